"""development aid (under gpurun): which transient kernel the automatic choice ends up with on a Monte-Carlo batch,
its miss statistics and the time of the full .TRAN.  usage: tools/spec_probe.py netlist instances [CSIM_JIT]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 3:
    os.environ["CSIM_JIT"] = sys.argv[3]
import torch
import bench
from circuitsimulator_b200 import capi, shard
from circuitsimulator_b200.netlist import Netlist

name, B = sys.argv[1], int(sys.argv[2])
nl = Netlist(bench.load_netlist_text(name))
plan = nl.plan(device=0)
dev = torch.device("cuda", 0)
probes = nl.plot_probes if len(nl.plot_probes) else np.arange(2, dtype=np.int32)
setup = nl.tran_setup(probes=probes, out_stride=1, n_steps=nl.n_steps)
params = torch.from_numpy(shard.shard_params(nl, B, 0, 1, bench.SEED)).to(dev)
for rep in range(int(os.environ.get('PROBE_REPS', '3'))):
    h, st = capi.new_diag(B, dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    d = capi.dc(plan, params, h, st)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    t = capi.tran(plan, params, d["x"], setup, pivot_hash=h, status=st)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    info = plan.spec_info()
    print(f"rep {rep}: dc {1e3*(t1-t0):.1f} ms  tran {1e3*(t2-t1):.1f} ms  kind {info['kind']}  fallback {info['fallback_solves']} of "
          f"{info['solves']} solves  status_or {int(t['status'].max())}  newton {int(t['iters_total'].sum())}")
print("\n".join(l for l in info["log"].splitlines() if "warning" not in l and "^" not in l and l.strip() and "const double" not in l and "Remark" not in l)[:3000])
