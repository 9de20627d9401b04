"""Development aid (under gpurun): time the thread-per-instance transient kernel for several CSIM_TPI_OPTS masks and
check that every mask gives the same bits (final state, pivot hash, Newton totals).
usage: tools/tpi_opts_sweep.py netlist steps instances mask [mask ...]     (mask 'd' = library defaults)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
name, ns, B = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
g, topo = support.load_golden(name)
P = support.perturb(topo.params, topo, B, 20261104)
params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
dt = float(g["dt"])
m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([1, 2], np.int32), out_stride=10)
ref = None
for mask in sys.argv[4:]:
    if mask == "d":
        os.environ.pop("CSIM_TPI_OPTS", None)
    else:
        os.environ["CSIM_TPI_OPTS"] = mask
    plan = support.make_plan(topo)
    plan.jit_mode(3)
    best = 1e30
    for rep in range(3):
        d = capi.dc(plan, params)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t = capi.tran(plan, params, d["x"], setup, pivot_hash=d["hash"], status=d["status"])
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    info = plan.spec_info()
    cur = (t["x"].cpu().numpy().view(np.uint64), t["hash"].cpu().numpy(), t["iters_total"].cpu().numpy())
    same = "-" if ref is None else all(np.array_equal(a, b) for a, b in zip(ref, cur))
    if ref is None:
        ref = cur
    tot = int(t["iters_total"].sum())
    print(f"opts {mask:>6s} {name} B={B} steps={ns}: tran {best:9.2f} ms  {B * ns / best / 1e3:8.2f} M inst-tp/s  "
          f"{best * 1e3 / (tot / B):7.3f} us/newton-iter  redone {info['fallback_solves']}/{info['solves']}  same bits: {same}",
          flush=True)
    plan.close()
