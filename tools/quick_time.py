"""Quick device timing of the fused DC+TRAN path and the stamp kernel (development aid, not the bench)."""
import sys, os, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
name = sys.argv[1] if len(sys.argv) > 1 else "dbmixer"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
ns = int(sys.argv[3]) if len(sys.argv) > 3 else 500
g, topo = support.load_golden(name)
plan = support.make_plan(topo)
P = support.perturb(topo.params, topo, B, 20261104)
params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
dt = float(g["dt"])
m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([1, 2], np.int32))
e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
for rep in range(2):
    e[0].record(); d = capi.dc(plan, params); e[1].record()
    x = d["x"].clone()
    e[2].record(); t = capi.tran(plan, params, x, setup, pivot_hash=d["hash"], status=d["status"]); e[3].record()
    torch.cuda.synchronize()
    tot = int(t["iters_total"].sum())
    print(f"{name} B={B} steps={ns}: dc {e[0].elapsed_time(e[1]):.2f} ms, tran {e[2].elapsed_time(e[3]):.2f} ms, "
          f"{tot/B/ns:.2f} it/step, {B*ns/e[2].elapsed_time(e[3])*1e3/1e6:.2f} M inst-timepoints/s, "
          f"{e[2].elapsed_time(e[3])*1e3/(tot/B):.2f} us per newton iter (latency), status or={int(t['status'].max())}")
# stamp kernel
N = topo.n_unknowns
x = d["x"]; xp = x.clone()
nnz = plan.pattern(1)[1].size
vals = torch.empty((nnz, B), dtype=torch.float64, device="cuda"); rhs = torch.empty((N, B), dtype=torch.float64, device="cuda")
sv = torch.ones((len(topo.source_elems()), B), dtype=torch.float64, device="cuda")
for Bs in (B, 65536, 1 << 20):
    if Bs != B:
        reps = Bs // B
        params2 = params.repeat(1, reps).contiguous(); x2 = x.repeat(1, reps).contiguous(); xp2 = x2.clone()
        sv2 = sv.repeat(1, reps).contiguous()
        vals = torch.empty((nnz, Bs), dtype=torch.float64, device="cuda"); rhs = torch.empty((N, Bs), dtype=torch.float64, device="cuda")
    else:
        params2, x2, xp2, sv2 = params, x, xp, sv
    for rep in range(3):
        e[0].record(); capi.stamp_csr(plan, 1, params2, x2, xp2, sv2, dt=dt, gmin=1e-6, out=(vals, rhs)); e[1].record()
        torch.cuda.synchronize()
    ms = e[0].elapsed_time(e[1])
    byt = 8 * (topo.n_params + N + N + nnz + N) * Bs
    print(f"stamp_csr TRAN B={Bs}: {ms*1e3:.1f} us, {byt/ms/1e6:.1f} GB/s algorithmic")
