"""Development aid (under gpurun): time csim_dev_tran launch by launch in automatic mode and print the variant statistics.
usage: tools/tran_tpi_probe.py netlist instances launches [steps]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
name, B, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
g, topo = support.load_golden(name)
ns = int(sys.argv[4]) if len(sys.argv) > 4 else support.n_steps_of(g)
P = O.monte_carlo(topo, B, 20261104 + 3)
params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
plan = support.make_plan(topo)
dt = float(g["dt"])
m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([1, 2], np.int32))
d = capi.dc(plan, params)
prev = (0, 0)
for k in range(L):
    x = d["x"].clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    t = capi.tran(plan, params, x, setup)
    e1.record()
    torch.cuda.synchronize()
    i = plan.spec_info()
    fb, sv = i["fallback_solves"] - prev[0], i["solves"] - prev[1]
    if sv <= 0:
        fb, sv = i["fallback_solves"], i["solves"]
    prev = (i["fallback_solves"], i["solves"])
    nv = i["log"].split(" predicted")[0].split()[-1]
    print(f"launch {k}: {e0.elapsed_time(e1):8.2f} ms  variants {nv}  redone {fb} of {sv} ({100.0 * fb / max(sv, 1):.3f} %)", flush=True)
