"""Development aid (under gpurun): time csim_dev_dc launch by launch and print the DC kernel's variant statistics.
usage: tools/dc_tpi_probe.py netlist instances launches"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
name, B, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
g, topo = support.load_golden(name)
P = O.monte_carlo(topo, B, 20261104 + 3)
params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
plan = support.make_plan(topo)
prev = (0, 0)
for k in range(L):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    d = capi.dc(plan, params)
    e1.record()
    torch.cuda.synchronize()
    i = plan.dc_spec_info()
    print(f"launch {k}: {e0.elapsed_time(e1):8.2f} ms  variants {i['n_variants']}  redone {i['fallback_solves']} of {i['solves']} "
          f"({100.0 * i['fallback_solves'] / max(i['solves'], 1):.2f} %)  iters/instance {float(d['iters'].sum()) / B:.1f}", flush=True)
print(plan.dc_spec_info()["log"][:3000])
