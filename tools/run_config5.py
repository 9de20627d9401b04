#!/usr/bin/env python
"""BASELINE.json config 5 at full size: generated 200-stage inverter/RC ladder (N = 404 unknowns),
.TRAN 1e-11 2e-9 (200 steps), 4,096 Monte-Carlo instances, on the large-matrix path (dense_kernels.cu:
one CTA per instance, blocked LU; CSIM_BIG_KIND=warp selects big_kernels.cu instead).
The CPU side is the oracle's dense restatement on a bounded sample (O(N^3) per Newton iteration).
Prints one JSON line.   usage: tools/run_config5.py [--stages 200] [--instances 4096] [--steps 0]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--stages", type=int, default=200)
ap.add_argument("--instances", type=int, default=4096)
ap.add_argument("--steps", type=int, default=0)
ap.add_argument("--cpu-steps", type=int, default=2)
args = ap.parse_args()

import __graft_entry__ as g  # noqa: E402
g.build()
import torch  # noqa: E402
from circuitsimulator_b200 import capi  # noqa: E402
from circuitsimulator_b200.netlist import Netlist, ladder_netlist  # noqa: E402

nl = Netlist(ladder_netlist(args.stages))
B = args.instances
ns = args.steps or nl.n_steps
P = nl.monte_carlo(B, 20261104 + 5)
plan = nl.plan()
setup = nl.tran_setup(n_steps=ns)
params = torch.from_numpy(P).cuda()


def run():
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"], status=d["status"])
    return d, t


e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
torch.cuda.synchronize()
e[0].record()
d = capi.dc(plan, params)
e[1].record()
t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"], status=d["status"])
e[2].record()
torch.cuda.synchronize()
dc_ms, tran_ms = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])
newton = int(t["iters_total"].sum().item())
dc_solves = int(d["iters"].sum().item())
N = nl.n_unknowns
line = {
    "config": f"{args.stages}-stage inverter/RC ladder, N = {N}, {B} Monte-Carlo instances, {ns} steps of "
              f"{setup.dt:g} s, large-matrix path ("
              + ("warp per instance, zero skipping" if os.environ.get("CSIM_BIG_KIND", "").startswith("w")
                 else "CTA per instance, blocked dense LU") + ", matrix in HBM/L2)",
    "dc_ms": dc_ms, "tran_ms": tran_ms, "instance_timepoints_per_s": B * ns / ((dc_ms + tran_ms) / 1e3),
    "tran_newton_iterations": newton, "dc_newton_iterations": dc_solves,
    "solves_per_s": (newton + dc_solves) / ((dc_ms + tran_ms) / 1e3),
    "status_or": int(t["status"].max().item()),
}
# FP64 pipe roofline of the dense factorisation: (N^3 - N)/3 multiply-subtract pairs per solve (the reference executes
# exactly these; FMA is forbidden), against the measured 18.46 T separate mul/add per second (profiles/r1_microbench_fp64.txt)
pairs = (N ** 3 - N) // 3
line["roofline"] = {"kernel": "tran_dense_kernel / dc_dense_kernel", "bound": "fp64 pipe (separate mul and add, no FMA)",
                    "achieved": line["solves_per_s"] * 2 * pairs / 1e12, "peak": 18.46, "unit": "Tflop/s",
                    "frac": line["solves_per_s"] * 2 * pairs / 18.46e12, "pairs_per_solve": pairs}
if args.cpu_steps > 0:
    from oracle import oraclelib as O
    topo = O.Topology(nl.n_unknowns, nl.n_node_eq, nl.elem_type, nl.elem_flags, nl.elem_nodes, nl.elem_branch,
                      nl.elem_param, nl.n_params, nl.params, nl.node_names, nl.elem_names)
    m, grp, tab = O.make_source_tables(topo, nl.params, setup.dt, args.cpu_steps)
    t0 = time.perf_counter()
    od = O.dc(topo, P[:, 0])
    t1 = time.perf_counter()
    ot = O.tran(topo, P[:, 0], setup.dt, 0, args.cpu_steps, m, grp, tab, od["x"], want_wave=False)
    t2 = time.perf_counter()
    per_solve = (t2 - t0) / (int(od["iters"].sum()) + ot["total"])
    line["cpu_port"] = {"seconds_per_solve": per_solve, "cores": 1,
                        "instance_timepoints_per_s_per_core_extrapolated": ns / (per_solve * (dc_solves + newton) / B),
                        "sample": f"instance 0: DC ({int(od['iters'].sum())} solves) + {args.cpu_steps} steps "
                                  f"({ot['total']} solves), oracle C restatement (dense O(N^3))"}
    # parity of what was timed: instance 0's first steps equal the oracle's
    assert np.array_equal(d["x"][:, 0].cpu().numpy().view(np.uint64), od["x"].view(np.uint64)), "DC parity"
    line["parity_checked"] = "instance 0 DC operating point bit-identical to the oracle"
print(json.dumps(line))
