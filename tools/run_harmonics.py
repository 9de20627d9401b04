#!/usr/bin/env python
"""Throughput of csim_dev_harmonics at config 4's size: 16,384 signals (one probe of every sweep point) x one
period of 100,000 samples, n_harm harmonics.  HBM roofline: 8 bytes per sample read once.  Prints one JSON line.
usage: tools/run_harmonics.py [--signals 16384] [--samples 100000] [--n-harm 3]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--signals", type=int, default=16384)
ap.add_argument("--samples", type=int, default=100000)
ap.add_argument("--n-harm", type=int, default=3)
args = ap.parse_args()
import __graft_entry__ as g  # noqa: E402
g.build()
import torch  # noqa: E402
from circuitsimulator_b200 import capi  # noqa: E402

x = torch.rand((args.signals, args.samples), dtype=torch.float64, device="cuda")
tables = capi.twiddle_tables(args.samples)
for _ in range(3):
    capi.harmonics(x, args.samples, args.n_harm, tables)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 5
e0.record()
for _ in range(reps):
    re, im = capi.harmonics(x, args.samples, args.n_harm, tables)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
peak = 6451.8
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
gbs = x.numel() * 8 / (ms / 1e3) / 1e9
print(json.dumps({"kernel": "harmonics_kernel", "signals": args.signals, "samples": args.samples, "n_harm": args.n_harm,
                  "ms": ms, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                         "bytes_per_launch": x.numel() * 8}}))
