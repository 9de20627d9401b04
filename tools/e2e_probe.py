"""development aid (under gpurun): wall time of csim_batch_set_params + csim_batch_tran (host buffers) on the bench
workload for several CSIM_WAVE_CHUNKS settings, beside the device time of the same call.
usage: tools/e2e_probe.py [instances] [chunks ...]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from circuitsimulator_b200 import capi, shard
from circuitsimulator_b200.netlist import Netlist

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
chunk_list = sys.argv[2:] or ["1", "0", "16"]
nl = Netlist(bench.load_netlist_text("dbmixer"))
plan = nl.plan(device=0)
probes = nl.plot_probes if len(nl.plot_probes) else np.arange(2, dtype=np.int32)
setup = nl.tran_setup(probes=probes, out_stride=1, n_steps=nl.n_steps)
ph = torch.from_numpy(shard.shard_params(nl, B, 0, 1, bench.SEED)).pin_memory().numpy()
N = nl.n_unknowns


def pinned(shape, dtype):
    return torch.empty(shape, dtype=dtype).pin_memory().numpy()


hout = dict(x_dc=pinned((N, B), torch.float64), x_final=pinned((N, B), torch.float64),
            wave=pinned((B, len(probes), max(setup.n_out, 1)), torch.float64), iters_total=pinned((B,), torch.int32),
            iters_step=None, hash=pinned((B,), torch.int64).view(np.uint64), status=pinned((B,), torch.int32).view(np.uint32))
batch = capi.Batch(plan, B)
ref = None
for ch in chunk_list * 2:
    os.environ["CSIM_WAVE_CHUNKS"] = ch
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    batch.set_params(ph)
    batch.tran(setup, out=hout)
    wall = time.perf_counter() - t0
    s = float(hout["wave"].sum())
    ref = s if ref is None else ref
    print(f"chunks {ch:>3s}: wall {wall*1e3:8.1f} ms  device (DC + transient) {batch.last_kernel_ms:8.1f} ms  "
          f"gap {wall*1e3 - batch.last_kernel_ms:6.1f} ms  {B * nl.n_steps / wall / 1e6:6.1f} M/s  same={s == ref}")
