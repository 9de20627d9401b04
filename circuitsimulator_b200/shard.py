"""Multi-GPU plumbing: the batch is sharded in contiguous blocks of instances, one process per GPU,
no collective on the data path (instances never interact — SURVEY.md §8(e)); the only exchange is the
final gather of probe waveforms and per-instance diagnostics to rank 0 (NCCL over NVLink on the GPU
box, gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """[first, first + count) of the global batch owned by `rank`: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(int(n_total), int(world))
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def shard_params(netlist, n_total: int, rank: int, world: int, seed: int, rel: float = 0.05) -> np.ndarray:
    """This rank's [P, count] slice of ONE global Monte-Carlo batch (same draws whatever the world size)."""
    first, count = shard_range(n_total, rank, world)
    return netlist.monte_carlo(count, seed, rel=rel, first=first)


def gather_instances(t, dst: int = 0, dim: int = 0):
    """Gathers a per-instance tensor (instances along `dim`, possibly different counts per rank) to `dst`,
    in rank order — i.e. in global instance order for contiguous shards.  Returns the concatenation on
    `dst` and None elsewhere.  Works on CUDA tensors with the nccl backend and CPU tensors with gloo."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return t
    world, rank = dist.get_world_size(), dist.get_rank()
    n = torch.tensor([t.shape[dim]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    t = t.movedim(dim, 0).contiguous()
    if rank == dst:
        parts = [torch.empty((c,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device) for c in counts]
        parts[dst] = t
        ops = [dist.P2POp(dist.irecv, parts[r], r) for r in range(world) if r != dst and counts[r] > 0]
        for q in (dist.batch_isend_irecv(ops) if ops else []):
            q.wait()
        return torch.cat(parts, 0).movedim(0, dim)
    if t.shape[0] > 0:
        for q in dist.batch_isend_irecv([dist.P2POp(dist.isend, t, dst)]):
            q.wait()
    return None
