"""Multi-GPU host logic for view-sharded volume / TF gradient steps.

The reference has no multi-GPU code at all; rays and views are independent
(renderer/renderer_kernels.cuh:100-105), so the path shards over the view/batch dimension:
view v goes to rank v mod G, the volume and TF are replicated, every rank accumulates its views'
gradients into private adj_volume / adj_tf tensors, and ONE exchange step per optimisation step
sums them (NCCL all-reduce over NVLink on GPUs, gloo in the CPU tests). Per-ray outputs
(images, adj_stepSize, adj_camera*) stay on the owning rank.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Indices of the views owned by ``rank`` (round-robin, so every rank gets a spread of
    view directions and therefore a similar number of samples)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of size {world_size}")
    return list(range(rank, n_views, world_size))


def batches(indices: Sequence[int], batch: int) -> Iterable[List[int]]:
    """Launch-sized groups of a rank's views (one kernel launch renders ``batch`` views)."""
    if batch <= 0:
        raise ValueError("batch must be positive")
    for i in range(0, len(indices), batch):
        yield list(indices[i:i + batch])


def allreduce_gradients(tensors, group=None):
    """Sums the replicated-parameter gradients (adj_volume, adj_tf) over all ranks, in place.

    One collective per tensor, issued back to back on the current stream (NCCL) so they pipeline;
    a no-op when torch.distributed is not initialised (single GPU)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    handles = [dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group, async_op=True)
               for t in tensors if t is not None]
    for h in handles:
        h.wait()
