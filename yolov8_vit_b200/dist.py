"""Multi-GPU plumbing of the hot path: images are independent, so the batch is
sharded contiguously across ranks with NO collective on the data path; the only exchange is the
final gather of the fixed-size per-image result records (NCCL over NVLink on the GPU box; the same
code runs over gloo on CPU tensors in the tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_images: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) of rank `rank`; the first n % world ranks take one extra image."""
    base, extra = divmod(n_images, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_records(records: torch.Tensor, n_images: int | None = None) -> torch.Tensor:
    """All-gather of per-rank records [B_local, max_det, 8] -> [sum B_local, max_det, 8] in global
    image order.  Ragged shards (n_images % world != 0) are padded to the largest shard for the
    collective and trimmed afterwards."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return records
    world, rank = dist.get_world_size(), dist.get_rank()
    if n_images is None:
        n_images = records.shape[0] * world
    sizes = [shard_range(n_images, r, world) for r in range(world)]
    bmax = max(hi - lo for lo, hi in sizes)
    send = records
    if records.shape[0] < bmax:
        pad = records.new_zeros((bmax - records.shape[0],) + tuple(records.shape[1:]))
        send = torch.cat([records, pad], 0)
    out = records.new_empty((world * bmax,) + tuple(records.shape[1:]))
    dist.all_gather_into_tensor(out, send.contiguous())
    if all(hi - lo == bmax for lo, hi in sizes):
        return out
    return torch.cat([out[r * bmax: r * bmax + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], 0)
