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


def balanced_ranges(counts, world: int) -> list[tuple[int, int]]:
    """Crop-count-balanced contiguous partition (SURVEY.md §8 rows e / f3): `counts[i]` is the number of
    detections the detector reported for image i (`num_dets` of the engine tuple, known before the second
    stage starts; `解读.md:83-86`), and the ViT cost of an image is proportional to it.  Rank k gets the
    images whose inclusive prefix sum of `counts + eps` falls in (k, k+1] * total / world -- a prefix sum
    and `world - 1` binary searches, no collective; every rank computes the same answer from the same
    counts.  Images stay in global order, so `gather_records(..., ranges=...)` concatenates the shards.
    The eps (one image costs at least its decode + NMS) keeps empty images spread over the ranks."""
    c = torch.as_tensor(counts, dtype=torch.float64).flatten() + 0.25
    n = c.numel()
    if n == 0:
        return [(0, 0)] * world
    pre = torch.cumsum(c, 0)
    targets = pre[-1] * torch.arange(1, world, dtype=torch.float64) / world
    # first image whose inclusive prefix exceeds the target closes the shard if that is nearer the target
    cut = torch.searchsorted(pre, targets, right=False)
    bounds = [0]
    for k in range(world - 1):
        j = int(cut[k])
        lo_side = float(pre[j - 1]) if j > 0 else 0.0
        hi_side = float(pre[min(j, n - 1)])
        b = j + 1 if (hi_side - float(targets[k])) < (float(targets[k]) - lo_side) else j
        bounds.append(min(n, max(bounds[-1], b)))
    bounds.append(n)
    return [(bounds[k], bounds[k + 1]) for k in range(world)]


def gather_records(records: torch.Tensor, n_images: int | None = None,
                   ranges: list[tuple[int, int]] | None = None) -> torch.Tensor:
    """All-gather of per-rank records [B_local, max_det, 8] -> [sum B_local, max_det, 8] in global
    image order.  Ragged shards (n_images % world != 0, or the crop-balanced `ranges` of
    `balanced_ranges`) are padded to the largest shard for the collective and trimmed afterwards."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return records
    world, rank = dist.get_world_size(), dist.get_rank()
    if n_images is None:
        n_images = records.shape[0] * world
    sizes = ranges if ranges is not None else [shard_range(n_images, r, world) for r in range(world)]
    if len(sizes) != world or records.shape[0] != sizes[rank][1] - sizes[rank][0]:
        raise ValueError(f"gather_records: rank {rank} holds {records.shape[0]} images, its range is {sizes[rank]}")
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
