"""world_size-2 gloo test (CPU) of the N>1 host logic: contiguous sharding + final record gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from yolov8_vit_b200.dist import balanced_ranges, gather_records, shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 256, 513):
        for w in (1, 2, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_balanced_ranges():
    g = torch.Generator().manual_seed(7)
    for w in (1, 2, 4, 8):
        for n in (0, 1, 5, 64, 513):
            counts = torch.randint(0, 100, (n,), generator=g)
            r = balanced_ranges(counts, w)
            assert len(r) == w and r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] and r[i][0] <= r[i][1] for i in range(w - 1))
            if n >= 8 * w:
                load = [float(counts[a:b].sum()) for a, b in r]
                even = [float(counts[a:b].sum()) for a, b in (shard_range(n, k, w) for k in range(w))]
                ideal = float(counts.sum()) / w
                assert max(load) <= ideal + 100            # never more than one image over the ideal
                assert max(load) <= max(even) + 1e-9 or max(load) <= ideal + 50
    # skewed batch: all detections in the first quarter of the images
    counts = [100] * 16 + [0] * 48
    r = balanced_ranges(counts, 4)
    load = [sum(counts[a:b]) for a, b in r]
    assert max(load) <= 500 and max(sum(counts[a:b]) for a, b in (shard_range(64, k, 4) for k in range(4))) == 1600
    assert balanced_ranges([0] * 8, 4) == [(0, 2), (2, 4), (4, 6), (6, 8)]


def _worker(rank, world, port, n_images, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(n_images * 3 * 8, dtype=torch.float32).reshape(n_images, 3, 8)
    if n_images < 0 or n_images == 9:                      # crop-balanced ragged ranges
        ranges = balanced_ranges([50, 50, 50, 1, 1, 1, 1, 1, 1], world)
        lo, hi = ranges[rank]
        got = gather_records(full[lo:hi].clone(), ranges=ranges)
        q.put((rank, bool(torch.equal(got, full)) and ranges[0][1] - ranges[0][0] < ranges[1][1] - ranges[1][0]))
        dist.destroy_process_group()
        return
    lo, hi = shard_range(n_images, rank, world)
    got = gather_records(full[lo:hi].clone(), n_images)
    q.put((rank, bool(torch.equal(got, full))))
    dist.destroy_process_group()


def _run(n_images):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=120) for _ in ps]
    [p.join(timeout=60) for p in ps]
    assert all(ok for _, ok in res), res


def test_gather_records_even():
    _run(8)


def test_gather_records_ragged():
    _run(7)


def test_gather_records_balanced():
    _run(9)
