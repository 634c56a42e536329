"""world_size-2 gloo test (CPU) of the N>1 host logic: contiguous sharding + final record gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from yolov8_vit_b200.dist import gather_records, shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 256, 513):
        for w in (1, 2, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def _worker(rank, world, port, n_images, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_images, rank, world)
    full = torch.arange(n_images * 3 * 8, dtype=torch.float32).reshape(n_images, 3, 8)
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
