"""-m "not gpu": the N>1 path on CPU — world_size-2 gloo run of the sample sharding + timing plumbing."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from multimeditron_b200 import sharding


def test_shard_rows_partition():
    for n in (0, 1, 7, 8, 64, 257):
        for w in (1, 2, 3, 8):
            seen = []
            for r in range(w):
                seen += list(sharding.shard_rows(n, r, w))
            assert seen == list(range(n))
            sizes = [len(sharding.shard_rows(n, r, w)) for r in range(w)]
            assert max(sizes) - min(sizes) <= 1


def test_shard_batch_rows_travel_with_their_images():
    per_row = [4, 0, 1, 2, 4, 4, 0, 3]
    got = []
    for r in range(3):
        rows, imgs = sharding.shard_batch(None, per_row, r, 3)
        assert len(imgs) == sum(per_row[rows.start: rows.stop])
        got += list(imgs)
    assert got == list(range(sum(per_row)))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rk, lrk, ws = sharding.env_rank_world()
    rows = sharding.shard_rows(10, rk, ws)
    # every rank processes only its rows; no data-path collective; only the timing reduction is collective
    local = torch.arange(10)[rows.start: rows.stop].sum().item()
    t = sharding.max_over_ranks(1.0 + rank)
    total = sharding.sum_over_ranks(float(len(rows)))
    dist.barrier()
    q.put((rank, local, t, total))
    dist.destroy_process_group()


def test_world_size_2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [10, 35]       # rows 0..4 and 5..9
    assert all(r[2] == 2.0 for r in res)          # max over ranks
    assert all(r[3] == 10.0 for r in res)         # all rows owned exactly once
