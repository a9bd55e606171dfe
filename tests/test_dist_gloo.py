"""CPU tests of the N>1 host logic (world_size 2, gloo): channel sharding, the one gather of the sharded rows onto
rank 0 (uneven shards), and the max-over-ranks timing reduction. The GPU box runs the same code over NCCL."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, units, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pebblesdr_b200.dist import shard_range, gather_rows, max_over_ranks
        c0, c1 = shard_range(units, world, rank)
        # row r of the full result is (r + 1) * (1 + 2j) * [0..7]
        local = torch.stack([(r + 1) * (1 + 2j) * torch.arange(8, dtype=torch.float64).to(torch.complex128) for r in range(c0, c1)]) \
            if c1 > c0 else torch.zeros(0, 8, dtype=torch.complex128)
        full = gather_rows(local, units, dst=0)
        ok = True
        if rank == 0:
            want = torch.stack([(r + 1) * (1 + 2j) * torch.arange(8, dtype=torch.float64).to(torch.complex128) for r in range(units)])
            ok = full is not None and full.shape == want.shape and bool(torch.equal(full, want))
        else:
            ok = full is None
        t = max_over_ranks(10.0 + rank, torch.device("cpu"))
        q.put((rank, ok and t == 10.0 + world - 1, (c0, c1)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("units", [5, 8])
def test_shard_gather_world2(units):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, units, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=90) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    ranges = [rg for _, _, rg in res]
    assert ranges[0][0] == 0 and ranges[0][1] == ranges[1][0] and ranges[1][1] == units   # contiguous cover


def test_shard_range_properties():
    from pebblesdr_b200.dist import shard_range, channel_frequencies
    for units in (1, 7, 64, 1024):
        for world in (1, 2, 4, 8):
            rs = [shard_range(units, world, r) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == units
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1
    f = channel_frequencies(1024)
    assert f[0] == -30.69e6 and abs(f[1023] - 30.69e6) < 1.0
