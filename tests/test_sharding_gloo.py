"""N > 1 host logic on CPU: two processes over gloo exercise the scatter / gather plumbing and the
shard arithmetic bench.py uses across GPUs (no compute -- ciphertext shards are independent)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_shard_bounds_cover_batch_exactly():
    from gemini_seal_b200.shard import shard_bounds
    for batch in (0, 1, 7, 64, 65, 1000):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(batch, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and 0 <= (b - a) - (d - c) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gemini_seal_b200.shard import gather_batch, scatter_batch
    B, shape = 3, (2, 4, 16)
    full = torch.arange(world * B * 2 * 4 * 16, dtype=torch.int64).reshape((world * B,) + shape) if rank == 0 else None
    mine = scatter_batch(full, (B,) + shape, torch.int64, "cpu")
    expect = torch.arange(world * B * 2 * 4 * 16, dtype=torch.int64).reshape((world * B,) + shape)[rank * B:(rank + 1) * B]
    ok = torch.equal(mine, expect)
    back = gather_batch(mine + rank)  # a rank-dependent "result"
    if rank == 0:
        want = torch.cat([expect_r + r for r, expect_r in enumerate(
            torch.arange(world * B * 2 * 4 * 16, dtype=torch.int64).reshape((world * B,) + shape).chunk(world))])
        ok = ok and torch.equal(back, want)
    else:
        ok = ok and back is None
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_scatter_gather_two_ranks_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert results == {0: True, 1: True}
