"""N>1 host logic on CPU: world-size-2 gloo.  Utterance sharding needs no collective on the data path, sharded ==
unsharded results, every rank takes the same direction-dropout decisions, timing is max-over-ranks."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from test_wrappers import _classes, _oracle_backend


def test_shard_rows_partition():
    from paper_accurate_fast_cheap_b200.sharding import shard_rows
    for B in (0, 1, 7, 8, 16, 64):
        for G in (1, 2, 4, 8):
            rows = [shard_rows(B, r, G) for r in range(G)]
            assert rows[0][0] == 0 and rows[-1][1] == B
            assert all(rows[i][1] == rows[i + 1][0] for i in range(G - 1))
            assert max(e - b for b, e in rows) - min(e - b for b, e in rows) <= 1


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from paper_accurate_fast_cheap_b200 import sharding
        w, classes = _classes()

        class MP:                                   # minimal monkeypatch context for the oracle backend
            def setattr(self, obj, name, val):
                setattr(obj, name, val)
        _oracle_backend(MP(), w)
        torch.manual_seed(4242)
        m = classes["drop_both"](64, 64, 12, "v6", "bi", 2048, False, 3)
        B, T = 5, 40                                # uneven shards: 3 + 2 rows
        x = torch.randn(B, T, 64, generator=torch.Generator().manual_seed(7))
        m.eval()
        with torch.no_grad():
            full, _ = m(x, x, x)
            mine, _ = m(sharding.shard_batch(x, rank, world), None, None)
        got = sharding.gather_batch(mine, B)
        ok_fwd = oracle.rel_err(got, full) <= 1e-6
        # training: same seed on every rank (train.py:71) -> identical direction decisions, step after step
        m.train()
        torch.manual_seed(5)
        decisions = [m._directions() for _ in range(12)]
        flat = torch.tensor([[int(a), int(b)] for a, b in decisions])
        both = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(both, flat)
        ok_drop = all(torch.equal(both[0], o) for o in both) and len({tuple(d) for d in decisions}) == 3
        t = sharding.max_over_ranks(1.0 + rank)
        if rank == 0:
            out.put((bool(ok_fwd), bool(ok_drop), t))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_equals_unsharded():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    ok_fwd, ok_drop, t = out.get(timeout=10)
    assert ok_fwd, "sharded forward differs from the unsharded one"
    assert ok_drop, "ranks disagree on direction dropout (or a branch was never taken)"
    assert t == 2.0
