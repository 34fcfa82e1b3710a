"""Multi-GPU host logic on CPU: world_size 2 over gloo (no GPU).  The partition / all_gather / join code of
vote-saver-protocol_b200/sharded.py runs for real; the local multiexp and the point sum are the oracle's (the CUDA
library is exercised by the -m gpu tests and by bench.py under torchrun)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sharded():
    importlib.import_module("vote-saver-protocol_b200")
    return importlib.import_module("vote-saver-protocol_b200.sharded")


def test_shard_range_covers_every_point_once():
    sh = _sharded()
    for n in (0, 1, 7, 8, 1000, (1 << 20) + 3):
        for world in (1, 2, 3, 8):
            edges = [sh.shard_range(n, g, world) for g in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[g][1] == edges[g + 1][0] for g in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.shard_range(10, 2, 2)


def test_ballot_indices_partition():
    sh = _sharded()
    for k in (0, 1, 5, 4096):
        for world in (1, 2, 4, 8):
            allidx = np.concatenate([sh.ballot_indices(k, g, world) for g in range(world)])
            assert sorted(allidx.tolist()) == list(range(k))


def _worker(rank: int, world: int, port: int, n: int, out_dir: str):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import cref, fixtures as fx
        import helpers as h
        sh = _sharded()
        bases, scalars = h.random_g1(n, seed=5), fx.random_fr(n, seed=6)
        scalars[3] = 0  # a zero and a one scalar, and a point at infinity, inside rank 0's range
        scalars[4] = 0
        scalars[4, 0] = 1
        bases[5] = 0
        lo, hi = sh.shard_range(n, rank, world)

        def point_sum(pts):
            acc = np.zeros(12, dtype=np.uint64)
            for p in pts:
                acc = cref.g1_add(acc, p)
            return acc

        got = sh.sharded_multiexp(lambda: cref.msm_g1(np.ascontiguousarray(bases[lo:hi]),
                                                      np.ascontiguousarray(scalars[lo:hi])), point_sum)
        want = cref.msm_g1(bases, scalars)
        assert np.array_equal(got, want), "rank %d: sharded multiexp differs from the single-range one" % rank

        # an empty shard (n < world) contributes the point at infinity
        tiny = sh.sharded_multiexp(lambda: cref.msm_g1(np.ascontiguousarray(bases[:1]), np.ascontiguousarray(scalars[:1]))
                                   if rank == 0 else np.zeros(12, dtype=np.uint64), point_sum)
        assert np.array_equal(tiny, cref.msm_g1(np.ascontiguousarray(bases[:1]), np.ascontiguousarray(scalars[:1])))

        # proof-per-GPU: every rank "proves" its ballots (row = ballot id pattern), results come back in ballot order
        k = 7
        mine = sh.ballot_indices(k, rank, world)
        local = np.stack([np.full(48, 1000 + i, dtype=np.uint64) for i in mine]) if len(mine) else np.zeros((0, 48), np.uint64)
        allp = sh.gather_ballot_results(local, k)
        assert allp.shape == (k, 48) and all(int(allp[i, 0]) == 1000 + i for i in range(k))
        with open(os.path.join(out_dir, "ok%d" % rank), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_sharded_multiexp_and_ballot_gather_world2_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), 301, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / ("ok%d" % r)).exists() for r in range(world))
