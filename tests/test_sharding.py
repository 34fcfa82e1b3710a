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


def test_slice_proving_key_partitions_every_query_and_shares_sum_to_the_whole():
    """point-range slicing of a proving key (sharded.slice_proving_key): the slices tile A / B / H / L exactly, and the
    oracle's multiexps over the slices (with the scalar ranges vsp_prove_partial uses) sum to the full multiexps"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import cref, fixtures as fx
    import helpers as h
    sh = _sharded()
    cs, primary, aux = h.synthetic_r1cs(61, 6, seed=3)
    arr = h.synthetic_pk_arrays(cs, 4)
    _, m = cref.domain_plan(cs.num_constraints + cs.num_inputs + 1)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    w = np.concatenate([cref.ints_to_limbs([1], 4), P, A])
    H = cref.witness_map(fx.fill_pk(cref.ProvingKeyDesc(), arr).cs, P, A)
    ni = cs.num_inputs
    full = {"A": cref.msm_g1(arr["A_query"], w), "H": cref.msm_g1(np.ascontiguousarray(arr["H_query"][:m - 1]), H[:m - 1].copy()),
            "L": cref.msm_g1(arr["L_query"], A), "B": cref.msm_g1(arr["B_g1"], np.ascontiguousarray(w[arr["B_idx"]]))}
    for world in (1, 2, 3, 5):
        acc = {k: np.zeros(12, dtype=np.uint64) for k in full}
        seen = {k: [] for k in "ABHL"}
        for rank in range(world):
            sl, rng = sh.slice_proving_key(arr, m, rank, world)
            for k in "ABHL":
                seen[k].append(rng[k])
            a_lo, a_hi = rng["A"]
            h_lo, h_hi = rng["H"]
            l_lo, l_hi = rng["L"]
            parts = {"A": (sl["A_query"][:a_hi - a_lo], w[a_lo:a_hi]),
                     "H": (sl["H_query"][:h_hi - h_lo], H[h_lo:h_hi]),
                     "L": (sl["L_query"][:l_hi - l_lo], w[ni + 1 + l_lo: ni + 1 + l_hi]),
                     "B": (sl["B_g1"][:sl["B_count"]], w[sl["B_idx"][:sl["B_count"]]])}
            for k, (b, s) in parts.items():
                if len(s):
                    acc[k] = cref.g1_add(acc[k], cref.msm_g1(np.ascontiguousarray(b), np.ascontiguousarray(s)))
        for k in "ABHL":
            assert seen[k][0][0] == 0 and all(seen[k][i][1] == seen[k][i + 1][0] for i in range(world - 1))
        assert seen["A"][-1][1] == cs.num_variables + 1 and seen["H"][-1][1] == m - 1
        for k in full:
            assert np.array_equal(acc[k], full[k]), (world, k)


class _StubPhase:
    """stands in for vote_phase.VotePhase on a GPU-less box: a ballot's "blobs" name the ballot and the scalars it was given"""

    def __init__(self, rank):
        self.rank, self.calls = rank, 0

    def generate_votes(self, ballots, rs=None, rnd=None, rerandomize=True):
        self.calls += 1
        assert rs is None or rs.shape == (len(ballots), 3, 4)
        return [(b"proof%d" % idx, b"vote%d" % vote, sk + b"|rs%d" % (0 if rs is None else int(rs[j, 0, 0])),
                 b"rank%d" % self.rank) for j, (idx, vote, sk) in enumerate(ballots)]


def _vote_worker(rank: int, world: int, port: int, out_dir: str):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import importlib
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        vp = importlib.import_module("vote-saver-protocol_b200.vote_phase")
        k = 7
        ballots = [(i, (3 * i) % 25, b"sk%d" % i) for i in range(k)]
        rs = np.zeros((k, 3, 4), dtype=np.uint64)
        rs[:, 0, 0] = np.arange(100, 100 + k)
        out = vp.generate_votes_sharded(_StubPhase(rank), ballots, rs=rs)
        assert len(out) == k
        for i, (proof, pin, ct, sn) in enumerate(out):
            # ballot order, each ballot made by rank i mod world with ITS scalars
            assert (proof, pin, ct, sn) == (b"proof%d" % i, b"vote%d" % ((3 * i) % 25), b"sk%d|rs%d" % (i, 100 + i), b"rank%d" % (i % world))
        assert vp.generate_votes_sharded(_StubPhase(rank), []) == []
        with open(os.path.join(out_dir, "vote_ok%d" % rank), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_vote_phase_ballots_shard_over_ranks_world2_gloo(tmp_path):
    """vote_phase.generate_votes_sharded: ballots i mod G per rank, blobs back in ballot order on every rank, per-ballot
    scalars following their ballot (so the output is the same at every world size); without a process group it is the
    single-process call"""
    import importlib
    world = 2
    mp.spawn(_vote_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / ("vote_ok%d" % r)).exists() for r in range(world))
    vp = importlib.import_module("vote-saver-protocol_b200.vote_phase")
    single = vp.generate_votes_sharded(_StubPhase(0), [(i, i, b"s") for i in range(3)])
    assert [b[0] for b in single] == [b"proof0", b"proof1", b"proof2"] and all(b[3] == b"rank0" for b in single)
