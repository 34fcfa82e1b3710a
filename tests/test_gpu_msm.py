"""GPU parity: algebra::multiexp over G1 / G2 through the C ABI against the oracle's BDLO12 restatement and the
reference's own point-addition known answers (libs/pubkey/test/bls.cpp).  Bit-exact canonical affine output."""
import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import (G1_GEN_LIMBS, expected_generated_msm, generated_base_scalars, golden, hexlist, random_g1,
                     random_g2)

pytestmark = pytest.mark.gpu


def _ones(n):
    s = np.zeros((n, 4), dtype=np.uint64)
    s[:, 0] = 1
    return s


def test_g1_aggregate_known_answer(vsp, ctx):
    """bls.cpp:858-998: etalon_agg_sig = sum of ten G1 points = all-ones multiexp"""
    kat = golden("curve_kat.json")["g1_aggregate"]
    pts = cref.g1_to_limbs([(int(p["x"], 16), int(p["y"], 16)) for p in kat["points"]])
    want = cref.g1_to_limbs([(int(kat["sum"]["x"], 16), int(kat["sum"]["y"], 16))])[0]
    assert np.array_equal(vsp.multiexp(ctx, pts, _ones(10)), want)
    assert np.array_equal(vsp.points_sum(ctx, pts), want)


def test_g2_aggregate_known_answer(vsp, ctx):
    """bls.cpp:669-856: the same over G2"""
    kat = golden("curve_kat.json")["g2_aggregate"]
    f2 = lambda v: tuple(hexlist(v))
    pts = cref.g2_to_limbs([(f2(p["x"]), f2(p["y"])) for p in kat["points"]])
    want = cref.g2_to_limbs([(f2(kat["sum"]["x"]), f2(kat["sum"]["y"]))])[0]
    assert np.array_equal(vsp.multiexp(ctx, pts, _ones(10)), want)
    assert np.array_equal(vsp.points_sum(ctx, pts), want)


@pytest.mark.parametrize("n", [0, 1, 2, 3, 17, 100, 1000, 1 << 12, (1 << 14) + 5])
def test_g1_multiexp_matches_oracle(vsp, ctx, n):
    bases = random_g1(n, seed=n + 1) if n else np.zeros((0, 12), np.uint64)
    scalars = fx.random_fr(n, seed=n + 2) if n else np.zeros((0, 4), np.uint64)
    want = cref.msm_g1(bases, scalars) if n else np.zeros(12, np.uint64)
    assert np.array_equal(vsp.multiexp(ctx, bases, scalars), want)


@pytest.mark.parametrize("n", [1, 2, 9, 300, 1 << 10])
def test_g2_multiexp_matches_oracle(vsp, ctx, n):
    bases = random_g2(n, seed=n + 3)
    scalars = fx.random_fr(n, seed=n + 4)
    assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g2(bases, scalars))


def test_g1_multiexp_edge_scalars_and_points(vsp, ctx):
    """zero / one / r-1 scalars, repeated bases (bucket doubling), P and -P in one bucket, infinity among the bases"""
    n = 600
    bases = random_g1(n, seed=77)
    bases[10:20] = bases[10]          # the same point ten times: equal-point additions inside a bucket
    neg = cref.g1_mul(o.R - 1, bases[30])
    bases[31] = neg                   # P and -P
    bases[40] = 0                     # point at infinity as a base
    scalars = fx.random_fr(n, seed=78)
    scalars[0:5] = 0
    scalars[5:10] = _ones(5)
    scalars[10:20] = scalars[10]
    scalars[30] = scalars[31] = scalars[32]
    scalars[50] = cref.ints_to_limbs([o.R - 1], 4)[0]
    scalars[51] = cref.ints_to_limbs([o.R - 2], 4)[0]
    scalars[52] = cref.ints_to_limbs([1 << 254], 4)[0]
    assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g1(bases, scalars))
    # everything cancels -> the point at infinity (all-zero)
    b2 = np.stack([bases[30], neg])
    s2 = np.stack([scalars[60], scalars[60]])
    assert not vsp.multiexp(ctx, b2, s2).any()
    # all scalars zero
    assert not vsp.multiexp(ctx, bases, np.zeros((n, 4), np.uint64)).any()


@pytest.mark.parametrize("n", [1 << 12, 1 << 15])
def test_g1_witness_like_scalars(vsp, ctx, n):
    """90 % zero / one scalars: what multiexp_with_mixed_addition special-cases in the reference"""
    bases = random_g1(n, seed=5)
    scalars = fx.witness_like_fr(n, seed=6)
    assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g1(bases, scalars, mixed=True))


def test_resident_bases_prefix_and_errors(vsp, ctx):
    n = 500
    bases = random_g1(n, seed=11)
    scalars = fx.random_fr(n, seed=12)
    h = vsp.G1Bases(ctx, bases)
    assert np.array_equal(h.multiexp(scalars), cref.msm_g1(bases, scalars))
    assert np.array_equal(h.multiexp(scalars[:123]), cref.msm_g1(bases[:123], scalars[:123]))
    with pytest.raises(ValueError):
        h.multiexp(fx.random_fr(n + 1, seed=1))
    with pytest.raises(ValueError):
        vsp.multiexp(ctx, bases, scalars[:10])
    h.close()


def test_generated_bases_match_oracle(vsp, ctx):
    """vsp_g1_bases_generate: bases[i] = (splitmix64(seed + i) | 1) * G"""
    n, seed = 300, 42
    h = vsp.G1Bases.generate(ctx, n, seed)
    want = cref.g1_fixed_base_mul64(G1_GEN_LIMBS, generated_base_scalars(n, seed))
    assert np.array_equal(h.read(), want)
    assert np.array_equal(h.read(100, 7), want[100:107])


@pytest.mark.parametrize("log_n", [16, 20])
def test_g1_multiexp_checksum_full_size(vsp, ctx, log_n):
    """BASELINE config 2 sizes: sum s_i (k_i G) == (sum s_i k_i mod r) G, independent of n"""
    n, seed = 1 << log_n, 1234
    h = vsp.G1Bases.generate(ctx, n, seed)
    scalars = fx.random_fr(n, seed=1)
    assert np.array_equal(h.multiexp(scalars), expected_generated_msm(n, seed, scalars))
    # linearity: msm(a) + msm(b) == msm(a + b mod r) on a witness-like second vector
    if log_n == 16:
        b = fx.witness_like_fr(n, seed=2)
        ab = cref.ints_to_limbs([(x + y) % o.R for x, y in zip(cref.limbs_to_ints(scalars), cref.limbs_to_ints(b))], 4)
        lhs = vsp.points_sum(ctx, np.stack([h.multiexp(scalars), h.multiexp(b)]))
        assert np.array_equal(lhs, h.multiexp(ab))


def test_g1_multiexp_one_hot_bucket_at_scale(vsp, ctx):
    """2^21 points under a 0/1 witness: every non-zero digit lands in ONE bucket (32,768 tasks), which exercises the
    sliced two-level collapse beyond its minimum slice length and the warp-aggregated counters of the global sort;
    then the same points with small scalars (values 1..8: a handful of hot buckets).  Checksum as above."""
    n, seed = 1 << 21, 77
    h = vsp.G1Bases.generate(ctx, n, seed)
    k = generated_base_scalars(n, seed).astype(object)
    rng = np.random.default_rng(5)
    bits = np.zeros((n, 4), dtype=np.uint64)
    bits[:, 0] = rng.integers(0, 2, size=n, dtype=np.uint64)
    small = np.zeros((n, 4), dtype=np.uint64)
    small[:, 0] = rng.integers(1, 9, size=n, dtype=np.uint64)
    for sc in (bits, _ones(n), small):
        acc = int((k * sc[:, 0].astype(object)).sum()) % o.R
        assert np.array_equal(h.multiexp(sc), cref.g1_mul(acc, G1_GEN_LIMBS))
    h.close()


@pytest.mark.parametrize("affine", ["1", "0"])
def test_both_accumulate_kernels_match_oracle(vsp, ctx, monkeypatch, affine):
    """VSP_MSM_AFFINE forces the batched-affine bucket accumulation (1: shared division-step inversions, lanes walking
    their task queues) or the XYZZ one (0) whatever the launch size; by default the device-resident task count picks.
    Random G1 / G2 multiexps, the exceptional additions (the same base ten times, P and -P, infinite bases, a finite base
    with x = 0), witness-like scalars and a 2^16-point checksum, each against the oracle."""
    monkeypatch.setenv("VSP_MSM_AFFINE", affine)
    for n in (1, 2, 17, 1000, (1 << 14) + 5):
        bases, scalars = random_g1(n, seed=n + 11), fx.random_fr(n, seed=n + 12)
        assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g1(bases, scalars)), n
    for n in (1, 9, 300, 1 << 10):
        bases, scalars = random_g2(n, seed=n + 13), fx.random_fr(n, seed=n + 14)
        assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g2(bases, scalars)), n
    n = 700
    bases = random_g1(n, seed=81)
    bases[10:20] = bases[10]
    bases[31] = cref.g1_mul(o.R - 1, bases[30])
    bases[40] = 0
    bases[41] = cref.g1_to_limbs([(0, 2)])[0]   # on y^2 = x^3 + 4 with x = 0: finite, though all-zero x looks like infinity
    bases[42] = cref.g1_to_limbs([(0, o.P - 2)])[0]
    scalars = fx.random_fr(n, seed=82)
    scalars[0:5] = 0
    scalars[5:10] = _ones(5)
    scalars[10:20] = scalars[10]
    scalars[30] = scalars[31] = scalars[32]
    scalars[41] = scalars[42] = scalars[43]     # (0, 2) and (0, -2) meet each other and a third point in every bucket
    small = scalars.copy()
    small[:, 1:] = 0
    small[:, 0] &= np.uint64(3)                  # a handful of crowded buckets: long lane queues, many equal-point meetings
    for sc in (scalars, small, _ones(n)):
        assert np.array_equal(vsp.multiexp(ctx, bases, sc), cref.msm_g1(bases, sc))
        h = vsp.G1Bases(ctx, bases)              # the window-table path over the same inputs
        assert np.array_equal(h.multiexp(sc), cref.msm_g1(bases, sc))
        h.close()
    g2 = random_g2(64, seed=83)
    g2[3:9] = g2[3]
    g2[20] = 0
    s2 = fx.random_fr(64, seed=84)
    s2[3:9] = s2[3]
    assert np.array_equal(vsp.multiexp(ctx, g2, s2), cref.msm_g2(g2, s2))
    nb = 1 << 15
    bases, scalars = random_g1(nb, seed=5), fx.witness_like_fr(nb, seed=6)
    assert np.array_equal(vsp.multiexp(ctx, bases, scalars), cref.msm_g1(bases, scalars, mixed=True))
    n, seed = 1 << 16, 4321
    h = vsp.G1Bases.generate(ctx, n, seed)
    scalars = fx.random_fr(n, seed=3)
    assert np.array_equal(h.multiexp(scalars), expected_generated_msm(n, seed, scalars))
    h.close()
