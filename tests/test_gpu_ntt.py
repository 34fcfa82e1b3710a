"""GPU parity: evaluation domains over Fr (fft / inverse_fft / multiply_by_coset / divide_by_z_on_coset) through the
C ABI against the oracle, the reference's known answers (libs/math/test/polynomial_dfs.cpp) and the properties of
libs/math/test/evaluation_domain.cpp.  Bit-exact."""
import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import golden, fr_ints, run_fft_kat

pytestmark = pytest.mark.gpu


def test_fft_known_answers(vsp, ctx):
    """polynomial_dfs.cpp:50-113, 387-581: size-8 / size-16 FFT, iFFT and dfs-multiplication known answers"""
    def fft(m, v):
        return fr_ints(vsp.make_evaluation_domain(ctx, m).fft(cref.ints_to_limbs(v, 4)))

    def ifft(m, v):
        return fr_ints(vsp.make_evaluation_domain(ctx, m).inverse_fft(cref.ints_to_limbs(v, 4)))

    cases = golden("fr_fft_kat.json")["cases"]
    assert len(cases) >= 7
    for case in cases:
        got, want = run_fft_kat(case, fft, ifft)
        assert got == want, case["name"]


@pytest.mark.parametrize("log_n", [1, 2, 3, 4, 5, 7, 9, 10, 11, 12, 13, 14, 16, 18, 20])
def test_basic_radix2_fft_matches_oracle(vsp, ctx, log_n):
    m = 1 << log_n
    a = fx.random_fr(m, seed=100 + log_n)
    dom = vsp.make_evaluation_domain(ctx, m)
    assert dom.kind == vsp.DOMAIN_BASIC_RADIX2 and dom.m == m
    f = dom.fft(a)
    assert np.array_equal(f, cref.domain_fft(m, a, inverse=False))
    assert np.array_equal(dom.inverse_fft(a), cref.domain_fft(m, a, inverse=True))
    assert np.array_equal(dom.inverse_fft(f), a)  # evaluation_domain.cpp:103-121


@pytest.mark.parametrize("log_n", [21, 22, 24])
def test_large_fft_properties(vsp, ctx, log_n):
    """sizes the oracle is slow at: ifft(fft(a)) == a, and linearity fft(a + b) == fft(a) + fft(b) on a sample"""
    m = 1 << log_n
    a = fx.random_fr(m, seed=7)
    dom = vsp.make_evaluation_domain(ctx, m)
    f = dom.fft(a)
    assert np.array_equal(dom.inverse_fft(f), a)
    # a delta at position k transforms to the geometric sequence omega^(k j): checks the root and the output order
    k = 5
    d = np.zeros((m, 4), dtype=np.uint64)
    d[k, 0] = 1
    fd = fr_ints(dom.fft(d)[:64])
    w = o.unity_root(m)
    assert fd == [pow(w, k * j, o.R) for j in range(64)]


@pytest.mark.parametrize("m_req", [3, 5, 6, 12, 24, 33, 40, 96, 1000, 1024 + 64, (1 << 14) + (1 << 10), (1 << 16) + 4,
                                   (1 << 17) + (1 << 16)])
def test_step_and_rounded_domains_match_oracle(vsp, ctx, m_req):
    """make_evaluation_domain.hpp:51-106: non-powers of two use step_radix2 (or round up); results match the oracle"""
    kind, m = cref.domain_plan(m_req)
    dom = vsp.make_evaluation_domain(ctx, m_req)
    assert (dom.kind, dom.m) == (kind, m)
    a = fx.random_fr(m, seed=m_req)
    assert np.array_equal(dom.fft(a), cref.domain_fft(m, a, inverse=False))
    assert np.array_equal(dom.inverse_fft(a), cref.domain_fft(m, a, inverse=True))
    assert np.array_equal(dom.inverse_fft(dom.fft(a)), a)
    z = dom.divide_by_z_on_coset(a)
    assert np.array_equal(z, cref.divide_by_z_on_coset(m, a))


def test_fft_equals_naive_evaluation(vsp, ctx):
    """evaluation_domain.cpp:68-100: fft(f)[i] == f(omega^i) for m = 4"""
    m = 4
    f = [2, 5, 3, 8]
    dom = vsp.make_evaluation_domain(ctx, m)
    got = fr_ints(dom.fft(cref.ints_to_limbs(f, 4)))
    w = o.unity_root(m)
    want = [sum(c * pow(w, i * j, o.R) for j, c in enumerate(f)) % o.R for i in range(m)]
    assert got == want


def test_short_input_is_zero_padded_long_input_rejected(vsp, ctx):
    """basic_radix2_domain.hpp:69-79: a is resized to m; longer input throws std::invalid_argument"""
    dom = vsp.make_evaluation_domain(ctx, 16)
    a = fx.random_fr(11, seed=3)
    padded = np.zeros((16, 4), dtype=np.uint64)
    padded[:11] = a
    assert np.array_equal(dom.fft(a), cref.domain_fft(16, padded))
    with pytest.raises(ValueError):
        dom.fft(fx.random_fr(17, seed=4))
    # step_radix2_domain.hpp:112-114: inverse_fft requires a.size() == m
    step = vsp.make_evaluation_domain(ctx, 24)
    with pytest.raises(ValueError):
        step.inverse_fft(fx.random_fr(20, seed=5))


def test_domain_selection_errors(vsp):
    with pytest.raises(ValueError):
        vsp.domain_plan(1)
    with pytest.raises(ValueError):
        vsp.domain_plan(0)


@pytest.mark.parametrize("n", [1, 2, 7, 8, 1000, 1 << 12, (1 << 15) + 3])
def test_multiply_by_coset_matches_oracle(vsp, ctx, n):
    """coset.hpp:37-43: a[i] *= g^i, a[0] untouched"""
    a = fx.random_fr(n, seed=n)
    for g in (7, pow(7, -1, o.R), 0x1234567890ABCDEF1234567890ABCDEF):
        got = vsp.multiply_by_coset(ctx, a, g)
        assert np.array_equal(got, cref.multiply_by_coset(a, g))
    assert np.array_equal(vsp.multiply_by_coset(ctx, a, 7)[0], a[0])


def test_coset_round_trip(vsp, ctx):
    """evaluation_domain.cpp:124-144: coset FFT then coset iFFT with the multiplicative generator is the identity"""
    m = 1 << 10
    a = fx.random_fr(m, seed=9)
    dom = vsp.make_evaluation_domain(ctx, m)
    g = 7
    c = dom.fft(vsp.multiply_by_coset(ctx, a, g))
    back = vsp.multiply_by_coset(ctx, dom.inverse_fft(c), pow(g, -1, o.R))
    assert np.array_equal(back, a)


@pytest.mark.parametrize("m", [2, 8, 1 << 10, 1 << 16])
def test_divide_by_z_on_coset_basic(vsp, ctx, m):
    a = fx.random_fr(m, seed=m + 1)
    dom = vsp.make_evaluation_domain(ctx, m)
    assert np.array_equal(dom.divide_by_z_on_coset(a), cref.divide_by_z_on_coset(m, a))
