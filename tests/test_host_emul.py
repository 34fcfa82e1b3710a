"""CPU tier: the device field / curve templates (csrc/fp.cuh, csrc/ec.cuh) compiled with g++ under VSP_HOST_EMUL —
PTX mad.lo.cc / madc.hi.cc / addc / subc emulated bit for bit — checked against Python big-int arithmetic.  Covers
the limb schedules (even/odd-column Montgomery product, final reduction) and the XYZZ formulas with their
exceptional cases (doubling inside add, P + (-P), infinity operands) without a GPU."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

from oracle import pyref as o
from oracle.cref import g1_from_limbs, g1_to_limbs, g2_from_limbs, g2_to_limbs, ints_to_limbs, limbs_to_ints

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul")
P = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint64))


@pytest.fixture(scope="module")
def emul():
    subprocess.check_call(["make", "-C", HERE, "-s"])
    return C.CDLL(os.path.join(HERE, "libemul.so"))


@pytest.mark.parametrize("name,mod,limbs", [("fr", o.R, 4), ("fq", o.P, 6)])
def test_field_ops(emul, name, mod, limbs):
    fn = getattr(emul, "emul_%s_op" % name)
    rng = random.Random(limbs)
    edge = [0, 1, 2, mod - 1, mod - 2, (1 << (32 * limbs)) % mod, mod >> 1]
    pairs = [(a, b) for a in edge for b in edge] + [(rng.randrange(mod), rng.randrange(mod)) for _ in range(400)]
    for i, (a, b) in enumerate(pairs):
        A, B, O = ints_to_limbs([a], limbs), ints_to_limbs([b], limbs), np.zeros((1, limbs), np.uint64)
        want = [a * b % mod, (a + b) % mod, (a - b) % mod, (-a) % mod, pow(a, mod - 2, mod), a * a % mod,
                pow(a, mod - 2, mod),  # op 6: the binary-Euclid inverse (fp_inv_gcd), every pair
                pow(a, mod - 2, mod)]  # op 7: the constant-time division-step inverse (fp_inv_safegcd), every pair
        for op in range(8):
            if op == 4 and i % 40:
                continue
            fn(op, P(A), P(B), P(O))
            assert limbs_to_ints(O)[0] == want[op], (name, op, hex(a), hex(b))


@pytest.mark.parametrize("name,mod,limbs", [("fr", o.R, 4), ("fq", o.P, 6)])
def test_raw_montgomery_product(emul, name, mod, limbs):
    fn = getattr(emul, "emul_%s_montmul_raw" % name)
    rinv = pow(1 << (64 * limbs), -1, mod)
    rng = random.Random(7)
    for _ in range(300):
        a, b = rng.randrange(mod), rng.randrange(mod)
        A, B, O = ints_to_limbs([a], limbs), ints_to_limbs([b], limbs), np.zeros((1, limbs), np.uint64)
        fn(P(A), P(B), P(O))
        assert limbs_to_ints(O)[0] == a * b * rinv % mod


@pytest.mark.parametrize("name,mod,limbs", [("fr", o.R, 4), ("fq", o.P, 6)])
def test_raw_montgomery_square(emul, name, mod, limbs):
    """fp_sqr has its own limb schedule (cross products once, doubled; sliding-frame reduction): raw < p inputs whose
    32-bit limbs are biased to all-ones / zero / one so every carry path of the chains fires"""
    fn = getattr(emul, "emul_%s_montsqr_raw" % name)
    rinv = pow(1 << (64 * limbs), -1, mod)
    rng = random.Random(11)
    pool = [0xffffffff, 0, 1, 0x80000000, 0x7fffffff, 0xfffffffe]
    vals = [0, 1, mod - 1, mod - 2, mod >> 1, (1 << (32 * limbs)) % mod]
    while len(vals) < 1500:
        ws = [rng.choice(pool) if rng.random() < 0.7 else rng.getrandbits(32) for _ in range(2 * limbs)]
        v = sum(w << (32 * i) for i, w in enumerate(ws))
        if len(vals) % 3 == 0:
            v = rng.randrange(mod)
        vals.append(v % mod if v >= mod and rng.random() < 0.5 else (v if v < mod else v & ((1 << (mod.bit_length() - 1)) - 1)))
    for a in vals:
        A, O = ints_to_limbs([a], limbs), np.zeros((1, limbs), np.uint64)
        fn(P(A), P(O))
        assert limbs_to_ints(O)[0] == a * a * rinv % mod, (name, hex(a))


@pytest.mark.parametrize("grp", ["g1", "g2"])
def test_curve_ops(emul, grp):
    fn = getattr(emul, "emul_%s_op" % grp)
    cv, gen, tl, fl, w = ((o.G1, o.g1_generator(), g1_to_limbs, g1_from_limbs, 12) if grp == "g1" else
                          (o.G2, o.g2_generator(), g2_to_limbs, g2_from_limbs, 24))
    rng = random.Random(2)
    pts = [cv.to_affine(cv.mul(rng.randrange(1, o.R), gen)) for _ in range(6)]
    neg3 = cv.to_affine(cv.neg(cv.from_affine(pts[3])))
    cases = [(pts[0], pts[1]), (pts[2], pts[2]), (pts[3], neg3), (None, pts[4]), (pts[4], None), (None, None),
             (pts[5], pts[0])]
    for a, b in cases:
        A, B, O = tl([a]), tl([b]), np.zeros((1, w), np.uint64)
        ja = cv.from_affine(a) if a else cv.inf()
        jb = cv.from_affine(b) if b else cv.inf()
        for op in range(5):
            k = rng.randrange(1, 70000) if op == 3 else 0
            fn(op, P(A), P(B), C.c_uint32(k), P(O))
            want = {0: cv.add(ja, jb), 1: cv.add(ja, jb), 2: cv.dbl(ja), 3: cv.mul(k, ja), 4: cv.dbl(ja)}[op]
            want = None if cv.is_inf(want) else cv.to_affine(want)
            assert fl(O)[0] == want, (grp, op)


@pytest.mark.parametrize("name,mod,limbs", [("fr", o.R, 4), ("fq", o.P, 6)])
def test_division_step_inverse_stress(emul, name, mod, limbs):
    """fp_inv_safegcd (the inverse shared by a batch of affine bucket additions): powers of two, values next to 0 and p,
    sparse and dense bit patterns, and a few thousand random elements, against pow(a, -1, p)"""
    fn = getattr(emul, "emul_%s_op" % name)
    rng = random.Random(99 + limbs)
    vals = [pow(2, k, mod) for k in range(0, 64 * limbs + 3, 7)] + [mod - pow(2, k, mod) for k in range(0, 64 * limbs, 29)]
    vals += [(1 << k) - 1 for k in range(1, 64 * limbs - 3, 31)] + [rng.getrandbits(rng.randrange(1, 64 * limbs)) % mod for _ in range(500)]
    vals += [rng.randrange(mod) for _ in range(2500)]
    A, O = np.zeros((1, limbs), np.uint64), np.zeros((1, limbs), np.uint64)
    for a in vals:
        a %= mod
        A[:] = ints_to_limbs([a], limbs)
        fn(7, P(A), P(A), P(O))
        assert limbs_to_ints(O)[0] == (pow(a, -1, mod) if a else 0), hex(a)


@pytest.mark.parametrize("name,mod,limbs", [("fr", o.R, 4), ("fq", o.P, 6)])
def test_double_width_product_and_shared_reduction(emul, name, mod, limbs):
    """fp_mul_wide + fp_redc (lazy reduction): operands up to 2p - 2 (an unreduced sum), limb patterns that fire every carry
    path, and a b - c d on one reduction, against Python integers"""
    fn = getattr(emul, "emul_%s_wide" % name)
    rinv = pow(1 << (64 * limbs), -1, mod)
    rng = random.Random(17 + limbs)
    pat = [0, 0xffffffff, 1, 0x80000000]

    def biased(bound):
        while True:
            v = sum(rng.choice(pat + [rng.getrandbits(32)]) << (32 * i) for i in range(2 * limbs))
            v %= bound
            if v:
                return v

    Z = np.zeros((1, limbs), np.uint64)
    for it in range(600):
        a, b = (biased(2 * mod - 1), biased(2 * mod - 1)) if it % 2 else (rng.randrange(mod), rng.randrange(mod))
        O = np.zeros((1, limbs), np.uint64)
        fn(0, P(ints_to_limbs([a], limbs)), P(ints_to_limbs([b], limbs)), P(Z), P(Z), P(O))
        assert limbs_to_ints(O)[0] == a * b * rinv % mod, (hex(a), hex(b))
        a, b, c, d = (biased(mod) for _ in range(4)) if it % 2 else (rng.randrange(mod) for _ in range(4))
        fn(1, P(ints_to_limbs([a], limbs)), P(ints_to_limbs([b], limbs)), P(ints_to_limbs([c], limbs)), P(ints_to_limbs([d], limbs)), P(O))
        assert limbs_to_ints(O)[0] == (a * b - c * d) * rinv % mod
    for a, b, c, d in ((mod - 1, mod - 1, 0, 0), (0, 0, mod - 1, mod - 1), (mod - 1, mod - 1, mod - 1, mod - 1), (1, 1, mod - 1, mod - 1)):
        O = np.zeros((1, limbs), np.uint64)
        fn(1, P(ints_to_limbs([a], limbs)), P(ints_to_limbs([b], limbs)), P(ints_to_limbs([c], limbs)), P(ints_to_limbs([d], limbs)), P(O))
        assert limbs_to_ints(O)[0] == (a * b - c * d) * rinv % mod


def test_fq2_lazy_products(emul):
    """Fq2Ops::mul / sqr / mul_sub_mul (three double-width products, two reductions) on raw Montgomery-domain limbs"""
    p = o.P
    rinv = pow(1 << 384, -1, p)
    rng = random.Random(23)
    edge = [0, 1, p - 1, p - 2, (1 << 384) % p]

    def f2mul(x, y):
        return ((x[0] * y[0] - x[1] * y[1]) % p, (x[0] * y[1] + x[1] * y[0]) % p)

    def pack(x):
        return np.concatenate([ints_to_limbs([x[0]], 6)[0], ints_to_limbs([x[1]], 6)[0]])

    cases = [tuple((rng.choice(edge), rng.choice(edge)) for _ in range(4)) for _ in range(60)]
    cases += [tuple((rng.randrange(p), rng.randrange(p)) for _ in range(4)) for _ in range(300)]
    for a, b, c, d in cases:
        O = np.zeros(12, np.uint64)
        A, B, Cc, D = pack(a), pack(b), pack(c), pack(d)
        emul.emul_fq2_raw(0, P(A), P(B), P(Cc), P(D), P(O))
        want = f2mul(a, b)
        assert tuple(limbs_to_ints(O.reshape(2, 6))) == (want[0] * rinv % p, want[1] * rinv % p)
        emul.emul_fq2_raw(1, P(A), P(B), P(Cc), P(D), P(O))
        want = f2mul(a, a)
        assert tuple(limbs_to_ints(O.reshape(2, 6))) == (want[0] * rinv % p, want[1] * rinv % p)
        emul.emul_fq2_raw(2, P(A), P(B), P(Cc), P(D), P(O))
        ab, cd = f2mul(a, b), f2mul(c, d)
        assert tuple(limbs_to_ints(O.reshape(2, 6))) == ((ab[0] - cd[0]) * rinv % p, (ab[1] - cd[1]) * rinv % p)


def test_scratch_cache_size_classes(emul):
    """csrc/scratch_class.h, the size classes of the context's scratch cache: a class holds its request with at most 1/8 of
    slack, is its own class (so a released block is found again by the same request), never shrinks with the request, and
    there are eight of them per power of two"""
    f = emul.emul_scratch_size_class
    f.restype, f.argtypes = C.c_uint64, [C.c_uint64]
    rng = random.Random(5)
    sizes = [1, 16, 511, 512, 513, 1023, 1024, 1025, (1 << 20) - 1, 1 << 20, (1 << 20) + 1, 3 << 30, (1 << 33) + 5, 1 << 40]
    sizes += [rng.randrange(1, 1 << rng.randrange(10, 42)) for _ in range(20000)]
    for n in sizes:
        c = f(n)
        assert c >= n and c >= 512 and c % 64 == 0, n
        assert n <= 512 or c * 8 <= n * 9, (n, c)   # at most 1/8 above the request
        assert f(c) == c, (n, c)                     # a class is a fixed point
    assert all(f(n) <= f(n + 1) for n in range(1, 70000))
    for top in (12, 20, 31, 36):
        classes = {f(n) for n in range((1 << top) + 1, (1 << (top + 1)) + 1, max(1, 1 << (top - 9)))}
        assert len(classes) == 8, top
