"""CPU tier: the pairing ALGORITHM the GPU uses (csrc/pairing.cu), restated over the oracle's field arithmetic and checked
against the oracle's textbook pairing — sparse scaled lines over the twist (no Fq12 inversions, slopes depend on the G2
argument only), the Frobenius constants and the exact x-chain of the final exponentiation.  The GPU tier then compares
the CUDA implementation bit for bit with the oracle (tests/test_gpu_pairing.py)."""
import random

from oracle import pyref as o

P, R, F2 = o.P, o.R, o._Fq2
X = -o.BLS_X


def _line(c0, c1, py):
    z = (0, 0)
    return ((c0, c1, z), (z, (py % P, 0), z))  # c0 + c1 w^2 + py w^3


def _slope_table(q_aff):
    """the 68 (lambda, lambda Tx - Ty) pairs of a G2 point: 63 doublings + 5 additions (k_g2_line_table)"""
    qx, qy = q_aff
    tx, ty = qx, qy
    out = []
    for bit in bin(o.BLS_X)[3:]:
        lam = F2.mul(F2.mul((3, 0), F2.sqr(tx)), F2.inv(F2.add(ty, ty)))
        out.append((lam, F2.sub(F2.mul(lam, tx), ty)))
        nx = F2.sub(F2.sqr(lam), F2.add(tx, tx))
        tx, ty = nx, F2.sub(F2.mul(lam, F2.sub(tx, nx)), ty)
        if bit == "1":
            lam = F2.mul(F2.sub(qy, ty), F2.inv(F2.sub(qx, tx)))
            out.append((lam, F2.sub(F2.mul(lam, tx), ty)))
            nx = F2.sub(F2.sub(F2.sqr(lam), tx), qx)
            tx, ty = nx, F2.sub(F2.mul(lam, F2.sub(tx, nx)), ty)
    return out


def _multi_miller(pairs):
    """shared squarings over a product of pairings (k_multi_miller)"""
    tabs = [(p, _slope_table(q)) for p, q in pairs]
    f, k = o.F12_ONE, 0
    for bit in bin(o.BLS_X)[3:]:
        f = o.f12mul(f, f)
        for _ in range(2 if bit == "1" else 1):
            for (px, py), tab in tabs:
                lam, c0 = tab[k]
                f = o.f12mul(f, _line(c0, F2.neg(F2.mul(lam, (px, 0))), py))
            k += 1
    assert k == 68
    return o.f12conj(f)


def _f2pow(a, e):
    r = (1, 0)
    for b in bin(e)[2:]:
        r = F2.mul(r, r)
        if b == "1":
            r = F2.mul(r, a)
    return r


_G1C = [_f2pow((1, 1), i * (P - 1) // 6) for i in range(6)]
_G2C = [_f2pow((1, 1), i * (P * P - 1) // 6) for i in range(6)]


def _coeffs(f):
    c0, c1 = f
    return [c0[0], c1[0], c0[1], c1[1], c0[2], c1[2]]


def _from_coeffs(a):
    return ((a[0], a[2], a[4]), (a[1], a[3], a[5]))


def _frob(f):
    return _from_coeffs([F2.mul((a[0], (-a[1]) % P), _G1C[i]) for i, a in enumerate(_coeffs(f))])


def _frob2(f):
    return _from_coeffs([F2.mul(a, _G2C[i]) for i, a in enumerate(_coeffs(f))])


def _final_exp_x_chain(f):
    e0 = (X - 1) ** 2 // 3
    powx = lambda a: o.f12conj(o.f12pow(a, -X))
    t = o.f12mul(o.f12conj(f), o.f12inv(f))
    f1 = o.f12mul(_frob2(t), t)
    a = o.f12pow(f1, e0)
    a = o.f12mul(o.f12mul(a, a), a)  # f1^((x - 1)^2): crypto3's exponent is 3 x the exact one (oracle/pyref.py)
    b = o.f12mul(powx(a), _frob(a))
    c = o.f12mul(o.f12mul(powx(powx(b)), _frob2(b)), o.f12conj(b))
    return o.f12mul(c, o.f12mul(o.f12mul(f1, f1), f1))


def test_hard_part_identity_and_frobenius_constants():
    assert (X - 1) ** 2 % 3 == 0
    assert (P ** 4 - P ** 2 + 1) // R == ((X - 1) ** 2 // 3) * (X + P) * (X * X + P * P - 1) + 1
    assert all(g[1] == 0 for g in _G2C)  # the p^2 constants lie in Fq
    rng = random.Random(5)
    f = tuple(tuple((rng.randrange(P), rng.randrange(P)) for _ in range(3)) for _ in range(2))
    assert _frob(f) == o.f12pow(f, P)
    assert _frob2(f) == o.f12pow(f, P * P)


def test_sparse_line_miller_loop_and_x_chain_match_the_textbook_pairing():
    rng = random.Random(1)
    pts = []
    for _ in range(2):
        p = o.G1.to_affine(o.G1.mul(rng.randrange(1, R), o.g1_generator()))
        q = o.G2.to_affine(o.G2.mul(rng.randrange(1, R), o.g2_generator()))
        pts.append((p, q))
    want = o.F12_ONE
    for p, q in pts:
        want = o.f12mul(want, o.miller_loop(p, q))
    want = o.final_exponentiation(want)
    assert _final_exp_x_chain(_multi_miller(pts)) == want
