"""CPU tier: pin the oracle (C restatement + Python big-int restatement) on the reference's own known answers and on
each other, and run the self-consistency gate for the unpinned parts (generate -> prove -> verify; SAVER round trip).

Golden vectors: tests/golden/*.json, extracted from /root/reference by tests/golden/make_golden.py:
  libs/math/test/polynomial_dfs.cpp:50-113, 387-581      Fr FFT / iFFT / dfs multiplication
  libs/pubkey/test/bls.cpp:111-303, 669-998              G1 / G2 points, encodings, ten-point aggregates
"""
import random

import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import (fr_ints, golden, hexlist, random_g1, random_g2, run_fft_kat, synthetic_pk_arrays, synthetic_r1cs,
                     synthetic_saver_pub)


# ---- Fr FFT ------------------------------------------------------------------------------------
def _c_fft(m, v):
    return fr_ints(cref.domain_fft(m, cref.ints_to_limbs(v, 4)))


def _c_ifft(m, v):
    return fr_ints(cref.domain_fft(m, cref.ints_to_limbs(v, 4), inverse=True))


def _py_fft(m, v):
    a = list(v) + [0] * (m - len(v))
    o.make_evaluation_domain(m).fft(a)
    return a


def _py_ifft(m, v):
    a = list(v)
    o.make_evaluation_domain(m).inverse_fft(a)
    return a


@pytest.mark.parametrize("impl", ["c", "py"])
def test_fr_fft_known_answers(impl):
    fft, ifft = (_c_fft, _c_ifft) if impl == "c" else (_py_fft, _py_ifft)
    cases = golden("fr_fft_kat.json")["cases"]
    assert len(cases) == 7
    for case in cases:
        got, want = run_fft_kat(case, fft, ifft)
        assert got == want, case["name"]


@pytest.mark.parametrize("m_req", [2, 4, 8, 64, 1 << 10, 3, 6, 12, 24, 40, 1000, 1024 + 64])
def test_c_and_python_domains_agree(m_req):
    kind, m = cref.domain_plan(m_req)
    pk, pm = o.domain_plan(m_req)
    assert pm == m
    a = fx.random_fr(m, seed=m_req)
    ints = fr_ints(a)
    dom = o.make_evaluation_domain(m_req)
    f = list(ints)
    dom.fft(f)
    assert fr_ints(cref.domain_fft(m, a)) == f
    g = list(ints)
    dom.inverse_fft(g)
    assert fr_ints(cref.domain_fft(m, a, inverse=True)) == g
    z = list(ints)
    dom.divide_by_z_on_coset(z)
    assert fr_ints(cref.divide_by_z_on_coset(m, a)) == z
    c = list(ints)
    o.multiply_by_coset(c, 7)
    assert fr_ints(cref.multiply_by_coset(a, 7)) == c


def test_domain_selection_rules():
    """make_evaluation_domain.hpp:51-106 for s = 32"""
    assert cref.domain_plan(8) == (1, 8)
    assert cref.domain_plan(24) == (3, 24)          # big + 2^j exactly -> step
    assert cref.domain_plan(5) == (3, 5)            # small_m = 1 = 2^0 is a legal step domain (type_traits.hpp:74-85)
    assert cref.domain_plan(1000) == (1, 1024)      # big + rounded_small = 512 + 512 -> basic
    assert cref.domain_plan(1100) == (3, 1024 + 128)
    assert cref.domain_plan(5000) == (3, 4096 + 1024)
    with pytest.raises(ValueError):
        cref.domain_plan(1)


# ---- curves ------------------------------------------------------------------------------------
def test_g1_g2_aggregate_known_answers():
    kat = golden("curve_kat.json")
    ones = np.zeros((10, 4), np.uint64)
    ones[:, 0] = 1
    g1 = kat["g1_aggregate"]
    pts = cref.g1_to_limbs([(int(p["x"], 16), int(p["y"], 16)) for p in g1["points"]])
    want = cref.g1_to_limbs([(int(g1["sum"]["x"], 16), int(g1["sum"]["y"], 16))])[0]
    assert np.array_equal(cref.msm_g1(pts, ones), want)
    acc = pts[0]
    for p in pts[1:]:
        acc = cref.g1_add(acc, p)
    assert np.array_equal(acc, want)
    g2 = kat["g2_aggregate"]
    f2 = lambda v: tuple(hexlist(v))
    pts2 = cref.g2_to_limbs([(f2(p["x"]), f2(p["y"])) for p in g2["points"]])
    want2 = cref.g2_to_limbs([(f2(g2["sum"]["x"]), f2(g2["sum"]["y"]))])[0]
    assert np.array_equal(cref.msm_g2(pts2, ones), want2)
    # Python oracle too
    acc = o.G1.inf()
    for p in g1["points"]:
        acc = o.G1.add(acc, o.G1.from_affine((int(p["x"], 16), int(p["y"], 16))))
    assert o.G1.to_affine(acc) == (int(g1["sum"]["x"], 16), int(g1["sum"]["y"], 16))


def test_g1_serialization_known_answers():
    """bls.cpp:111-196: Jacobian p3 normalises onto the curve; p4's blst-generated encodings"""
    kat = golden("curve_kat.json")["g1_serialization"]
    x, y, z = hexlist(kat["p3_jacobian"])
    aff = o.G1.to_affine((x, y, z))
    assert (aff[1] * aff[1] - aff[0] ** 3 - 4) % o.P == 0
    got = cref.g1_jacobian_to_affine(cref.ints_to_limbs([x, y, z], 6).reshape(-1))
    assert cref.g1_from_limbs(got)[0] == aff
    p4 = tuple(hexlist(kat["p4_affine"])[:2])
    assert o.g1_uncompressed(o.G1.from_affine(p4)).hex() == kat["p4_octets"]
    assert o.g1_compress(o.G1.from_affine(p4)).hex() == kat["p4_octets_compressed"]


def test_g2_serialization_known_answers():
    kat = golden("curve_kat.json")["g2_serialization"]
    v = hexlist(kat["p4_affine"])
    p4 = ((v[0], v[1]), (v[2], v[3]))
    assert o.g2_uncompressed(o.G2.from_affine(p4)).hex() == kat["p4_octets"]
    assert o.g2_compress(o.G2.from_affine(p4)).hex() == kat["p4_octets_compressed"]


@pytest.mark.parametrize("n", [1, 5, 64, 700])
def test_c_multiexp_agrees_with_python_and_naive(n):
    bases = random_g1(n, seed=n)
    scalars = fx.random_fr(n, seed=n + 1)
    got = cref.msm_g1(bases, scalars)
    if n <= 64:
        pts = [o.G1.from_affine(p) for p in cref.g1_from_limbs(bases)]
        want = o.multiexp_bdlo12(o.G1, pts, fr_ints(scalars))
        assert cref.g1_from_limbs(got)[0] == o.G1.to_affine(want)
    # chunked (MULTICORE-style) and mixed-addition variants give the same point
    assert np.array_equal(cref.msm_g1(bases, scalars, chunks=4), got)
    assert np.array_equal(cref.msm_g1(bases, scalars, mixed=True), got)
    acc = np.zeros(12, np.uint64)
    for i in range(min(n, 64)):
        acc = cref.g1_add(acc, cref.g1_mul(fr_ints(scalars[i:i + 1])[0], bases[i]))
    if n <= 64:
        assert np.array_equal(acc, got)


def test_c_g2_multiexp_agrees_with_naive():
    n = 20
    bases = random_g2(n, seed=3)
    scalars = fx.random_fr(n, seed=4)
    acc = np.zeros(24, np.uint64)
    for i in range(n):
        acc = cref.g2_add(acc, cref.g2_mul(fr_ints(scalars[i:i + 1])[0], bases[i]))
    assert np.array_equal(cref.msm_g2(bases, scalars), acc)


# ---- prover self-consistency (the gate for the UPSTREAM-RECALL parts) -----------------------------
def test_generate_prove_verify_and_c_prover_agrees():
    """r1cs_gg_ppzksnark.cpp:62-164 style: an oracle proof verifies under the oracle's pairing check, and the C
    prover emits the same bytes as the Python one for fixed r, s."""
    rng = random.Random(11)
    cs, primary, aux = o.r1cs_example_with_field_input(20, 4, rng)
    pk, vk = o.groth16_generate(cs, rng)
    r, s = rng.randrange(o.R), rng.randrange(o.R)
    proof = o.groth16_prove(pk, primary, aux, r, s)
    assert o.groth16_verify(vk, primary, proof)
    bad = list(primary)
    bad[0] = (bad[0] + 1) % o.R
    assert not o.groth16_verify(vk, bad, proof)
    arr = fx.pk_arrays(pk)
    got = cref.groth16_prove(fx.fill_pk(cref.ProvingKeyDesc(), arr), cref.ints_to_limbs(primary, 4),
                             cref.ints_to_limbs(aux, 4), r, s)
    assert np.array_equal(got, fx.proof_to_limbs(proof))


def test_reference_made_proof_in_data_bin_verifies():
    """bin/cli/src/data.bin holds the one crypto3-made Groth16 proof of the tree, followed by its verification key
    (tests/golden/make_golden.py::databin).  It pins, against bytes the reference wrote: the compressed point decoders, the
    192-byte proof layout, the verification key's field order, the orientation of the verification equation and the value
    convention of pair_reduced (alpha_g1_beta_g2 is compared coefficient for coefficient, not through a decision)."""
    g = golden("databin_groth16.json")
    pr = bytes.fromhex(g["proof"])
    A, B, C = o.g1_decompress(pr[:48]), o.g2_decompress(pr[48:144]), o.g1_decompress(pr[144:])
    gamma, delta = o.g2_decompress(bytes.fromhex(g["gamma_g2"])), o.g2_decompress(bytes.fromhex(g["delta_g2"]))
    first = o.g1_decompress(bytes.fromhex(g["gamma_ABC_first"]))
    J1, J2 = o.G1.from_affine, o.G2.from_affine
    for pt in (A, C, first):
        assert o.G1.on_curve_affine(pt) and o.G1.in_subgroup(J1(pt))
    for pt in (B, gamma, delta):
        assert o.G2.on_curve_affine(pt) and o.G2.in_subgroup(J2(pt))
    # the decoders invert the encoders bit for bit
    assert o.g1_compress(J1(A)) + o.g2_compress(J2(B)) + o.g1_compress(J1(C)) == pr
    gt = o.gt_from_coeffs([int(x, 16) for x in g["alpha_g1_beta_g2"]])
    proof = (J1(A), J2(B), J1(C))
    assert o.groth16_verify_gt(gt, J2(gamma), J2(delta), J1(first), [], [], proof)
    assert not o.groth16_verify_gt(gt, J2(gamma), J2(delta), J1(first), [], [], (J1(C), J2(B), J1(A)))
    # the exact (p^12 - 1) / r power does NOT reproduce the stored element: crypto3's reduced pairing is its cube
    f = o.F12_ONE
    for p_, q_ in ((A, B), (o.G1.to_affine(o.G1.neg(J1(first))), gamma), (o.G1.to_affine(o.G1.neg(J1(C))), delta)):
        f = o.f12mul(f, o.miller_loop(p_, q_))
    x = o.final_exponentiation_exact(f)
    assert x != gt and o.f12mul(o.f12mul(x, x), x) == gt


def test_step_domain_prover_verifies():
    rng = random.Random(12)
    cs, primary, aux = o.r1cs_example_with_field_input(19, 4, rng)  # m_req = 24 -> step_radix2
    assert o.domain_plan(24)[1] == 24
    pk, vk = o.groth16_generate(cs, rng)
    proof = o.groth16_prove(pk, primary, aux, 5, 6)
    assert o.groth16_verify(vk, primary, proof)


def test_c_witness_map_agrees_with_python():
    for nc, ni in ((10, 5), (18, 5), (60, 3)):
        cs, primary, aux = synthetic_r1cs(nc, ni, seed=nc)
        _, H, _ = o.witness_map(cs, primary, aux)
        arr = fx.r1cs_arrays(cs)
        got = cref.witness_map(fx.fill_r1cs(cref.R1CSDesc(), arr), cref.ints_to_limbs(primary, 4),
                               cref.ints_to_limbs(aux, 4))
        assert fr_ints(got) == H


def test_saver_c_agrees_with_python():
    """elgamal_verifiable.hpp:341-352, 632-667 on a synthetic key (shape only)"""
    msg_size = 7  # elgamal_verifiable.cpp:339, 369 uses a 7-slot message
    pub_arr, gabc = synthetic_saver_pub(msg_size, seed=1)
    gabc = np.ascontiguousarray(gabc[:msg_size])
    rng = random.Random(2)
    msg_i = [rng.randrange(16) for _ in range(msg_size)]
    r = rng.randrange(o.R)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    got = cref.saver_ciphertext(opub, gabc, cref.ints_to_limbs(msg_i, 4), r)
    J = lambda a: [o.G1.from_affine(p) if p else o.G1.inf() for p in cref.g1_from_limbs(a)]
    pub = o.SaverPublicKey(J(pub_arr["delta_g1"])[0], J(pub_arr["delta_s_g1"]), J(pub_arr["t_g1"]), [],
                           J(pub_arr["delta_sum_s_g1"])[0], J(pub_arr["gamma_inverse_sum_s_g1"])[0])
    want = o.saver_ciphertext(pub, J(gabc), msg_i, r)
    assert cref.g1_from_limbs(got) == [o.G1.to_affine(p) for p in want]


def test_saver_encrypt_verifies_and_survives_rerandomisation():
    """elgamal_verifiable.cpp:353-489 style round trip on the oracle: keygen -> encrypt (ciphertext + encrypted_input
    proof, g_C += r_saver * gamma_inverse_sum_s_g1) -> verify_encryption true; a tampered ciphertext is rejected; the
    same after rerandomize.  This is the gate for the UPSTREAM-RECALL encrypted_input adjustment."""
    rng = random.Random(31)
    msg_size = 3
    cs, primary, aux = o.r1cs_example_with_field_input(12, 5, rng)
    pk, vk = o.groth16_generate(cs, rng)
    pub, _sk, _svk = o.saver_keygen(vk, msg_size, [rng.randrange(1, o.R) for _ in range(3 * msg_size + 2)])
    r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
    plain = list(primary[:msg_size])
    ct, proof = o.saver_encrypt(pk, vk, pub, plain, primary, aux, r_saver, r, s)
    rest = list(primary[msg_size:])
    assert o.saver_verify_encryption(vk, pub, ct, rest, proof)
    bad = list(ct)
    bad[1] = o.G1.add(bad[1], o.g1_generator())
    assert not o.saver_verify_encryption(vk, pub, bad, rest, proof)
    rnd = [rng.randrange(1, o.R) for _ in range(3)]
    ct2, proof2 = o.saver_rerandomize(pub, pk.delta_g2, ct, proof, rnd)
    assert ct2 != ct
    assert o.saver_verify_encryption(vk, pub, ct2, rest, proof2)


def test_saver_decrypt_round_trip_on_the_oracle():
    """elgamal_verifiable.cpp:442-489 style on the oracle: keygen -> ciphertext -> decrypt recovers the message and
    verify_decryption accepts it (elgamal_verifiable.hpp:401-441, 548-575 restated from source); a wrong message or a
    wrong decryption proof is rejected, and a message outside the brute-force range fails like the reference asserts."""
    rng = random.Random(41)
    msg_size = 2
    cs, primary, aux = o.r1cs_example_with_field_input(8, 3, rng)
    _pk, vk = o.groth16_generate(cs, rng)
    pub, rho, svk = o.saver_keygen(vk, msg_size, [rng.randrange(1, o.R) for _ in range(3 * msg_size + 2)])
    plain = [3, 0]
    ct = o.saver_ciphertext(pub, vk.gamma_ABC_rest, plain, rng.randrange(1, o.R))
    m, c0 = o.saver_decrypt(ct, rho, svk, vk.gamma_ABC_rest, block_bits=2)
    assert m == plain
    assert o.saver_verify_decryption(ct, plain, svk, vk.gamma_ABC_rest, c0)
    assert not o.saver_verify_decryption(ct, [2, 0], svk, vk.gamma_ABC_rest, c0)
    assert not o.saver_verify_decryption(ct, plain, svk, vk.gamma_ABC_rest, o.G1.add(c0, o.g1_generator()))
    ct6 = o.saver_ciphertext(pub, vk.gamma_ABC_rest, [6, 0], rng.randrange(1, o.R))
    with pytest.raises(ValueError):
        o.saver_decrypt(ct6, rho, svk, vk.gamma_ABC_rest, block_bits=2)
