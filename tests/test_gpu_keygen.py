"""GPU key generation against the oracle: the Groth16 generator (zk::generate, call site
bin/cli/include/nil/vote_saver/common.hpp:916-917) and the SAVER key pair (generate_keypair_op::process,
libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:218-271), bit for bit on fixed randomness, plus the
reference tests' own gate (libs/marshalling/zk/test/r1cs_gg_ppzksnark.cpp:160-163, libs/pubkey/test/elgamal_verifiable.cpp:
442-489): a proof made end to end on the GPU from a GPU-generated key passes the oracle's pairing check."""
import importlib
import random

import numpy as np
import pytest

from helpers import fr_ints, synthetic_r1cs
from oracle import cref, fixtures as fx, pyref as o

pytestmark = pytest.mark.gpu
vsp = importlib.import_module("vote-saver-protocol_b200")


@pytest.fixture(scope="module")
def ctx():
    with vsp.Context(0) as c:
        yield c


def _aff1(p):
    return o.G1.to_affine(p)


def _toxic(seed):
    rng = random.Random(seed)
    return [rng.randrange(1, o.R) for _ in range(5)]


def _vk_from_arrays(vk):
    J1 = lambda a: [o.G1.from_affine(p) if p else o.G1.inf() for p in cref.g1_from_limbs(a)]
    J2 = lambda a: [o.G2.from_affine(p) if p else o.G2.inf() for p in cref.g2_from_limbs(a)]
    gabc = J1(vk["gamma_ABC_g1"])
    return o.VerificationKey(J1(vk["alpha_g1"])[0], J2(vk["beta_g2"])[0], J2(vk["gamma_g2"])[0], J2(vk["delta_g2"])[0],
                             J1(vk["gamma_g1"])[0], J1(vk["delta_g1"])[0], gabc[0], gabc[1:])


def test_generator_multiples_match_oracle(ctx):
    """scalar * G1::one() / G2::one(): edge scalars and random ones against the oracle's double-and-add"""
    sc = fx.random_fr(61, seed=5)
    edge = cref.ints_to_limbs([0, 1, 2, 255, 256, o.R - 1, 1 << 254, (1 << 248) - 1], 4)
    sc = np.concatenate([edge, sc])
    g1 = cref.g1_to_limbs([o.G1.to_affine(o.g1_generator())])[0]
    g2 = cref.g2_to_limbs([o.G2.to_affine(o.g2_generator())])[0]
    got1 = vsp.g1_generator_mul_batch(ctx, sc)
    got2 = vsp.g2_generator_mul_batch(ctx, sc[:24])
    for i, k in enumerate(fr_ints(sc)):
        assert np.array_equal(got1[i], cref.g1_mul(k, g1)), i
        if i < 24:
            assert np.array_equal(got2[i], cref.g2_mul(k, g2)), i
    assert vsp.g1_generator_mul_batch(ctx, sc[:0]).shape == (0, 12)


def test_variable_base_multiples_match_oracle(ctx):
    from helpers import random_g1
    pts = random_g1(9, seed=3)
    sc = fx.random_fr(9, seed=4)
    got = vsp.g1_mul_batch(ctx, pts, sc)
    for i, k in enumerate(fr_ints(sc)):
        assert np.array_equal(got[i], cref.g1_mul(k, pts[i]))
    # the G1 jobs split k = k1 + k2 * lambda, lambda = z^2 - 1 (endomorphism): scalars around every boundary of the split
    z2 = 0xd201000000010000 ** 2
    lam = z2 - 1
    assert (lam * lam + lam + 1) % o.R == 0
    edge = [0, 1, 2, lam, z2, z2 - 2, z2 + 1, 2 * z2 - 1, (1 << 128) - 1, 1 << 128, (1 << 128) + 1, (o.R - 1) // z2 * z2,
            (o.R - 1) // z2 * z2 - 1, o.R - 1, o.R - 2, (o.R - 1) // 2, lam * lam % o.R, (1 << 254) + 12345]
    epts = random_g1(len(edge), seed=6)
    got = vsp.g1_mul_batch(ctx, epts, cref.ints_to_limbs(edge, 4))
    for i, k in enumerate(edge):
        assert np.array_equal(got[i], cref.g1_mul(k, epts[i])), hex(k)


def _check_keypair(ctx, cs, toxic):
    pk_want, vk_want = o.groth16_generate(cs, None, toxic=toxic)
    pk, vk = vsp.generate(ctx, fx.r1cs_arrays(cs), toxic)
    want = fx.pk_arrays(pk_want)
    for name in ("alpha_g1", "beta_g1", "beta_g2", "delta_g1", "delta_g2", "A_query", "H_query"):
        assert np.array_equal(pk[name], want[name]), name
    assert pk["B_count"] == want["B_count"] and pk["H_count"] == want["H_count"]
    nb = pk["B_count"]
    assert np.array_equal(pk["B_idx"][:nb], want["B_idx"][:nb])
    assert np.array_equal(pk["B_g1"][:nb], want["B_g1"][:nb]) and np.array_equal(pk["B_g2"][:nb], want["B_g2"][:nb])
    if cs.auxiliary_input_size:
        assert np.array_equal(pk["L_query"], want["L_query"])
    assert np.array_equal(vk["gamma_g2"], cref.g2_to_limbs([o.G2.to_affine(vk_want.gamma_g2)]))
    assert np.array_equal(vk["gamma_g1"], cref.g1_to_limbs([_aff1(vk_want.gamma_g1)]))
    assert np.array_equal(vk["gamma_ABC_g1"],
                          cref.g1_to_limbs([_aff1(p) for p in [vk_want.gamma_ABC_first] + vk_want.gamma_ABC_rest]))
    return pk, vk


@pytest.mark.parametrize("nc,ni,binary", [(27, 4, False), (19, 4, False), (40, 7, True), (31, 0, False)])
def test_generator_matches_oracle(ctx, nc, ni, binary):
    """basic (m = 32, 64) and step (m = 24, 48) domains, with and without primary inputs"""
    cs, _primary, _aux = synthetic_r1cs(nc, ni, seed=nc, binary=binary)
    _check_keypair(ctx, cs, _toxic(nc))


def test_generator_t_inside_the_domain(ctx):
    """t = omega^3: evaluate_all_lagrange_polynomials returns the unit vector (basic_radix2_domain_aux.hpp:97-110), Z(t) = 0"""
    cs, _p, _a = synthetic_r1cs(27, 4, seed=2)
    tox = _toxic(9)
    tox[0] = pow(o.unity_root(32), 3, o.R)
    pk, _vk = _check_keypair(ctx, cs, tox)
    assert not pk["H_query"].any()
    cs2, _p, _a = synthetic_r1cs(19, 4, seed=2)  # step domain 16 + 8: t in the small coset
    tox[0] = o.make_evaluation_domain(24).get_domain_element(16 + 5)
    _check_keypair(ctx, cs2, tox)


def test_generator_rejects_bad_randomness(ctx):
    cs, _p, _a = synthetic_r1cs(10, 2, seed=1)
    tox = _toxic(1)
    tox[4] = 0  # delta must be invertible
    with pytest.raises(ValueError):
        vsp.generate(ctx, fx.r1cs_arrays(cs), tox)


@pytest.mark.parametrize("nc,ni", [(59, 4), (300, 6), (4065, 30)])
def test_gpu_key_gpu_proof_verifies(ctx, nc, ni):
    """generate -> prove on the GPU, verify with the oracle's pairing check; a wrong primary input is rejected
    (r1cs_gg_ppzksnark.cpp:160-163).  4065 + 30 + 1 = 2^12: a key the Python oracle could not generate in test time."""
    cs, primary, aux = synthetic_r1cs(nc, ni, seed=nc)
    arr, vk = vsp.generate(ctx, fx.r1cs_arrays(cs), _toxic(nc + 1))
    pk = vsp.ProvingKey(ctx, arr)
    rng = random.Random(nc)
    r, s = rng.randrange(o.R), rng.randrange(o.R)
    proof = pk.prove(cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4), r, s)
    pk.close()
    J1 = lambda a: o.G1.from_affine(cref.g1_from_limbs(a.reshape(1, 12))[0])
    J2 = lambda a: o.G2.from_affine(cref.g2_from_limbs(a.reshape(1, 24))[0])
    pr = (J1(proof[:12]), J2(proof[12:36]), J1(proof[36:]))
    ovk = _vk_from_arrays(vk)
    assert o.groth16_verify(ovk, primary, pr)
    bad = list(primary)
    bad[0] = (bad[0] + 1) % o.R
    assert not o.groth16_verify(ovk, bad, pr)


def test_saver_keygen_matches_oracle_and_encryption_verifies(ctx):
    """keys from the GPU generator + GPU SAVER keygen, ballot encrypted and proved on the GPU, then the oracle's
    verify_encryption (elgamal_verifiable.hpp:486-490) accepts it and rejects a tampered ciphertext"""
    msg_size = 3
    cs, primary, aux = synthetic_r1cs(26, 5, seed=8)
    arr, vk = vsp.generate(ctx, fx.r1cs_arrays(cs), _toxic(21))
    ovk = _vk_from_arrays(vk)
    rng = random.Random(22)
    rnd = [rng.randrange(1, o.R) for _ in range(3 * msg_size + 2)]
    pub_arr, rho, svk = vsp.saver_generate_keypair(ctx, vk, msg_size, rnd)
    opub, orho, osvk = o.saver_keygen(ovk, msg_size, rnd)
    want = fx.saver_pub_arrays(opub)
    for name in ("delta_g1", "delta_s_g1", "t_g1", "t_g2", "delta_sum_s_g1", "gamma_inverse_sum_s_g1"):
        assert np.array_equal(pub_arr[name], want[name]), name
    assert rho == orho
    G2L = lambda pts: cref.g2_to_limbs([o.G2.to_affine(p) for p in pts])
    assert np.array_equal(svk["rho_g2"], G2L([osvk.rho_g2]))
    assert np.array_equal(svk["rho_sv_g2"], G2L(osvk.rho_sv_g2))
    assert np.array_equal(svk["rho_rhov_g2"], G2L(osvk.rho_rhov_g2))
    with pytest.raises(ValueError):
        vsp.saver_generate_keypair(ctx, vk, msg_size, rnd[:-1])

    pk = vsp.ProvingKey(ctx, arr)
    pub = vsp.SaverPublicKey(pub_arr)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
    gabc_rest = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + msg_size])
    ct, proof = pk.encrypt(pub, gabc_rest, np.ascontiguousarray(P[:msg_size]), P, A, r_saver, r, s)
    pk.close()
    J1 = lambda a: [o.G1.from_affine(p) if p else o.G1.inf() for p in cref.g1_from_limbs(a)]
    pr = (J1(proof[:12].reshape(1, 12))[0], o.G2.from_affine(cref.g2_from_limbs(proof[12:36].reshape(1, 24))[0]),
          J1(proof[36:].reshape(1, 12))[0])
    oct = J1(ct)
    rest = list(primary[msg_size:])
    assert o.saver_verify_encryption(ovk, opub, oct, rest, pr)
    bad = list(oct)
    bad[1] = o.G1.add(bad[1], o.g1_generator())
    assert not o.saver_verify_encryption(ovk, opub, bad, rest, pr)
