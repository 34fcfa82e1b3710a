"""GPU pairing side against the oracle: reduced pairings as bit-exact Fq12 values
(algebra::pair_reduced, call sites libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:415-424), bilinearity at
sizes the oracle cannot reach, and the two verifiers (zk::verify, libs/marshalling/zk/test/r1cs_gg_ppzksnark.cpp:160-163;
verify_encryption_op::process, elgamal_verifiable.hpp:486-490) deciding exactly as the oracle does on honest, tampered
and rerandomised ballots (the flow of libs/pubkey/test/elgamal_verifiable.cpp:442-489)."""
import importlib
import random

import numpy as np
import pytest

from helpers import synthetic_r1cs
from oracle import cref, fixtures as fx, pyref as o

pytestmark = pytest.mark.gpu
vsp = importlib.import_module("vote-saver-protocol_b200")


@pytest.fixture(scope="module")
def ctx():
    with vsp.Context(0) as c:
        yield c


def _gt_limbs(f):
    """pyref Fq12 ((a0, a1, a2), (b0, b1, b2)) -> 72 limbs in the ABI's tower order"""
    flat = [c for half in f for e2 in half for c in e2]
    return cref.ints_to_limbs(flat, 6).reshape(-1)


G1L = lambda pts: cref.g1_to_limbs([o.G1.to_affine(p) for p in pts])
G2L = lambda pts: cref.g2_to_limbs([o.G2.to_affine(p) for p in pts])


def test_reduced_pairing_matches_oracle_bit_for_bit(ctx):
    rng = random.Random(3)
    ps = [o.G1.mul(rng.randrange(1, o.R), o.g1_generator()) for _ in range(3)]
    qs = [o.G2.mul(rng.randrange(1, o.R), o.g2_generator()) for _ in range(3)]
    gt, one = vsp.pairing_products(ctx, G1L(ps), G2L(qs))
    for i in range(3):
        want = o.final_exponentiation(o.miller_loop(o.G1.to_affine(ps[i]), o.G2.to_affine(qs[i])))
        assert np.array_equal(gt[i], _gt_limbs(want)), i
    assert not one.any()
    # a product of three pairings in one check
    gt3, _ = vsp.pairing_products(ctx, G1L(ps), G2L(qs), pairs=3)
    f = o.F12_ONE
    for p, q in zip(ps, qs):
        f = o.f12mul(f, o.miller_loop(o.G1.to_affine(p), o.G2.to_affine(q)))
    assert np.array_equal(gt3[0], _gt_limbs(o.final_exponentiation(f)))


def test_pairing_bilinearity_and_infinity(ctx):
    """e(aP, bQ) e(-abP, Q) = 1 for many random (a, b); a pair with a point at infinity contributes 1"""
    n = 64
    a, b = fx.random_fr(n, seed=1), fx.random_fr(n, seed=2)
    ai, bi = cref.limbs_to_ints(a), cref.limbs_to_ints(b)
    ab = cref.ints_to_limbs([(-x * y) % o.R for x, y in zip(ai, bi)], 4)
    g2 = G2L([o.g2_generator()])[0]
    aP, abP = vsp.g1_generator_mul_batch(ctx, a), vsp.g1_generator_mul_batch(ctx, ab)
    bQ = vsp.g2_generator_mul_batch(ctx, b)
    g1 = np.stack([aP, abP], axis=1).reshape(-1, 12)
    q = np.stack([bQ, np.tile(g2, (n, 1))], axis=1).reshape(-1, 24)
    _gt, one = vsp.pairing_products(ctx, g1, q, pairs=2)
    assert one.all()
    g1[1] = vsp.g1_generator_mul_batch(ctx, a[:1])[0]  # break check 0
    _gt, one = vsp.pairing_products(ctx, g1, q, pairs=2)
    assert not one[0] and one[1:].all()
    z1, z2 = np.zeros((1, 12), np.uint64), np.zeros((1, 24), np.uint64)
    _gt, one = vsp.pairing_products(ctx, np.concatenate([z1, aP[:1]]), np.concatenate([bQ[:1], z2]))
    assert one.all()


def _keys_and_ballots(ctx, nc, ni, msg_size, k, seed):
    cs, _p, _a = synthetic_r1cs(nc, ni, seed=seed)
    rng = random.Random(seed)
    arr, vk = vsp.generate(ctx, fx.r1cs_arrays(cs), [rng.randrange(1, o.R) for _ in range(5)])
    pub_arr, _rho, _svk = vsp.saver_generate_keypair(ctx, vk, msg_size, [rng.randrange(1, o.R) for _ in range(3 * msg_size + 2)])
    return cs, arr, vk, pub_arr, rng


def test_verify_batch_decides_like_the_oracle(ctx):
    nc, ni, k = 40, 6, 5
    cs, arr, vk, _pub, rng = _keys_and_ballots(ctx, nc, ni, 2, k, seed=5)
    pk = vsp.ProvingKey(ctx, arr)
    prim, proofs = [], []
    for b in range(k):
        _cs, primary, aux = synthetic_r1cs(nc, ni, seed=5)  # same circuit, same satisfying assignment
        proofs.append(pk.prove(cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4), rng.randrange(o.R), rng.randrange(o.R)))
        prim.append(cref.ints_to_limbs(primary, 4))
    pk.close()
    prim, proofs = np.stack(prim), np.stack(proofs)
    assert vsp.verify_batch(ctx, vk, prim, proofs).all()
    prim[1, 0, 0] ^= np.uint64(1)            # wrong primary input
    proofs[2, 36:48] = proofs[3, 0:12]       # C replaced by another valid G1 point
    proofs[4, 12:36] = proofs[0, 12:36]      # B from another proof
    got = vsp.verify_batch(ctx, vk, prim, proofs)
    assert got.tolist() == [True, False, False, True, False]
    # the oracle agrees on one accept and one reject
    from test_gpu_keygen import _vk_from_arrays
    ovk = _vk_from_arrays(vk)
    J1 = lambda a: o.G1.from_affine(cref.g1_from_limbs(a.reshape(1, 12))[0])
    J2 = lambda a: o.G2.from_affine(cref.g2_from_limbs(a.reshape(1, 24))[0])
    for b in (0, 1):
        pr = (J1(proofs[b, :12]), J2(proofs[b, 12:36]), J1(proofs[b, 36:]))
        assert o.groth16_verify(ovk, cref.limbs_to_ints(prim[b]), pr) == bool(got[b])


def test_verify_encryption_batch_decides_like_the_oracle(ctx):
    nc, ni, msg_size, k = 33, 7, 4, 6
    cs, arr, vk, pub_arr, rng = _keys_and_ballots(ctx, nc, ni, msg_size, k, seed=9)
    pk = vsp.ProvingKey(ctx, arr)
    pub = vsp.SaverPublicKey(pub_arr)
    _cs, primary, aux = synthetic_r1cs(nc, ni, seed=9)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    gabc_rest = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + msg_size])
    cts, proofs = [], []
    for b in range(k):
        r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
        ct, proof = pk.encrypt(pub, gabc_rest, np.ascontiguousarray(P[:msg_size]), P, A, r_saver, r, s)
        cts.append(ct)
        proofs.append(proof)
    # ballot 5: a rerandomised copy of ballot 0 (rerandomize_op, elgamal_verifiable.hpp:632-667) still verifies
    rnd = [rng.randrange(1, o.R) for _ in range(3)]
    cts[5], proofs[5] = vsp.rerandomize(ctx, pub, arr["delta_g2"], rnd, cts[0], proofs[0])
    pk.close()
    cts, proofs = np.stack(cts), np.stack(proofs)
    rest = np.tile(P[msg_size:], (k, 1, 1))
    assert vsp.verify_encryption_batch(ctx, vk, pub, cts, rest, proofs).all()
    g = G1L([o.g1_generator()])[0]
    cts[1, 1] = cref.g1_add(cts[1, 1], g)        # tampered message slot
    cts[2, msg_size + 1] = cts[3, msg_size + 1]  # sum element of another ballot
    rest[4, 0, 0] ^= np.uint64(1)                # wrong public part of the primary input
    got = vsp.verify_encryption_batch(ctx, vk, pub, cts, rest, proofs)
    assert got.tolist() == [True, False, False, True, False, True]
    from test_gpu_keygen import _vk_from_arrays
    ovk = _vk_from_arrays(vk)
    J1s = lambda a: [o.G1.from_affine(p) if p else o.G1.inf() for p in cref.g1_from_limbs(a)]
    J2s = lambda a: [o.G2.from_affine(p) if p else o.G2.inf() for p in cref.g2_from_limbs(a)]
    opub = o.SaverPublicKey(J1s(pub_arr["delta_g1"])[0], J1s(pub_arr["delta_s_g1"]), J1s(pub_arr["t_g1"]), J2s(pub_arr["t_g2"]),
                            J1s(pub_arr["delta_sum_s_g1"])[0], J1s(pub_arr["gamma_inverse_sum_s_g1"])[0])
    for b in (0, 1, 5):
        pr = (J1s(proofs[b, :12].reshape(1, 12))[0], J2s(proofs[b, 12:36].reshape(1, 24))[0], J1s(proofs[b, 36:].reshape(1, 12))[0])
        want = o.saver_verify_encryption(ovk, opub, J1s(cts[b]), cref.limbs_to_ints(rest[b]), pr)
        assert want == bool(got[b]), b
    with pytest.raises(ValueError):
        vsp.verify_encryption_batch(ctx, vk, pub, cts, rest[:, :1], proofs)  # gamma_ABC count mismatch


@pytest.mark.xfail(strict=False, reason="vsp_saver_verify_encryption_gt_batch (a host wrapper over the verifier's GT branch) was "
                                        "added after the round's GPU minutes were spent and has not run on hardware yet")
def test_verify_encryption_with_the_key_in_gt_form_decides_like_the_group_form(ctx):
    """the verification key as it is serialised (alpha_g1_beta_g2 = pair_reduced(alpha_g1, beta_g2), verification_key.hpp:128-152)
    gives the same decisions as the key with alpha_g1 and beta_g2, on honest, rerandomised and tampered ballots"""
    nc, ni, msg_size, k = 33, 7, 4, 5
    cs, arr, vk, pub_arr, rng = _keys_and_ballots(ctx, nc, ni, msg_size, k, seed=19)
    pk = vsp.ProvingKey(ctx, arr)
    pub = vsp.SaverPublicKey(pub_arr)
    _cs, primary, aux = synthetic_r1cs(nc, ni, seed=19)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    gabc_rest = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + msg_size])
    cts, proofs = [], []
    for b in range(k):
        r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
        ct, proof = pk.encrypt(pub, gabc_rest, np.ascontiguousarray(P[:msg_size]), P, A, r_saver, r, s)
        cts.append(ct)
        proofs.append(proof)
    cts[4], proofs[4] = vsp.rerandomize(ctx, pub, arr["delta_g2"], [rng.randrange(1, o.R) for _ in range(3)], cts[0], proofs[0])
    pk.close()
    cts, proofs = np.stack(cts), np.stack(proofs)
    rest = np.tile(P[msg_size:], (k, 1, 1))
    gt, _ = vsp.pairing_products(ctx, vk["alpha_g1"], vk["beta_g2"])
    gt_vk = {"alpha_g1_beta_g2": gt[0], "gamma_g2": vk["gamma_g2"], "delta_g2": vk["delta_g2"], "gamma_ABC_g1": vk["gamma_ABC_g1"]}
    assert vsp.verify_encryption_gt_batch(ctx, gt_vk, pub, cts, rest, proofs).all()
    cts[1, 1] = cref.g1_add(cts[1, 1], G1L([o.g1_generator()])[0])  # tampered message slot
    cts[2, msg_size + 1] = cts[3, msg_size + 1]                       # sum element of another ballot
    rest[3, 0, 0] ^= np.uint64(1)                                     # wrong public part of the primary input
    got = vsp.verify_encryption_gt_batch(ctx, gt_vk, pub, cts, rest, proofs)
    assert got.tolist() == vsp.verify_encryption_batch(ctx, vk, pub, cts, rest, proofs).tolist() == [True, False, False, False, True]
    wrong_gt = dict(gt_vk, alpha_g1_beta_g2=vsp.pairing_products(ctx, vk["alpha_g1"], vk["gamma_g2"])[0][0])
    assert not vsp.verify_encryption_gt_batch(ctx, wrong_gt, pub, cts, rest, proofs).any()


def test_tally_decrypt_and_verify_decryption(ctx):
    """the tally flow of bin/cli (common.hpp: sum the ballots' ciphertexts, decrypt_op, verify_decryption_op): keys, ballots
    and the tally on the GPU; small case cross-checked against the oracle's restatement of elgamal_verifiable.hpp:401-441,
    548-575"""
    msg_size, voters = 3, 5
    cs, arr, vk, pub_arr, rng = _keys_and_ballots(ctx, 20, 5, msg_size, voters, seed=13)
    rnd = [rng.randrange(1, o.R) for _ in range(3 * msg_size + 2)]
    pub_arr, rho, svk = vsp.saver_generate_keypair(ctx, vk, msg_size, rnd)
    pub = vsp.SaverPublicKey(pub_arr)
    gabc_rest = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + msg_size])
    votes = [[1, 0, 0], [0, 1, 0], [1, 0, 0], [1, 0, 0], [0, 0, 1]]
    total = None
    for v in votes:
        ct = vsp.saver_ciphertext(ctx, pub, gabc_rest, cref.ints_to_limbs(v, 4), rng.randrange(1, o.R))
        total = ct if total is None else np.stack([cref.g1_add(total[i], ct[i]) for i in range(msg_size + 2)])
    m, c0 = vsp.decrypt(ctx, total, rho, svk, gabc_rest, block_bits=3)
    assert m == [3, 1, 1]
    assert vsp.verify_decryption(ctx, total, m, svk, gabc_rest, c0)
    assert not vsp.verify_decryption(ctx, total, [3, 1, 0], svk, gabc_rest, c0)
    assert not vsp.verify_decryption(ctx, total, m, svk, gabc_rest, cref.g1_add(c0, G1L([o.g1_generator()])[0]))
    with pytest.raises(ValueError, match="Decryption failed"):
        vsp.decrypt(ctx, total, rho, svk, gabc_rest, block_bits=1)  # 3 votes do not fit exp <= 2
    # the oracle agrees on a two-slot ciphertext (one Python pairing costs about a second)
    from test_gpu_keygen import _vk_from_arrays
    ovk = _vk_from_arrays(vk)
    J1s = lambda a: [o.G1.from_affine(p) if p else o.G1.inf() for p in cref.g1_from_limbs(a)]
    J2s = lambda a: [o.G2.from_affine(p) if p else o.G2.inf() for p in cref.g2_from_limbs(a)]
    osvk = o.SaverVerificationKey(J2s(svk["rho_g2"])[0], J2s(svk["rho_sv_g2"])[:1], J2s(svk["rho_rhov_g2"])[:1])
    small = np.concatenate([total[:2], total[-1:]])  # slot 0 only: ct_0, ct_1, (sum element unused by decrypt)
    om, oc0 = o.saver_decrypt(J1s(small), rho, osvk, ovk.gamma_ABC_rest, block_bits=2)
    gm, gc0 = vsp.decrypt(ctx, small, rho, svk, gabc_rest, block_bits=2)
    assert om == gm == [3]
    assert np.array_equal(gc0, G1L([oc0])[0])
    assert o.saver_verify_decryption(J1s(small), om, osvk, ovk.gamma_ABC_rest, oc0)
    assert vsp.verify_decryption(ctx, small, gm, svk, gabc_rest, gc0)


def test_empty_batches_and_infinity_proof_elements(ctx):
    """empty inputs are no-ops; a proof whose A is the point at infinity drops e(A, B) from the product exactly as the
    oracle's pairing does (miller_loop of an infinite argument is 1), so the check fails unless the rest is 1 too"""
    cs, arr, vk, pub_arr, rng = _keys_and_ballots(ctx, 14, 3, 1, 1, seed=21)
    assert vsp.verify_batch(ctx, vk, np.zeros((0, 3, 4), np.uint64), np.zeros((0, 48), np.uint64)).shape == (0,)
    gt, one = vsp.pairing_products(ctx, np.zeros((0, 12), np.uint64), np.zeros((0, 24), np.uint64))
    assert gt.shape == (0, 72) and one.shape == (0,)
    pk = vsp.ProvingKey(ctx, arr)
    _cs, primary, aux = synthetic_r1cs(14, 3, seed=21)
    proof = pk.prove(cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4), 5, 6)
    pk.close()
    P = cref.ints_to_limbs(primary, 4)[None]
    assert vsp.verify_batch(ctx, vk, P, proof[None]).tolist() == [True]
    bad = proof.copy()
    bad[:12] = 0
    assert vsp.verify_batch(ctx, vk, P, bad[None]).tolist() == [False]
    from test_gpu_keygen import _vk_from_arrays
    J2 = lambda a: o.G2.from_affine(cref.g2_from_limbs(a.reshape(1, 24))[0])
    J1 = lambda a: o.G1.from_affine(cref.g1_from_limbs(a.reshape(1, 12))[0])
    assert not o.groth16_verify(_vk_from_arrays(vk), primary, (o.G1.inf(), J2(bad[12:36]), J1(bad[36:])))
    # SAVER key pair and decryption with an empty message
    pub0, rho0, svk0 = vsp.saver_generate_keypair(ctx, vk, 0, [7, 9])
    assert pub0["t_g2"].shape[0] == 1 and rho0 == 7
    ct0 = vsp.saver_ciphertext(ctx, vsp.SaverPublicKey(pub0), np.zeros((1, 12), np.uint64), np.zeros((0, 4), np.uint64), 11)
    m0, c0 = vsp.decrypt(ctx, ct0, rho0, svk0, np.zeros((1, 12), np.uint64))
    assert m0 == [] and vsp.verify_decryption(ctx, ct0, [], svk0, np.zeros((1, 12), np.uint64), c0)


def test_reference_made_proof_verifies_on_the_gpu(ctx):
    """bin/cli/src/data.bin (tests/golden/databin_groth16.json): the reference's own proof, decoded by the product codec,
    is accepted under the reference's own verification key (alpha_g1_beta_g2 as a GT value) — and the GPU's pair_reduced
    product reproduces that stored Fq12 element limb for limb, which pins its value convention on crypto3's."""
    from helpers import golden
    codec = importlib.import_module("vote-saver-protocol_b200.codec")
    g = golden("databin_groth16.json")
    proof = codec.proof_from_bytes(bytes.fromhex(g["proof"]))
    gamma, delta = codec.g2_decompress(bytes.fromhex(g["gamma_g2"])), codec.g2_decompress(bytes.fromhex(g["delta_g2"]))
    first = codec.g1_decompress(bytes.fromhex(g["gamma_ABC_first"]))
    gt = cref.ints_to_limbs([int(x, 16) for x in g["alpha_g1_beta_g2"]], 6).reshape(-1)
    none = np.zeros((1, 0, 4), np.uint64)
    assert vsp.verify_gt_batch(ctx, gt, gamma, delta, first, none, proof[None]).tolist() == [True]
    swapped = proof.copy()
    swapped[0:12], swapped[36:48] = proof[36:48], proof[0:12]
    assert vsp.verify_gt_batch(ctx, gt, gamma, delta, first, none, swapped[None]).tolist() == [False]
    bad_gt = gt.copy()
    bad_gt[0] ^= np.uint64(1)
    assert vsp.verify_gt_batch(ctx, bad_gt, gamma, delta, first, none, proof[None]).tolist() == [False]
    # e(A, B) e(-first, gamma) e(-C, delta) as a value
    neg = lambda p: cref.g1_to_limbs([o.G1.to_affine(o.G1.neg(o.G1.from_affine(cref.g1_from_limbs(p.reshape(1, 12))[0])))])[0]
    g1 = np.stack([proof[:12], neg(first), neg(proof[36:])])
    g2 = np.stack([proof[12:36], gamma, delta])
    got, _ = vsp.pairing_products(ctx, g1, g2, pairs=3)
    assert np.array_equal(got[0], gt)


def test_ill_formed_proof_elements_are_rejected(ctx):
    """proof.is_well_formed() / the decoders' guarantees: an element off the curve, outside the order-r subgroup or with an
    unreduced coordinate makes the ballot invalid, whatever the pairing of such limbs would give; the oracle agrees"""
    nc, ni, k = 24, 4, 7
    cs, arr, vk, _pub, rng = _keys_and_ballots(ctx, nc, ni, 2, k, seed=21)
    pk = vsp.ProvingKey(ctx, arr)
    _cs, primary, aux = synthetic_r1cs(nc, ni, seed=21)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    proofs = np.stack([pk.prove(P, A, rng.randrange(o.R), rng.randrange(o.R)) for _ in range(k)])
    pk.close()
    prim = np.tile(P, (k, 1, 1))
    assert vsp.verify_batch(ctx, vk, prim, proofs).all()
    proofs[1, 6] ^= np.uint64(1)                       # A.y off the curve
    proofs[2, 12] ^= np.uint64(2)                      # B.x.c0 off the twist
    pmod = cref.ints_to_limbs([o.P], 6)[0]
    with np.errstate(over="ignore"):
        carry, t = 0, proofs[3, 36:42].copy()          # C.x + p: same residue, unreduced limbs
        for i in range(6):
            v = int(t[i]) + int(pmod[i]) + carry
            t[i], carry = np.uint64(v & (2 ** 64 - 1)), v >> 64
    if carry == 0:
        proofs[3, 36:42] = t
    # a curve point outside G1: the cofactor is not 1, so a random x with a square x^3 + 4 almost never lands in the subgroup
    x = 7
    while True:
        y = o.fq_sqrt((x ** 3 + 4) % o.P)
        if y is not None and not o.G1.in_subgroup(o.G1.from_affine((x, y))):
            break
        x += 1
    proofs[4, 0:12] = cref.g1_to_limbs([(x, y)])[0]
    # the same on the twist for B
    x0 = 1
    while True:
        yy = o.fq2_sqrt(o._Fq2.add(o._Fq2.mul(o._Fq2.sqr((x0, 0)), (x0, 0)), (4, 4)))
        if yy is not None and not o.G2.in_subgroup(o.G2.from_affine(((x0, 0), yy))):
            break
        x0 += 1
    proofs[5, 12:36] = cref.g2_to_limbs([((x0, 0), yy)])[0]
    got = vsp.verify_batch(ctx, vk, prim, proofs)
    assert got.tolist() == [True, False, False, carry != 0, False, False, True]
    from test_gpu_keygen import _vk_from_arrays
    ovk = _vk_from_arrays(vk)
    J1 = lambda a: o.G1.from_affine(cref.g1_from_limbs(a.reshape(1, 12))[0])
    J2 = lambda a: o.G2.from_affine(cref.g2_from_limbs(a.reshape(1, 24))[0])
    for b in (1, 4, 5, 6):
        pr = (J1(proofs[b, :12]), J2(proofs[b, 12:36]), J1(proofs[b, 36:]))
        assert o.groth16_verify(ovk, primary, pr) == bool(got[b])
