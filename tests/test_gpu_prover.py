"""GPU parity: r1cs_to_qap::witness_map, zk::prove (basic and encrypted_input modes), batched proving, and the SAVER
ciphertext / rerandomisation through the C ABI against the oracle.  Bit-exact (every group element canonical affine).

Whole-proof parity is against the oracle's restatement (PARITY UNPINNED upstream: the reference has no proof-byte
fixtures); tests/test_oracle.py holds the generate -> prove -> verify self-consistency gate."""
import random

import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import synthetic_pk_arrays, synthetic_r1cs, synthetic_saver_pub

pytestmark = pytest.mark.gpu


def _instance(nc, ni, seed, binary=False):
    cs, primary, aux = synthetic_r1cs(nc, ni, seed, binary)
    arr = synthetic_pk_arrays(cs, seed + 1)
    return cs, cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4) if aux else np.zeros((0, 4), np.uint64), arr


def _oracle_pk(arr):
    return fx.fill_pk(cref.ProvingKeyDesc(), arr)


# (constraints, inputs): m_req = nc + ni + 1 -> 16 basic; 24 step; 128 basic; 1024+64 step (exact); 5000 -> rounds to 5120 step
CASES = [(10, 5, False), (18, 5, False), (97, 30, False), (1057, 30, False), (4969, 30, True)]


@pytest.mark.parametrize("nc,ni,binary", CASES)
def test_witness_map_matches_oracle(vsp, ctx, nc, ni, binary):
    cs, primary, aux, arr = _instance(nc, ni, seed=nc, binary=binary)
    pk = vsp.ProvingKey(ctx, arr)
    want = cref.witness_map(_oracle_pk(arr).cs, primary, aux)
    got = pk.witness_map(primary, aux)
    assert pk.m == want.shape[0] - 1
    assert np.array_equal(got, want)
    pk.close()


@pytest.mark.parametrize("nc,ni,binary", CASES)
def test_prove_matches_oracle(vsp, ctx, nc, ni, binary):
    cs, primary, aux, arr = _instance(nc, ni, seed=nc + 100, binary=binary)
    pk = vsp.ProvingKey(ctx, arr)
    rng = random.Random(nc)
    r, s = rng.randrange(o.R), rng.randrange(o.R)
    want = cref.groth16_prove(_oracle_pk(arr), primary, aux, r, s)
    assert np.array_equal(pk.prove(primary, aux, r, s), want)
    pk.close()


def test_prove_encrypted_input_mode_and_degenerate_randomness(vsp, ctx):
    cs, primary, aux, arr = _instance(97, 30, seed=900)
    pub_arr, _ = synthetic_saver_pub(25, seed=901)
    pub = vsp.SaverPublicKey(pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    opk = _oracle_pk(arr)
    rng = random.Random(5)
    for r, s, rs in ((rng.randrange(o.R), rng.randrange(o.R), rng.randrange(o.R)), (0, 0, 0), (1, o.R - 1, 1)):
        want = cref.groth16_prove(opk, primary, aux, r, s, rs, pub_arr["gamma_inverse_sum_s_g1"])
        assert np.array_equal(pk.prove(primary, aux, r, s, pub, rs), want)
    with pytest.raises(ValueError):
        pk.prove(primary, aux, o.R, 1)  # randomness must be reduced mod r
    pk.close()


def test_prove_batch_equals_single_proofs(vsp, ctx):
    nc, ni, k = 97, 30, 5
    cs, _, _ = synthetic_r1cs(nc, ni, 0)
    arr = synthetic_pk_arrays(cs, 3)
    pub_arr, _ = synthetic_saver_pub(25, seed=4)
    pub = vsp.SaverPublicKey(pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    opk = _oracle_pk(arr)
    prim, aux, rs = [], [], []
    rng = random.Random(9)
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=b)  # same constraint shape, different witness
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    got = pk.prove_batch(np.stack(prim), np.stack(aux), np.stack(rs), pub)
    for b in range(k):
        r, s, rsv = cref.limbs_to_ints(rs[b])
        want = cref.groth16_prove(opk, prim[b], aux[b], r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"])
        assert np.array_equal(got[b], want), b
    pk.close()


def test_saver_encrypt_and_rerandomize_match_oracle(vsp, ctx):
    """elgamal_verifiable.hpp:322-357 and :632-667 with msg_size = 25 (common.hpp:163)"""
    nc, ni, msg_size = 97, 30, 25
    cs, primary, aux = synthetic_r1cs(nc, ni, seed=21, binary=True)
    arr = synthetic_pk_arrays(cs, 22)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(msg_size, seed=23)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:msg_size])
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    msg = np.ascontiguousarray(P[:msg_size])
    pub = vsp.SaverPublicKey(pub_arr)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    rng = random.Random(31)
    r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
    ct, proof = pk.encrypt(pub, gamma_abc_rest, msg, P, A, r_saver, r, s)
    want_ct = cref.saver_ciphertext(opub, gamma_abc_rest, msg, r_saver)
    want_proof = cref.groth16_prove(_oracle_pk(arr), P, A, r, s, r_saver, pub_arr["gamma_inverse_sum_s_g1"])
    assert np.array_equal(ct, want_ct)
    assert np.array_equal(proof, want_proof)
    assert np.array_equal(vsp.saver_ciphertext(ctx, pub, gamma_abc_rest, msg, r_saver), want_ct)
    # "Plain text should be a prefix of primary input" (:336-339)
    bad = msg.copy()
    bad[3, 0] ^= 1
    with pytest.raises(ValueError):
        pk.encrypt(pub, gamma_abc_rest, bad, P, A, r_saver, r, s)
    rnd = [rng.randrange(1, o.R) for _ in range(3)]
    ct2, proof2 = vsp.rerandomize(ctx, pub, arr["delta_g2"], rnd, ct, proof)
    want_ct2, want_proof2 = cref.saver_rerandomize(opub, arr["delta_g2"], rnd, want_ct, want_proof)
    assert np.array_equal(ct2, want_ct2)
    assert np.array_equal(proof2, want_proof2)
    pk.close()


@pytest.mark.parametrize("binary", [True, False])
def test_encrypt_batch_matches_oracle(vsp, ctx, binary):
    """vsp_saver_encrypt_batch: k ballots, chunked (VSP_BATCH_CHUNK default 256 -> one chunk; k = 7 exercises ragged
    lanes), messages with 0 / 1 and general slots, each ciphertext and proof against the oracle"""
    nc, ni, msg_size, k = 97, 30, 25, 7
    cs, _, _ = synthetic_r1cs(nc, ni, 0, binary=binary)
    arr = synthetic_pk_arrays(cs, 41)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(msg_size, seed=42)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:msg_size])
    pub = vsp.SaverPublicKey(pub_arr)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    opk = _oracle_pk(arr)
    key = pub.resident(ctx, gamma_abc_rest)
    rng = random.Random(43)
    prim, aux, rs = [], [], []
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=b, binary=binary)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    cts, proofs = pk.encrypt_batch(key, np.stack(prim), np.stack(aux), np.stack(rs))
    for b in range(k):
        r, s, rsv = cref.limbs_to_ints(rs[b])
        msg = np.ascontiguousarray(prim[b][:msg_size])
        assert np.array_equal(cts[b], cref.saver_ciphertext(opub, gamma_abc_rest, msg, rsv)), b
        want = cref.groth16_prove(opk, prim[b], aux[b], r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"])
        assert np.array_equal(proofs[b], want), b
    key.close()
    pk.close()


def test_warp_per_row_r1cs_kernel_matches(vsp, ctx, monkeypatch):
    """the warp-per-row R1CS evaluation (long rows) and the thread-per-row one (short rows) give the same H"""
    cs, primary, aux, arr = _instance(1057, 30, 3)
    want = cref.witness_map(_oracle_pk(arr).cs, primary, aux)
    for force in (None, "1"):
        if force:
            monkeypatch.setenv("VSP_R1CS_WARP_ROWS", force)
        pk = vsp.ProvingKey(ctx, arr)
        assert np.array_equal(pk.witness_map(primary, aux), want)
        pk.close()


def test_prove_from_fast_proving_key_blob_to_proof_bytes(vsp, ctx):
    """the drop-in data path of the vote phase: r1cs_proving_key.bin bytes -> vsp_codec_fast_pk_read -> vsp_pk_load ->
    vsp_prove -> 192-byte proof.bin body, equal to the oracle's proof in the reference's compressed encoding"""
    import importlib
    codec = importlib.import_module("vote-saver-protocol_b200.codec")
    cs, primary, aux, arr = _instance(97, 30, seed=8, binary=True)
    blob = codec.write_fast_proving_key(arr)
    pk = vsp.ProvingKey(ctx, codec.read_fast_proving_key(blob))
    r, s = 0x1234567890abcdef, o.R - 5
    proof = pk.prove(primary, aux, r, s)
    want = cref.groth16_prove(_oracle_pk(arr), primary, aux, r, s)
    assert np.array_equal(proof, want)
    gA, gB, gC = cref.g1_from_limbs(want[None, 0:12])[0], cref.g2_from_limbs(want[None, 12:36])[0], \
        cref.g1_from_limbs(want[None, 36:48])[0]
    assert codec.proof_bytes(proof) == (o.g1_compress(o.G1.from_affine(gA)) + o.g2_compress(o.G2.from_affine(gB)) +
                                        o.g1_compress(o.G1.from_affine(gC)))
    pk.close()


def test_tuning_switches_do_not_change_results(vsp, ctx, monkeypatch):
    """DESIGN.md 4 'tuning switches': a 12-ballot batch (shared-memory sort, radix-4 NTT passes by default) proved again
    under the global-counter sort, the radix-2 pass kernel, another reduction segment length and another window width;
    every proof identical, and ballot 0 equal to the oracle's"""
    nc, ni, k = 1000, 30, 12
    cs, _, _ = synthetic_r1cs(nc, ni, 0, binary=True)
    arr = synthetic_pk_arrays(cs, 5)
    prim, aux, rs = [], [], []
    rng = random.Random(19)
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=b, binary=True)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R), rng.randrange(o.R), 0], 4))
    prim, aux, rs = np.stack(prim), np.stack(aux), np.stack(rs)

    def run():
        pk = vsp.ProvingKey(ctx, arr)
        out = pk.prove_batch(prim, aux, rs)
        pk.close()
        return out

    base = run()
    r, s, _ = cref.limbs_to_ints(rs[0])
    assert np.array_equal(base[0], cref.groth16_prove(_oracle_pk(arr), prim[0], aux[0], r, s))
    for name, val in (("VSP_MSM_GLOBAL_SORT", "1"), ("VSP_NTT_RADIX2", "1"), ("VSP_MSM_SEG_LOG", "2"), ("VSP_MSM_C", "9")):
        monkeypatch.setenv(name, val)
        assert np.array_equal(run(), base), name
        monkeypatch.delenv(name)


@pytest.mark.parametrize("chunk", [2, 3])
def test_multi_chunk_batches_match_oracle(vsp, ctx, chunk):
    """k = 7 ballots in chunks of 2 and 3 (vsp_ctx_set_batch_chunk): the double-buffered assembly tails, the witness
    prefetch's slot reuse, the ragged last chunk and join_tails, on the host-buffer and the device-resident entry points
    of prove_batch and encrypt_batch, every proof and cipher text against the oracle"""
    import torch
    nc, ni, msg_size, k = 97, 30, 25, 7
    cs, _, _ = synthetic_r1cs(nc, ni, 0, binary=True)
    arr = synthetic_pk_arrays(cs, 61)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(msg_size, seed=62)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:msg_size])
    pub = vsp.SaverPublicKey(pub_arr)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    opk = _oracle_pk(arr)
    rng = random.Random(63)
    prim, aux, rs = [], [], []
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=100 + b, binary=True)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    prim, aux, rs = np.stack(prim), np.stack(aux), np.stack(rs)
    want_plain, want_enc, want_ct = [], [], []
    for b in range(k):
        r, s, rsv = cref.limbs_to_ints(rs[b])
        want_plain.append(cref.groth16_prove(opk, prim[b], aux[b], r, s))
        want_enc.append(cref.groth16_prove(opk, prim[b], aux[b], r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"]))
        want_ct.append(cref.saver_ciphertext(opub, gamma_abc_rest, np.ascontiguousarray(prim[b][:msg_size]), rsv))
    want_plain, want_enc, want_ct = np.stack(want_plain), np.stack(want_enc), np.stack(want_ct)
    ctx.set_batch_chunk(chunk)
    try:
        pk = vsp.ProvingKey(ctx, arr)
        key = pub.resident(ctx, gamma_abc_rest)
        assert np.array_equal(pk.prove_batch(prim, aux, rs), want_plain)
        cts, proofs = pk.encrypt_batch(key, prim, aux, rs)
        assert np.array_equal(cts, want_ct) and np.array_equal(proofs, want_enc)
        # device-resident: (1 || primary || auxiliary) per ballot
        assign = np.zeros((k, ni + cs.num_variables - ni + 1, 4), dtype=np.uint64)
        assign[:, 0, 0] = 1
        assign[:, 1:ni + 1] = prim
        assign[:, ni + 1:] = aux
        d_assign = torch.from_numpy(assign.view(np.int64)).cuda()
        d_proofs = torch.zeros((k, 48), dtype=torch.int64, device="cuda")
        d_cts = torch.zeros((k, msg_size + 2, 12), dtype=torch.int64, device="cuda")
        pk.prove_batch_dev(k, d_assign.data_ptr(), rs, d_proofs.data_ptr())
        ctx.synchronize()
        assert np.array_equal(d_proofs.cpu().numpy().view(np.uint64), want_plain)
        pk.encrypt_batch_dev(key, k, d_assign.data_ptr(), rs, d_cts.data_ptr(), d_proofs.data_ptr())
        ctx.synchronize()
        assert np.array_equal(d_proofs.cpu().numpy().view(np.uint64), want_enc)
        assert np.array_equal(d_cts.cpu().numpy().view(np.uint64), want_ct)
        key.close()
        pk.close()
    finally:
        ctx.set_batch_chunk(0)


def test_long_constraint_rows_and_table_less_key(vsp, ctx, monkeypatch):
    """the reference's field-input example closes with one constraint over every variable (r1cs_examples.hpp:116-126):
    5,000 constraints put ~5,000 terms in that row, which the prover cuts over many blocks (k_r1cs_long_*); and the same
    key loaded without window tables (VSP_TABLE_MAX_LOG = 0, the policy of 2^23+ point queries) proves identically"""
    cs, primary, aux, arr = _instance(5000, 4, seed=12)
    want_h = cref.witness_map(_oracle_pk(arr).cs, primary, aux)
    r, s = 0xabcdef, o.R - 7
    want = cref.groth16_prove(_oracle_pk(arr), primary, aux, r, s)
    for table_max in (None, "0"):
        if table_max is not None:
            monkeypatch.setenv("VSP_TABLE_MAX_LOG", table_max)
        pk = vsp.ProvingKey(ctx, arr)
        assert np.array_equal(pk.witness_map(primary, aux), want_h)
        assert np.array_equal(pk.prove(primary, aux, r, s), want)
        pk.close()


@pytest.mark.parametrize("affine", ["1", "0"])
def test_batch_proofs_under_either_accumulate_kernel(vsp, ctx, monkeypatch, affine):
    """a 9-ballot encrypt batch with the bucket accumulation forced onto the batched-affine kernel and onto the XYZZ one"""
    monkeypatch.setenv("VSP_MSM_AFFINE", affine)
    nc, ni, msg_size, k = 1000, 30, 25, 9
    cs, _, _ = synthetic_r1cs(nc, ni, 0, binary=False)
    arr = synthetic_pk_arrays(cs, 71)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(msg_size, seed=72)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:msg_size])
    pub = vsp.SaverPublicKey(pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    opk = _oracle_pk(arr)
    key = pub.resident(ctx, gamma_abc_rest)
    rng = random.Random(73)
    prim, aux, rs = [], [], []
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=b, binary=False)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    _cts, proofs = pk.encrypt_batch(key, np.stack(prim), np.stack(aux), np.stack(rs))
    for b in (0, 4, 8):
        r, s, rsv = cref.limbs_to_ints(rs[b])
        assert np.array_equal(proofs[b], cref.groth16_prove(opk, prim[b], aux[b], r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"])), b
    key.close()
    pk.close()


def test_rerandomize_batch_matches_oracle(vsp, ctx):
    """vsp_saver_rerandomize_batch[_dev] (rerandomize_op::process, elgamal_verifiable.hpp:632-667, for k ballots): every
    cipher text and proof equal to the oracle's per-ballot restatement and to the one-ballot entry point; the device-
    resident variant in place on the outputs of encrypt_batch_dev"""
    import torch
    nc, ni, msg_size, k = 97, 30, 25, 6
    cs, _, _ = synthetic_r1cs(nc, ni, 0, binary=True)
    arr = synthetic_pk_arrays(cs, 81)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(msg_size, seed=82)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:msg_size])
    pub = vsp.SaverPublicKey(pub_arr)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    key = pub.resident(ctx, gamma_abc_rest)
    rng = random.Random(83)
    prim, aux, rs = [], [], []
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=200 + b, binary=True)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    prim, aux, rs = np.stack(prim), np.stack(aux), np.stack(rs)
    cts, proofs = pk.encrypt_batch(key, prim, aux, rs)
    rnd = np.stack([cref.ints_to_limbs([rng.randrange(1, o.R) for _ in range(3)], 4) for _ in range(k)])
    cts2, proofs2 = pk.rerandomize_batch(key, rnd, cts, proofs)
    for b in range(k):
        want_ct, want_pr = cref.saver_rerandomize(opub, arr["delta_g2"], cref.limbs_to_ints(rnd[b]), cts[b], proofs[b])
        assert np.array_equal(cts2[b], want_ct) and np.array_equal(proofs2[b], want_pr), b
    one_ct, one_pr = vsp.rerandomize(ctx, pub, arr["delta_g2"], cref.limbs_to_ints(rnd[2]), cts[2], proofs[2])
    assert np.array_equal(one_ct, cts2[2]) and np.array_equal(one_pr, proofs2[2])
    # device-resident, in place
    assign = np.zeros((k, cs.num_variables + 1, 4), dtype=np.uint64)
    assign[:, 0, 0] = 1
    assign[:, 1:ni + 1] = prim
    assign[:, ni + 1:] = aux
    d_assign = torch.from_numpy(assign.view(np.int64)).cuda()
    d_proofs = torch.zeros((k, 48), dtype=torch.int64, device="cuda")
    d_cts = torch.zeros((k, msg_size + 2, 12), dtype=torch.int64, device="cuda")
    pk.encrypt_batch_dev(key, k, d_assign.data_ptr(), rs, d_cts.data_ptr(), d_proofs.data_ptr())
    pk.rerandomize_batch_dev(key, k, rnd, d_cts.data_ptr(), d_proofs.data_ptr())
    ctx.synchronize()
    assert np.array_equal(d_cts.cpu().numpy().view(np.uint64), cts2)
    assert np.array_equal(d_proofs.cpu().numpy().view(np.uint64), proofs2)
    bad = rnd.copy()
    bad[1, 1] = 0  # z1 = 0 has no inverse
    with pytest.raises(ValueError):
        pk.rerandomize_batch(key, bad, cts, proofs)
    key.close()
    pk.close()


def test_plain_text_must_be_shorter_than_the_primary_input(vsp, ctx):
    """'Primary input should be longer than plain text.' (elgamal_verifiable.hpp:325-326): msg_size == num_inputs is
    rejected by the one-ballot and the batched entry points"""
    nc, ni = 40, 6
    cs, primary, aux, arr = _instance(nc, ni, seed=31, binary=True)
    pub_arr, gamma_abc_rest = synthetic_saver_pub(ni, seed=32)
    gamma_abc_rest = np.ascontiguousarray(gamma_abc_rest[:ni])
    pub = vsp.SaverPublicKey(pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    with pytest.raises(ValueError):
        pk.encrypt(pub, gamma_abc_rest, primary, primary, aux, 5, 6, 7)
    key = pub.resident(ctx, gamma_abc_rest)
    with pytest.raises(ValueError):
        pk.encrypt_batch(key, primary[None], aux[None], cref.ints_to_limbs([5, 6, 7], 4)[None])
    key.close()
    pk.close()


def test_scratch_cache_reaches_a_steady_state_and_trims(vsp, ctx):
    """the context's scratch cache (vsp_ctx_scratch_stats / vsp_ctx_trim): the first batch cudaMallocs its blocks, equal
    batches after it make no driver allocation and return the same proofs; trim hands every unused block back and the
    next batch still agrees"""
    nc, ni, k = 97, 30, 5
    cs, _, _ = synthetic_r1cs(nc, ni, 0)
    arr = synthetic_pk_arrays(cs, 3)
    pub_arr, _ = synthetic_saver_pub(25, seed=4)
    pub = vsp.SaverPublicKey(pub_arr)
    pk = vsp.ProvingKey(ctx, arr)
    rng = random.Random(19)
    prim, aux, rs = [], [], []
    for b in range(k):
        _, p, a = synthetic_r1cs(nc, ni, seed=b)
        prim.append(cref.ints_to_limbs(p, 4))
        aux.append(cref.ints_to_limbs(a, 4))
        rs.append(cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4))
    prim, aux, rs = np.stack(prim), np.stack(aux), np.stack(rs)
    first = pk.prove_batch(prim, aux, rs, pub)
    s0 = ctx.scratch_stats()
    assert s0["driver_allocations"] > 0 and s0["owned_bytes"] >= s0["unused_bytes"] > 0
    for _ in range(3):
        assert np.array_equal(pk.prove_batch(prim, aux, rs, pub), first)
    s1 = ctx.scratch_stats()
    assert s1["driver_allocations"] == s0["driver_allocations"] and s1["owned_bytes"] == s0["owned_bytes"]
    ctx.trim()
    s2 = ctx.scratch_stats()
    assert s2["unused_bytes"] == 0 and s2["owned_bytes"] == s0["owned_bytes"] - s0["unused_bytes"]
    assert np.array_equal(pk.prove_batch(prim, aux, rs, pub), first)
    assert ctx.scratch_stats()["driver_allocations"] > s2["driver_allocations"]
    pk.close()
