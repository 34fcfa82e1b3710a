"""CPU tier: the host half of the blobs-to-blobs vote phase (vote-saver-protocol_b200/vote_phase.py) — deserialisation of what
process_encrypted_input_mode_vote_phase reads (bin/cli/include/nil/vote_saver/common.hpp:996-1006), the reference's
root check, the sk bit array, and the four output blobs of serialize_data (common.hpp:462-485, 1149-1160).  The GPU half is
tests/test_gpu_voting.py."""
import importlib
import random
import struct

import numpy as np
import pytest

from oracle import cref, pyref as o

from helpers import golden, random_g1, random_g2, synthetic_pk_arrays, synthetic_r1cs

importlib.import_module("vote-saver-protocol_b200")
codec = importlib.import_module("vote-saver-protocol_b200.codec")
vp = importlib.import_module("vote-saver-protocol_b200.vote_phase")
vt = importlib.import_module("vote-saver-protocol_b200.voting")


def _tree(depth, rng):
    sks = [rng.getrandbits(255) for _ in range(1 << depth)]
    leaves = [vt.pedersen_hash_native(vt.digest_bits(s)) for s in sks]
    nodes, level = list(leaves), list(leaves)
    while len(level) > 1:
        level = [vt.pedersen_hash_native(vt.digest_bits(level[i]) + vt.digest_bits(level[i + 1])) for i in range(0, len(level), 2)]
        nodes += level
    return sks, leaves, nodes


def _blobs(depth=2, eid_bits=64, seed=5):
    rng = random.Random(seed)
    sks, leaves, nodes = _tree(depth, rng)
    eid = rng.getrandbits(eid_bits)
    g = golden("databin_groth16.json")
    g1, g2 = random_g1(40, seed), random_g2(30, seed + 1)
    n = 25
    pub = {"msg_size": n, "delta_g1": g1[0:1].copy(), "delta_sum_s_g1": g1[1:2].copy(), "gamma_inverse_sum_s_g1": g1[2:3].copy(),
           "delta_s_g1": random_g1(n, seed + 2), "t_g1": random_g1(n, seed + 3), "t_g2": g2[:n + 1].copy()}
    cs, _, _ = synthetic_r1cs(40, 30, seed=seed, binary=True)
    pk_arr = synthetic_pk_arrays(cs, seed)
    gt = cref.ints_to_limbs([int(x, 16) for x in g["alpha_g1_beta_g2"]], 6).reshape(-1)
    vk_blob = codec.groth16_vk_bytes(gt, g2[26], g2[27], g1[3], g1[4], random_g1(31, seed + 4))
    root_bits = vt.digest_bits(nodes[-1])
    blobs = {
        "merkle_tree_blob": codec.merkle_tree_bytes(cref.ints_to_limbs(nodes, 4), depth),
        "rt_blob": codec.primary_input_bytes(codec.bits_to_field_elements(root_bits)),
        "eid_blob": codec.primary_input_bytes(cref.ints_to_limbs([eid], 4)),
        "pk_eid_blob": codec.saver_pubkey_bytes(pub),
        "proving_key_blob": codec.write_fast_proving_key(pk_arr),
        "verification_key_blob": vk_blob,
    }
    return blobs, sks, leaves, nodes, eid, pub, pk_arr


def test_parse_vote_inputs_reads_every_blob_and_checks_the_root():
    blobs, sks, leaves, nodes, eid, pub, pk_arr = _blobs()
    inp = vp.parse_vote_inputs(2, 64, **blobs)
    assert inp.eid == eid and inp.tree.shape == (7, 4)
    assert cref.limbs_to_ints(inp.tree) == nodes
    assert cref.limbs_to_ints(inp.rt_packed) == [nodes[-1] & ((1 << 254) - 1), nodes[-1] >> 254]
    assert inp.pubkey["msg_size"] == 25 and np.array_equal(inp.pubkey["t_g2"], pub["t_g2"])
    assert int(inp.proving_key["cs"]["num_constraints"]) == 40 and np.array_equal(inp.proving_key["A_query"], pk_arr["A_query"])
    assert inp.verification_key["gamma_ABC_g1"].shape == (31, 12)
    # the copath the phase hands to the witness generator is the tree's
    for idx in range(4):
        assert cref.limbs_to_ints(codec.merkle_copath(inp.tree, 2, idx)) == vt.merkle_path_native(leaves, idx)
    # BOOST_ASSERT(rt_field == admin_rt_field), common.hpp:1025: a tree that is not the published one is refused
    other = dict(blobs, rt_blob=codec.primary_input_bytes(cref.ints_to_limbs([5, 0], 4)))
    with pytest.raises(ValueError, match="published root"):
        vp.parse_vote_inputs(2, 64, **other)
    with pytest.raises(ValueError):
        vp.parse_vote_inputs(3, 64, **blobs)  # 7 digests are not a depth-3 tree
    with pytest.raises(ValueError):
        vp.parse_vote_inputs(2, 65, **blobs)


def test_eid_bits_come_from_the_packed_element():
    """common.hpp:1010-1015: eid[i] = bit i of eid_field[i / 254]; bits past eid_bits are not part of the election id"""
    blobs, *_ = _blobs(eid_bits=64, seed=6)
    wide = dict(blobs, eid_blob=codec.primary_input_bytes(cref.ints_to_limbs([(1 << 70) | 0xabc], 4)))
    assert vp.parse_vote_inputs(2, 12, **wide).eid == 0xabc
    two = dict(blobs, eid_blob=codec.primary_input_bytes(cref.ints_to_limbs([1, 2], 4)))
    with pytest.raises(ValueError, match="field elements"):
        vp.parse_vote_inputs(2, 64, **two)


def test_secret_key_blob_is_the_bit_array_of_the_key():
    sk = random.Random(7).getrandbits(255) | 1 | (1 << 254)
    blob = codec.bitarray_bytes(vt.digest_bits(sk))
    assert len(blob) == 32 and blob[0] & 0x80 and blob[31] & 0x02  # sk[0] leads the first octet, sk[254] is bit 1 of the last
    assert cref.limbs_to_ints(vp.secret_key_from_blob(blob).reshape(1, 4))[0] == sk
    with pytest.raises(ValueError):
        vp.secret_key_from_blob(blob[:31])


def test_serialize_vote_writes_the_four_blobs_of_serialize_data():
    """proof | primary input from the eid on | cipher text | sn  (common.hpp:1149-1160)"""
    rng = random.Random(9)
    primary = cref.ints_to_limbs([0] * 7 + [1] + [0] * 17 + [rng.randrange(o.R) for _ in range(5)], 4)  # m | eid | sn(2) | rt(2)
    ct = random_g1(27, 10)
    proof = np.concatenate([random_g1(1, 11).reshape(-1), random_g2(1, 12).reshape(-1), random_g1(1, 13).reshape(-1)])
    proof_blob, pinput_blob, ct_blob, sn_blob = vp.serialize_vote(proof, primary, ct, 25, 1)
    assert proof_blob == codec.proof_bytes(proof) and len(proof_blob) == 192
    ints = cref.limbs_to_ints(primary)
    assert pinput_blob == struct.pack(">Q", 5) + b"".join(x.to_bytes(32, "big") for x in ints[25:])
    assert sn_blob == struct.pack(">Q", 2) + b"".join(x.to_bytes(32, "big") for x in ints[26:28])
    assert ct_blob == codec.ciphertext_bytes(ct) and np.array_equal(codec.ciphertext_from_bytes(ct_blob), ct)
    with pytest.raises(ValueError):
        vp.serialize_vote(proof, primary[:-1], ct, 25, 1)


class _HostWitness:
    """stands in for voting.WitnessGenerator (the CUDA generator) with the circuit module's host restatement, which the GPU
    generator equals bit for bit (tests/test_gpu_voting.py)"""

    def __init__(self, circuit):
        self.circuit, self.calls = circuit, []

    def generate(self, votes, eid, sks, indices, paths):
        self.calls.append((votes, eid, sks, indices, paths))
        rows = []
        for b in range(len(votes)):
            path = cref.limbs_to_ints(np.ascontiguousarray(paths[b]))[:self.circuit.depth]
            prim, aux = self.circuit.witness(int(votes[b]), int(eid), cref.limbs_to_ints(sks[b].reshape(1, 4))[0], int(indices[b]), path)
            rows.append(vt.ints_to_fr([1] + prim + aux))
        return np.stack(rows)

    def close(self):
        pass


class _EchoProver:
    """stands in for ProvingKey.encrypt_batch / rerandomize_batch: fixed points of the right shapes, and a record of what the
    phase passed in"""

    def __init__(self, msg_size):
        self.msg_size, self.seen = msg_size, {}

    def encrypt_batch(self, key, primary, auxiliary, rs):
        self.seen.update(primary=primary, auxiliary=auxiliary, rs=rs)
        k = rs.shape[0]
        cts = np.stack([random_g1(self.msg_size + 2, 100 + b) for b in range(k)])
        proofs = np.stack([np.concatenate([random_g1(1, 200 + b).reshape(-1), random_g2(1, 300 + b).reshape(-1),
                                           random_g1(1, 400 + b).reshape(-1)]) for b in range(k)])
        return cts, proofs

    def rerandomize_batch(self, key, rnd, cts, proofs):
        self.seen["rnd"] = rnd
        return cts[:, ::-1].copy(), proofs  # visibly different cipher texts

    def close(self):
        pass


def test_vote_phase_glue_with_host_stand_ins_for_the_three_gpu_calls():
    """everything VotePhase.generate_votes does around the witness generator, encrypt_batch and rerandomize_batch: voter index ->
    copath from the tree blob, sk blob -> key bits, eid element -> eid bits, the assignment split into primary | auxiliary, the
    scalars handed through, and the four blobs.  sn and rt in the published slice equal independent native computations, so
    every bit order on the way in is the circuit's."""
    blobs, sks, leaves, nodes, eid, pub, pk_arr = _blobs(depth=2, eid_bits=64, seed=15)
    inputs = vp.parse_vote_inputs(2, 64, **blobs)
    c = vt.VotingCircuit(depth=2)
    phase = object.__new__(vp.VotePhase)
    phase.ctx, phase.inputs, phase.circuit, phase.msg_size = None, inputs, c, c.msg_size
    phase.witness, phase.pk, phase.key, phase.pub = _HostWitness(c), _EchoProver(c.msg_size), None, None
    ballots = [(2, 7, codec.bitarray_bytes(vt.digest_bits(sks[2]))), (0, 24, codec.bitarray_bytes(vt.digest_bits(sks[0])))]
    rs = cref.ints_to_limbs(list(range(1, 7)), 4).reshape(2, 3, 4)
    rnd = cref.ints_to_limbs(list(range(11, 17)), 4).reshape(2, 3, 4)
    out = phase.generate_votes(ballots, rs=rs, rnd=rnd)
    assert np.array_equal(phase.pk.seen["rs"], rs) and np.array_equal(phase.pk.seen["rnd"], rnd)
    assert phase.pk.seen["primary"].shape == (2, c.num_inputs, 4) and phase.pk.seen["auxiliary"].shape == (2, c.num_aux, 4)
    cs = o.R1CS(c.num_inputs, c.num_aux, c.constraints)
    pack = lambda x: [x & ((1 << 254) - 1), x >> 254]
    for b, (idx, vote, _) in enumerate(ballots):
        proof_blob, pinput_blob, ct_blob, sn_blob = out[b]
        prim = cref.limbs_to_ints(phase.pk.seen["primary"][b])
        assert cs.is_satisfied(prim, cref.limbs_to_ints(phase.pk.seen["auxiliary"][b]))
        assert prim[:25] == [int(i == vote) for i in range(25)]
        sn = vt.pedersen_hash_native([(eid >> i) & 1 for i in range(64)] + vt.digest_bits(sks[idx]))
        assert cref.limbs_to_ints(codec.primary_input_from_bytes(pinput_blob)) == [eid] + pack(sn) + pack(nodes[-1])
        assert cref.limbs_to_ints(codec.primary_input_from_bytes(sn_blob)) == pack(sn)
        assert np.array_equal(codec.ciphertext_from_bytes(ct_blob), random_g1(27, 100 + b)[::-1])  # the rerandomised ones
        assert len(proof_blob) == 192
    # without rerandomisation the encrypt step's cipher texts go out; secrets are drawn when none are given
    out2 = phase.generate_votes(ballots[:1], rerandomize=False)
    assert np.array_equal(codec.ciphertext_from_bytes(out2[0][2]), random_g1(27, 100))
    drawn = cref.limbs_to_ints(phase.pk.seen["rs"].reshape(3, 4))
    assert len(set(drawn)) == 3 and all(0 <= x < o.R for x in drawn)
    with pytest.raises(ValueError, match="less than number of participants"):
        phase.generate_votes([(4, 0, ballots[0][2])])
    with pytest.raises(ValueError, match="options"):
        phase.generate_votes([(0, 25, ballots[0][2])])
    assert phase.generate_votes([]) == []


def test_tally_phase_glue_with_host_stand_ins(monkeypatch):
    """tally_phase.py around its three GPU calls (vsp_g1_sum, vsp_saver_decrypt, vsp_saver_verify_decryption), which are
    replaced by the oracle's point addition and by recorders: blobs are read in the reference's formats, the cipher texts are
    summed slot by slot, the keys reach decrypt / verify_decryption in the right slots, and the two output blobs are
    serialize_tally_phase_data's (common.hpp:511-522)"""
    tp = importlib.import_module("vote-saver-protocol_b200.tally_phase")
    vsp = importlib.import_module("vote-saver-protocol_b200")
    n, k = 25, 3
    cts = np.stack([random_g1(n + 2, 500 + b) for b in range(k)])
    g2 = random_g2(2 * n + 1, 60)
    rho = random.Random(61).randrange(o.R)
    g = golden("databin_groth16.json")
    gt = cref.ints_to_limbs([int(x, 16) for x in g["alpha_g1_beta_g2"]], 6).reshape(-1)
    gabc = random_g1(31, 62)
    vk_blob = codec.groth16_vk_bytes(gt, g2[0], g2[1], random_g1(1, 63)[0], random_g1(1, 64)[0], gabc)
    vk_eid_blob = codec.saver_vk_bytes(g2[0], g2[1:1 + n], g2[1 + n:])
    sk_blob = codec.saver_sk_bytes(cref.ints_to_limbs([rho], 4)[0])
    cts_blobs = [codec.ciphertext_bytes(c) for c in cts]
    want_agg = np.stack([cref.g1_add(cref.g1_add(cts[0][i], cts[1][i]), cts[2][i]) for i in range(n + 2)])
    seen = {}

    def points_sum(ctx, pts):
        acc = pts[0]
        for p in pts[1:]:
            acc = cref.g1_add(acc, p)
        return np.asarray(acc, dtype=np.uint64).reshape(12)

    def decrypt(ctx, ct, rho_, svk, gabc_rest, block_bits=4):
        seen["decrypt"] = (np.array(ct), rho_, svk, np.array(gabc_rest), block_bits)
        return [2, 0, 1] + [0] * (n - 3), random_g1(1, 65)[0]

    def verify_decryption(ctx, ct, plain, svk, gabc_rest, proof):
        seen["verify"] = (np.array(ct), list(plain), svk, np.array(gabc_rest), np.array(proof))
        return True

    monkeypatch.setattr(vsp, "points_sum", points_sum)
    monkeypatch.setattr(vsp, "decrypt", decrypt)
    monkeypatch.setattr(vsp, "verify_decryption", verify_decryption)
    dec_proof_blob, voting_res_blob = tp.tally_admin_phase(None, 2, cts_blobs, sk_blob, vk_eid_blob, None, vk_blob)
    ct, rho_, svk, rest, bits = seen["decrypt"]
    assert np.array_equal(ct, want_agg) and rho_ == rho and bits == 4
    assert np.array_equal(svk["rho_g2"], g2[0]) and np.array_equal(svk["rho_sv_g2"], g2[1:1 + n]) and np.array_equal(svk["rho_rhov_g2"], g2[1 + n:])
    assert np.array_equal(rest, gabc[1:1 + n])
    assert dec_proof_blob == o.g1_compress(o.G1.from_affine(cref.g1_from_limbs(random_g1(1, 65))[0]))
    assert voting_res_blob == struct.pack(">Q", n) + b"".join(x.to_bytes(32, "big") for x in [2, 0, 1] + [0] * (n - 3))
    assert tp.tally_voter_phase(None, 2, cts_blobs, vk_eid_blob, None, vk_blob, voting_res_blob, dec_proof_blob) is True
    ct, plain, svk, rest, proof = seen["verify"]
    assert np.array_equal(ct, want_agg) and plain == [2, 0, 1] + [0] * (n - 3) and np.array_equal(proof, random_g1(1, 65)[0])
    assert np.array_equal(rest, gabc[1:1 + n]) and np.array_equal(svk["rho_sv_g2"], g2[1:1 + n])
    # a result of the wrong length is refused before any GPU work; so are too many ballots and ragged cipher texts
    short = struct.pack(">Q", n - 1) + bytes(32 * (n - 1))
    seen.clear()
    assert tp.tally_voter_phase(None, 2, cts_blobs, vk_eid_blob, None, vk_blob, short, dec_proof_blob) is False and not seen
    with pytest.raises(ValueError, match="participants"):
        tp.read_ciphertexts(1, cts_blobs)
    with pytest.raises(ValueError, match="Wrong size"):
        tp.read_ciphertexts(2, cts_blobs[:2] + [codec.ciphertext_bytes(cts[2][:-1])])
    assert np.array_equal(tp.aggregate(None, cts[:1]), cts[0])


def test_init_voter_and_admin_data_phases_feed_the_vote_phase():
    """init_phase.py: voter key blobs (common.hpp:824-846) and the administrator's eid / rt / merkle_tree blobs (:943-985) on
    three voters of a depth-3 election — the unclaimed leaves are all-zero keys, as deserialize_voters_public_keys pads them —
    then read back by the vote phase's parser, whose root check passes; the tree equals a plain full-tree computation"""
    ip = importlib.import_module("vote-saver-protocol_b200.init_phase")
    rng = random.Random(31)
    sks = [rng.getrandbits(255) for _ in range(3)]
    keys = [ip.init_voter_phase(i, sk=s) for i, s in enumerate(sks)]
    for (pk_blob, sk_blob), s in zip(keys, sks):
        assert len(pk_blob) == 32 and len(sk_blob) == 32
        assert cref.limbs_to_ints(vp.secret_key_from_blob(sk_blob).reshape(1, 4))[0] == s
        assert list(codec.bitarray_from_bytes(pk_blob, 255)) == vt.digest_bits(vt.pedersen_hash_native(vt.digest_bits(s)))
    drawn_pk, drawn_sk = ip.init_voter_phase(0)
    assert drawn_sk != keys[0][1] and len(drawn_pk) == 32
    eid_blob, rt_blob, tree_blob = ip.init_admin_phase_generate_data(3, 64, [k[0] for k in keys], eid=0x1234)
    leaves = [vt.pedersen_hash_native(vt.digest_bits(s)) for s in sks] + [0] * 5
    nodes, level = list(leaves), list(leaves)
    while len(level) > 1:
        level = [vt.pedersen_hash_native(vt.digest_bits(level[i]) + vt.digest_bits(level[i + 1])) for i in range(0, len(level), 2)]
        nodes += level
    assert cref.limbs_to_ints(codec.merkle_tree_from_bytes(tree_blob, 3)) == nodes
    assert cref.limbs_to_ints(codec.primary_input_from_bytes(eid_blob)) == [0x1234]
    blobs, *_ = _blobs(depth=2, seed=33)
    blobs.update(merkle_tree_blob=tree_blob, rt_blob=rt_blob, eid_blob=eid_blob)
    inp = vp.parse_vote_inputs(3, 64, **blobs)
    assert inp.eid == 0x1234 and cref.limbs_to_ints(inp.tree)[-1] == nodes[-1]
    with pytest.raises(ValueError):
        ip.init_admin_phase_generate_data(1, 64, [k[0] for k in keys])
    with pytest.raises(ValueError):
        ip.init_admin_phase_generate_data(3, 8, [], eid=256)


def test_init_admin_keys_phase_glue_with_host_stand_ins(monkeypatch):
    """init_admin_phase_generate_keys around its three GPU calls (vsp_groth16_generate, vsp_saver_keygen, one pairing): the
    five blobs of serialize_initial_phase_admin_keys (common.hpp:330-366) carry what the generators returned, and the vote
    phase's parser reads them back"""
    ip = importlib.import_module("vote-saver-protocol_b200.init_phase")
    vsp = importlib.import_module("vote-saver-protocol_b200")
    cs, _, _ = synthetic_r1cs(40, 30, seed=3, binary=True)
    pk_arr = synthetic_pk_arrays(cs, 3)
    g1, g2 = random_g1(40, 70), random_g2(60, 71)
    vk = {"alpha_g1": g1[0:1], "beta_g2": g2[0:1], "gamma_g2": g2[1:2], "delta_g2": g2[2:3], "gamma_g1": g1[1:2], "delta_g1": g1[2:3],
          "gamma_ABC_g1": g1[3:34]}
    n = 25
    pub = {"msg_size": n, "delta_g1": g1[2:3].copy(), "delta_s_g1": random_g1(n, 72), "t_g1": random_g1(n, 73), "t_g2": g2[3:3 + n + 1].copy(),
           "delta_sum_s_g1": g1[34:35].copy(), "gamma_inverse_sum_s_g1": g1[35:36].copy()}
    svk = {"rho_g2": g2[30:31], "rho_sv_g2": g2[31:31 + n].copy(), "rho_rhov_g2": random_g2(n, 74)}
    gt = cref.ints_to_limbs([int(x, 16) for x in golden("databin_groth16.json")["alpha_g1_beta_g2"]], 6).reshape(1, 72)
    seen = {}
    monkeypatch.setattr(vsp, "generate", lambda ctx, arrays, toxic: (seen.update(cs=arrays, toxic=toxic), (pk_arr, vk))[1])
    monkeypatch.setattr(vsp, "saver_generate_keypair", lambda ctx, vk_, m, rnd: (seen.update(rnd=rnd, m=m), (pub, 12345, svk))[1])
    monkeypatch.setattr(vsp, "pairing_products", lambda ctx, a, b: (seen.update(pair=(a, b)), (gt, np.array([False])))[1])
    pk_blob, vk_blob, pub_blob, sk_blob, svk_blob = ip.init_admin_phase_generate_keys(None, 0, 8)
    c = vt.VotingCircuit(depth=0, eid_bits=8)
    assert int(seen["cs"]["num_constraints"]) == len(c.constraints) and int(seen["cs"]["num_inputs"]) == 30
    assert len(seen["toxic"]) == 5 and len(seen["rnd"]) == 3 * n + 2 and all(0 < x < o.R for x in seen["toxic"] + seen["rnd"])
    assert np.array_equal(seen["pair"][0], vk["alpha_g1"]) and np.array_equal(seen["pair"][1], vk["beta_g2"])
    back = codec.groth16_vk_from_bytes(vk_blob)
    assert np.array_equal(back["alpha_g1_beta_g2"], gt[0]) and np.array_equal(back["gamma_ABC_g1"], vk["gamma_ABC_g1"])
    assert np.array_equal(back["delta_g1"], vk["delta_g1"][0]) and np.array_equal(back["gamma_g1"], vk["gamma_g1"][0])
    assert np.array_equal(codec.read_fast_proving_key(pk_blob)["H_query"], pk_arr["H_query"])
    assert np.array_equal(codec.saver_pubkey_from_bytes(pub_blob)["t_g2"], pub["t_g2"])
    assert cref.limbs_to_ints(codec.saver_sk_from_bytes(sk_blob).reshape(1, 4))[0] == 12345
    a, b, cc = codec.saver_vk_from_bytes(svk_blob)
    assert np.array_equal(a, svk["rho_g2"][0]) and np.array_equal(b, svk["rho_sv_g2"]) and np.array_equal(cc, svk["rho_rhov_g2"])


def test_verify_votes_glue_with_a_host_stand_in(monkeypatch):
    """vote_phase.verify_votes around vsp_saver_verify_encryption_gt_batch: ballots whose blobs do not decode or do not fit the
    keys are False without reaching the GPU, the others go in one batch with the key in its serialised (GT) form"""
    vsp = importlib.import_module("vote-saver-protocol_b200")
    blobs, *_ = _blobs(seed=44)
    rng = random.Random(45)
    seen = {}

    class Pub:
        def __init__(self, arr):
            self.arrays, self.msg_size = arr, int(arr["msg_size"])

    def verify(ctx, vk, pub, cts, rests, proofs):
        seen.update(vk=vk, cts=np.array(cts), rests=np.array(rests), proofs=np.array(proofs))
        return np.array([True, False][:len(proofs)] + [True] * max(0, len(proofs) - 2))

    monkeypatch.setattr(vsp, "SaverPublicKey", Pub)
    monkeypatch.setattr(vsp, "verify_encryption_gt_batch", verify)

    def ballot(seed, n_ct=27, n_rest=5):
        proof = np.concatenate([random_g1(1, seed).reshape(-1), random_g2(1, seed).reshape(-1), random_g1(1, seed + 1).reshape(-1)])
        rest = cref.ints_to_limbs([rng.randrange(o.R) for _ in range(n_rest)], 4)
        return codec.proof_bytes(proof), codec.primary_input_bytes(rest), codec.ciphertext_bytes(random_g1(n_ct, seed + 2))

    good = [ballot(1), ballot(2), ballot(3)]
    broken_proof = (bytes(192), good[0][1], good[0][2])
    short_ct = ballot(4, n_ct=26)
    long_rest = ballot(5, n_rest=6)
    unreduced = (good[0][0], struct.pack(">Q", 5) + o.R.to_bytes(32, "big") + bytes(32 * 4), good[0][2])
    got = vp.verify_votes(None, blobs["verification_key_blob"], blobs["pk_eid_blob"],
                          [good[0], broken_proof, good[1], short_ct, long_rest, unreduced, good[2]])
    assert got == [True, False, False, False, False, False, True]
    assert seen["cts"].shape == (3, 27, 12) and seen["rests"].shape == (3, 5, 4) and seen["proofs"].shape == (3, 48)
    assert np.array_equal(seen["proofs"][1], codec.proof_from_bytes(good[1][0]))
    assert seen["vk"]["alpha_g1_beta_g2"].shape == (72,) and seen["vk"]["gamma_ABC_g1"].shape == (31, 12)
    seen.clear()
    assert vp.verify_votes(None, blobs["verification_key_blob"], blobs["pk_eid_blob"], [broken_proof]) == [False] and not seen
    assert vp.verify_votes(None, blobs["verification_key_blob"], blobs["pk_eid_blob"], []) == []


def test_vote_phase_refuses_a_key_for_another_circuit():
    """a proving key whose constraint system is not this package's ballot circuit (e.g. one made by the reference's own
    generator over crypto3-blueprint's components) is rejected before anything is uploaded"""
    blobs, *_ = _blobs(depth=2, seed=51)
    inputs = vp.parse_vote_inputs(2, 64, **blobs)  # the synthetic 40-constraint key of _blobs
    with pytest.raises(ValueError, match="is not the ballot circuit"):
        vp.VotePhase(None, inputs)
