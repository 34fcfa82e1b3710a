"""Run by tests/test_gpu_voting.py in a process of its own: the blobs-to-blobs vote phase (vote-saver-protocol_b200/vote_phase.py)
on the real depth-2 ballot circuit, BASELINE configs[0] — every input in the reference's byte formats (common.hpp:996-1006),
the blobs made by init_phase.py's three phases, the four output blobs parsed back and checked: the ballot is accepted by verify_encryption, a forged root is not, sn is the
native Pedersen hash of eid || sk, the public slice is eid | sn | rt, and the tally of four voters decrypts to their votes, directly and through
tally_phase.py's blobs.
Prints one JSON line."""
import importlib
import json
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

from oracle import cref, pyref as o  # noqa: E402  (the checker)

vsp = importlib.import_module("vote-saver-protocol_b200")
codec = importlib.import_module("vote-saver-protocol_b200.codec")
vp = importlib.import_module("vote-saver-protocol_b200.vote_phase")
vt = importlib.import_module("vote-saver-protocol_b200.voting")


def main():
    rng = random.Random(21)
    depth, eid_bits = 2, 64
    res = {}
    with vsp.Context(0) as ctx:
        ip = importlib.import_module("vote-saver-protocol_b200.init_phase")
        c = vt.VotingCircuit(depth=depth)
        # the initial phases (common.hpp:824-985), blobs out: administrator keys, voter keys, session data
        toxic = [rng.randrange(1, o.R) for _ in range(5)]
        keygen_rnd = [rng.randrange(1, o.R) for _ in range(3 * c.msg_size + 2)]
        pk_blob, vk_blob, pk_eid_blob, sk_eid_blob, vk_eid_blob = ip.init_admin_phase_generate_keys(ctx, depth, eid_bits, toxic=toxic,
                                                                                                  rnd=keygen_rnd)
        sks = [rng.getrandbits(255) for _ in range(1 << depth)]
        voter_keys = [ip.init_voter_phase(i, sk=s) for i, s in enumerate(sks)]
        eid = rng.getrandbits(eid_bits)
        eid_blob, rt_blob, tree_blob = ip.init_admin_phase_generate_data(depth, eid_bits, [k[0] for k in voter_keys], eid=eid)
        blobs = dict(merkle_tree_blob=tree_blob, rt_blob=rt_blob, eid_blob=eid_blob, pk_eid_blob=pk_eid_blob,
                     proving_key_blob=pk_blob, verification_key_blob=vk_blob)
        sk_blobs = [k[1] for k in voter_keys]
        nodes = cref.limbs_to_ints(codec.merkle_tree_from_bytes(tree_blob, depth))
        # the same keys as arrays, for the verifier calls below (the generators are deterministic in their secrets)
        pk_arr, vk = vsp.generate(ctx, c.r1cs_arrays(), toxic)
        pub_arr, rho, svk = vsp.saver_generate_keypair(ctx, vk, c.msg_size, keygen_rnd)
        gt, _ = vsp.pairing_products(ctx, vk["alpha_g1"], vk["beta_g2"])
        res["key_blobs_are_the_generated_keys"] = bool(
            pk_blob == codec.write_fast_proving_key(pk_arr) and pk_eid_blob == codec.saver_pubkey_bytes(pub_arr) and
            sk_eid_blob == codec.saver_sk_bytes(cref.ints_to_limbs([rho], 4)[0]))
        pub = vsp.SaverPublicKey(pub_arr)
        gabc = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + c.msg_size])

        # one ballot through the reference's own entry point shape
        voter, vote = 2, 7
        proof_blob, pinput_blob, ct_blob, sn_blob = vp.generate_vote(ctx, depth, eid_bits, voter, vote, blobs["merkle_tree_blob"],
                                                                    blobs["rt_blob"], blobs["eid_blob"], sk_blobs[voter],
                                                                    blobs["pk_eid_blob"], blobs["proving_key_blob"],
                                                                    blobs["verification_key_blob"])
        proof, ct = codec.proof_from_bytes(proof_blob), codec.ciphertext_from_bytes(ct_blob)
        rest = codec.primary_input_from_bytes(pinput_blob)
        res["blob_sizes"] = [len(proof_blob), len(pinput_blob), len(ct_blob), len(sn_blob)]
        res["sizes_ok"] = res["blob_sizes"] == [192, 8 + 32 * 5, 8 + 48 * (c.msg_size + 2), 8 + 32 * 2]
        res["accepted"] = bool(vsp.verify_encryption_batch(ctx, vk, pub, ct[None], rest[None], proof[None])[0])
        forged = rest.copy()
        forged[3, 0] ^= np.uint64(1)
        res["forged_root_rejected"] = not bool(vsp.verify_encryption_batch(ctx, vk, pub, ct[None], forged[None], proof[None])[0])
        sn = vt.pedersen_hash_native([(eid >> i) & 1 for i in range(eid_bits)] + vt.digest_bits(sks[voter]))
        pack = lambda x: [x & ((1 << 254) - 1), x >> 254]
        res["sn_is_native_hash"] = cref.limbs_to_ints(codec.primary_input_from_bytes(sn_blob)) == pack(sn)
        res["public_slice_is_eid_sn_rt"] = cref.limbs_to_ints(rest) == [eid] + pack(sn) + pack(nodes[-1])

        # the four voters in one call on resident keys, fixed scalars: twice the same blobs; then the tally
        inputs = vp.parse_vote_inputs(depth, eid_bits, **blobs)
        votes = [3, 3, 7, 0]
        ballots = [(i, votes[i], sk_blobs[i]) for i in range(4)]
        rs = np.stack([cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4) for _ in range(4)])
        rnd = np.stack([cref.ints_to_limbs([rng.randrange(1, o.R) for _ in range(3)], 4) for _ in range(4)])
        with vp.VotePhase(ctx, inputs) as phase:
            out = phase.generate_votes(ballots, rs=rs, rnd=rnd)
            res["deterministic_given_scalars"] = out == phase.generate_votes(ballots, rs=rs, rnd=rnd)
        cts = np.stack([codec.ciphertext_from_bytes(b[2]) for b in out])
        proofs = np.stack([codec.proof_from_bytes(b[0]) for b in out])
        rests = np.stack([codec.primary_input_from_bytes(b[1]) for b in out])
        res["batch_accepted"] = bool(vsp.verify_encryption_batch(ctx, vk, pub, cts, rests, proofs).all())
        # the verifier that only holds blobs: the key in its serialised (GT) form, vsp_saver_verify_encryption_gt_batch
        published = [(b[0], b[1], b[2]) for b in out]
        tampered = (published[1][0], published[0][1], published[1][2])  # voter 1's proof and cipher text with voter 0's sn
        res["verify_votes_from_blobs"] = vp.verify_votes(ctx, blobs["verification_key_blob"], blobs["pk_eid_blob"],
                                                         published + [tampered]) == [True, True, True, True, False]
        total = cts[0]
        for b in range(1, 4):
            total = np.stack([cref.g1_add(total[i], cts[b][i]) for i in range(c.msg_size + 2)])
        counts, c0 = vsp.decrypt(ctx, total, rho, svk, gabc, block_bits=4)
        res["tally"] = counts == [votes.count(i) for i in range(c.msg_size)]
        res["tally_proof"] = bool(vsp.verify_decryption(ctx, total, counts, svk, gabc, c0))
        # the tally phase on the same ballots, blobs in and out (common.hpp:1176-1296)
        tp = importlib.import_module("vote-saver-protocol_b200.tally_phase")
        ct_blobs = [b[2] for b in out]
        dec_proof_blob, voting_res_blob = tp.tally_admin_phase(ctx, depth, ct_blobs, sk_eid_blob, vk_eid_blob, None,
                                                               blobs["verification_key_blob"])
        res["tally_blobs"] = cref.limbs_to_ints(codec.primary_input_from_bytes(voting_res_blob)) == [votes.count(i) for i in range(c.msg_size)]
        res["tally_blobs_verify"] = tp.tally_voter_phase(ctx, depth, ct_blobs, vk_eid_blob, None, blobs["verification_key_blob"],
                                                         voting_res_blob, dec_proof_blob)
        wrong = codec.primary_input_bytes(cref.ints_to_limbs([votes.count(i) + (i == 5) for i in range(c.msg_size)], 4))
        res["wrong_result_rejected"] = not tp.tally_voter_phase(ctx, depth, ct_blobs, vk_eid_blob, None,
                                                                blobs["verification_key_blob"], wrong, dec_proof_blob)
        # the key as the vote phase deserialised it (alpha_g1_beta_g2 in GT form) is the generated one
        res["vk_blob_round_trip"] = bool(np.array_equal(inputs.verification_key["alpha_g1_beta_g2"], gt[0]) and
                                         np.array_equal(inputs.verification_key["gamma_ABC_g1"], vk["gamma_ABC_g1"]))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
