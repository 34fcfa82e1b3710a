"""GPU tier, SURVEY.md 8f-4: the batched CUDA witness generator (csrc/voting.cu) against the circuit module's host
restatement bit for bit, against the oracle's R1CS evaluator, and through the whole protocol loop on the real ballot circuit
— keys generated on the GPU from the circuit's R1CS, ballots witnessed, encrypted and proved on the GPU, accepted by the GPU
verifier and by the oracle's pairing check (the reference tests' own gate, libs/pubkey/test/elgamal_verifiable.cpp:442-489)."""
import importlib
import random

import numpy as np
import pytest

from oracle import cref, pyref as o

pytestmark = pytest.mark.gpu
vt = importlib.import_module("vote-saver-protocol_b200.voting")


def _limbs(x):
    return np.frombuffer(int(x).to_bytes(32, "little"), dtype=np.uint64).copy()


def _ballots(c, rng, k, leaves=None, sks=None):
    votes = [rng.randrange(c.msg_size) for _ in range(k)]
    if leaves is not None:
        idx = [rng.randrange(len(leaves)) for _ in range(k)]
        keys = [sks[i] for i in idx]
        paths = [vt.merkle_path_native(leaves, i) for i in idx]
    else:  # a deep tree: random sibling digests stand for the other 2^depth - 1 voters
        idx = [rng.randrange(1 << c.depth) for _ in range(k)]
        keys = [rng.getrandbits(255) for _ in range(k)]
        paths = [[rng.randrange(o.R) for _ in range(c.depth)] for _ in range(k)]
    return votes, idx, keys, paths


@pytest.mark.parametrize("depth,k", [(2, 5), (16, 3), (0, 2)])
def test_gpu_witness_equals_host_restatement_and_satisfies(vsp, ctx, depth, k):
    rng = random.Random(40 + depth)
    c = vt.VotingCircuit(depth=depth)
    sks = leaves = None
    if depth == 2:
        sks = [rng.getrandbits(255) for _ in range(4)]
        leaves = [vt.pedersen_hash_native(vt.digest_bits(s)) for s in sks]
    votes, idx, keys, paths = _ballots(c, rng, k, leaves, sks)
    eid = rng.getrandbits(64)
    gen = vt.WitnessGenerator(ctx, c)
    got = gen.generate(np.array(votes, np.uint32), eid, np.stack([_limbs(s) for s in keys]), np.array(idx, np.uint32),
                       np.array([[_limbs(d) for d in p] for p in paths], dtype=np.uint64).reshape(k, max(depth, 1) if depth else 0, 4)
                       if depth else np.zeros((k, 1, 4), np.uint64))
    cs = o.R1CS(c.num_inputs, c.num_aux, c.constraints)
    for b in range(k):
        prim, aux = c.witness(votes[b], eid, keys[b], idx[b], paths[b])
        want = vt.ints_to_fr([1] + prim + aux)
        assert np.array_equal(got[b], want), (depth, b, int(np.argmax((got[b] != want).any(axis=1))))
        full = cref.limbs_to_ints(got[b])
        assert cs.is_satisfied(full[1:c.num_inputs + 1], full[c.num_inputs + 1:])
    if leaves is not None:
        rt = vt.merkle_root_native(leaves)
        for b in range(k):
            full = cref.limbs_to_ints(got[b])
            assert full[29] + (full[30] << 254) == rt
    with pytest.raises(ValueError):
        gen.generate(np.array([c.msg_size], np.uint32), eid, _limbs(1)[None], np.array([0], np.uint32),
                     np.zeros((1, max(depth, 1), 4), np.uint64))
    gen.close()


def test_ballots_on_the_real_circuit_prove_and_verify(vsp, ctx):
    """BASELINE configs[0]: 4 voters (tree depth 2), msg_size 25, eid 64 bits: Groth16 and SAVER keys generated on the GPU from
    the circuit's R1CS, 4 ballots witnessed + encrypted + proved on the GPU, every one accepted by verify_encryption_batch, a
    ballot whose public root is forged rejected, one ballot put through the oracle's pairing check, and the tally decrypted"""
    import torch
    rng = random.Random(77)
    c = vt.VotingCircuit(depth=2)
    arr = c.r1cs_arrays()
    pk_arr, vk = vsp.generate(ctx, arr, [rng.randrange(1, o.R) for _ in range(5)])
    pub_arr, rho, svk = vsp.saver_generate_keypair(ctx, vk, c.msg_size, [rng.randrange(1, o.R) for _ in range(3 * c.msg_size + 2)])
    sks = [rng.getrandbits(255) for _ in range(4)]
    leaves = [vt.pedersen_hash_native(vt.digest_bits(s)) for s in sks]
    k = 4
    votes = [3, 3, 7, 0]
    idx = [0, 1, 2, 3]
    eid = rng.getrandbits(64)
    gen = vt.WitnessGenerator(ctx, c)
    d_assign = torch.zeros((k, c.num_vars + 1, 4), dtype=torch.int64, device="cuda")
    gen.generate_dev(k, np.array(votes, np.uint32), eid, np.stack([_limbs(s) for s in sks]), np.array(idx, np.uint32),
                     np.array([[_limbs(d) for d in vt.merkle_path_native(leaves, i)] for i in idx], dtype=np.uint64), d_assign.data_ptr())
    pk = vsp.ProvingKey(ctx, pk_arr)
    pub = vsp.SaverPublicKey(pub_arr)
    gabc = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + c.msg_size])
    key = pub.resident(ctx, gabc)
    rs = np.stack([cref.ints_to_limbs([rng.randrange(o.R) for _ in range(3)], 4) for _ in range(k)])
    d_cts = torch.zeros((k, c.msg_size + 2, 12), dtype=torch.int64, device="cuda")
    d_proofs = torch.zeros((k, 48), dtype=torch.int64, device="cuda")
    pk.encrypt_batch_dev(key, k, d_assign.data_ptr(), rs, d_cts.data_ptr(), d_proofs.data_ptr())
    ctx.synchronize()
    cts, proofs = d_cts.cpu().numpy().view(np.uint64), d_proofs.cpu().numpy().view(np.uint64)
    assign = d_assign.cpu().numpy().view(np.uint64)
    rest = np.ascontiguousarray(assign[:, 1 + c.msg_size:1 + c.num_inputs])  # eid | sn | rt
    assert vsp.verify_encryption_batch(ctx, vk, pub, cts, rest, proofs).all()
    forged = rest.copy()
    forged[1, 3, 0] ^= np.uint64(1)  # another Merkle root
    assert vsp.verify_encryption_batch(ctx, vk, pub, cts, forged, proofs).tolist() == [True, False, True, True]
    # the oracle's verifier on ballot 2
    from test_gpu_keygen import _vk_from_arrays
    ovk = _vk_from_arrays(vk)
    J1 = lambda a: o.G1.from_affine(cref.g1_from_limbs(np.ascontiguousarray(a).reshape(1, 12))[0])
    J2 = lambda a: o.G2.from_affine(cref.g2_from_limbs(np.ascontiguousarray(a).reshape(1, 24))[0])
    opub = o.SaverPublicKey(J1(pub_arr["delta_g1"]), [J1(p) for p in pub_arr["delta_s_g1"]], [J1(p) for p in pub_arr["t_g1"]],
                            [J2(p) for p in pub_arr["t_g2"]], J1(pub_arr["delta_sum_s_g1"]), J1(pub_arr["gamma_inverse_sum_s_g1"]))
    pr = (J1(proofs[2, :12]), J2(proofs[2, 12:36]), J1(proofs[2, 36:]))
    assert o.saver_verify_encryption(ovk, opub, [J1(p) for p in cts[2]], cref.limbs_to_ints(rest[2]), pr)
    # tally: the sum of the cipher texts decrypts to the vote counts
    total = cts[0]
    for b in range(1, k):
        total = np.stack([cref.g1_add(total[i], cts[b][i]) for i in range(c.msg_size + 2)])
    counts, c0 = vsp.decrypt(ctx, total, rho, svk, gabc, block_bits=4)
    assert counts == [votes.count(i) for i in range(c.msg_size)]
    assert vsp.verify_decryption(ctx, total, counts, svk, gabc, c0)
    key.close()
    pk.close()
    gen.close()


@pytest.mark.xfail(strict=False, reason="vote_phase.py's GPU half was written after the round's GPU minutes were spent and has "
                                        "not run on hardware yet; its host half is covered by tests/test_vote_phase.py")
def test_vote_phase_from_blobs_to_blobs():
    """generate_vote (bin/cli/src/wasm.cpp:112-147) / process_encrypted_input_mode_vote_phase (common.hpp:985-1166) through
    vote-saver-protocol_b200/vote_phase.py: tests/vote_phase_check.py in a process of its own (a fault there cannot reach the
    tests collected after this one)"""
    import json
    import os
    import subprocess
    import sys
    script = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vote_phase_check.py")
    run = subprocess.run([sys.executable, script], capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stderr[-2000:]
    res = json.loads(run.stdout.strip().splitlines()[-1])
    for key in ("sizes_ok", "accepted", "forged_root_rejected", "sn_is_native_hash", "public_slice_is_eid_sn_rt",
                "deterministic_given_scalars", "batch_accepted", "tally", "tally_proof", "vk_blob_round_trip", "tally_blobs",
                "tally_blobs_verify", "wrong_result_rejected", "key_blobs_are_the_generated_keys", "verify_votes_from_blobs"):
        assert res[key] is True, (key, res)
