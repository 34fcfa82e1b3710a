"""The vote phase from blobs to blobs: the host-side mirror of `process_encrypted_input_mode_vote_phase`
(/root/reference/bin/cli/include/nil/vote_saver/common.hpp:985-1166) and of its C export `generate_vote`
(/root/reference/bin/cli/src/wasm.cpp:112-147) — the same inputs in the reference's own byte formats, the same four output
blobs, with everything between deserialisation and serialisation on the GPU:

    merkle_tree | rt | eid | sk | pk_eid | proving_key | verification_key          (common.hpp:996-1006)
      -> copath of the voter's leaf                                                 (:1024-1027, merkle/proof.hpp:65-91)
      -> witness of the ballot circuit                                              (:1054-1128)   vsp_voting_witness_batch_dev
      -> encrypt<verifiable_encryption>(m, {d(), pk_eid, keypair, primary, aux})    (:1131-1136)   vsp_saver_encrypt_batch
      -> rerandomize(rnd[3], cipher_text, {pk_eid, keypair, proof})                 (:1139-1146)   vsp_saver_rerandomize_batch
      -> proof | primary input after the message slots | cipher text | sn           (:1149-1160, serialize_data :462-485)

`parse_vote_inputs` and `serialize_vote` are plain host code over the C ABI's codecs (no GPU); `VotePhase` keeps the keys,
the circuit's tables and the fixed-base tables resident and takes any number of ballots per call, which is what the
reference's per-vote call cannot do.  The ballot circuit is this package's restatement (voting.py: parity with
crypto3-blueprint's wire values unpinned), so the proving key must have been generated for it (vsp_groth16_generate over
VotingCircuit.r1cs_arrays()); a key made by the reference's own generator describes crypto3's circuit and is rejected by
the size check below.

Secret scalars (the SAVER r, Groth16's r and s, the three rerandomisation values) come from the operating system
(`secrets`), as the reference draws them from random::algebraic_random_device (common.hpp:1131); tests pass them in.
"""
from __future__ import annotations

import secrets
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import codec, voting

FR_MODULUS = voting.Q


@dataclass
class VoteInputs:
    """what the vote phase deserialises once per election (common.hpp:996-1006)"""
    tree_depth: int
    eid_bits: int
    tree: np.ndarray            # (2^(depth+1) - 1, 4) digests, leaves first, root last
    eid: int                    # the election id, bit i = eid[i]
    rt_packed: np.ndarray       # the administrator's root as packed field elements (2, 4)
    pubkey: dict                # arrays of vsp_saver_pubkey_t
    proving_key: dict           # arrays of vsp_proving_key_t
    verification_key: dict      # groth16_vk_from_bytes


def _limbs_to_int(a) -> int:
    return int.from_bytes(np.ascontiguousarray(a, dtype=np.uint64).tobytes(), "little")


def _digest_bits(limbs) -> np.ndarray:
    x = _limbs_to_int(limbs)
    return np.array([(x >> i) & 1 for i in range(voting.DIGEST_BITS)], dtype=np.uint8)


def parse_vote_inputs(tree_depth: int, eid_bits: int, merkle_tree_blob: bytes, rt_blob: bytes, eid_blob: bytes,
                      pk_eid_blob: bytes, proving_key_blob: bytes, verification_key_blob: bytes) -> VoteInputs:
    """common.hpp:996-1006, 1010-1015, 1024-1025: deserialise, rebuild the eid bits from its packed elements, and check
    that the tree's root is the root the administrator published (the reference's BOOST_ASSERT(rt_field == admin_rt_field))"""
    if eid_bits < 1 or eid_bits > 64:
        raise ValueError("eid_bits must be within 1..64 (the witness generator takes the election id as one 64-bit word)")
    tree = codec.merkle_tree_from_bytes(merkle_tree_blob, tree_depth)
    rt_packed = codec.primary_input_from_bytes(rt_blob)
    eid_field = codec.primary_input_from_bytes(eid_blob)
    if eid_field.shape[0] != (eid_bits + voting.CHUNK_BITS - 1) // voting.CHUNK_BITS:
        raise ValueError("eid blob holds %d field elements, eid_bits = %d needs %d" % (
            eid_field.shape[0], eid_bits, (eid_bits + voting.CHUNK_BITS - 1) // voting.CHUNK_BITS))
    eid_all = _limbs_to_int(eid_field[0])
    eid = eid_all & ((1 << eid_bits) - 1)  # bit_test(eid_field[i / chunk_size], i % chunk_size), i < eid_bits
    want_rt = codec.bits_to_field_elements(_digest_bits(tree[-1]))
    if rt_packed.shape != want_rt.shape or not np.array_equal(rt_packed, want_rt):
        raise ValueError("the Merkle tree's root is not the published root (rt)")
    return VoteInputs(tree_depth, eid_bits, tree, eid, rt_packed, codec.saver_pubkey_from_bytes(pk_eid_blob),
                      codec.read_fast_proving_key(proving_key_blob), codec.groth16_vk_from_bytes(verification_key_blob))


def secret_key_from_blob(sk_blob: bytes) -> np.ndarray:
    """deserialize_bitarray<secret_key_bits>(sk_blob) (common.hpp:1000-1001) as the 4-limb integer whose bit i is sk[i]"""
    bits = codec.bitarray_from_bytes(sk_blob, voting.DIGEST_BITS)
    x = sum(int(b) << i for i, b in enumerate(bits))
    return np.frombuffer(x.to_bytes(32, "little"), dtype=np.uint64).copy()


def serialize_vote(proof, primary, ct, msg_size: int, eid_elements: int) -> Tuple[bytes, bytes, bytes, bytes]:
    """common.hpp:1149-1160 + serialize_data (:462-485): proof | primary input from the eid on | cipher text | the sn slice.
    primary: the ballot's full primary input (ni, 4); ct: (msg_size + 2, 12); proof: 48 limbs."""
    p = np.ascontiguousarray(primary, dtype=np.uint64).reshape(-1, 4)
    sn_offset = msg_size + eid_elements
    rt_offset = sn_offset + (voting.DIGEST_BITS + voting.CHUNK_BITS - 1) // voting.CHUNK_BITS
    if p.shape[0] != rt_offset + (voting.DIGEST_BITS + voting.CHUNK_BITS - 1) // voting.CHUNK_BITS:
        raise ValueError("primary input has %d elements, expected m | eid | sn | rt" % p.shape[0])
    return (codec.proof_bytes(proof), codec.primary_input_bytes(p[msg_size:]), codec.ciphertext_bytes(ct),
            codec.primary_input_bytes(p[sn_offset:rt_offset]))


def _random_fr(n: int) -> np.ndarray:
    return voting.ints_to_fr([secrets.randbelow(FR_MODULUS) for _ in range(n)])


class VotePhase:
    """One election's resident state on one GPU: proving key and its window tables, the SAVER key's fixed-base tables, the
    ballot circuit's witness program.  `generate_votes` then turns (voter index, vote, sk blob) triples into the reference's
    four blobs each."""

    def __init__(self, ctx, inputs: VoteInputs):
        import importlib
        vsp = importlib.import_module(__package__)
        self.ctx, self.inputs = ctx, inputs
        self.circuit = voting.VotingCircuit(depth=inputs.tree_depth, eid_bits=inputs.eid_bits,
                                            msg_size=int(inputs.pubkey["msg_size"]))
        cs = inputs.proving_key["cs"]
        if (int(cs["num_inputs"]), int(cs["num_aux"]), int(cs["num_constraints"])) != (
                self.circuit.num_inputs, self.circuit.num_aux, len(self.circuit.constraints)):
            raise ValueError("the proving key's constraint system (%d inputs, %d auxiliary, %d constraints) is not the ballot "
                             "circuit of depth %d (%d, %d, %d)" % (cs["num_inputs"], cs["num_aux"], cs["num_constraints"],
                                                                   inputs.tree_depth, self.circuit.num_inputs,
                                                                   self.circuit.num_aux, len(self.circuit.constraints)))
        gabc = inputs.verification_key["gamma_ABC_g1"]
        if gabc.shape[0] != self.circuit.num_inputs + 1:
            raise ValueError("the verification key's gamma_ABC_g1 does not match the circuit's primary input")
        self.msg_size = self.circuit.msg_size
        self.pk = vsp.ProvingKey(ctx, inputs.proving_key)
        self.pub = vsp.SaverPublicKey(inputs.pubkey)
        self.key = self.pub.resident(ctx, np.ascontiguousarray(gabc[1:1 + self.msg_size]))
        self.witness = voting.WitnessGenerator(ctx, self.circuit)

    def close(self):
        for obj in (getattr(self, "witness", None), getattr(self, "key", None), getattr(self, "pk", None)):
            if obj is not None:
                obj.close()
        self.witness = self.key = self.pk = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def generate_votes(self, ballots: Sequence[Tuple[int, int, bytes]], rs=None, rnd=None,
                       rerandomize: bool = True) -> List[Tuple[bytes, bytes, bytes, bytes]]:
        """ballots: (voter_idx, vote, sk_blob) each.  rs (k, 3, 4) = (r, s, r_saver) and rnd (k, 3, 4) = (r, z1, z2) are the
        secret scalars; left out, they are drawn from the operating system.  -> (proof_blob, pinput_blob, ct_blob, sn_blob)
        per ballot, as serialize_data writes them."""
        k = len(ballots)
        if k == 0:
            return []
        n_voters = 1 << self.inputs.tree_depth
        for idx, vote, _ in ballots:
            if not 0 <= idx < n_voters:
                raise ValueError("Voter index should be less than number of participants!")  # common.hpp:1018
            if not 0 <= vote < self.msg_size:
                raise ValueError("vote %d is not one of the %d options" % (vote, self.msg_size))
        votes = np.array([b[1] for b in ballots], dtype=np.uint32)
        indices = np.array([b[0] for b in ballots], dtype=np.uint32)
        sks = np.stack([secret_key_from_blob(b[2]) for b in ballots])
        depth = self.inputs.tree_depth
        paths = np.stack([codec.merkle_copath(self.inputs.tree, depth, b[0]) for b in ballots]) if depth else \
            np.zeros((k, 1, 4), dtype=np.uint64)
        assign = self.witness.generate(votes, self.inputs.eid, sks, indices, paths)  # (k, num_vars + 1, 4)
        ni = self.circuit.num_inputs
        primary = np.ascontiguousarray(assign[:, 1:1 + ni])
        auxiliary = np.ascontiguousarray(assign[:, 1 + ni:])
        rs = _random_fr(3 * k).reshape(k, 3, 4) if rs is None else np.ascontiguousarray(rs, dtype=np.uint64).reshape(k, 3, 4)
        cts, proofs = self.pk.encrypt_batch(self.key, primary, auxiliary, rs)
        if rerandomize:
            rnd = _random_fr(3 * k).reshape(k, 3, 4) if rnd is None else \
                np.ascontiguousarray(rnd, dtype=np.uint64).reshape(k, 3, 4)
            cts, proofs = self.pk.rerandomize_batch(self.key, rnd, cts, proofs)
        n_eid = (self.inputs.eid_bits + voting.CHUNK_BITS - 1) // voting.CHUNK_BITS
        return [serialize_vote(proofs[b], primary[b], cts[b], self.msg_size, n_eid) for b in range(k)]


def generate_votes_sharded(phase, ballots: Sequence[Tuple[int, int, bytes]], rs=None, rnd=None, rerandomize: bool = True,
                           group=None) -> List[Tuple[bytes, bytes, bytes, bytes]]:
    """proof-per-GPU sharding of a ballot list (BASELINE configs[3]): one process per GPU under torch.distributed,
    every rank holding its own VotePhase on its own GPU; rank g takes ballots i = g (mod G) (sharded.ballot_indices), no
    collective touches the data path, and the four blobs of every ballot come back in ballot order on every rank (one
    all_gather_object of the ranks' blob lists).  rs / rnd, when given, are indexed by ballot, so the result does not depend
    on the world size.  Without an initialised process group this is phase.generate_votes.  Under the NCCL backend the
    caller has made its rank's GPU the current CUDA device (torch.cuda.set_device), as all_gather_object needs."""
    from . import sharded
    rank, world = sharded._world(group)
    k = len(ballots)
    mine = sharded.ballot_indices(k, rank, world)
    pick = lambda a: None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.uint64).reshape(k, 3, 4)[mine])
    local = phase.generate_votes([ballots[int(i)] for i in mine], rs=pick(rs), rnd=pick(rnd), rerandomize=rerandomize)
    if world == 1:
        return local
    gathered = [None] * world
    sharded._dist().all_gather_object(gathered, local, group=group)
    out: List[Optional[Tuple[bytes, bytes, bytes, bytes]]] = [None] * k
    for g in range(world):
        for j, i in enumerate(sharded.ballot_indices(k, g, world)):
            out[int(i)] = gathered[g][j]
    return out  # type: ignore[return-value]


def verify_votes(ctx, verification_key_blob: bytes, pk_eid_blob: bytes,
                 ballots: Sequence[Tuple[bytes, bytes, bytes]]) -> List[bool]:
    """verify_encryption<encryption_scheme_type>(cipher_text, {pk_eid, vk, proof, primary input after the message slots}) for
    ballots given as (proof_blob, pinput_blob, ct_blob) — the vote phase's own check (common.hpp:1163-1169) and what the
    verifier of the published ballots runs — from the key blobs alone: the verification key stays in the form it is
    serialised in (alpha_g1_beta_g2 as a GT value; vsp_saver_verify_encryption_gt_batch).  A ballot whose blobs do not
    decode, or whose sizes do not fit the keys, is False."""
    import importlib
    vsp = importlib.import_module(__package__)
    vk = codec.groth16_vk_from_bytes(verification_key_blob)
    pub = vsp.SaverPublicKey(codec.saver_pubkey_from_bytes(pk_eid_blob))
    n = pub.msg_size
    n_rest = vk["gamma_ABC_g1"].shape[0] - 1 - n
    if n_rest < 0:
        raise ValueError("the verification key's gamma_ABC_g1 is shorter than the public key's message size")
    ok = [False] * len(ballots)
    rows, proofs, cts, rests = [], [], [], []
    for b, (proof_blob, pinput_blob, ct_blob) in enumerate(ballots):
        try:
            proof, ct, rest = codec.proof_from_bytes(proof_blob), codec.ciphertext_from_bytes(ct_blob), \
                codec.primary_input_from_bytes(pinput_blob)
        except ValueError:
            continue
        if ct.shape[0] != n + 2 or rest.shape[0] != n_rest or any(
                int.from_bytes(np.ascontiguousarray(r).tobytes(), "little") >= FR_MODULUS for r in rest):
            continue
        rows.append(b)
        proofs.append(proof)
        cts.append(ct)
        rests.append(rest)
    if rows:
        got = vsp.verify_encryption_gt_batch(ctx, vk, pub, np.stack(cts), np.stack(rests), np.stack(proofs))
        for b, g in zip(rows, got):
            ok[b] = bool(g)
    return ok


def generate_vote(ctx, tree_depth: int, eid_bits: int, voter_idx: int, vote: int, merkle_tree_blob: bytes, rt_blob: bytes,
                  eid_blob: bytes, sk_blob: bytes, pk_eid_blob: bytes, proving_key_blob: bytes,
                  verification_key_blob: bytes, rs: Optional[np.ndarray] = None,
                  rnd: Optional[np.ndarray] = None) -> Tuple[bytes, bytes, bytes, bytes]:
    """`generate_vote` of bin/cli/src/wasm.cpp:112-147, argument for argument: one ballot, keys loaded for this call only.
    -> (proof_blob, pinput_blob, ct_blob, sn_blob)"""
    inputs = parse_vote_inputs(tree_depth, eid_bits, merkle_tree_blob, rt_blob, eid_blob, pk_eid_blob, proving_key_blob,
                               verification_key_blob)
    with VotePhase(ctx, inputs) as phase:
        return phase.generate_votes([(voter_idx, vote, sk_blob)], rs=rs, rnd=rnd)[0]
