"""The initial phases from blobs to blobs: host-side mirrors of `process_encrypted_input_mode_init_voter_phase`,
`..._init_admin_phase_generate_keys` and `..._init_admin_phase_generate_data`
(/root/reference/bin/cli/include/nil/vote_saver/common.hpp:824-846, 849-941, 943-985).  Together with vote_phase.py and
tally_phase.py they are the whole `process_encrypted_input_mode` sequence of bin/cli (src/main.cpp:37-120), every artefact
in the reference's byte formats.

    voter   sk = 255 random bits, pk = hash<merkle_hash_type>(sk)            -> voter_pk | voter_sk bit-array blobs
    admin   R1CS of the ballot circuit -> zk::generate -> generate_keypair    -> r1cs_proving_key | r1cs_verification_key |
                                                                                public_key | secret_key | verification_key
    admin   Merkle tree over the voters' public keys, eid, rt                -> eid | rt | merkle_tree blobs

Key generation runs on the GPU (vsp_groth16_generate, vsp_saver_keygen, one pairing for alpha_g1_beta_g2); the voter's
hash and the tree are the circuit module's native Pedersen hash on the host (one hash per voter and per non-default tree
node: subtrees that hold no voter collapse to one value per level).  Random values come from `secrets`, where the reference
uses std::rand for sk / eid (common.hpp:809-821, 963-965) and algebraic_random_device for key material (:921-925).
"""
from __future__ import annotations

import importlib
import secrets
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import codec, voting


def _vsp():
    return importlib.import_module(__package__)


def _limbs(x: int) -> np.ndarray:
    return np.frombuffer(int(x).to_bytes(32, "little"), dtype=np.uint64).copy()


def init_voter_phase(voter_idx: int, sk: Optional[int] = None) -> Tuple[bytes, bytes]:
    """-> (voter_pk_blob, voter_sk_blob): 32-octet bit arrays (serialize_initial_phase_voter_data)"""
    del voter_idx  # only logged by the reference
    if sk is None:
        sk = secrets.randbits(voting.DIGEST_BITS)
    if not 0 <= sk < (1 << voting.DIGEST_BITS):
        raise ValueError("a secret key has %d bits" % voting.DIGEST_BITS)
    pk = voting.pedersen_hash_native(voting.digest_bits(sk))
    return codec.bitarray_bytes(voting.digest_bits(pk)), codec.bitarray_bytes(voting.digest_bits(sk))


def voters_tree(tree_depth: int, public_keys: Sequence[int]) -> List[int]:
    """make_merkle_tree over 2^tree_depth leaves, the voters' public keys first and all-zero keys after them
    (deserialize_voters_public_keys, common.hpp:638-655): every node, leaves first, root last"""
    n = 1 << tree_depth
    if len(public_keys) > n:
        raise ValueError("more public keys than participants")
    level, default, width, nodes = list(public_keys), 0, n, []
    while True:
        nodes += level + [default] * (width - len(level))
        if width == 1:
            return nodes
        nxt = []
        for i in range(0, len(level), 2):
            right = level[i + 1] if i + 1 < len(level) else default
            nxt.append(voting.pedersen_hash_native(voting.digest_bits(level[i]) + voting.digest_bits(right)))
        default = voting.pedersen_hash_native(voting.digest_bits(default) * 2)
        level, width = nxt, width // 2


def init_admin_phase_generate_data(tree_depth: int, eid_bits: int, public_keys_blobs: Sequence[bytes],
                                   eid: Optional[int] = None) -> Tuple[bytes, bytes, bytes]:
    """-> (eid_blob, rt_blob, merkle_tree_blob)   (serialize_initial_phase_admin_data, common.hpp:368-386)"""
    if eid is None:
        eid = secrets.randbits(eid_bits)
    if not 0 <= eid < (1 << eid_bits):
        raise ValueError("eid does not fit eid_bits")
    pks = []
    for blob in public_keys_blobs:
        bits = codec.bitarray_from_bytes(blob, voting.DIGEST_BITS)
        pks.append(sum(int(b) << i for i, b in enumerate(bits)))
    nodes = voters_tree(tree_depth, pks)
    tree = np.stack([_limbs(x) for x in nodes])
    eid_field = codec.bits_to_field_elements([(eid >> i) & 1 for i in range(eid_bits)])
    rt_field = codec.bits_to_field_elements(voting.digest_bits(nodes[-1]))
    return codec.primary_input_bytes(eid_field), codec.primary_input_bytes(rt_field), codec.merkle_tree_bytes(tree, tree_depth)


def init_admin_phase_generate_keys(ctx, tree_depth: int, eid_bits: int, toxic: Optional[Sequence[int]] = None,
                                   rnd: Optional[Sequence[int]] = None, msg_size: int = 25):
    """-> (r1cs_proving_key_blob, r1cs_verification_key_blob, public_key_blob, secret_key_blob, verification_key_blob)
    (serialize_initial_phase_admin_keys, common.hpp:330-366).  toxic = (t, alpha, beta, gamma, delta) and rnd (3 msg_size + 2
    scalars) are the generators' secrets; left out, they are drawn from the operating system."""
    vsp = _vsp()
    q = voting.Q
    circuit = voting.VotingCircuit(depth=tree_depth, eid_bits=eid_bits, msg_size=msg_size)
    toxic = [1 + secrets.randbelow(q - 1) for _ in range(5)] if toxic is None else list(toxic)
    rnd = [1 + secrets.randbelow(q - 1) for _ in range(3 * msg_size + 2)] if rnd is None else list(rnd)
    pk_arr, vk = vsp.generate(ctx, circuit.r1cs_arrays(), toxic)
    pub_arr, rho, svk = vsp.saver_generate_keypair(ctx, vk, msg_size, rnd)
    gt, _ = vsp.pairing_products(ctx, vk["alpha_g1"], vk["beta_g2"])  # alpha_g1_beta_g2 as the key stores it
    return (codec.write_fast_proving_key(pk_arr),
            codec.groth16_vk_bytes(np.asarray(gt).reshape(-1)[:72], vk["gamma_g2"], vk["delta_g2"], vk["delta_g1"], vk["gamma_g1"],
                                   vk["gamma_ABC_g1"]),
            codec.saver_pubkey_bytes(pub_arr), codec.saver_sk_bytes(_limbs(rho)),
            codec.saver_vk_bytes(svk["rho_g2"], svk["rho_sv_g2"], svk["rho_rhov_g2"]))
