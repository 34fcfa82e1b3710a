"""The tally phase from blobs to blobs: the host-side mirror of `process_encrypted_input_mode_tally_admin_phase` and
`process_encrypted_input_mode_tally_voter_phase` (/root/reference/bin/cli/include/nil/vote_saver/common.hpp:1176-1236,
1239-1296) over the C ABI — the consumers of what the vote phase (vote_phase.py) produces:

    administrator   cts blobs, sk_eid, vk_eid, vk_crs  ->  sum of the cipher texts slot by slot (:1207-1215)
                    -> decrypt<verifiable_encryption>(ct_agg, {sk_eid, vk_eid, keypair})   (:1219-1223)   vsp_saver_decrypt
                    -> dec_proof_blob (the G1 point rho * ct_agg[0], compressed) | voting_res_blob (count | Fr...)
                       (serialize_tally_phase_data, :511-522)
    voter           cts blobs, vk_eid, vk_crs, voting_res, dec_proof  ->  the same sum
                    -> verify_decryption(ct_agg, voting_result, {vk_eid, keypair, dec_proof})   (:1283-1285)
                       vsp_saver_verify_decryption

The reference also deserialises the proving key in both calls (`pk_crs_blob`, :1189-1190, :1251-1252) but uses only the
verification key's gamma_ABC_g1.rest from the key pair (elgamal_verifiable.hpp:401-441, 548-575); the argument is accepted
and ignored here.  dec_proof_blob is written by `nil::marshalling::pack<endianness>(dec.second, status)` (:515): the
marshalling library's type inference (in the un-vendored `marshalling` core) picks the type for a G1 value; it is taken here
to be the one curve_element every other G1 point on this path uses, i.e. 48 compressed bytes.  The slot-by-slot sums run on the GPU (vsp_g1_sum), one call per cipher-text slot.
"""
from __future__ import annotations

import importlib
from typing import Optional, Sequence, Tuple

import numpy as np

from . import codec


def _vsp():
    return importlib.import_module(__package__)


def read_ciphertexts(tree_depth: int, cts_blobs: Sequence[bytes]) -> np.ndarray:
    """deserialize_ct over every ballot (common.hpp:1196-1201, 1262-1267): (k, msg_size + 2, 12).  The reference asserts
    cts_blobs.size() <= 2^tree_depth and equal sizes ("Wrong size of the ct!")."""
    if len(cts_blobs) == 0:
        raise ValueError("no cipher texts to tally")
    if len(cts_blobs) > (1 << tree_depth):
        raise ValueError("more cipher texts than participants")
    cts = [codec.ciphertext_from_bytes(b) for b in cts_blobs]
    if any(c.shape != cts[0].shape for c in cts) or cts[0].shape[0] < 2:
        raise ValueError("Wrong size of the ct!")
    return np.stack(cts)


def aggregate(ctx, cts: np.ndarray) -> np.ndarray:
    """ct_agg[i] = sum over ballots of ct[i] (common.hpp:1207-1215): (msg_size + 2, 12)"""
    vsp = _vsp()
    if cts.shape[0] == 1:
        return np.ascontiguousarray(cts[0])
    return np.stack([vsp.points_sum(ctx, np.ascontiguousarray(cts[:, i])).reshape(12) for i in range(cts.shape[1])])


def serialize_tally(counts: Sequence[int], dec_proof) -> Tuple[bytes, bytes]:
    """serialize_tally_phase_data (common.hpp:511-522) -> (dec_proof_blob, voting_res_blob)"""
    vals = np.frombuffer(b"".join(int(c).to_bytes(32, "little") for c in counts), dtype=np.uint64).reshape(-1, 4) \
        if len(counts) else np.zeros((0, 4), dtype=np.uint64)
    return codec.g1_compress(dec_proof), codec.primary_input_bytes(vals)


def tally_admin_phase(ctx, tree_depth: int, cts_blobs: Sequence[bytes], sk_eid_blob: bytes, vk_eid_blob: bytes,
                      pk_crs_blob: Optional[bytes], vk_crs_blob: bytes, block_bits: int = 4) -> Tuple[bytes, bytes]:
    """process_encrypted_input_mode_tally_admin_phase -> (dec_proof_blob, voting_res_blob)"""
    vsp = _vsp()
    rho = int.from_bytes(codec.saver_sk_from_bytes(sk_eid_blob).tobytes(), "little")
    rho_g2, rho_sv_g2, rho_rhov_g2 = codec.saver_vk_from_bytes(vk_eid_blob)
    vk = codec.groth16_vk_from_bytes(vk_crs_blob)
    cts = read_ciphertexts(tree_depth, cts_blobs)
    n = cts.shape[1] - 2
    if rho_sv_g2.shape[0] != n or vk["gamma_ABC_g1"].shape[0] < n + 1:
        raise ValueError("Cipher text size should match the verification key (exclusive of 2 elements in CT).")
    svk = {"rho_g2": rho_g2, "rho_sv_g2": rho_sv_g2, "rho_rhov_g2": rho_rhov_g2}
    gabc = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + n])
    counts, c0 = vsp.decrypt(ctx, aggregate(ctx, cts), rho, svk, gabc, block_bits=block_bits)
    return serialize_tally(counts, c0)


def tally_voter_phase(ctx, tree_depth: int, cts_blobs: Sequence[bytes], vk_eid_blob: bytes, pk_crs_blob: Optional[bytes],
                      vk_crs_blob: bytes, voting_res_blob: bytes, dec_proof_blob: bytes) -> bool:
    """process_encrypted_input_mode_tally_voter_phase -> the verification's answer (the reference asserts it)"""
    vsp = _vsp()
    rho_g2, rho_sv_g2, rho_rhov_g2 = codec.saver_vk_from_bytes(vk_eid_blob)
    vk = codec.groth16_vk_from_bytes(vk_crs_blob)
    result = codec.primary_input_from_bytes(voting_res_blob)
    dec_proof = codec.g1_decompress(dec_proof_blob)
    cts = read_ciphertexts(tree_depth, cts_blobs)
    n = cts.shape[1] - 2
    if result.shape[0] != n or rho_sv_g2.shape[0] != n or vk["gamma_ABC_g1"].shape[0] < n + 1:
        return False
    counts = [int.from_bytes(np.ascontiguousarray(r).tobytes(), "little") for r in result]
    svk = {"rho_g2": rho_g2, "rho_sv_g2": rho_sv_g2, "rho_rhov_g2": rho_rhov_g2}
    gabc = np.ascontiguousarray(vk["gamma_ABC_g1"][1:1 + n])
    return bool(vsp.verify_decryption(ctx, aggregate(ctx, cts), counts, svk, gabc, dec_proof))
