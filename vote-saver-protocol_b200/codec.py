"""The reference's wire formats either side of the ballot-proof path, through the C ABI's vsp_codec_* entry points
(host only — no GPU needed): compressed proof / cipher text / primary-input blobs as bin/cli's vote phase writes them
(bin/cli/include/nil/vote_saver/common.hpp:462-485, 1154-1160) and the "fast" proving key it reads (common.hpp:1004-1006;
libs/marshalling/zk/include/nil/crypto3/marshalling/zk/types/r1cs_gg_ppzksnark/fast_proving_key.hpp:66-103)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _u64(a, cols=None) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return a.reshape(-1, cols) if cols else a


def _check(rc: int, what: str):
    if rc != 0:
        raise ValueError("%s: malformed input (vsp error %d)" % (what, rc))


def _out(n: int):
    return (C.c_uint8 * n)()


def g1_compress(point) -> bytes:
    p, out = _u64(point).reshape(-1), _out(48)
    _check(_lib.load().vsp_codec_g1_compress(p.ctypes.data, out), "g1_compress")
    return bytes(out)


def g2_compress(point) -> bytes:
    p, out = _u64(point).reshape(-1), _out(96)
    _check(_lib.load().vsp_codec_g2_compress(p.ctypes.data, out), "g2_compress")
    return bytes(out)


def proof_bytes(proof) -> bytes:
    """g_A | g_B | g_C (48 limbs) -> the 192-byte proof.bin body"""
    p, out = _u64(proof).reshape(-1), _out(192)
    if p.shape[0] != 48:
        raise ValueError("a proof is 48 limbs")
    _check(_lib.load().vsp_codec_proof_write(p.ctypes.data, out), "proof_write")
    return bytes(out)


def g1_decompress(octets: bytes) -> np.ndarray:
    """48 compressed octets -> 12 canonical affine limbs (all-zero = infinity); ValueError where the reference's decoder fails"""
    if len(octets) != 48:
        raise ValueError("a compressed G1 point is 48 bytes")
    out = np.zeros(12, dtype=np.uint64)
    _check(_lib.load().vsp_codec_g1_decompress(bytes(octets), out.ctypes.data), "g1_decompress")
    return out


def g2_decompress(octets: bytes) -> np.ndarray:
    if len(octets) != 96:
        raise ValueError("a compressed G2 point is 96 bytes")
    out = np.zeros(24, dtype=np.uint64)
    _check(_lib.load().vsp_codec_g2_decompress(bytes(octets), out.ctypes.data), "g2_decompress")
    return out


def proof_from_bytes(octets: bytes) -> np.ndarray:
    """the 192-byte proof.bin body -> g_A | g_B | g_C (48 limbs)"""
    if len(octets) != 192:
        raise ValueError("a proof is 192 bytes")
    out = np.zeros(48, dtype=np.uint64)
    _check(_lib.load().vsp_codec_proof_read(bytes(octets), out.ctypes.data), "proof_read")
    return out


def ciphertext_bytes(ct) -> bytes:
    c = _u64(ct, 12)
    out = _out(8 + 48 * c.shape[0])
    _check(_lib.load().vsp_codec_ciphertext_write(c.ctypes.data, c.shape[0], out), "ciphertext_write")
    return bytes(out)


def primary_input_bytes(values) -> bytes:
    v = _u64(values, 4)
    out = _out(8 + 32 * v.shape[0])
    _check(_lib.load().vsp_codec_primary_input_write(v.ctypes.data, v.shape[0], out), "primary_input_write")
    return bytes(out)


def primary_input_from_bytes(blob: bytes) -> np.ndarray:
    n = C.c_uint64(0)
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    lib = _lib.load()
    _check(lib.vsp_codec_primary_input_read(buf, len(blob), C.byref(n), None, 0), "primary_input_read")
    out = np.zeros((n.value, 4), dtype=np.uint64)
    _check(lib.vsp_codec_primary_input_read(buf, len(blob), C.byref(n), out.ctypes.data, n.value), "primary_input_read")
    return out


def ciphertext_from_bytes(blob: bytes) -> np.ndarray:
    """ct.bin -> n x 12 limbs (deserialize_ct, common.hpp:776)"""
    n = C.c_uint64(0)
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    lib = _lib.load()
    _check(lib.vsp_codec_ciphertext_read(buf, len(blob), C.byref(n), None, 0), "ciphertext_read")
    out = np.zeros((n.value, 12), dtype=np.uint64)
    _check(lib.vsp_codec_ciphertext_read(buf, len(blob), C.byref(n), out.ctypes.data, n.value), "ciphertext_read")
    return out


def _two_pass_write(call) -> bytes:
    n = C.c_size_t(0)
    _check(call(None, 0, C.byref(n)), "size query")
    out = _out(n.value)
    _check(call(out, n.value, C.byref(n)), "write")
    return bytes(out)


def saver_pubkey_bytes(pub: dict) -> bytes:
    """the arrays dict SaverPublicKey takes (field names of vsp_saver_pubkey_t) -> the pk_eid blob (common.hpp:357)"""
    arr = {k: (_u64(v) if isinstance(v, np.ndarray) else v) for k, v in pub.items()}
    desc = fill_desc(_lib.SaverPubkeyDesc(), arr)
    lib = _lib.load()
    return _two_pass_write(lambda out, cap, n: lib.vsp_codec_saver_pubkey_write(C.byref(desc), out, cap, n))


def saver_pubkey_from_bytes(blob: bytes) -> dict:
    """the pk_eid blob (deserialize_pk_eid, common.hpp:691) -> the arrays dict SaverPublicKey takes"""
    lib = _lib.load()
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    n = C.c_uint64(0)
    _check(lib.vsp_codec_saver_pubkey_read(buf, len(blob), C.byref(n), None), "saver_pubkey_read")
    k = n.value
    arr = {"msg_size": k, "delta_g1": np.zeros((1, 12), dtype=np.uint64),
           "delta_s_g1": np.zeros((max(k, 1), 12), dtype=np.uint64), "t_g1": np.zeros((max(k, 1), 12), dtype=np.uint64),
           "t_g2": np.zeros((k + 1, 24), dtype=np.uint64), "delta_sum_s_g1": np.zeros((1, 12), dtype=np.uint64),
           "gamma_inverse_sum_s_g1": np.zeros((1, 12), dtype=np.uint64)}
    desc = fill_desc(_lib.SaverPubkeyDesc(), arr)
    _check(lib.vsp_codec_saver_pubkey_read(buf, len(blob), None, C.byref(desc)), "saver_pubkey_read")
    arr["delta_s_g1"], arr["t_g1"] = arr["delta_s_g1"][:k], arr["t_g1"][:k]
    return arr


def saver_vk_bytes(rho_g2, rho_sv_g2, rho_rhov_g2) -> bytes:
    a, b, c = _u64(rho_g2).reshape(-1), _u64(rho_sv_g2, 24), _u64(rho_rhov_g2, 24)
    if a.shape[0] != 24 or b.shape != c.shape:
        raise ValueError("rho_g2 is one G2 point; rho_sv_g2 and rho_rhov_g2 hold msg_size G2 points each")
    lib = _lib.load()
    return _two_pass_write(lambda out, cap, n: lib.vsp_codec_saver_vk_write(b.shape[0], a.ctypes.data, b.ctypes.data,
                                                                            c.ctypes.data, out, cap, n))


def saver_vk_from_bytes(blob: bytes):
    """-> (rho_g2 [24], rho_sv_g2 [n, 24], rho_rhov_g2 [n, 24])   (deserialize_vk_eid, common.hpp:706)"""
    lib = _lib.load()
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    n = C.c_uint64(0)
    _check(lib.vsp_codec_saver_vk_read(buf, len(blob), C.byref(n), None, None, None, 0), "saver_vk_read")
    k = n.value
    a, b, c = (np.zeros(24, dtype=np.uint64), np.zeros((max(k, 1), 24), dtype=np.uint64),
               np.zeros((max(k, 1), 24), dtype=np.uint64))
    _check(lib.vsp_codec_saver_vk_read(buf, len(blob), None, a.ctypes.data, b.ctypes.data, c.ctypes.data, k), "saver_vk_read")
    return a, b[:k], c[:k]


def saver_sk_bytes(rho) -> bytes:
    r, out = _u64(rho).reshape(-1), _out(32)
    _check(_lib.load().vsp_codec_saver_sk_write(r.ctypes.data, out), "saver_sk_write")
    return bytes(out)


def saver_sk_from_bytes(blob: bytes) -> np.ndarray:
    out = np.zeros(4, dtype=np.uint64)
    _check(_lib.load().vsp_codec_saver_sk_read(bytes(blob), len(blob), out.ctypes.data), "saver_sk_read")
    return out


def groth16_vk_bytes(alpha_g1_beta_g2, gamma_g2, delta_g2, delta_g1, gamma_g1, gamma_ABC_g1) -> bytes:
    """the extended verification key blob (common.hpp:351): alpha_g1_beta_g2 72 limbs (vsp_pairing_products' order),
    gamma_ABC_g1 = first | rest as (1 + n) x 12 limbs"""
    gt, g2a, g2b = _u64(alpha_g1_beta_g2).reshape(-1), _u64(gamma_g2).reshape(-1), _u64(delta_g2).reshape(-1)
    g1a, g1b, abc = _u64(delta_g1).reshape(-1), _u64(gamma_g1).reshape(-1), _u64(gamma_ABC_g1, 12)
    if gt.shape[0] != 72 or g2a.shape[0] != 24 or g2b.shape[0] != 24 or g1a.shape[0] != 12 or g1b.shape[0] != 12:
        raise ValueError("groth16 verification key: wrong limb counts")
    lib = _lib.load()
    return _two_pass_write(lambda out, cap, n: lib.vsp_codec_groth16_vk_write(
        gt.ctypes.data, g2a.ctypes.data, g2b.ctypes.data, g1a.ctypes.data, g1b.ctypes.data, abc.shape[0], abc.ctypes.data,
        out, cap, n))


def groth16_vk_from_bytes(blob: bytes) -> dict:
    """deserialize_vk_crs (common.hpp:734) -> {alpha_g1_beta_g2 [72], gamma_g2, delta_g2 [24], delta_g1, gamma_g1 [12],
    gamma_ABC_g1 [(1 + n), 12]}: the arguments of vsp_groth16_verify_gt_batch, and gamma_ABC_g1[1:] is the gamma_abc_rest
    the SAVER entry points take"""
    lib = _lib.load()
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    n = C.c_uint64(0)
    _check(lib.vsp_codec_groth16_vk_read(buf, len(blob), C.byref(n), None, None, None, None, None, None, 0), "groth16_vk_read")
    vk = {"alpha_g1_beta_g2": np.zeros(72, dtype=np.uint64), "gamma_g2": np.zeros(24, dtype=np.uint64),
          "delta_g2": np.zeros(24, dtype=np.uint64), "delta_g1": np.zeros(12, dtype=np.uint64),
          "gamma_g1": np.zeros(12, dtype=np.uint64), "gamma_ABC_g1": np.zeros((n.value, 12), dtype=np.uint64)}
    _check(lib.vsp_codec_groth16_vk_read(buf, len(blob), None, *(vk[k].ctypes.data for k in (
        "alpha_g1_beta_g2", "gamma_g2", "delta_g2", "delta_g1", "gamma_g1", "gamma_ABC_g1")), n.value), "groth16_vk_read")
    return vk


def bitarray_bytes(bits) -> bytes:
    """serialize_bitarray (common.hpp:574-591): 0 / 1 values -> ceil(n / 8) octets, first bit in the top of the first octet"""
    b = np.ascontiguousarray(bits, dtype=np.uint8).reshape(-1)
    out = _out((b.shape[0] + 7) // 8)
    _check(_lib.load().vsp_codec_bitarray_write(b.ctypes.data, b.shape[0], out), "bitarray_write")
    return bytes(out)


def bitarray_from_bytes(blob: bytes, n_bits: int) -> np.ndarray:
    out = np.zeros(n_bits, dtype=np.uint8)
    _check(_lib.load().vsp_codec_bitarray_read(bytes(blob), len(blob), n_bits, out.ctypes.data), "bitarray_read")
    return out


def bits_to_field_elements(bits) -> np.ndarray:
    """get_multi_field_element_from_bits (common.hpp:557-570): -> ceil(n / 254) x 4 limbs"""
    b = np.ascontiguousarray(bits, dtype=np.uint8).reshape(-1)
    out = np.zeros(((b.shape[0] + 253) // 254, 4), dtype=np.uint64)
    _check(_lib.load().vsp_codec_bits_to_field_elements(b.ctypes.data, b.shape[0], out.ctypes.data), "bits_to_field_elements")
    return out


def merkle_tree_bytes(digests, depth: int) -> bytes:
    d = _u64(digests, 4)
    if d.shape[0] != (2 << depth) - 1:
        raise ValueError("a depth-%d tree holds %d digests" % (depth, (2 << depth) - 1))
    out = _out(32 * d.shape[0])
    _check(_lib.load().vsp_codec_merkle_tree_write(d.ctypes.data, depth, out), "merkle_tree_write")
    return bytes(out)


def merkle_tree_from_bytes(blob: bytes, depth: int) -> np.ndarray:
    """deserialize_merkle_tree (common.hpp:614-636) -> (2^(depth + 1) - 1) x 4 limbs, leaves first, root last"""
    if not 0 <= depth <= 40 or len(blob) != 32 * ((2 << depth) - 1):
        raise ValueError("a depth-%d tree blob is %d digests of 32 bytes" % (depth, (2 << depth) - 1 if 0 <= depth <= 40 else 0))
    out = np.zeros(((2 << depth) - 1, 4), dtype=np.uint64)
    _check(_lib.load().vsp_codec_merkle_tree_read(bytes(blob), len(blob), depth, out.ctypes.data), "merkle_tree_read")
    return out


def merkle_copath(digests, depth: int, index: int) -> np.ndarray:
    """merkle_proof(tree, index) for arity 2 -> depth x 4 limbs, leaf level first (a `paths` row of the witness generator)"""
    d = _u64(digests, 4)
    if d.shape[0] != (2 << depth) - 1:
        raise ValueError("a depth-%d tree holds %d digests" % (depth, (2 << depth) - 1))
    out = np.zeros((depth, 4), dtype=np.uint64)
    _check(_lib.load().vsp_codec_merkle_copath(d.ctypes.data, depth, index, out.ctypes.data), "merkle_copath")
    return out


def read_fast_proving_key(blob: bytes) -> dict:
    """r1cs_proving_key.bin -> the arrays dict ProvingKey(ctx, arrays) takes (field names of vsp_proving_key_t)"""
    lib = _lib.load()
    buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
    sz = _lib.FastPkSizes()
    _check(lib.vsp_codec_fast_pk_sizes(buf, len(blob), C.byref(sz)), "fast_pk_sizes")
    nc = sz.num_constraints
    cs = {"num_inputs": sz.num_inputs, "num_aux": sz.num_aux, "num_constraints": nc}
    for k, nnz in (("a", sz.nnz_a), ("b", sz.nnz_b), ("c", sz.nnz_c)):
        cs[k + "_row"] = np.zeros(nc + 1, dtype=np.uint64)
        cs[k + "_col"] = np.zeros(max(nnz, 1), dtype=np.uint32)
        cs[k + "_val"] = np.zeros((max(nnz, 1), 4), dtype=np.uint64)
    arr = {"cs": cs}
    for name, width in (("alpha_g1", 12), ("beta_g1", 12), ("beta_g2", 24), ("delta_g1", 12), ("delta_g2", 24)):
        arr[name] = np.zeros((1, width), dtype=np.uint64)
    arr["A_query"] = np.zeros((max(sz.A_count, 1), 12), dtype=np.uint64)
    arr["B_count"] = sz.B_count
    arr["B_idx"] = np.zeros(max(sz.B_count, 1), dtype=np.uint32)
    arr["B_g2"] = np.zeros((max(sz.B_count, 1), 24), dtype=np.uint64)
    arr["B_g1"] = np.zeros((max(sz.B_count, 1), 12), dtype=np.uint64)
    arr["H_count"] = sz.H_count
    arr["H_query"] = np.zeros((max(sz.H_count, 1), 12), dtype=np.uint64)
    arr["L_query"] = np.zeros((max(sz.L_count, 1), 12), dtype=np.uint64)
    if sz.A_count != sz.num_inputs + sz.num_aux + 1 or sz.L_count != sz.num_aux:
        raise ValueError("proving key: A_query / L_query sizes do not match the constraint system")
    desc = fill_desc(_lib.ProvingKeyDesc(), arr)
    _check(lib.vsp_codec_fast_pk_read(buf, len(blob), C.byref(desc)), "fast_pk_read")
    for k, nnz in (("a", sz.nnz_a), ("b", sz.nnz_b), ("c", sz.nnz_c)):
        cs[k + "_col"], cs[k + "_val"] = cs[k + "_col"][:max(nnz, 1)], cs[k + "_val"][:max(nnz, 1)]
    arr["_sizes"] = {f: getattr(sz, f) for f, _ in sz._fields_}
    return arr


def write_fast_proving_key(arrays: dict) -> bytes:
    lib = _lib.load()
    arrays = {k: v for k, v in arrays.items() if not k.startswith("_")}
    desc = fill_desc(_lib.ProvingKeyDesc(), arrays)
    n = C.c_size_t(0)
    _check(lib.vsp_codec_fast_pk_write(C.byref(desc), None, 0, C.byref(n)), "fast_pk_write")
    out = _out(n.value)
    _check(lib.vsp_codec_fast_pk_write(C.byref(desc), out, n.value, C.byref(n)), "fast_pk_write")
    return bytes(out)


def fill_desc(desc, arr: dict):
    """populate a ctypes descriptor from a dict of numpy arrays / ints with the same field names"""
    for name, ftype in desc._fields_:
        v = arr[name]
        if isinstance(v, dict):
            fill_desc(getattr(desc, name), v)
        elif isinstance(v, np.ndarray):
            setattr(desc, name, v.ctypes.data_as(ftype))
        else:
            setattr(desc, name, int(v))
    return desc
