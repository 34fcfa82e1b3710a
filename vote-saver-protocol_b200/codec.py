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
