"""ctypes front-end of the C ORACLE (oracle/liboracle.so, built from vsp_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Arrays are numpy uint64, canonical little-endian limbs: Fr (n, 4); G1 affine (n, 12); G2 affine (n, 24).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)


class R1CSDesc(C.Structure):
    """orc_r1cs_t == vsp_r1cs_t (include/vsp_b200.h)"""
    _fields_ = [("num_inputs", C.c_uint64), ("num_aux", C.c_uint64), ("num_constraints", C.c_uint64),
                ("a_row", u64p), ("b_row", u64p), ("c_row", u64p),
                ("a_col", u32p), ("b_col", u32p), ("c_col", u32p),
                ("a_val", u64p), ("b_val", u64p), ("c_val", u64p)]


class ProvingKeyDesc(C.Structure):
    """orc_proving_key_t == vsp_proving_key_t"""
    _fields_ = [("cs", R1CSDesc),
                ("alpha_g1", u64p), ("beta_g1", u64p), ("beta_g2", u64p), ("delta_g1", u64p), ("delta_g2", u64p),
                ("A_query", u64p), ("B_count", C.c_uint64), ("B_idx", u32p), ("B_g2", u64p), ("B_g1", u64p),
                ("H_count", C.c_uint64), ("H_query", u64p), ("L_query", u64p)]


class SaverPubkeyDesc(C.Structure):
    """orc_saver_pubkey_t == vsp_saver_pubkey_t"""
    _fields_ = [("msg_size", C.c_uint64), ("delta_g1", u64p), ("delta_s_g1", u64p), ("t_g1", u64p), ("t_g2", u64p),
                ("delta_sum_s_g1", u64p), ("gamma_inverse_sum_s_g1", u64p)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = [os.path.join(_HERE, f) for f in ("vsp_oracle.c", "curve_tmpl.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_init()
        _LIB.orc_g1_bases_load.restype = C.c_void_p
        _LIB.orc_bases_free.argtypes = [C.c_void_p]
        _LIB.orc_msm_g1_loaded.argtypes = [C.c_uint64, C.c_void_p, u64p, C.c_uint64, C.c_int, u64p]
    return _LIB


def P(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def P32(a: np.ndarray):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u32p)


# ---- int <-> limb helpers ---------------------------------------------------------------------
def ints_to_limbs(xs, limbs: int) -> np.ndarray:
    nb = 8 * limbs
    buf = b"".join(int(x).to_bytes(nb, "little") for x in xs)
    return np.frombuffer(buf, dtype=np.uint64).reshape(len(xs), limbs).copy()


def limbs_to_ints(a: np.ndarray):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    flat = a.reshape(-1, a.shape[-1])
    nb = 8 * a.shape[-1]
    raw = flat.tobytes()
    return [int.from_bytes(raw[i * nb:(i + 1) * nb], "little") for i in range(flat.shape[0])]


def g1_to_limbs(aff_points) -> np.ndarray:
    """list of affine (x, y) int tuples or None -> (n, 12)"""
    out = []
    for p in aff_points:
        out += [0, 0] if p is None else [p[0], p[1]]
    return ints_to_limbs(out, 6).reshape(len(aff_points), 12)


def g1_from_limbs(a: np.ndarray):
    v = limbs_to_ints(np.ascontiguousarray(a).reshape(-1, 6))
    pts = []
    for i in range(0, len(v), 2):
        pts.append(None if v[i] == 0 and v[i + 1] == 0 else (v[i], v[i + 1]))
    return pts


def g2_to_limbs(aff_points) -> np.ndarray:
    out = []
    for p in aff_points:
        out += [0, 0, 0, 0] if p is None else [p[0][0], p[0][1], p[1][0], p[1][1]]
    return ints_to_limbs(out, 6).reshape(len(aff_points), 24)


def g2_from_limbs(a: np.ndarray):
    v = limbs_to_ints(np.ascontiguousarray(a).reshape(-1, 6))
    pts = []
    for i in range(0, len(v), 4):
        pts.append(None if not any(v[i:i + 4]) else ((v[i], v[i + 1]), (v[i + 2], v[i + 3])))
    return pts


# ---- oracle calls -----------------------------------------------------------------------------
def domain_plan(m: int):
    kind, size = C.c_int(0), C.c_uint64(0)
    if lib().orc_domain_plan(C.c_uint64(m), C.byref(kind), C.byref(size)):
        raise ValueError("no radix-2 family domain for %d" % m)
    return kind.value, size.value


def domain_fft(m: int, a: np.ndarray, inverse: bool = False) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    out = np.zeros((m, 4), dtype=np.uint64)
    rc = lib().orc_domain_fft(C.c_uint64(m), int(inverse), P(a), C.c_uint64(a.shape[0]), P(out))
    if rc:
        raise ValueError("orc_domain_fft rc=%d" % rc)
    return out


def basic_radix2_fft(a: np.ndarray, omega: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    w = ints_to_limbs([omega], 4)
    if lib().orc_fr_basic_radix2_fft(C.c_uint64(a.shape[0]), P(w), P(a)):
        raise ValueError("expected n == (1u << logn)")
    return a


def multiply_by_coset(a: np.ndarray, g: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_multiply_by_coset(C.c_uint64(a.shape[0]), P(ints_to_limbs([g], 4)), P(a))
    return a


def divide_by_z_on_coset(m: int, a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64).copy()
    if lib().orc_divide_by_z_on_coset(C.c_uint64(m), P(a)):
        raise ValueError("bad domain")
    return a


def msm_g1(bases: np.ndarray, scalars: np.ndarray, chunks: int = 1, mixed: bool = False) -> np.ndarray:
    bases = np.ascontiguousarray(bases, dtype=np.uint64)
    scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
    out = np.zeros(12, dtype=np.uint64)
    lib().orc_msm_g1(C.c_uint64(scalars.shape[0]), P(bases), P(scalars), C.c_uint64(chunks), int(mixed), P(out))
    return out


def msm_g2(bases: np.ndarray, scalars: np.ndarray, chunks: int = 1, mixed: bool = False) -> np.ndarray:
    bases = np.ascontiguousarray(bases, dtype=np.uint64)
    scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
    out = np.zeros(24, dtype=np.uint64)
    lib().orc_msm_g2(C.c_uint64(scalars.shape[0]), P(bases), P(scalars), C.c_uint64(chunks), int(mixed), P(out))
    return out


def g1_fixed_base_mul64(base: np.ndarray, k: np.ndarray) -> np.ndarray:
    k = np.ascontiguousarray(k, dtype=np.uint64)
    out = np.zeros((k.shape[0], 12), dtype=np.uint64)
    lib().orc_g1_fixed_base_mul64(P(np.ascontiguousarray(base, dtype=np.uint64)), C.c_uint64(k.shape[0]), P(k), P(out))
    return out


def g2_fixed_base_mul64(base: np.ndarray, k: np.ndarray) -> np.ndarray:
    k = np.ascontiguousarray(k, dtype=np.uint64)
    out = np.zeros((k.shape[0], 24), dtype=np.uint64)
    lib().orc_g2_fixed_base_mul64(P(np.ascontiguousarray(base, dtype=np.uint64)), C.c_uint64(k.shape[0]), P(k), P(out))
    return out


def g1_mul(k: int, pt: np.ndarray) -> np.ndarray:
    out = np.zeros(12, dtype=np.uint64)
    lib().orc_g1_mul(P(ints_to_limbs([k], 4)), P(np.ascontiguousarray(pt, dtype=np.uint64)), P(out))
    return out


def g2_mul(k: int, pt: np.ndarray) -> np.ndarray:
    out = np.zeros(24, dtype=np.uint64)
    lib().orc_g2_mul(P(ints_to_limbs([k], 4)), P(np.ascontiguousarray(pt, dtype=np.uint64)), P(out))
    return out


def g1_add(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.zeros(12, dtype=np.uint64)
    lib().orc_g1_add(P(np.ascontiguousarray(a, dtype=np.uint64)), P(np.ascontiguousarray(b, dtype=np.uint64)), P(out))
    return out


def g2_add(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.zeros(24, dtype=np.uint64)
    lib().orc_g2_add(P(np.ascontiguousarray(a, dtype=np.uint64)), P(np.ascontiguousarray(b, dtype=np.uint64)), P(out))
    return out


def g1_jacobian_to_affine(xyz: np.ndarray) -> np.ndarray:
    out = np.zeros(12, dtype=np.uint64)
    lib().orc_g1_jacobian_to_affine(P(np.ascontiguousarray(xyz, dtype=np.uint64)), P(out))
    return out


def witness_map(cs_desc: R1CSDesc, primary: np.ndarray, aux: np.ndarray) -> np.ndarray:
    _, m = domain_plan(cs_desc.num_constraints + cs_desc.num_inputs + 1)
    out = np.zeros((m + 1, 4), dtype=np.uint64)
    mo = C.c_uint64(0)
    rc = lib().orc_witness_map(C.byref(cs_desc), P(primary), P(aux), P(out), C.byref(mo))
    if rc:
        raise ValueError("orc_witness_map rc=%d" % rc)
    return out


def groth16_prove(pk_desc: ProvingKeyDesc, primary, aux, r: int, s: int, r_saver=None, gamma_inverse_sum_s_g1=None,
                  chunks: int = 1) -> np.ndarray:
    out = np.zeros(48, dtype=np.uint64)
    rs = ints_to_limbs([r, s], 4)
    if r_saver is None:
        a5, a6 = None, None
    else:
        a5 = P(ints_to_limbs([r_saver], 4))
        a6 = P(np.ascontiguousarray(gamma_inverse_sum_s_g1, dtype=np.uint64))
    rc = lib().orc_groth16_prove(C.byref(pk_desc), P(primary), P(aux), P(rs[0:1].copy()), P(rs[1:2].copy()), a5, a6,
                                 C.c_uint64(chunks), P(out))
    if rc:
        raise ValueError("orc_groth16_prove rc=%d" % rc)
    return out


def saver_ciphertext(pub: SaverPubkeyDesc, gamma_abc_rest: np.ndarray, msg: np.ndarray, r: int) -> np.ndarray:
    out = np.zeros((pub.msg_size + 2, 12), dtype=np.uint64)
    lib().orc_saver_ciphertext(C.byref(pub), P(gamma_abc_rest), P(msg), P(ints_to_limbs([r], 4)), P(out))
    return out


def saver_rerandomize(pub: SaverPubkeyDesc, delta_g2: np.ndarray, rnd, ct: np.ndarray, proof: np.ndarray):
    ct_out = np.zeros_like(ct)
    proof_out = np.zeros(48, dtype=np.uint64)
    lib().orc_saver_rerandomize(C.byref(pub), P(delta_g2), P(ints_to_limbs(list(rnd), 4)), P(ct), P(proof), P(ct_out),
                                P(proof_out))
    return ct_out, proof_out
