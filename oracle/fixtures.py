"""Bridges between the Python oracle's objects and the flat numpy layouts the C ABIs take.  TEST INFRASTRUCTURE ONLY.

`*_arrays` return plain dicts of numpy arrays (kept alive by the caller); `fill_*` populate a ctypes Structure of
either ABI (oracle/cref.py or the product binding) from such a dict — both ABIs share field names.
"""
from __future__ import annotations

import ctypes as C
import random

import numpy as np

from . import pyref as o
from .cref import ints_to_limbs, g1_to_limbs, g2_to_limbs


def r1cs_arrays(cs: o.R1CS) -> dict:
    arr = {"num_inputs": cs.num_inputs, "num_aux": cs.auxiliary_input_size, "num_constraints": cs.num_constraints}
    for k, name in enumerate("abc"):
        row = [0]
        col = []
        val = []
        for con in cs.constraints:
            for idx, coeff in con[k]:
                col.append(idx)
                val.append(coeff % o.R)
            row.append(len(col))
        arr[name + "_row"] = np.array(row, dtype=np.uint64)
        arr[name + "_col"] = np.array(col if col else [0], dtype=np.uint32)
        arr[name + "_val"] = ints_to_limbs(val, 4) if val else np.zeros((1, 4), dtype=np.uint64)
    return arr


def _ptr(a: np.ndarray, field_type):
    return a.ctypes.data_as(field_type)


def fill_r1cs(desc, arr: dict):
    for name, ftype in desc._fields_:
        v = arr[name]
        if isinstance(v, np.ndarray):
            setattr(desc, name, _ptr(v, ftype))
        else:
            setattr(desc, name, v)
    return desc


def pk_arrays(pk: o.ProvingKey) -> dict:
    aff1 = lambda p: o.G1.to_affine(p)
    aff2 = lambda p: o.G2.to_affine(p)
    b_idx = [i for i, e in enumerate(pk.B_query) if e is not None]
    arr = {
        "cs": r1cs_arrays(pk.cs),
        "alpha_g1": g1_to_limbs([aff1(pk.alpha_g1)]), "beta_g1": g1_to_limbs([aff1(pk.beta_g1)]),
        "beta_g2": g2_to_limbs([aff2(pk.beta_g2)]), "delta_g1": g1_to_limbs([aff1(pk.delta_g1)]),
        "delta_g2": g2_to_limbs([aff2(pk.delta_g2)]),
        "A_query": g1_to_limbs([aff1(p) for p in pk.A_query]),
        "B_count": len(b_idx),
        "B_idx": np.array(b_idx if b_idx else [0], dtype=np.uint32),
        "B_g2": g2_to_limbs([aff2(pk.B_query[i][0]) for i in b_idx]) if b_idx else np.zeros((1, 24), np.uint64),
        "B_g1": g1_to_limbs([aff1(pk.B_query[i][1]) for i in b_idx]) if b_idx else np.zeros((1, 12), np.uint64),
        "H_count": len(pk.H_query),
        "H_query": g1_to_limbs([aff1(p) for p in pk.H_query]),
        "L_query": g1_to_limbs([aff1(p) for p in pk.L_query]) if pk.L_query else np.zeros((1, 12), np.uint64),
    }
    return arr


def fill_pk(desc, arr: dict):
    for name, ftype in desc._fields_:
        v = arr[name]
        if name == "cs":
            fill_r1cs(desc.cs, v)
        elif isinstance(v, np.ndarray):
            setattr(desc, name, _ptr(v, ftype))
        else:
            setattr(desc, name, v)
    return desc


def saver_pub_arrays(pub: o.SaverPublicKey) -> dict:
    aff1 = lambda p: o.G1.to_affine(p)
    return {
        "msg_size": len(pub.delta_s_g1),
        "delta_g1": g1_to_limbs([aff1(pub.delta_g1)]),
        "delta_s_g1": g1_to_limbs([aff1(p) for p in pub.delta_s_g1]),
        "t_g1": g1_to_limbs([aff1(p) for p in pub.t_g1]),
        "t_g2": g2_to_limbs([o.G2.to_affine(p) for p in pub.t_g2]),
        "delta_sum_s_g1": g1_to_limbs([aff1(pub.delta_sum_s_g1)]),
        "gamma_inverse_sum_s_g1": g1_to_limbs([aff1(pub.gamma_inverse_sum_s_g1)]),
    }


def fill_struct(desc, arr: dict):
    for name, ftype in desc._fields_:
        v = arr[name]
        if isinstance(v, np.ndarray):
            setattr(desc, name, _ptr(v, ftype))
        else:
            setattr(desc, name, v)
    return desc


def proof_to_limbs(proof) -> np.ndarray:
    gA, gB, gC = proof
    return np.concatenate([g1_to_limbs([o.G1.to_affine(gA)]).ravel(), g2_to_limbs([o.G2.to_affine(gB)]).ravel(),
                           g1_to_limbs([o.G1.to_affine(gC)]).ravel()])


# ---- synthetic, seeded workloads (no reference needed at run time) ------------------------------
def random_fr(n: int, seed: int) -> np.ndarray:
    """uniform in [0, r): 64-bit PRNG limbs, rejection on the top limb."""
    rng = np.random.default_rng(seed)
    out = rng.integers(0, 1 << 64, size=(n, 4), dtype=np.uint64)
    top = np.uint64(o.R >> 192)
    # r's top limb is 0x73eda753299d7d48: clear bit 63 then reject >= top (keeps uniformity up to 2^-190 bias)
    out[:, 3] &= np.uint64((1 << 63) - 1)
    bad = out[:, 3] >= top
    while bad.any():
        out[bad, 3] = rng.integers(0, 1 << 63, size=int(bad.sum()), dtype=np.uint64)
        bad = out[:, 3] >= top
    return out


def witness_like_fr(n: int, seed: int, frac_full: float = 0.1) -> np.ndarray:
    """90 % zero/one, 10 % uniform (SURVEY.md 8d config 2 'witness-like')."""
    rng = np.random.default_rng(seed)
    out = np.zeros((n, 4), dtype=np.uint64)
    out[:, 0] = rng.integers(0, 2, size=n, dtype=np.uint64)
    full = rng.random(n) < frac_full
    out[full] = random_fr(int(full.sum()), seed + 1)
    return out
