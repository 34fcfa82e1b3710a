"""Shared test inputs: seeded synthetic points, R1CS instances and proving keys (all CPU, via the oracle)."""
from __future__ import annotations

import json
import os
import random

import numpy as np

from oracle import cref, fixtures as fx, pyref as o

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MASK64 = (1 << 64) - 1


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def generated_base_scalars(n: int, seed: int, first: int = 0) -> np.ndarray:
    """k_i of vsp_g1_bases_generate: splitmix64(seed + i) | 1"""
    with np.errstate(over="ignore"):
        idx = np.arange(first, first + n, dtype=np.uint64) + np.uint64(seed)
    return splitmix64(idx) | np.uint64(1)


G1_GEN_LIMBS = cref.g1_to_limbs([o.G1_GEN])
G2_GEN_LIMBS = cref.g2_to_limbs([o.G2_GEN])


def random_g1(n: int, seed: int) -> np.ndarray:
    k = splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(seed * 1000003 + 17)) | np.uint64(1)
    return cref.g1_fixed_base_mul64(G1_GEN_LIMBS, k)


def random_g2(n: int, seed: int) -> np.ndarray:
    k = splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(seed * 1000003 + 29)) | np.uint64(1)
    return cref.g2_fixed_base_mul64(G2_GEN_LIMBS, k)


def fr_ints(a: np.ndarray):
    return cref.limbs_to_ints(a)


def expected_generated_msm(n: int, seed: int, scalars: np.ndarray) -> np.ndarray:
    """sum_i s_i * (k_i G) = (sum_i s_i k_i mod r) G — a size-independent checksum for generated bases"""
    k = generated_base_scalars(n, seed)
    acc = 0
    s = fr_ints(scalars)
    for ki, si in zip(k.tolist(), s):
        acc += ki * si
    return cref.g1_mul(acc % o.R, G1_GEN_LIMBS)


def synthetic_r1cs(num_constraints: int, num_inputs: int, seed: int, binary: bool = False):
    rng = random.Random(seed)
    gen = o.r1cs_example_with_binary_input if binary else o.r1cs_example_with_field_input
    return gen(num_constraints, num_inputs, rng)


def synthetic_pk_arrays(cs: o.R1CS, seed: int) -> dict:
    """A proving key of the right *shape* over random bases (k_i G): enough for oracle parity of the prover, not a
    verifying key.  B_query is sparse where no constraint's B touches the variable."""
    _, m = cref.domain_plan(cs.num_constraints + cs.num_inputs + 1)
    nv = cs.num_variables
    used_b = sorted({idx for (_a, b, _c) in cs.constraints for idx, _ in b})
    g1 = random_g1(5 + (nv + 1) + len(used_b) + (m - 1) + (nv - cs.num_inputs), seed)
    g2 = random_g2(2 + len(used_b), seed)
    pos = [0]

    def take(n):
        r = g1[pos[0]: pos[0] + n]
        pos[0] += n
        return np.ascontiguousarray(r)

    arr = {"cs": fx.r1cs_arrays(cs)}
    arr["alpha_g1"], arr["beta_g1"], arr["delta_g1"] = take(1), take(1), take(1)
    arr["beta_g2"], arr["delta_g2"] = np.ascontiguousarray(g2[0:1]), np.ascontiguousarray(g2[1:2])
    arr["A_query"] = take(nv + 1)
    arr["B_count"] = len(used_b)
    arr["B_idx"] = np.array(used_b if used_b else [0], dtype=np.uint32)
    arr["B_g1"] = take(len(used_b)) if used_b else np.zeros((1, 12), np.uint64)
    arr["B_g2"] = np.ascontiguousarray(g2[2:]) if used_b else np.zeros((1, 24), np.uint64)
    arr["H_count"] = m - 1
    arr["H_query"] = take(m - 1)
    na = nv - cs.num_inputs
    arr["L_query"] = take(na) if na else np.zeros((1, 12), np.uint64)
    return arr


def synthetic_saver_pub(msg_size: int, seed: int):
    g1 = random_g1(3 + 2 * msg_size + msg_size, seed + 7)
    g2 = random_g2(msg_size + 1, seed + 7)
    arr = {
        "msg_size": msg_size,
        "delta_g1": np.ascontiguousarray(g1[0:1]),
        "delta_sum_s_g1": np.ascontiguousarray(g1[1:2]),
        "gamma_inverse_sum_s_g1": np.ascontiguousarray(g1[2:3]),
        "delta_s_g1": np.ascontiguousarray(g1[3:3 + msg_size]),
        "t_g1": np.ascontiguousarray(g1[3 + msg_size:3 + 2 * msg_size]),
        "t_g2": g2,
    }
    gamma_abc_rest = np.ascontiguousarray(g1[3 + 2 * msg_size:])
    return arr, gamma_abc_rest


def hexlist(xs):
    return [int(x, 16) for x in xs]


def run_fft_kat(case, fft, ifft):
    """Replay one tests/golden/fr_fft_kat.json case with the given transforms.
    fft(m, ints) / ifft(m, ints) -> list of m ints (basic domain of size m; short input zero-padded).
    Returns (got, want) lists of ints."""
    op = case["op"]
    if op == "fft":  # polynomial_dfs.cpp:50-71 from_coefficients: coefficients -> evaluations on size n
        return fft(case["n"], hexlist(case["in"])), hexlist(case["out"])
    if op == "ifft":  # polynomial_dfs.cpp:77-111 coefficients(): evaluations -> coefficients, trailing zeros trimmed
        got = ifft(case["n"], hexlist(case["in"]))
        want = hexlist(case["out"])
        assert all(v == 0 for v in got[len(want):])
        return got[:len(want)], want
    if op == "dfs_mul":  # polynomial_dfs.hpp:445-463 operator*, resize = iFFT(k) -> zero-pad -> FFT(s) (:346-358)
        a, b, want = hexlist(case["a"]), hexlist(case["b"]), hexlist(case["c"])
        s = 1
        while s < max(len(a), len(b), case["a_deg"] + case["b_deg"] + 1):
            s *= 2

        def resize(v):
            if len(v) == s:
                return v
            if len(v) == 1:
                return v * s
            return fft(s, ifft(len(v), v) + [0] * (s - len(v)))

        ra, rb = resize(a), resize(b)
        return [x * y % o.R for x, y in zip(ra, rb)], want
    raise ValueError(op)
