"""Synthetic, seeded ballot workloads for bench.py and the profiling probes (numpy only — no oracle imports).

The real `encrypted_input_voting` circuit lives in the un-vendored crypto3-blueprint submodule, so the
ballot circuit is a stand-in with the same *interface shape*: 30 primary inputs of which the first 25 are the 0/1
message slots (bin/cli/include/nil/vote_saver/common.hpp:163, 863-876), an all-bits witness, and the constraint shape
of the reference's own synthetic generator `generate_r1cs_example_with_binary_input`
(libs/marshalling/zk/test/detail/r1cs_examples.hpp:154-210): constraint i is  (2 x_u) * x_v = x_u + x_v - x_new,
i.e. x_new = x_u XOR x_v.
"""
from __future__ import annotations

import numpy as np

FR_MODULUS = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001


def fr_limbs(x: int) -> np.ndarray:
    return np.frombuffer(int(x % FR_MODULUS).to_bytes(32, "little"), dtype=np.uint64).copy()


def ballot_r1cs(num_constraints: int, num_inputs: int = 30, seed: int = 0) -> dict:
    """CSR arrays of the XOR-chain R1CS (field names of vsp_r1cs_t) plus the wiring needed to make witnesses."""
    rng = np.random.default_rng(seed)
    nc, ni = num_constraints, num_inputs
    # u, v index the variables that exist before constraint i: ni inputs + i earlier outputs (1-based columns)
    hi = ni + np.arange(nc, dtype=np.int64)
    u = (rng.random(nc) * hi).astype(np.int64)
    v = (rng.random(nc) * hi).astype(np.int64)
    new = ni + np.arange(nc, dtype=np.int64)  # 0-based variable index of the output
    two, one, minus_one = fr_limbs(2), fr_limbs(1), fr_limbs(-1)
    same = u == v
    # A: one term (u+1, 2);  B: one term (v+1, 1)
    a_row = np.arange(nc + 1, dtype=np.uint64)
    b_row = np.arange(nc + 1, dtype=np.uint64)
    a_col = (u + 1).astype(np.uint32)
    b_col = (v + 1).astype(np.uint32)
    a_val = np.tile(two, (nc, 1))
    b_val = np.tile(one, (nc, 1))
    # C: (u+1, 1), (v+1, 1), (new+1, -1)   or   (u+1, 2), (new+1, -1) when u == v
    c_cnt = np.where(same, 2, 3)
    c_row = np.zeros(nc + 1, dtype=np.uint64)
    c_row[1:] = np.cumsum(c_cnt)
    nnz = int(c_row[-1])
    c_col = np.zeros(nnz, dtype=np.uint32)
    c_val = np.zeros((nnz, 4), dtype=np.uint64)
    start = c_row[:-1].astype(np.int64)
    c_col[start] = (u + 1).astype(np.uint32)
    c_val[start] = np.where(same[:, None], two[None, :], one[None, :])
    ns = ~same
    c_col[start[ns] + 1] = (v[ns] + 1).astype(np.uint32)
    c_val[start[ns] + 1] = one
    last = start + c_cnt - 1
    c_col[last] = (new + 1).astype(np.uint32)
    c_val[last] = minus_one
    return {"num_inputs": ni, "num_aux": nc, "num_constraints": nc,
            "a_row": a_row, "b_row": b_row, "c_row": c_row, "a_col": a_col, "b_col": b_col, "c_col": c_col,
            "a_val": np.ascontiguousarray(a_val), "b_val": np.ascontiguousarray(b_val), "c_val": c_val,
            "_u": u, "_v": v}


def ballot_witness(cs: dict, seed: int):
    """(primary (ni, 4), auxiliary (nc, 4)) uint64 limbs: random input bits, outputs by the XOR chain."""
    rng = np.random.default_rng(seed + 1000003)
    ni, nc = cs["num_inputs"], cs["num_constraints"]
    full = np.zeros(ni + nc, dtype=np.uint8)
    full[:ni] = rng.integers(0, 2, size=ni, dtype=np.uint8)
    u, v = cs["_u"], cs["_v"]
    f = full.tolist()
    ul, vl = u.tolist(), v.tolist()
    for i in range(nc):
        f[ni + i] = f[ul[i]] ^ f[vl[i]]
    full = np.array(f, dtype=np.uint64)
    out = np.zeros((ni + nc, 4), dtype=np.uint64)
    out[:, 0] = full
    return out[:ni].copy(), out[ni:].copy()


def b_query_support(cs: dict) -> np.ndarray:
    """sorted variable indices with a non-zero B polynomial (the sparse B_query's index set)"""
    return np.unique(cs["b_col"]).astype(np.uint32)


def proving_key_arrays(vsp, ctx, cs: dict, m: int, seed: int = 7) -> dict:
    """Shape-only proving key: every query is a seeded synthetic point k_i * G built on the device
    (vsp_g1_bases_generate / vsp_g2_points_generate), so proofs are checkable against the oracle but not verifying."""
    ni, na = cs["num_inputs"], cs["num_aux"]
    nv = ni + na
    b_idx = b_query_support(cs)
    nb = len(b_idx)
    n_g1 = 3 + (nv + 1) + nb + (m - 1) + na
    h = vsp.G1Bases.generate(ctx, n_g1, seed, table=False)
    g1 = h.read()
    h.close()
    g2 = vsp.generate_g2_points(ctx, 2 + nb, seed + 1)
    pos = [0]

    def take(n):
        r = np.ascontiguousarray(g1[pos[0]: pos[0] + n])
        pos[0] += n
        return r

    arr = {"cs": {k: v for k, v in cs.items() if not k.startswith("_")}}
    arr["alpha_g1"], arr["beta_g1"], arr["delta_g1"] = take(1), take(1), take(1)
    arr["beta_g2"], arr["delta_g2"] = np.ascontiguousarray(g2[0:1]), np.ascontiguousarray(g2[1:2])
    arr["A_query"] = take(nv + 1)
    arr["B_count"] = nb
    arr["B_idx"] = b_idx
    arr["B_g1"] = take(nb)
    arr["B_g2"] = np.ascontiguousarray(g2[2:])
    arr["H_count"] = m - 1
    arr["H_query"] = take(m - 1)
    arr["L_query"] = take(na)
    return arr


def saver_pubkey_arrays(vsp, ctx, msg_size: int = 25, seed: int = 11):
    h = vsp.G1Bases.generate(ctx, 3 + 3 * msg_size, seed, table=False)
    g1 = h.read()
    h.close()
    g2 = vsp.generate_g2_points(ctx, msg_size + 1, seed + 1)
    arr = {"msg_size": msg_size,
           "delta_g1": np.ascontiguousarray(g1[0:1]),
           "delta_sum_s_g1": np.ascontiguousarray(g1[1:2]),
           "gamma_inverse_sum_s_g1": np.ascontiguousarray(g1[2:3]),
           "delta_s_g1": np.ascontiguousarray(g1[3:3 + msg_size]),
           "t_g1": np.ascontiguousarray(g1[3 + msg_size:3 + 2 * msg_size]),
           "t_g2": g2}
    return arr, np.ascontiguousarray(g1[3 + 2 * msg_size:])


def random_fr(n: int, seed: int) -> np.ndarray:
    """uniform in [0, r) up to a 2^-190 bias: 64-bit limbs with the top limb rejection-sampled below r's top limb"""
    rng = np.random.default_rng(seed)
    out = rng.integers(0, 1 << 64, size=(n, 4), dtype=np.uint64)
    top = np.uint64(FR_MODULUS >> 192)
    out[:, 3] &= np.uint64((1 << 63) - 1)
    bad = out[:, 3] >= top
    while bad.any():
        out[bad, 3] = rng.integers(0, 1 << 63, size=int(bad.sum()), dtype=np.uint64)
        bad = out[:, 3] >= top
    return out


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE configs[4]: the reference's synthetic field-input system (generate_r1cs_example_with_field_input,
# libs/marshalling/zk/test/detail/r1cs_examples.hpp:75-144), sized so that the evaluation domain is exactly 2^log_m
# ---------------------------------------------------------------------------------------------------------------------
def field_r1cs(num_constraints: int, num_inputs: int) -> dict:
    """CSR arrays (vsp_r1cs_t field names).  Variables x_1 .. x_nv, nv = num_constraints + 2 (index 0 is the constant 1):
    constraint i < nc - 1:  i odd  x_{i+1} * x_{i+2} = x_{i+3};   i even  (x_{i+1} + x_{i+2}) * 1 = x_{i+3}
    constraint nc - 1:      (sum_{j<nv} x_j) * (sum_{j<nv} x_j) = x_nv"""
    nc, ni = int(num_constraints), int(num_inputs)
    assert ni <= nc + 2 and nc >= 2
    nv = nc + 2
    one = fr_limbs(1)
    i = np.arange(nc - 1, dtype=np.int64)
    odd = (i % 2) == 1
    # A: odd rows (i + 1); even rows (i + 1), (i + 2); last row 1 .. nv - 1
    a_cnt = np.where(odd, 1, 2)
    a_row = np.zeros(nc + 1, dtype=np.uint64)
    a_row[1:nc] = np.cumsum(a_cnt)
    a_row[nc] = a_row[nc - 1] + np.uint64(nv - 1)
    a_col = np.zeros(int(a_row[nc]), dtype=np.uint32)
    st = a_row[:nc - 1].astype(np.int64)
    a_col[st] = (i + 1).astype(np.uint32)
    a_col[st[~odd] + 1] = (i[~odd] + 2).astype(np.uint32)
    a_col[int(a_row[nc - 1]):] = np.arange(1, nv, dtype=np.uint32)
    # B: odd rows (i + 2); even rows the constant (0); last row 1 .. nv - 1
    b_row = np.zeros(nc + 1, dtype=np.uint64)
    b_row[1:nc] = np.arange(1, nc, dtype=np.uint64)
    b_row[nc] = np.uint64(nc - 1 + nv - 1)
    b_col = np.zeros(int(b_row[nc]), dtype=np.uint32)
    b_col[:nc - 1] = np.where(odd, i + 2, 0).astype(np.uint32)
    b_col[nc - 1:] = np.arange(1, nv, dtype=np.uint32)
    # C: (i + 3); last row nv
    c_row = np.arange(nc + 1, dtype=np.uint64)
    c_col = np.concatenate([(i + 3).astype(np.uint32), np.array([nv], dtype=np.uint32)])

    def ones(n):
        return np.ascontiguousarray(np.tile(one, (n, 1)))

    return {"num_inputs": ni, "num_aux": nv - ni, "num_constraints": nc,
            "a_row": a_row, "b_row": b_row, "c_row": c_row, "a_col": a_col, "b_col": b_col, "c_col": c_col,
            "a_val": ones(len(a_col)), "b_val": ones(len(b_col)), "c_val": ones(len(c_col))}


def field_witness(cs: dict, seed: int):
    """(primary (ni, 4), auxiliary (na, 4)): two seeded field elements, then a + b / a * b alternately, then the square of
    the sum of everything before it (r1cs_examples.hpp:84-126)"""
    nc, ni = cs["num_constraints"], cs["num_inputs"]
    a, b = (int.from_bytes(x.tobytes(), "little") for x in random_fr(2, seed))
    vals = [a, b]
    app, q = vals.append, FR_MODULUS
    for i in range(nc - 1):
        t = a * b % q if i & 1 else (a + b) % q
        app(t)
        a, b = b, t
    vals.append(pow(sum(vals) % q, 2, q))
    raw = b"".join(v.to_bytes(32, "little") for v in vals)
    full = np.frombuffer(raw, dtype=np.uint64).reshape(-1, 4)
    return np.ascontiguousarray(full[:ni]), np.ascontiguousarray(full[ni:])


def key_layout(cs: dict, m: int) -> dict:
    """where each query of proving_key_arrays sits in the seeded G1 / G2 point streams (so that a rank can generate only
    its own slice of a query: point j of stream `seed` is k_j * G with k_j = splitmix64(seed + j) | 1)"""
    ni, na = cs["num_inputs"], cs["num_aux"]
    nv = ni + na
    nb = len(b_query_support(cs))
    a0 = 3
    b0 = a0 + nv + 1
    h0 = b0 + nb
    l0 = h0 + m - 1
    return {"nv": nv, "nb": nb, "A": (a0, nv + 1), "B_g1": (b0, nb), "H": (h0, m - 1), "L": (l0, na), "B_g2": (2, nb)}


def proving_key_slice_arrays(vsp, ctx, cs: dict, m: int, seed: int, ranges: dict) -> dict:
    """the slice of proving_key_arrays(vsp, ctx, cs, m, seed) that sharded.slice_proving_key would cut for `ranges`
    ({"A": (lo, hi), "B": ..., "H": ..., "L": ...}), generated directly: no rank ever holds the whole key"""
    lay = key_layout(cs, m)

    def g1(off, lo, hi):
        if hi <= lo:
            return np.zeros((1, 12), dtype=np.uint64)
        h = vsp.G1Bases.generate(ctx, hi - lo, seed + off + lo, table=False)
        out = h.read()
        h.close()
        return out

    head = g1(0, 0, 3)
    g2_head = vsp.generate_g2_points(ctx, 2, seed + 1)
    blo, bhi = ranges["B"]
    b_idx = b_query_support(cs)
    arr = {"cs": {k: v for k, v in cs.items() if not k.startswith("_")},
           "alpha_g1": np.ascontiguousarray(head[0:1]), "beta_g1": np.ascontiguousarray(head[1:2]),
           "delta_g1": np.ascontiguousarray(head[2:3]),
           "beta_g2": np.ascontiguousarray(g2_head[0:1]), "delta_g2": np.ascontiguousarray(g2_head[1:2]),
           "A_query": g1(lay["A"][0], *ranges["A"]),
           "B_count": bhi - blo,
           "B_idx": np.ascontiguousarray(b_idx[blo:bhi]) if bhi > blo else np.zeros(1, dtype=np.uint32),
           "B_g1": g1(lay["B_g1"][0], blo, bhi),
           "B_g2": vsp.generate_g2_points(ctx, bhi - blo, seed + 1 + 2 + blo) if bhi > blo else np.zeros((1, 24), np.uint64),
           "H_count": ranges["H"][1] - ranges["H"][0],
           "H_query": g1(lay["H"][0], *ranges["H"]),
           "L_query": g1(lay["L"][0], *ranges["L"])}
    return arr
