"""The ballot circuit of VoteSAVER's encrypted_input mode: R1CS builder, variable layout and batched witness generation.

Replaces, on the prover's side, what bin/cli builds from crypto3-blueprint components on every vote
(/root/reference/bin/cli/include/nil/vote_saver/common.hpp:1054-1107 re-synthesises the R1CS, :1110-1128 generates the
witness in nine steps): `components::encrypted_input_voting<2, pedersen<jubjub, sha2<256>, ...>, ...>` over

    m_block (msg_size one-hot vote bits)         eid_block (eid_bits)          sk_block (255-bit voter secret)
    pk  = PedersenHash(sk)                                    -- hash<merkle_hash_type>(sk), common.hpp:833
    rt  = Merkle root over the voters' pk, arity 2            -- containers/merkle/proof.hpp:66-92, address bits
    sn  = PedersenHash(eid || sk)                             -- common.hpp:1044-1047
    primary input = m (25) | pack(eid) | pack(sn) | pack(rt), chunk_size = 254 bits  -- common.hpp:861-876: 30 elements

PARITY UNPINNED.  crypto3-blueprint and crypto3-hash are un-vendored submodules: neither the component
sources nor the find_group_hash generators exist in /root/reference, and the reference prints the circuit's size only at
run time (common.hpp:911-912).  This module restates the construction those components are ports of — the Zcash Sapling
Pedersen hash over Jubjub as ethsnarks / bellman arithmetise it (3-bit signed windows looked up in Montgomery form, 63
windows per generator, Montgomery additions inside a segment, Edwards additions between segments, the u-coordinate as the
digest) — with its own deterministic generators (SHA-256 try-and-increment, below).  It has the reference circuit's
interface (the 30 primary inputs, their order and packing) and shape; its constraint count and wire values are this
module's, not crypto3's.  What holds it: every witness satisfies the R1CS (checked by the oracle's own evaluator), sn and
rt equal an independent native computation, and generate -> prove -> verify accepts (tests/test_voting.py,
tests/test_gpu_voting.py).

Layout (variable 0 is the constant 1; indices below are 1-based variable numbers):
    primary   m[msg_size] | eid_packed[ceil(eid_bits/254)] | sn_packed[2] | rt_packed[2]
    auxiliary eid bits | sk bits | address bits[depth] | PedersenHash(sk) | PedersenHash(eid || sk) |
              per level: sibling bits[255], left[255], right[255], PedersenHash(left || right)
A Pedersen hash over n input bits (W = ceil(n/3) windows, S = ceil(W/63) segments) allocates, in this order:
    per window w: c_w = b0 b1, y_w = ylc (1 - 2 b2), and unless w opens a segment lambda_w, x_w, yy_w (Montgomery sum)
    per segment s: u_s, v_s (its Edwards form);  per segment s >= 1: U, A, B, C, u3, v3 (Edwards sum so far)
    255 digest bits (little-endian bits of the final u)
and 5 W + 5 S + 250 constraints.  The CUDA generator (csrc/voting.cu) walks the same layout from the tables this module
hands it, one thread per (ballot, hash, segment).
"""
from __future__ import annotations

import ctypes as C
import hashlib
from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import numpy as np

Q = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001  # Fr of BLS12-381 = Jubjub's base field
JUBJUB_D = (-10240 * pow(10241, -1, Q)) % Q   # -u^2 + v^2 = 1 + d u^2 v^2
MONT_A = 40962                                # y^2 = x^3 + A x^2 + x
JUBJUB_ORDER = 6554484396890773809930967563523245729705921265872317281365359162392183254199  # prime subgroup, cofactor 8
DIGEST_BITS = 255
CHUNK_BITS = 254                              # field_type::value_bits - 1 (common.hpp:861)
WINDOWS_PER_SEGMENT = 63


def _sqrt(a: int):
    """a square root mod Q (Tonelli-Shanks; Q - 1 = 2^32 t) or None"""
    a %= Q
    if a == 0:
        return 0
    if pow(a, (Q - 1) // 2, Q) != 1:
        return None
    s, t = 32, (Q - 1) >> 32
    z = 7  # a non-residue: the field's multiplicative generator
    c, x, b, m = pow(z, t, Q), pow(a, (t + 1) // 2, Q), pow(a, t, Q), s
    while b != 1:
        i, bb = 0, b
        while bb != 1:
            bb = bb * bb % Q
            i += 1
        g = pow(c, 1 << (m - i - 1), Q)
        x, b, c, m = x * g % Q, b * g * g % Q, g * g % Q, i
    return x


MONT_SCALE = min(_sqrt(-(MONT_A + 2)), Q - _sqrt(-(MONT_A + 2)))  # (x, y) -> u = scale x / y


# ---- Jubjub, twisted Edwards (complete addition law) ---------------------------------------------------------------
def ed_add(p, q):
    (u1, v1), (u2, v2) = p, q
    t = JUBJUB_D * u1 * u2 % Q * v1 % Q * v2 % Q
    return ((u1 * v2 + v1 * u2) * pow(1 + t, -1, Q) % Q, (v1 * v2 + u1 * u2) * pow(1 - t, -1, Q) % Q)


def _proj_add(p, q):
    """projective twisted Edwards addition, a = -1 (add-2008-bbjlp): no inversion"""
    (x1, y1, z1), (x2, y2, z2) = p, q
    a = z1 * z2 % Q
    b = a * a % Q
    c = x1 * x2 % Q
    d = y1 * y2 % Q
    e = JUBJUB_D * c % Q * d % Q
    f, g = (b - e) % Q, (b + e) % Q
    return (a * f % Q * ((x1 + y1) * (x2 + y2) - c - d) % Q, a * g % Q * (d + c) % Q, f * g % Q)


def ed_mul(k: int, p):
    """k p by double-and-add in projective coordinates, one inversion at the end"""
    acc, pp = (0, 1, 1), (p[0], p[1], 1)
    for bit in bin(k)[2:] if k else "":
        acc = _proj_add(acc, acc)
        if bit == "1":
            acc = _proj_add(acc, pp)
    zi = pow(acc[2], -1, Q)
    return (acc[0] * zi % Q, acc[1] * zi % Q)


def ed_neg(p):
    return ((-p[0]) % Q, p[1])


def ed_to_mont(p):
    u, v = p
    x = (1 + v) * pow(1 - v, -1, Q) % Q
    return x, MONT_SCALE * x % Q * pow(u, -1, Q) % Q


def pedersen_generator(i: int):
    """segment generator I_i: SHA-256("VoteSAVER_PH" | i | counter) as a v-coordinate, the root u with the hash's top bit as
    its parity, times the cofactor; first counter that gives a point of the prime-order subgroup"""
    ctr = 0
    while True:
        h = hashlib.sha256(b"VoteSAVER_PH" + i.to_bytes(4, "little") + ctr.to_bytes(4, "little")).digest()
        ctr += 1
        v = int.from_bytes(h, "little") % Q
        u = _sqrt((v * v - 1) * pow(1 + JUBJUB_D * v * v, -1, Q))
        if u is None or u == 0:
            continue
        if (u & 1) != (h[31] >> 7):
            u = Q - u
        p = ed_mul(8, (u, v))
        if p != (0, 1) and ed_mul(JUBJUB_ORDER, p) == (0, 1):
            return p


_GENERATORS: Dict[int, Tuple[int, int]] = {}


def generator(i: int):
    if i not in _GENERATORS:
        _GENERATORS[i] = pedersen_generator(i)
    return _GENERATORS[i]


def pedersen_hash_native(bits: Sequence[int]) -> int:
    """sum_i <M_i> I_i in Edwards coordinates, <M_i> = sum_j enc(m_j) 2^(4 j), enc = (1 + b0 + 2 b1)(1 - 2 b2); the
    u-coordinate of the sum.  Independent of the circuit's Montgomery walk: scalar multiples, not window tables."""
    bits = list(bits) + [0] * (-len(bits) % 3)
    acc = (0, 1)
    for seg in range(0, len(bits) // 3, WINDOWS_PER_SEGMENT):
        k = 0
        for j, w in enumerate(range(seg, min(seg + WINDOWS_PER_SEGMENT, len(bits) // 3))):
            b0, b1, b2 = bits[3 * w:3 * w + 3]
            k += (1 + b0 + 2 * b1) * (1 - 2 * b2) << (4 * j)
        acc = ed_add(acc, ed_mul(k % JUBJUB_ORDER, generator(seg // WINDOWS_PER_SEGMENT)))
    return acc[0]


def digest_bits(x: int) -> List[int]:
    return [(x >> i) & 1 for i in range(DIGEST_BITS)]


def merkle_root_native(leaves: Sequence[int]) -> int:
    """arity-2 Pedersen tree over leaf digests (integers); len(leaves) a power of two"""
    level = list(leaves)
    while len(level) > 1:
        level = [pedersen_hash_native(digest_bits(level[i]) + digest_bits(level[i + 1])) for i in range(0, len(level), 2)]
    return level[0]


def merkle_path_native(leaves: Sequence[int], index: int) -> List[int]:
    """sibling digests from the leaf level up (containers/merkle/proof.hpp:66-92 for arity 2)"""
    level, path = list(leaves), []
    while len(level) > 1:
        path.append(level[index ^ 1])
        level = [pedersen_hash_native(digest_bits(level[i]) + digest_bits(level[i + 1])) for i in range(0, len(level), 2)]
        index >>= 1
    return path


class SparseVoterTree:
    """A 2^depth-leaf voter tree in which only the first len(sks) leaves hold real public keys (pk_i = PedersenHash(sk_i))
    and every other leaf holds the same default digest: the subtrees made of defaults only collapse to one value per level,
    so a consistent depth-16 tree costs a few hundred hashes instead of 65,535.  Gives the root and every voter's copath."""

    def __init__(self, depth: int, sks: Sequence[int], default_leaf: int = 0):
        n = len(sks)
        assert n <= (1 << depth)
        self.depth = depth
        level = [pedersen_hash_native(digest_bits(s)) for s in sks]
        self.leaves = list(level)
        default = default_leaf
        self.levels = []  # (real nodes of the level, the all-default node of the level)
        for _ in range(depth):
            self.levels.append((level, default))
            nxt = []
            for i in range(0, len(level), 2):
                right = level[i + 1] if i + 1 < len(level) else default
                nxt.append(pedersen_hash_native(digest_bits(level[i]) + digest_bits(right)))
            default = pedersen_hash_native(digest_bits(default) * 2)
            level = nxt
        self.root = level[0] if level else default

    def path(self, index: int) -> List[int]:
        out = []
        for level, default in self.levels:
            sib = index ^ 1
            out.append(level[sib] if sib < len(level) else default)
            index >>= 1
        return out


# ---- the circuit ---------------------------------------------------------------------------------------------------
LinComb = List[Tuple[int, int]]


@dataclass
class HashInstance:
    in_vars: List[int]   # variable index of every input bit
    var_base: int        # first internal variable
    out_base: int        # first digest-bit variable
    windows: int
    segments: int


def window_offset(w: int) -> int:
    """internal-variable offset of window w inside a hash instance: 2 per window + 3 per window that does not open a segment"""
    return 5 * w - 3 * ((w + WINDOWS_PER_SEGMENT - 1) // WINDOWS_PER_SEGMENT)


@dataclass
class VotingCircuit:
    depth: int
    eid_bits: int = 64
    msg_size: int = 25
    constraints: List[Tuple[LinComb, LinComb, LinComb]] = field(default_factory=list)
    num_vars: int = 0
    hashes: List[HashInstance] = field(default_factory=list)

    def __post_init__(self):
        self._build()

    # -- allocation
    def _alloc(self, n: int = 1) -> int:
        first = self.num_vars + 1
        self.num_vars += n
        return first

    def _c(self, a: LinComb, b: LinComb, c: LinComb):
        self.constraints.append((a, b, c))

    def _boolean(self, v: int):
        self._c([(v, 1)], [(0, 1), (v, -1)], [])

    def window_table(self, seg: int, j: int):
        """Montgomery coordinates of k 2^(4 j) I_seg, k = 1..4"""
        key = (seg, j)
        if key not in self._tables:
            base = ed_mul(pow(2, 4 * j, JUBJUB_ORDER), generator(seg))
            pts, acc = [], (0, 1)
            for _k in range(4):
                acc = ed_add(acc, base)
                pts.append(ed_to_mont(acc))
            self._tables[key] = pts
        return self._tables[key]

    def _pedersen(self, in_vars: List[int]) -> HashInstance:
        n = len(in_vars)
        W = (n + 2) // 3
        S = (W + WINDOWS_PER_SEGMENT - 1) // WINDOWS_PER_SEGMENT
        var_base = self._alloc(window_offset(W) + 2 * S + 6 * (S - 1))
        out_base = self._alloc(DIGEST_BITS)
        h = HashInstance(list(in_vars), var_base, out_base, W, S)
        self.hashes.append(h)
        bit = lambda i: [(in_vars[i], 1)] if i < n else []  # padding bits are the constant 0
        seg_uv = []
        for w in range(W):
            seg, j = divmod(w, WINDOWS_PER_SEGMENT)
            (x1, y1), (x2, y2), (x3, y3), (x4, y4) = self.window_table(seg, j)
            b0, b1, b2 = bit(3 * w), bit(3 * w + 1), bit(3 * w + 2)
            o = var_base + window_offset(w)
            c_v, y_v = o, o + 1
            self._c(b0, b1, [(c_v, 1)])
            lc = lambda k1, k2, k3, k4: ([(0, k1)] + [(v, (k2 - k1) * co) for v, co in b0] + [(v, (k3 - k1) * co) for v, co in b1]
                                         + [(c_v, k4 - k3 - k2 + k1)])
            xlc, ylc = lc(x1, x2, x3, x4), lc(y1, y2, y3, y4)
            self._c(ylc, [(0, 1)] + [(v, -2 * co) for v, co in b2], [(y_v, 1)])
            if j == 0:
                acc_x, acc_y = xlc, [(y_v, 1)]
            else:
                lam, xs, ys = o + 2, o + 3, o + 4
                neg = lambda l: [(v, -co) for v, co in l]
                self._c([(lam, 1)], xlc + neg(acc_x), [(y_v, 1)] + neg(acc_y))
                self._c([(lam, 1)], [(lam, 1)], [(0, MONT_A)] + acc_x + xlc + [(xs, 1)])
                self._c([(lam, 1)], acc_x + [(xs, -1)], acc_y + [(ys, 1)])
                acc_x, acc_y = [(xs, 1)], [(ys, 1)]
            if j == WINDOWS_PER_SEGMENT - 1 or w == W - 1:
                u_v = var_base + window_offset(W) + 2 * seg
                v_v = u_v + 1
                self._c(acc_y, [(u_v, 1)], [(v, MONT_SCALE * co) for v, co in acc_x])
                self._c(acc_x + [(0, 1)], [(v_v, 1)], acc_x + [(0, -1)])
                seg_uv.append((u_v, v_v))
        cu, cv = seg_uv[0]
        for s in range(1, S):
            o = var_base + window_offset(W) + 2 * S + 6 * (s - 1)
            U, A, B, Cc, u3, v3 = range(o, o + 6)
            su, sv = seg_uv[s]
            self._c([(cu, 1), (cv, 1)], [(su, 1), (sv, 1)], [(U, 1)])
            self._c([(sv, 1)], [(cu, 1)], [(A, 1)])
            self._c([(su, 1)], [(cv, 1)], [(B, 1)])
            self._c([(A, JUBJUB_D)], [(B, 1)], [(Cc, 1)])
            self._c([(0, 1), (Cc, 1)], [(u3, 1)], [(A, 1), (B, 1)])
            # -u^2 + v^2 curve: v3 = (v1 v2 + u1 u2) / (1 - C), and v1 v2 + u1 u2 = U - A - B with U = (u1 + v1)(u2 + v2)
            self._c([(0, 1), (Cc, -1)], [(v3, 1)], [(U, 1), (A, -1), (B, -1)])
            cu, cv = u3, v3
        for i in range(DIGEST_BITS):
            self._boolean(out_base + i)
        self._c([(out_base + i, pow(2, i, Q)) for i in range(DIGEST_BITS)], [(0, 1)], [(cu, 1)])
        return h

    def _pack(self, bit_vars: List[int], packed_first: int):
        for k in range(0, len(bit_vars), CHUNK_BITS):
            chunk = bit_vars[k:k + CHUNK_BITS]
            self._c([(v, 1 << i) for i, v in enumerate(chunk)], [(0, 1)], [(packed_first + k // CHUNK_BITS, 1)])

    def _build(self):
        self._tables = {}
        n_eid_packed = (self.eid_bits + CHUNK_BITS - 1) // CHUNK_BITS
        n_dig_packed = (DIGEST_BITS + CHUNK_BITS - 1) // CHUNK_BITS
        self.m = self._alloc(self.msg_size)
        self.eid_packed = self._alloc(n_eid_packed)
        self.sn_packed = self._alloc(n_dig_packed)
        self.rt_packed = self._alloc(n_dig_packed)
        self.num_inputs = self.num_vars
        self.eid = self._alloc(self.eid_bits)
        self.sk = self._alloc(DIGEST_BITS)
        self.address = self._alloc(self.depth)
        for i in range(self.msg_size):
            self._boolean(self.m + i)
        self._c([(self.m + i, 1) for i in range(self.msg_size)], [(0, 1)], [(0, 1)])  # exactly one vote
        for i in range(self.eid_bits):
            self._boolean(self.eid + i)
        for i in range(DIGEST_BITS):
            self._boolean(self.sk + i)
        for i in range(self.depth):
            self._boolean(self.address + i)
        sk_vars = list(range(self.sk, self.sk + DIGEST_BITS))
        self.h_pk = self._pedersen(sk_vars)
        self.h_sn = self._pedersen(list(range(self.eid, self.eid + self.eid_bits)) + sk_vars)
        cur = self.h_pk.out_base
        self.levels = []
        for lvl in range(self.depth):
            sib = self._alloc(DIGEST_BITS)
            left = self._alloc(DIGEST_BITS)
            right = self._alloc(DIGEST_BITS)
            a = self.address + lvl
            for i in range(DIGEST_BITS):
                self._boolean(sib + i)
                # address bit 1: the running digest is the right child
                self._c([(a, 1)], [(sib + i, 1), (cur + i, -1)], [(left + i, 1), (cur + i, -1)])
                self._c([(a, 1)], [(cur + i, 1), (sib + i, -1)], [(right + i, 1), (sib + i, -1)])
            h = self._pedersen(list(range(left, left + DIGEST_BITS)) + list(range(right, right + DIGEST_BITS)))
            self.levels.append((sib, left, right, h))
            cur = h.out_base
        self.root_bits = cur
        self._pack(list(range(self.eid, self.eid + self.eid_bits)), self.eid_packed)
        self._pack(list(range(self.h_sn.out_base, self.h_sn.out_base + DIGEST_BITS)), self.sn_packed)
        self._pack(list(range(self.root_bits, self.root_bits + DIGEST_BITS)), self.rt_packed)
        self.num_aux = self.num_vars - self.num_inputs
        self.num_constraints = len(self.constraints)

    # -- R1CS in the C ABI's CSR form (vsp_r1cs_t field names)
    def r1cs_arrays(self) -> dict:
        arr = {"num_inputs": self.num_inputs, "num_aux": self.num_aux, "num_constraints": self.num_constraints}
        for k, name in enumerate("abc"):
            rows, cols, vals = [0], [], bytearray()
            for con in self.constraints:
                merged: Dict[int, int] = {}
                for idx, co in con[k]:
                    merged[idx] = (merged.get(idx, 0) + co) % Q
                for idx in sorted(merged):
                    if merged[idx]:
                        cols.append(idx)
                        vals += merged[idx].to_bytes(32, "little")
                rows.append(len(cols))
            arr[name + "_row"] = np.array(rows, dtype=np.uint64)
            arr[name + "_col"] = np.array(cols if cols else [0], dtype=np.uint32)
            arr[name + "_val"] = (np.frombuffer(bytes(vals), dtype=np.uint64).reshape(-1, 4).copy() if cols
                                  else np.zeros((1, 4), dtype=np.uint64))
        return arr

    # -- host restatement of the witness generation (what csrc/voting.cu computes per ballot), Python integers
    def _pedersen_witness(self, h: HashInstance, w: List[int]):
        n = len(h.in_vars)
        bit = lambda i: w[h.in_vars[i]] if i < n else 0
        inv = lambda x: pow(x, -1, Q)
        seg_uv = []
        for k in range(h.windows):
            seg, j = divmod(k, WINDOWS_PER_SEGMENT)
            tab = self.window_table(seg, j)
            b0, b1, b2 = bit(3 * k), bit(3 * k + 1), bit(3 * k + 2)
            px, py = tab[b0 + 2 * b1]
            o = h.var_base + window_offset(k)
            w[o] = b0 * b1
            w[o + 1] = py if not b2 else (Q - py) % Q
            if j == 0:
                ax, ay = px, w[o + 1]
            else:
                lam = (w[o + 1] - ay) * inv(px - ax) % Q
                xs = (lam * lam - MONT_A - ax - px) % Q
                ys = (lam * (ax - xs) - ay) % Q
                w[o + 2], w[o + 3], w[o + 4] = lam, xs, ys
                ax, ay = xs, ys
            if j == WINDOWS_PER_SEGMENT - 1 or k == h.windows - 1:
                u = MONT_SCALE * ax % Q * inv(ay) % Q
                v = (ax - 1) * inv(ax + 1) % Q
                o2 = h.var_base + window_offset(h.windows) + 2 * seg
                w[o2], w[o2 + 1] = u, v
                seg_uv.append((u, v))
        cu, cv = seg_uv[0]
        for s in range(1, h.segments):
            o = h.var_base + window_offset(h.windows) + 2 * h.segments + 6 * (s - 1)
            su, sv = seg_uv[s]
            U = (cu + cv) * (su + sv) % Q
            A = sv * cu % Q
            B = su * cv % Q
            Cc = JUBJUB_D * A % Q * B % Q
            u3 = (A + B) * inv(1 + Cc) % Q
            v3 = (U - A - B) * inv(1 - Cc) % Q
            w[o:o + 6] = [U, A, B, Cc, u3, v3]
            cu, cv = u3, v3
        for i in range(DIGEST_BITS):
            w[h.out_base + i] = (cu >> i) & 1
        return cu

    def witness(self, vote: int, eid: int, sk: int, voter_index: int, path: Sequence[int]):
        """full assignment for one ballot as Python integers: (primary, auxiliary).  path: sibling digests, leaf level first."""
        assert 0 <= vote < self.msg_size and len(path) == self.depth
        w = [0] * (self.num_vars + 1)
        w[0] = 1
        w[self.m + vote] = 1
        for i in range(self.eid_bits):
            w[self.eid + i] = (eid >> i) & 1
        for i in range(DIGEST_BITS):
            w[self.sk + i] = (sk >> i) & 1
        for i in range(self.depth):
            w[self.address + i] = (voter_index >> i) & 1
        self._pedersen_witness(self.h_pk, w)
        self._pedersen_witness(self.h_sn, w)
        cur = self.h_pk.out_base
        for lvl, (sib, left, right, h) in enumerate(self.levels):
            a = w[self.address + lvl]
            for i in range(DIGEST_BITS):
                s_i, c_i = (path[lvl] >> i) & 1, w[cur + i]
                w[sib + i] = s_i
                w[left + i], w[right + i] = (s_i, c_i) if a else (c_i, s_i)
            self._pedersen_witness(h, w)
            cur = h.out_base

        def pack(first_bit: int, nbits: int, dst: int):
            for k in range(0, nbits, CHUNK_BITS):
                w[dst + k // CHUNK_BITS] = sum(w[first_bit + k + i] << i for i in range(min(CHUNK_BITS, nbits - k)))

        pack(self.eid, self.eid_bits, self.eid_packed)
        pack(self.h_sn.out_base, DIGEST_BITS, self.sn_packed)
        pack(self.root_bits, DIGEST_BITS, self.rt_packed)
        return w[1:self.num_inputs + 1], w[self.num_inputs + 1:]

    # -- tables for the CUDA generator
    def device_program(self) -> dict:
        """flat arrays csrc/voting.cu walks: per hash (in_off, n_bits, var_base, out_base, windows, segments), the input-bit
        variable indices, per level (sib, left, right, hash index), the window tables (canonical Fr, (segment, window, k) ->
        x | y) and the scalar layout fields"""
        hashes, in_idx = [], []
        for h in self.hashes:
            hashes.append([len(in_idx), len(h.in_vars), h.var_base, h.out_base, h.windows, h.segments])
            in_idx += h.in_vars
        max_seg = max(h.segments for h in self.hashes)
        tab = bytearray()
        for seg in range(max_seg):
            for j in range(WINDOWS_PER_SEGMENT):
                for (x, y) in self.window_table(seg, j):
                    tab += x.to_bytes(32, "little") + y.to_bytes(32, "little")
        for cst in (MONT_A, JUBJUB_D, MONT_SCALE):  # the curve constants ride behind the window tables
            tab += cst.to_bytes(32, "little")
        levels = [[sib, left, right, self.hashes.index(h)] for sib, left, right, h in self.levels]
        return {"hashes": np.array(hashes, dtype=np.uint32), "in_idx": np.array(in_idx, dtype=np.uint32),
                "levels": np.array(levels if levels else [[0, 0, 0, 0]], dtype=np.uint32),
                "tables": np.frombuffer(bytes(tab), dtype=np.uint64).reshape(-1, 4).copy(),
                "scalars": np.array([self.depth, self.eid_bits, self.msg_size, self.num_vars, self.num_inputs, self.m,
                                     self.eid_packed, self.sn_packed, self.rt_packed, self.eid, self.sk, self.address,
                                     self.hashes.index(self.h_pk), self.hashes.index(self.h_sn), max_seg], dtype=np.uint32)}


def ints_to_fr(vals: Sequence[int]) -> np.ndarray:
    return np.frombuffer(b"".join(int(v).to_bytes(32, "little") for v in vals), dtype=np.uint64).reshape(-1, 4).copy()


class WitnessGenerator:
    """Batched witness generation on the GPU (vsp_voting_*): the circuit's tables resident on the device, k ballots per
    call, full assignments (1 || primary || auxiliary) left in device memory for vsp_saver_encrypt_batch_dev."""

    def __init__(self, ctx, circuit: VotingCircuit):
        self.ctx, self.circuit = ctx, circuit
        prog = circuit.device_program()
        self._keep = prog
        h = C.c_void_p()
        p = lambda a: C.c_void_p(a.ctypes.data)
        ctx.check(ctx._lib.vsp_voting_program_create(ctx._h, p(prog["scalars"]), len(prog["scalars"]), p(prog["hashes"]),
                                                     len(prog["hashes"]), p(prog["in_idx"]), len(prog["in_idx"]),
                                                     p(prog["levels"]), circuit.depth, p(prog["tables"]),
                                                     len(prog["tables"]), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            self.ctx._lib.vsp_voting_program_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def generate_dev(self, k: int, votes, eid: int, sks, indices, paths, d_assignments: int):
        """votes (k,) uint32; sks (k, 4) uint64 (255-bit secrets); indices (k,) uint32 voter positions; paths (k, depth, 4)
        uint64 sibling digests, leaf level first; d_assignments: device pointer to k x (num_vars + 1) x 4 uint64.
        Stream-ordered on the context's stream once the small inputs are uploaded."""
        v = np.ascontiguousarray(votes, dtype=np.uint32)
        s = np.ascontiguousarray(sks, dtype=np.uint64).reshape(k, 4)
        ix = np.ascontiguousarray(indices, dtype=np.uint32)
        pt = np.ascontiguousarray(paths, dtype=np.uint64).reshape(k, max(self.circuit.depth, 1), 4) \
            if self.circuit.depth else np.zeros((k, 1, 4), dtype=np.uint64)
        p = lambda a: C.c_void_p(a.ctypes.data)
        self.ctx.check(self.ctx._lib.vsp_voting_witness_batch_dev(self.ctx._h, self._h, k, p(v), C.c_uint64(int(eid)), p(s),
                                                                  p(ix), p(pt), C.c_void_p(d_assignments)))

    def generate(self, votes, eid: int, sks, indices, paths) -> np.ndarray:
        """the same with the assignments copied back: (k, num_vars + 1, 4) uint64"""
        import torch
        k = len(votes)
        out = torch.zeros((k, self.circuit.num_vars + 1, 4), dtype=torch.int64, device="cuda:%d" % self.ctx.device)
        self.generate_dev(k, votes, eid, sks, indices, paths, out.data_ptr())
        self.ctx.synchronize()
        return out.cpu().numpy().view(np.uint64)
