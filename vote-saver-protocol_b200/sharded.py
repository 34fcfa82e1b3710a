"""Multi-GPU sharding of the ballot-proof path over torch.distributed (one process per GPU; NCCL over NVLink 5 /
NVSwitch on the GPU box, gloo in the CPU tests).  SURVEY.md 8e / DESIGN.md 7:

  * batched ballots shard proof-per-GPU: ballot i -> rank i mod G, proving key replicated, no data-path collective;
    proofs (48 limbs) and ciphertexts are gathered at the end if the caller wants them in one place;
  * one large multiexp shards by contiguous point range: rank g runs the whole bucket pipeline on bases
    [g N / G, (g + 1) N / G) and emits ONE affine point; the G points are exchanged by a single all_gather and summed
    locally (NCCL has no elliptic-curve reduction, so this cannot be an all_reduce).

The local multiexp and the point sum are injectable so the partition / exchange / join logic is testable on CPU with
world_size 2 (tests/test_sharding.py supplies the oracle's functions there; the product default is the CUDA library).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """contiguous point range [lo, hi) of rank `rank`: sizes differ by at most one, earlier ranks take the remainder"""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    q, r = divmod(n, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def ballot_indices(k: int, rank: int, world: int) -> np.ndarray:
    """ballots of rank `rank` under proof-per-GPU sharding: i = rank (mod world)"""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    return np.arange(rank, k, world, dtype=np.int64)


def _dist():
    import torch.distributed as dist
    return dist


def _world(group=None) -> Tuple[int, int]:
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def all_gather_rows(local: np.ndarray, group=None, device: Optional[str] = None) -> np.ndarray:
    """all_gather of one fixed-shape uint64 array per rank -> array of shape (world,) + local.shape.
    `device`: "cuda" moves the payload through the GPU (required by NCCL); None picks it from the backend."""
    import torch
    dist = _dist()
    rank, world = _world(group)
    local = np.ascontiguousarray(local, dtype=np.uint64)
    if world == 1:
        return local[None].copy()
    if device is None:
        device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.from_numpy(local.view(np.int64).copy()).to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    return np.stack([o.cpu().numpy().view(np.uint64) for o in out])


def sharded_multiexp(local_multiexp: Callable[[], np.ndarray], point_sum: Callable[[np.ndarray], np.ndarray],
                     group=None, device: Optional[str] = None) -> np.ndarray:
    """Point-range sharded multiexp.  local_multiexp() -> this rank's partial result as a canonical affine point
    (12 limbs G1 / 24 limbs G2, all-zero = infinity); point_sum(points (world, L)) -> their sum, same format.
    Every rank returns the full result (the join is replicated: G - 1 point additions)."""
    part = np.ascontiguousarray(local_multiexp(), dtype=np.uint64).reshape(-1)
    if part.shape[0] not in (12, 24):
        raise ValueError("a partial multiexp result must be one affine G1 (12 limbs) or G2 (24 limbs) point")
    parts = all_gather_rows(part, group, device)
    return np.ascontiguousarray(point_sum(parts), dtype=np.uint64).reshape(-1)


class ShardedBases:
    """This rank's contiguous slice of a large bases vector, resident on its GPU with its window table."""

    def __init__(self, vsp, ctx, n_total: int, *, bases_local: Optional[np.ndarray] = None, seed: Optional[int] = None,
                 group=None):
        self.vsp, self.ctx, self.group = vsp, ctx, group
        self.rank, self.world = _world(group)
        self.n_total = n_total
        self.lo, self.hi = shard_range(n_total, self.rank, self.world)
        if bases_local is not None:
            b = np.ascontiguousarray(bases_local, dtype=np.uint64)
            if b.shape[0] != self.hi - self.lo:
                raise ValueError("bases_local must hold exactly this rank's range")
            self.g2 = b.ndim == 2 and b.shape[1] == 24
            self.handle = (vsp.G2Bases if self.g2 else vsp.G1Bases)(ctx, b)
        else:
            # synthetic sweep bases: k_i * G with k_i = splitmix64(seed + i); every rank
            # generates only its own range, with the global index as the stream position
            self.g2 = False
            self.handle = vsp.G1Bases.generate(ctx, self.hi - self.lo, int(seed) + self.lo)

    def multiexp(self, scalars_local) -> np.ndarray:
        """scalars_local: this rank's (hi - lo, 4) canonical Fr.  Returns the full sum on every rank."""
        s = np.ascontiguousarray(scalars_local, dtype=np.uint64).reshape(-1, 4)
        if s.shape[0] != self.hi - self.lo:
            raise ValueError("scalars_local must hold exactly this rank's range")
        return sharded_multiexp(lambda: self.handle.multiexp(s), lambda pts: self.vsp.points_sum(self.ctx, pts),
                                self.group)

    def close(self):
        self.handle.close()


def gather_ballot_results(local: np.ndarray, k: int, group=None, device: Optional[str] = None) -> np.ndarray:
    """Inverse of ballot_indices: every rank passes the results of its ballots (rows in its own order); returns the
    (k, ...) array in ballot order on every rank.  Ranks with fewer ballots are padded for the exchange."""
    rank, world = _world(group)
    local = np.ascontiguousarray(local, dtype=np.uint64)
    per = -(-k // world)
    pad = np.zeros((per,) + local.shape[1:], dtype=np.uint64)
    pad[: local.shape[0]] = local
    allr = all_gather_rows(pad, group, device)
    out = np.zeros((k,) + local.shape[1:], dtype=np.uint64)
    for g in range(world):
        idx = ballot_indices(k, g, world)
        out[idx] = allr[g][: len(idx)]
    return out


# ---------------------------------------------------------------------------------------------------------------------
# one proof over G GPUs: every query sharded by contiguous point range (BASELINE configs[4])
# ---------------------------------------------------------------------------------------------------------------------
PARTIAL_G1 = ((0, 12), (12, 24), (24, 36), (36, 48))  # A | B1 | H | L inside the 72-limb partial vector
PARTIAL_G2 = (48, 72)  # B2


def slice_proving_key(arrays: dict, m: int, rank: int, world: int):
    """(arrays', slice) for rank `rank`: the same proving key with each of A_query, B_query, H_query, L_query cut to
    this rank's contiguous range (constraint system and the fixed elements replicated).  `arrays` follows
    vsp_proving_key_t's field names; m is the evaluation-domain size (H_query has m - 1 points)."""
    cs = arrays["cs"]
    nv, na = int(cs["num_inputs"]) + int(cs["num_aux"]), int(cs["num_aux"])
    nb = int(arrays["B_count"])
    rng = {"A": shard_range(nv + 1, rank, world), "B": shard_range(nb, rank, world),
           "H": shard_range(m - 1, rank, world), "L": shard_range(na, rank, world)}

    def cut(name, lo, hi, width):
        a = np.ascontiguousarray(arrays[name], dtype=np.uint64).reshape(-1, width)[lo:hi]
        return np.ascontiguousarray(a) if hi > lo else np.zeros((1, width), dtype=np.uint64)

    out = dict(arrays)
    out["A_query"] = cut("A_query", *rng["A"], 12)
    out["H_query"] = cut("H_query", *rng["H"], 12)
    out["H_count"] = rng["H"][1] - rng["H"][0]
    out["L_query"] = cut("L_query", *rng["L"], 12)
    lo, hi = rng["B"]
    out["B_count"] = hi - lo
    out["B_g1"] = cut("B_g1", lo, hi, 12)
    out["B_g2"] = cut("B_g2", lo, hi, 24)
    idx = np.ascontiguousarray(arrays["B_idx"], dtype=np.uint32)[lo:hi]
    out["B_idx"] = np.ascontiguousarray(idx) if hi > lo else np.zeros(1, dtype=np.uint32)
    return out, rng


def sum_partials(parts: np.ndarray, sum_g1: Callable[[np.ndarray], np.ndarray],
                 sum_g2: Callable[[np.ndarray], np.ndarray]) -> np.ndarray:
    """component-wise sum of the (world, 72) partial vectors -> 72 limbs"""
    parts = np.ascontiguousarray(parts, dtype=np.uint64).reshape(-1, 72)
    out = np.zeros(72, dtype=np.uint64)
    for lo, hi in PARTIAL_G1:
        out[lo:hi] = sum_g1(np.ascontiguousarray(parts[:, lo:hi]))
    lo, hi = PARTIAL_G2
    out[lo:hi] = sum_g2(np.ascontiguousarray(parts[:, lo:hi]))
    return out


def key_ranges(cs: dict, m: int, n_b: int, rank: int, world: int) -> dict:
    """the contiguous range of each query rank `rank` holds (what slice_proving_key cuts)"""
    nv, na = int(cs["num_inputs"]) + int(cs["num_aux"]), int(cs["num_aux"])
    return {"A": shard_range(nv + 1, rank, world), "B": shard_range(n_b, rank, world),
            "H": shard_range(m - 1, rank, world), "L": shard_range(na, rank, world)}


def vector_owners(world: int):
    """rank that extends aA, aB, aC to the coset in a sharded proof: three different ranks from world 3 up, (0, 1, 0) at 2"""
    if world >= 3:
        return (0, 1, 2)
    return (0, 1 % max(world, 1), 0)


PARTIAL_LIMBS = 144  # raw XYZZ shares of vsp_prove_partial_dev: evA | evB1 | evH | evL (24 limbs each) | evB2 (48)


class ShardedProver:
    """Groth16 prover for one large circuit over all ranks of `group`: this rank loads its slice of the proving key and
    proves its share; the shares meet in ONE all_gather, every rank sums them and assembles the proof.

    prove()      host arrays in and out, canonical affine shares (72 limbs) over the host — the portable form the CPU
                 tests drive over gloo;
    prove_dev()  everything stays on the GPU: vsp_prove_partial_dev leaves 144 limbs of raw XYZZ shares in device memory,
                 one all_gather_into_tensor (NCCL over NVLink) moves world x 1,152 bytes, vsp_prove_assemble_dev sums
                 them and writes the proof to device memory.  No host round trip, nothing synchronises."""

    def __init__(self, vsp, ctx, arrays: dict, group=None, sliced: bool = False, ranges: Optional[dict] = None):
        """`arrays`: the whole proving key (cut here), or with sliced=True this rank's slice already (`ranges` as
        key_ranges gives them) so that no rank ever has to hold the whole key."""
        self.vsp, self.ctx, self.group = vsp, ctx, group
        self.rank, self.world = _world(group)
        cs = arrays["cs"]
        _, self.m = vsp.domain_plan(int(cs["num_constraints"]) + int(cs["num_inputs"]) + 1)
        if sliced:
            self.ranges = ranges
        else:
            arrays, self.ranges = slice_proving_key(arrays, self.m, self.rank, self.world)
        self.pk = vsp.ProvingKey(ctx, arrays, slice=self.ranges)
        self._bufs = None
        self._abc = None
        # share the witness map's three coset extensions between ranks (prove_dev); off: every rank computes all three
        self.split_witness_map = True

    def _global_rank(self, group_rank: int) -> int:
        dist = _dist()
        return dist.get_global_rank(self.group, group_rank) if self.group is not None else group_rank

    def prove_dev(self, d_assignment, r: int, s: int, pubkey=None, r_saver=None):
        """d_assignment: torch int64 CUDA tensor holding (1 || primary || auxiliary) as canonical Fr limbs.  Returns the
        proof as a torch int64 CUDA tensor of 48 limbs (valid in stream order on the context's stream)."""
        import torch
        if self._bufs is None:
            dev = d_assignment.device
            self._bufs = (torch.zeros(PARTIAL_LIMBS, dtype=torch.int64, device=dev),
                          torch.zeros(self.world * PARTIAL_LIMBS, dtype=torch.int64, device=dev),
                          torch.zeros(48, dtype=torch.int64, device=dev))
        part, gathered, proof = self._bufs
        if self.world > 1 and self.split_witness_map:
            # the three coset extensions of the witness map go to (up to) three ranks; each owner broadcasts its vector
            # (m Fr over NCCL / NVLink) while every rank's A / L / B multiexps already run on the side streams
            if self._abc is None:
                self._abc = torch.empty((3, self.m * 4), dtype=torch.int64, device=d_assignment.device)
            owners = vector_owners(self.world)
            mask = sum(1 << v for v in range(3) if owners[v] == self.rank)
            self.pk.prove_partial_begin_dev(d_assignment.data_ptr(), mask, self._abc.data_ptr(), part.data_ptr())
            for v in range(3):
                _dist().broadcast(self._abc[v], src=self._global_rank(owners[v]), group=self.group)
            self.pk.prove_partial_finish_dev(self._abc.data_ptr(), part.data_ptr())
        else:
            self.pk.prove_partial_dev(d_assignment.data_ptr(), part.data_ptr())
        if self.world > 1:
            _dist().all_gather_into_tensor(gathered, part, group=self.group)
        else:
            gathered.copy_(part)
        self.pk.prove_assemble_dev(self.world, gathered.data_ptr(), r, s, proof.data_ptr(), pubkey, r_saver)
        return proof

    def prove(self, primary, auxiliary, r: int, s: int, pubkey=None, r_saver=None) -> np.ndarray:
        part = self.pk.prove_partial(primary, auxiliary)
        parts = all_gather_rows(part, self.group)
        sums = sum_partials(parts, lambda p: self.vsp.points_sum(self.ctx, p), lambda p: self.vsp.points_sum(self.ctx, p))
        return self.pk.prove_assemble(sums, r, s, pubkey, r_saver)

    def close(self):
        self.pk.close()
