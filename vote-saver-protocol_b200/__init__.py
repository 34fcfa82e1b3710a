"""vote-saver-protocol_b200 — host-side mirror of the crypto3 entry points of VoteSAVER's ballot-proof hot path,
running on one B200 through the C ABI of libvsp_b200.so (include/vsp_b200.h).

Reference interfaces mirrored (paths relative to NilFoundation/vote-saver-protocol @ ae4f646c):
  make_evaluation_domain(m) -> domain.fft / inverse_fft / divide_by_z_on_coset
        libs/math/include/nil/crypto3/math/algorithms/make_evaluation_domain.hpp:51-106,
        libs/math/include/nil/crypto3/math/domains/evaluation_domain.hpp:78-118
  multiply_by_coset(a, g)                       libs/math/include/nil/crypto3/math/coset.hpp:37-43
  multiexp(bases, scalars)                      call site libs/containers/include/nil/crypto3/container/sparse_vector.hpp:258-261
  prove(pk, primary, auxiliary, r, s[, pubkey, r_saver])
        zk::prove<r1cs_gg_ppzksnark<..., encrypted_input>>  call site libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:353
  encrypt / rerandomize                         elgamal_verifiable.hpp:322-357, 632-667
  generate(cs, toxic)                           zk::generate<...>  call site bin/cli/include/nil/vote_saver/common.hpp:916-917
  saver_generate_keypair(vk, msg_size, rnd)     generate_keypair_op::process, elgamal_verifiable.hpp:218-271
  pairing_products / verify_batch / verify_encryption_batch
        algebra::pair_reduced (elgamal_verifiable.hpp:415-424), zk::verify, verify_encryption_op::process (:486-490)

Data: numpy uint64 arrays of canonical little-endian limbs — Fr (n, 4); G1 affine (n, 12); G2 affine (n, 24);
the point at infinity is all-zero.  Error behaviour follows the reference: what throws std::invalid_argument there
raises ValueError here.  Everything computes on the GPU; without the built library or a CUDA device the calls raise.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import (DOMAIN_BASIC_RADIX2, DOMAIN_EXTENDED_RADIX2, DOMAIN_STEP_RADIX2, KeypairDesc, PkSliceDesc,
                   ProvingKeyDesc, R1CSDesc, SaverKeypairDesc, SaverPubkeyDesc, VerificationKeyDesc, VspError)

__all__ = ["Context", "EvaluationDomain", "make_evaluation_domain", "multiply_by_coset", "multiexp", "G1Bases",
           "G2Bases", "ProvingKey", "SaverPublicKey", "ResidentSaverKey", "domain_plan", "VspError", "library_path",
           "generate", "saver_generate_keypair", "pairing_products", "verify_batch", "verify_gt_batch", "verify_encryption_batch", "verify_encryption_gt_batch",
           "decrypt", "verify_decryption", "g1_mul_batch", "g1_generator_mul_batch", "g2_generator_mul_batch"]

FR_MULTIPLICATIVE_GENERATOR = 7


def library_path() -> str:
    return _lib.LIB_PATH


def _a(x, cols=None) -> np.ndarray:
    a = np.ascontiguousarray(x, dtype=np.uint64)
    if cols is not None and (a.ndim == 0 or a.shape[-1] != cols):
        a = a.reshape(-1, cols)
    return a


def _p(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)


def _fr(x: int) -> np.ndarray:
    return np.frombuffer(int(x).to_bytes(32, "little"), dtype=np.uint64).copy()


def domain_plan(min_size: int):
    """(kind, m) of make_evaluation_domain<Fr>(min_size); ValueError where the reference throws."""
    kind, m = C.c_int(0), C.c_uint64(0)
    if _lib.load().vsp_domain_plan(int(min_size), C.byref(kind), C.byref(m)) != 0:
        raise ValueError("no radix-2 family evaluation domain for size %d" % min_size)
    return kind.value, m.value


class Context:
    """One GPU + one CUDA stream (vsp_ctx).  `stream`: a raw cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._lib = _lib.load()
        h = C.c_void_p()
        rc = self._lib.vsp_ctx_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h))
        if rc != 0:
            raise VspError(rc, "vsp_ctx_create(device=%d) failed — a CUDA device is required, there is no CPU path" % device)
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.vsp_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def check(self, rc: int):
        if rc == 0:
            return
        text = (self._lib.vsp_last_error(self._h) or b"").decode("utf-8", "replace")
        if rc in (_lib.VSP_E_INVALID, _lib.VSP_E_DOMAIN):
            raise ValueError(text or "invalid argument")
        raise VspError(rc, text)

    def set_batch_chunk(self, ballots: int):
        """ballots per chunk of the batched prover (0 = default); results do not depend on it"""
        self.check(self._lib.vsp_ctx_set_batch_chunk(self._h, int(ballots)))

    @property
    def last_warning(self) -> str:
        return (self._lib.vsp_last_warning(self._h) or b"").decode("utf-8", "replace")

    def synchronize(self):
        self.check(self._lib.vsp_ctx_synchronize(self._h))

    def trim(self):
        """every unused block of the context's scratch cache back to the driver (synchronises the device)"""
        self.check(self._lib.vsp_ctx_trim(self._h))

    def scratch_stats(self) -> dict:
        """{"owned_bytes", "unused_bytes", "driver_allocations"} of the context's scratch cache"""
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self.check(self._lib.vsp_ctx_scratch_stats(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"owned_bytes": a.value, "unused_bytes": b.value, "driver_allocations": c.value}

    @property
    def launch_count(self) -> int:
        return int(self._lib.vsp_ctx_launch_count(self._h))

    KCLASSES = ("msm_accumulate_g1", "msm_accumulate_g2", "msm_sort", "msm_reduce", "ntt", "qap", "fixed_base", "other")

    def profile_enable(self, on: bool = True):
        self.check(self._lib.vsp_ctx_profile_enable(self._h, int(on)))

    def profile_read(self) -> dict:
        """{class: {"launches", "ms", "units"}} of device time per kernel class since profile_enable()"""
        out = {}
        for k, name in enumerate(self.KCLASSES):
            n, ms, u = C.c_uint64(0), C.c_double(0), C.c_double(0)
            self.check(self._lib.vsp_ctx_profile_read(self._h, k, C.byref(n), C.byref(ms), C.byref(u)))
            out[name] = {"launches": n.value, "ms": ms.value, "units": u.value}
        return out

    def accumulate_mads_per_add(self) -> float:
        """32x32->64 multiply-adds the G1 bucket-accumulation kernel issues per bucket addition (a property of the build)"""
        out = C.c_double(0)
        self.check(self._lib.vsp_accumulate_mads_per_add(self._h, C.byref(out)))
        return out.value

    def measure_dfma_rate(self, iters: int = 4096) -> float:
        out = C.c_double(0)
        self.check(self._lib.vsp_measure_dfma_rate(self._h, iters, C.byref(out)))
        return out.value

    def measure_imad_rate(self, iters: int = 4096) -> float:
        out = C.c_double(0)
        self.check(self._lib.vsp_measure_imad_rate(self._h, iters, C.byref(out)))
        return out.value


class EvaluationDomain:
    """evaluation_domain<Fr> of size m as make_evaluation_domain selects it (basic_radix2 or step_radix2)."""

    def __init__(self, ctx: Context, min_size: int):
        self.ctx = ctx
        self.kind, self.m = domain_plan(min_size)
        if self.kind == DOMAIN_EXTENDED_RADIX2:
            raise ValueError("extended_radix2 domain (log m = 33) is not executable")

    def fft(self, a) -> np.ndarray:
        return self._run(a, 0)

    def inverse_fft(self, a) -> np.ndarray:
        return self._run(a, 1)

    def _run(self, a, inverse: int) -> np.ndarray:
        a = _a(a, 4)
        out = np.empty((self.m, 4), dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_fr_fft(self.ctx._h, self.m, inverse, _p(a), a.shape[0], _p(out)))
        return out

    def divide_by_z_on_coset(self, a) -> np.ndarray:
        a = _a(a, 4).copy()
        if a.shape[0] != self.m:
            raise ValueError("expected a.size() == m")
        self.ctx.check(self.ctx._lib.vsp_fr_divide_by_z_on_coset(self.ctx._h, self.m, _p(a)))
        return a

    # device-resident variants (raw device pointers, e.g. tensor.data_ptr()); stream-ordered, no synchronisation
    def fft_dev(self, d_ptr: int, inverse: bool = False):
        self.ctx.check(self.ctx._lib.vsp_fr_fft_dev(self.ctx._h, self.m, int(inverse), C.c_void_p(d_ptr)))

    def fft_noalloc_dev(self, d_ptr: int, inverse: bool = False):
        """fft_dev without any allocation (static device scratch; basic radix-2 domains up to 2^16; no reference
        counterpart, see include/vsp_b200.h).  Calls that overlap on one GPU race on the shared scratch."""
        self.ctx.check(self.ctx._lib.vsp_fr_fft_noalloc_dev(self.ctx._h, self.m, int(inverse), C.c_void_p(d_ptr)))


def make_evaluation_domain(ctx: Context, min_size: int) -> EvaluationDomain:
    return EvaluationDomain(ctx, min_size)


def multiply_by_coset(ctx: Context, a, g: int = FR_MULTIPLICATIVE_GENERATOR) -> np.ndarray:
    a = _a(a, 4).copy()
    gg = _fr(g)
    ctx.check(ctx._lib.vsp_fr_multiply_by_coset(ctx._h, a.shape[0], _p(gg), _p(a)))
    return a


class _Bases:
    _G2 = False

    def __init__(self, ctx: Context, bases=None, *, table: bool = True, _handle=None, _n=0):
        """table=False: the points only — multiexps take the one-shot path with the bases already resident"""
        self.ctx = ctx
        lib = ctx._lib
        if _handle is not None:
            self._h, self.n = _handle, _n
            return
        b = _a(bases, 24 if self._G2 else 12)
        h = C.c_void_p()
        create = lib.vsp_g2_bases_create_ex if self._G2 else lib.vsp_g1_bases_create_ex
        ctx.check(create(ctx._h, b.shape[0], _p(b), int(table), C.byref(h)))
        self._h, self.n = h, b.shape[0]

    def close(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            (self.ctx._lib.vsp_g2_bases_destroy if self._G2 else self.ctx._lib.vsp_g1_bases_destroy)(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def multiexp_dev(self, n_scalars: int, d_scalars: int, d_out: int):
        """scalars and the affine result (12 / 24 limbs) stay on the device; stream-ordered, no synchronisation."""
        fn = self.ctx._lib.vsp_g2_msm_bases_dev if self._G2 else self.ctx._lib.vsp_g1_msm_bases_dev
        self.ctx.check(fn(self.ctx._h, self._h, n_scalars, C.c_void_p(d_scalars), C.c_void_p(d_out)))

    def multiexp(self, scalars) -> np.ndarray:
        s = _a(scalars, 4)
        out = np.zeros(24 if self._G2 else 12, dtype=np.uint64)
        fn = self.ctx._lib.vsp_g2_msm_bases if self._G2 else self.ctx._lib.vsp_g1_msm_bases
        self.ctx.check(fn(self.ctx._h, self._h, s.shape[0], _p(s), 0, _p(out)))
        return out


class G1Bases(_Bases):
    """Device-resident G1 bases (a proving-key query)."""
    _G2 = False

    @classmethod
    def generate(cls, ctx: Context, n: int, seed: int, table: bool = True) -> "G1Bases":
        """Synthetic bases k_i * G with k_i = splitmix64(seed + i) | 1, built on the device.
        table=False keeps the points only: multiexps over them run the one-shot (table-less) path."""
        h = C.c_void_p()
        ctx.check(ctx._lib.vsp_g1_bases_generate_ex(ctx._h, n, seed, int(table), C.byref(h)))
        return cls(ctx, _handle=h, _n=n)

    def read(self, first: int = 0, count: int | None = None) -> np.ndarray:
        count = self.n - first if count is None else count
        out = np.zeros((count, 12), dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_g1_bases_read(self.ctx._h, self._h, first, count, _p(out)))
        return out



class G2Bases(_Bases):
    _G2 = True


def generate_g2_points(ctx: Context, n: int, seed: int) -> np.ndarray:
    """n synthetic G2 points k_i * G2 (k_i = splitmix64(seed + i) | 1) as canonical affine limbs, built on the device"""
    out = np.zeros((n, 24), dtype=np.uint64)
    ctx.check(ctx._lib.vsp_g2_points_generate(ctx._h, n, seed, _p(out)))
    return out


def multiexp(ctx: Context, bases, scalars) -> np.ndarray:
    """algebra::multiexp<BDLO12>(bases, scalars): G1 if bases are (n, 12), G2 if (n, 24).  Host buffers in, affine out."""
    b = np.ascontiguousarray(bases, dtype=np.uint64)
    s = _a(scalars, 4)
    g2 = b.ndim == 2 and b.shape[1] == 24
    b = _a(b, 24 if g2 else 12)
    if b.shape[0] != s.shape[0]:
        raise ValueError("bases and scalars differ in length")
    out = np.zeros(24 if g2 else 12, dtype=np.uint64)
    fn = ctx._lib.vsp_g2_msm if g2 else ctx._lib.vsp_g1_msm
    ctx.check(fn(ctx._h, b.shape[0], _p(b), _p(s), _p(out)))
    return out


def points_sum(ctx: Context, points) -> np.ndarray:
    p = np.ascontiguousarray(points, dtype=np.uint64)
    g2 = p.shape[-1] == 24
    out = np.zeros(24 if g2 else 12, dtype=np.uint64)
    fn = ctx._lib.vsp_g2_sum if g2 else ctx._lib.vsp_g1_sum
    ctx.check(fn(ctx._h, p.shape[0], _p(p), _p(out)))
    return out


def _fill(desc, arr: dict):
    """populate a ctypes descriptor from a dict of numpy arrays / ints with the same field names"""
    for name, ftype in desc._fields_:
        v = arr[name]
        if isinstance(v, dict):
            _fill(getattr(desc, name), v)
        elif isinstance(v, np.ndarray):
            setattr(desc, name, v.ctypes.data_as(ftype))
        else:
            setattr(desc, name, v)
    return desc


class SaverPublicKey:
    """elgamal_verifiable public key (elgamal_verifiable.hpp:152-157) as flat arrays."""

    def __init__(self, arrays: dict):
        self.arrays = arrays  # keeps the numpy buffers alive
        self.desc = _fill(SaverPubkeyDesc(), arrays)
        self.msg_size = int(arrays["msg_size"])
        self._resident = None

    def resident(self, ctx: "Context", gamma_abc_rest) -> "ResidentSaverKey":
        """the key's fixed-base tables on the GPU (built once) for the batched encrypt path"""
        if self._resident is None or self._resident.ctx is not ctx or self._resident._h is None:
            self._resident = ResidentSaverKey(ctx, self, gamma_abc_rest)
        return self._resident


class ResidentSaverKey:
    """vsp_saver_pub: delta, delta_s, gamma_ABC rest, t and delta_sum_s as 8-bit fixed-base tables on the GPU."""

    def __init__(self, ctx: "Context", pubkey: SaverPublicKey, gamma_abc_rest):
        self.ctx, self.pubkey = ctx, pubkey
        g = _a(gamma_abc_rest, 12)
        if g.shape[0] < pubkey.msg_size:
            raise ValueError("gamma_abc_rest shorter than msg_size")
        h = C.c_void_p()
        ctx.check(ctx._lib.vsp_saver_pub_load(ctx._h, C.byref(pubkey.desc), _p(g), C.byref(h)))
        self._h = h
        self.msg_size = pubkey.msg_size

    def close(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            self.ctx._lib.vsp_saver_pub_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ProvingKey:
    """r1cs_gg_ppzksnark proving key resident on the GPU (fast_proving_key.hpp:66-103 field order)."""

    def __init__(self, ctx: Context, arrays: dict, slice: dict | None = None):
        """`slice`: {"A": (lo, hi), "B": ..., "H": ..., "L": ...} when `arrays` holds only a contiguous range of each
        query (point-range sharded proving, sharded.slice_proving_key); such a key serves prove_partial /
        prove_assemble only."""
        self.ctx = ctx
        self.arrays = arrays
        self.slice = slice
        desc = _fill(ProvingKeyDesc(), arrays)
        h = C.c_void_p()
        if slice is None:
            ctx.check(ctx._lib.vsp_pk_load(ctx._h, C.byref(desc), C.byref(h)))
        else:
            sd = PkSliceDesc(*[int(v) for q in "ABHL" for v in slice[q]])
            ctx.check(ctx._lib.vsp_pk_load_slice(ctx._h, C.byref(desc), C.byref(sd), C.byref(h)))
        self._h = h
        self.num_inputs = int(arrays["cs"]["num_inputs"])
        self.num_aux = int(arrays["cs"]["num_aux"])
        self.m = int(ctx._lib.vsp_pk_domain_size(h))

    def close(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            self.ctx._lib.vsp_pk_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def witness_map(self, primary, auxiliary) -> np.ndarray:
        """r1cs_to_qap::witness_map(cs, primary, auxiliary, 0, 0, 0).coefficients_for_H  (m + 1 values)"""
        p, a = _a(primary, 4), _a(auxiliary, 4)
        out = np.zeros((self.m + 1, 4), dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_witness_map(self.ctx._h, self._h, _p(p), _p(a), _p(out)))
        return out

    def prove(self, primary, auxiliary, r: int, s: int, pubkey: SaverPublicKey | None = None,
              r_saver: int | None = None) -> np.ndarray:
        """proof = g_A | g_B | g_C (48 limbs).  pubkey + r_saver select proving_mode::encrypted_input."""
        p, a = _a(primary, 4), _a(auxiliary, 4)
        out = np.zeros(48, dtype=np.uint64)
        rr, ss = _fr(r), _fr(s)
        if pubkey is not None:
            rs = _fr(r_saver)
            gi = pubkey.arrays["gamma_inverse_sum_s_g1"]
            rc = self.ctx._lib.vsp_prove(self.ctx._h, self._h, _p(p), _p(a), _p(rr), _p(ss), _p(rs), _p(gi), _p(out))
        else:
            rc = self.ctx._lib.vsp_prove(self.ctx._h, self._h, _p(p), _p(a), _p(rr), _p(ss), None, None, _p(out))
        self.ctx.check(rc)
        return out

    def prove_partial(self, primary, auxiliary) -> np.ndarray:
        """this key's (slice's) share of the five multiexps: A | B1 | H | L (G1) | B2 (G2) as 72 canonical limbs"""
        p, a = _a(primary, 4), _a(auxiliary, 4)
        out = np.zeros(72, dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_prove_partial(self.ctx._h, self._h, _p(p), _p(a), _p(out)))
        return out

    def prove_assemble(self, sums, r: int, s: int, pubkey: SaverPublicKey | None = None,
                       r_saver: int | None = None) -> np.ndarray:
        """the proof from the five multiexp results summed over all slices (layout of prove_partial)"""
        q = np.ascontiguousarray(sums, dtype=np.uint64).reshape(-1)
        if q.shape[0] != 72:
            raise ValueError("sums must be 72 limbs: A | B1 | H | L | B2")
        out = np.zeros(48, dtype=np.uint64)
        rr, ss = _fr(r), _fr(s)
        if pubkey is not None:
            rs = _fr(r_saver)
            gi = pubkey.arrays["gamma_inverse_sum_s_g1"]
            rc = self.ctx._lib.vsp_prove_assemble(self.ctx._h, self._h, _p(q), _p(rr), _p(ss), _p(rs), _p(gi), _p(out))
        else:
            rc = self.ctx._lib.vsp_prove_assemble(self.ctx._h, self._h, _p(q), _p(rr), _p(ss), None, None, _p(out))
        self.ctx.check(rc)
        return out

    def prove_partial_dev(self, d_assignment: int, d_partials: int):
        """device pointers: (1 || primary || auxiliary) in, 144 limbs of raw XYZZ shares out; stream-ordered"""
        self.ctx.check(self.ctx._lib.vsp_prove_partial_dev(self.ctx._h, self._h, C.c_void_p(d_assignment),
                                                           C.c_void_p(d_partials)))

    def prove_partial_begin_dev(self, d_assignment: int, vector_mask: int, d_abc: int, d_partials: int):
        """first half of prove_partial_dev: multiexps A / L / B started, coset extension of the vectors in vector_mask
        (bit 0 aA, 1 aB, 2 aC) left in d_abc (3 m Fr)"""
        self.ctx.check(self.ctx._lib.vsp_prove_partial_begin_dev(self.ctx._h, self._h, C.c_void_p(d_assignment), int(vector_mask),
                                                                 C.c_void_p(d_abc), C.c_void_p(d_partials)))

    def prove_partial_finish_dev(self, d_abc: int, d_partials: int):
        self.ctx.check(self.ctx._lib.vsp_prove_partial_finish_dev(self.ctx._h, self._h, C.c_void_p(d_abc), C.c_void_p(d_partials)))

    def prove_assemble_dev(self, world: int, d_gathered: int, r: int, s: int, d_proof: int,
                           pubkey: SaverPublicKey | None = None, r_saver: int | None = None):
        """d_gathered: world x 144 limbs (the all_gather of every rank's prove_partial_dev); proof left at d_proof"""
        rr, ss = _fr(r), _fr(s)
        if pubkey is not None:
            rs, gi = _fr(r_saver), pubkey.arrays["gamma_inverse_sum_s_g1"]
            rc = self.ctx._lib.vsp_prove_assemble_dev(self.ctx._h, self._h, world, C.c_void_p(d_gathered), _p(rr), _p(ss),
                                                      _p(rs), _p(gi), C.c_void_p(d_proof))
        else:
            rc = self.ctx._lib.vsp_prove_assemble_dev(self.ctx._h, self._h, world, C.c_void_p(d_gathered), _p(rr), _p(ss),
                                                      None, None, C.c_void_p(d_proof))
        self.ctx.check(rc)

    def prove_batch(self, primary, auxiliary, rs, pubkey: SaverPublicKey | None = None) -> np.ndarray:
        """k ballots: primary (k, ni, 4), auxiliary (k, na, 4), rs (k, 3, 4) = (r, s, r_saver).  -> (k, 48)"""
        p = np.ascontiguousarray(primary, dtype=np.uint64)
        a = np.ascontiguousarray(auxiliary, dtype=np.uint64)
        q = np.ascontiguousarray(rs, dtype=np.uint64)
        k = q.shape[0]
        out = np.zeros((k, 48), dtype=np.uint64)
        gi = _p(pubkey.arrays["gamma_inverse_sum_s_g1"]) if pubkey is not None else None
        self.ctx.check(self.ctx._lib.vsp_prove_batch(self.ctx._h, self._h, k, _p(p), _p(a), _p(q), gi, _p(out)))
        return out

    def prove_batch_dev(self, k: int, d_assignments: int, rs, d_proofs: int, pubkey: SaverPublicKey | None = None):
        """device-resident ballots: d_assignments -> k x (1 || primary || auxiliary) canonical Fr; d_proofs -> k x 48
        limbs.  Raw device pointers; stream-ordered, no synchronisation."""
        q = np.ascontiguousarray(rs, dtype=np.uint64)
        gi = _p(pubkey.arrays["gamma_inverse_sum_s_g1"]) if pubkey is not None else None
        self.ctx.check(self.ctx._lib.vsp_prove_batch_dev(self.ctx._h, self._h, k, C.c_void_p(d_assignments), _p(q), gi,
                                                         C.c_void_p(d_proofs)))

    def encrypt_batch(self, key: "ResidentSaverKey", primary, auxiliary, rs):
        """k ballots through encrypt_op::process: the message is the first msg_size primary inputs of each ballot.
        primary (k, ni, 4), auxiliary (k, na, 4), rs (k, 3, 4) = (r, s, r_saver) -> (cts (k, msg_size + 2, 12), proofs (k, 48))"""
        p = np.ascontiguousarray(primary, dtype=np.uint64)
        a = np.ascontiguousarray(auxiliary, dtype=np.uint64)
        q = np.ascontiguousarray(rs, dtype=np.uint64)
        k = q.shape[0]
        cts = np.zeros((k, key.msg_size + 2, 12), dtype=np.uint64)
        proofs = np.zeros((k, 48), dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_saver_encrypt_batch(self.ctx._h, self._h, key._h, k, _p(p), _p(a), _p(q), _p(cts),
                                                             _p(proofs)))
        return cts, proofs

    def encrypt_batch_dev(self, key: "ResidentSaverKey", k: int, d_assignments: int, rs, d_cts: int, d_proofs: int):
        """device-resident ballots (k x (1 || primary || auxiliary)); ciphertexts and proofs stay on the device"""
        q = np.ascontiguousarray(rs, dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_saver_encrypt_batch_dev(self.ctx._h, self._h, key._h, k, C.c_void_p(d_assignments),
                                                                 _p(q), C.c_void_p(d_cts), C.c_void_p(d_proofs)))

    def rerandomize_batch(self, key: "ResidentSaverKey", rnd, cts, proofs):
        """rerandomize_op::process for k ballots: rnd (k, 3, 4) = (r, z1, z2), cts (k, msg_size + 2, 12), proofs (k, 48)
        -> (cts', proofs')"""
        q = np.ascontiguousarray(rnd, dtype=np.uint64)
        k = q.shape[0]
        c = np.ascontiguousarray(cts, dtype=np.uint64).reshape(k, key.msg_size + 2, 12)
        pr = np.ascontiguousarray(proofs, dtype=np.uint64).reshape(k, 48)
        co, po = np.zeros_like(c), np.zeros_like(pr)
        self.ctx.check(self.ctx._lib.vsp_saver_rerandomize_batch(self.ctx._h, self._h, key._h, k, _p(q), _p(c), _p(pr), _p(co),
                                                                 _p(po)))
        return co, po

    def rerandomize_batch_dev(self, key: "ResidentSaverKey", k: int, rnd, d_cts: int, d_proofs: int, d_cts_out: int | None = None,
                              d_proofs_out: int | None = None):
        """the same on device-resident cipher texts and proofs (in place unless output pointers are given)"""
        q = np.ascontiguousarray(rnd, dtype=np.uint64)
        self.ctx.check(self.ctx._lib.vsp_saver_rerandomize_batch_dev(
            self.ctx._h, self._h, key._h, k, _p(q), C.c_void_p(d_cts), C.c_void_p(d_proofs),
            C.c_void_p(d_cts_out if d_cts_out is not None else d_cts),
            C.c_void_p(d_proofs_out if d_proofs_out is not None else d_proofs)))

    def encrypt(self, pubkey: SaverPublicKey, gamma_abc_rest, msg, primary, auxiliary, r_saver: int, r: int, s: int):
        """encrypt_op<elgamal_verifiable>::process -> (cipher_text (msg_size + 2, 12), proof (48,))"""
        g, mm, p, a = _a(gamma_abc_rest, 12), _a(msg, 4), _a(primary, 4), _a(auxiliary, 4)
        ct = np.zeros((pubkey.msg_size + 2, 12), dtype=np.uint64)
        proof = np.zeros(48, dtype=np.uint64)
        rs, rr, ss = _fr(r_saver), _fr(r), _fr(s)
        self.ctx.check(self.ctx._lib.vsp_saver_encrypt(self.ctx._h, self._h, C.byref(pubkey.desc), _p(g), _p(mm), _p(p),
                                                       _p(a), _p(rs), _p(rr), _p(ss), _p(ct), _p(proof)))
        return ct, proof


def saver_ciphertext(ctx: Context, pubkey: SaverPublicKey, gamma_abc_rest, msg, r_saver: int) -> np.ndarray:
    g, mm = _a(gamma_abc_rest, 12), _a(msg, 4)
    ct = np.zeros((pubkey.msg_size + 2, 12), dtype=np.uint64)
    rs = _fr(r_saver)
    ctx.check(ctx._lib.vsp_saver_ciphertext(ctx._h, C.byref(pubkey.desc), _p(g), _p(mm), _p(rs), _p(ct)))
    return ct


def rerandomize(ctx: Context, pubkey: SaverPublicKey, delta_g2, rnd, ct, proof):
    """rerandomize_op<elgamal_verifiable>::process; rnd = (r, z1, z2) -> (cipher_text', proof')"""
    d2, c, pr = _a(delta_g2), _a(ct, 12), _a(proof)
    q = np.concatenate([_fr(x) for x in rnd])
    ct_out = np.zeros_like(c)
    proof_out = np.zeros(48, dtype=np.uint64)
    ctx.check(ctx._lib.vsp_saver_rerandomize(ctx._h, C.byref(pubkey.desc), _p(d2), _p(q), _p(c), _p(pr), _p(ct_out),
                                             _p(proof_out)))
    return ct_out, proof_out


# ---- key generation --------------------------------------------------------------
def g1_mul_batch(ctx: Context, points, scalars) -> np.ndarray:
    """out[i] = scalars[i] * points[i] (variable base)"""
    p, sc = _a(points, 12), _a(scalars, 4)
    if p.shape[0] != sc.shape[0]:
        raise ValueError("points / scalars length mismatch")
    out = np.zeros((p.shape[0], 12), dtype=np.uint64)
    ctx.check(ctx._lib.vsp_g1_mul_batch(ctx._h, p.shape[0], _p(p), _p(sc), _p(out)))
    return out


def g1_generator_mul_batch(ctx: Context, scalars) -> np.ndarray:
    """out[i] = scalars[i] * G1::one()"""
    sc = _a(scalars, 4)
    out = np.zeros((sc.shape[0], 12), dtype=np.uint64)
    ctx.check(ctx._lib.vsp_g1_generator_mul_batch(ctx._h, sc.shape[0], _p(sc), _p(out)))
    return out


def g2_generator_mul_batch(ctx: Context, scalars) -> np.ndarray:
    """out[i] = scalars[i] * G2::one()"""
    sc = _a(scalars, 4)
    out = np.zeros((sc.shape[0], 24), dtype=np.uint64)
    ctx.check(ctx._lib.vsp_g2_generator_mul_batch(ctx._h, sc.shape[0], _p(sc), _p(out)))
    return out


def generate(ctx: Context, cs: dict, toxic):
    """zk::generate<r1cs_gg_ppzksnark<..., encrypted_input>>(constraint_system) with the generator's randomness
    toxic = (t, alpha, beta, gamma, delta) passed in.  `cs`: the R1CS arrays dict (vsp_r1cs_t field names).
    Returns (pk_arrays, vk_arrays): pk_arrays goes straight into ProvingKey(ctx, pk_arrays); vk_arrays holds
    alpha_g1, beta_g2, gamma_g2, delta_g2, gamma_g1, delta_g1, gamma_ABC_g1 (first | rest)."""
    ni, na, nc = int(cs["num_inputs"]), int(cs["num_aux"]), int(cs["num_constraints"])
    nvp1 = ni + na + 1
    _kind, m = domain_plan(nc + ni + 1)
    tox = np.concatenate([_fr(int(x)) for x in toxic])
    if tox.shape[0] != 20:
        raise ValueError("toxic = (t, alpha, beta, gamma, delta)")
    z1, z2 = (lambda n: np.zeros((max(n, 1), 12), np.uint64)), (lambda n: np.zeros((max(n, 1), 24), np.uint64))
    arr = {"alpha_g1": z1(1), "beta_g1": z1(1), "beta_g2": z2(1), "delta_g1": z1(1), "delta_g2": z2(1),
           "A_query": z1(nvp1), "B_count": 0, "B_idx": np.zeros(nvp1, np.uint32), "B_g2": z2(nvp1), "B_g1": z1(nvp1),
           "H_count": 0, "H_query": z1(m - 1), "L_query": z1(na),
           "gamma_g2": z2(1), "gamma_g1": z1(1), "gamma_ABC_g1": z1(ni + 1)}
    desc = _fill(KeypairDesc(), arr)
    csd = _fill(R1CSDesc(), cs)
    ctx.check(ctx._lib.vsp_groth16_generate(ctx._h, C.byref(csd), _p(tox), C.byref(desc)))
    nb = int(desc.B_count)
    pk = {"cs": cs, "alpha_g1": arr["alpha_g1"], "beta_g1": arr["beta_g1"], "beta_g2": arr["beta_g2"],
          "delta_g1": arr["delta_g1"], "delta_g2": arr["delta_g2"], "A_query": arr["A_query"], "B_count": nb,
          "B_idx": np.ascontiguousarray(arr["B_idx"][:max(nb, 1)]), "B_g2": np.ascontiguousarray(arr["B_g2"][:max(nb, 1)]),
          "B_g1": np.ascontiguousarray(arr["B_g1"][:max(nb, 1)]), "H_count": int(desc.H_count), "H_query": arr["H_query"],
          "L_query": arr["L_query"]}
    vk = {"alpha_g1": arr["alpha_g1"], "beta_g2": arr["beta_g2"], "gamma_g2": arr["gamma_g2"], "delta_g2": arr["delta_g2"],
          "gamma_g1": arr["gamma_g1"], "delta_g1": arr["delta_g1"], "gamma_ABC_g1": arr["gamma_ABC_g1"]}
    return pk, vk


def saver_generate_keypair(ctx: Context, vk: dict, msg_size: int, rnd):
    """generate_keypair_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:218-271): rnd = 3 msg_size + 2 scalars
    (rho, t_0, then s_i, v_i, t_i).  Returns (public-key arrays for SaverPublicKey, rho, verification-key arrays)."""
    n = int(msg_size)
    rnd = [int(x) for x in rnd]
    if len(rnd) < 3 * n + 2:
        raise ValueError("Too few numbers in the source of randomness.")  # the reference's assertion text (:221)
    r = np.concatenate([_fr(x) for x in rnd[:3 * n + 2]])
    gabc = _a(vk["gamma_ABC_g1"], 12)
    if gabc.shape[0] < n + 1:
        raise ValueError("gamma_ABC_g1 shorter than msg_size + 1")
    rest = np.ascontiguousarray(gabc[1:])
    z1, z2 = (lambda k: np.zeros((max(k, 1), 12), np.uint64)), (lambda k: np.zeros((max(k, 1), 24), np.uint64))
    arr = {"delta_s_g1": z1(n), "t_g1": z1(n), "t_g2": z2(n + 1), "delta_sum_s_g1": z1(1), "gamma_inverse_sum_s_g1": z1(1),
           "rho_g2": z2(1), "rho_sv_g2": z2(n), "rho_rhov_g2": z2(n), "rho": np.zeros(4, np.uint64)}
    desc = _fill(SaverKeypairDesc(), arr)
    ctx.check(ctx._lib.vsp_saver_keygen(ctx._h, n, _p(_a(vk["delta_g1"], 12)), _p(_a(vk["gamma_g1"], 12)), _p(rest), _p(r),
                                        C.byref(desc)))
    pub = {"msg_size": n, "delta_g1": _a(vk["delta_g1"], 12).copy(), "delta_s_g1": arr["delta_s_g1"], "t_g1": arr["t_g1"],
           "t_g2": arr["t_g2"], "delta_sum_s_g1": arr["delta_sum_s_g1"],
           "gamma_inverse_sum_s_g1": arr["gamma_inverse_sum_s_g1"]}
    svk = {"rho_g2": arr["rho_g2"], "rho_sv_g2": arr["rho_sv_g2"], "rho_rhov_g2": arr["rho_rhov_g2"]}
    rho = int.from_bytes(arr["rho"].tobytes(), "little")
    return pub, rho, svk


# ---- pairing side ----------------------------------------------------------------
def pairing_products(ctx: Context, g1, g2, pairs: int = 1):
    """prod_j e(g1[c][j], g2[c][j]) for every check c: returns (gt, is_one) with gt (n_checks, 72) canonical Fq12 limbs in
    tower order and is_one (n_checks,) bool.  algebra::pair_reduced, batched."""
    p1, p2 = _a(g1, 12), _a(g2, 24)
    if p1.shape[0] != p2.shape[0] or pairs < 1 or p1.shape[0] % pairs:
        raise ValueError("g1 / g2 must hold n_checks * pairs points each")
    n = p1.shape[0] // pairs
    gt = np.zeros((n, 72), dtype=np.uint64)
    one = np.zeros(n, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_pairing_products(ctx._h, n, pairs, _p(p1), _p(p2), _p(gt), _p(one)))
    return gt, one.astype(bool)


def _vk_desc(vk: dict):
    keep = {k: _a(vk[k], 24 if k.endswith("g2") else 12) for k in ("alpha_g1", "beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
    keep["gamma_ABC_count"] = keep["gamma_ABC_g1"].shape[0]
    return _fill(VerificationKeyDesc(), keep), keep


def verify_batch(ctx: Context, vk: dict, primary, proofs) -> np.ndarray:
    """zk::verify<r1cs_gg_ppzksnark<...>>(vk, primary_input, proof) for k proofs (basic mode): bool per proof.
    primary: (k, n_primary, 4); proofs: (k, 48)."""
    desc, keep = _vk_desc(vk)
    pr = _a(proofs, 48)
    k = pr.shape[0]
    x = np.ascontiguousarray(primary, dtype=np.uint64).reshape(k, -1, 4) if k else np.zeros((0, 0, 4), np.uint64)
    ok = np.zeros(k, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_groth16_verify_batch(ctx._h, C.byref(desc), k, _p(x), x.shape[1], _p(pr), _p(ok)))
    return ok.astype(bool)


def verify_gt_batch(ctx: Context, alpha_g1_beta_g2, gamma_g2, delta_g2, gamma_ABC_g1, primary, proofs) -> np.ndarray:
    """zk::verify with the verification key as the reference stores it (alpha_g1_beta_g2 as a GT value, 72 limbs)."""
    gt, g2, d2 = _a(alpha_g1_beta_g2).reshape(-1), _a(gamma_g2).reshape(-1, 24), _a(delta_g2).reshape(-1, 24)
    gabc = _a(gamma_ABC_g1).reshape(-1, 12)
    if gt.shape[0] != 72:
        raise ValueError("alpha_g1_beta_g2 must be 72 limbs")
    pr = _a(proofs, 48)
    k = pr.shape[0]
    x = np.ascontiguousarray(primary, dtype=np.uint64).reshape(k, -1, 4) if k else np.zeros((0, 0, 4), np.uint64)
    ok = np.zeros(k, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_groth16_verify_gt_batch(ctx._h, _p(gt), _p(g2), _p(d2), gabc.shape[0], _p(gabc), k, _p(x), x.shape[1],
                                                   _p(pr), _p(ok)))
    return ok.astype(bool)


def verify_encryption_batch(ctx: Context, vk: dict, pubkey: SaverPublicKey, cts, unencrypted_primary, proofs) -> np.ndarray:
    """verify_encryption_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:486-490) for k ballots: bool per ballot.
    cts: (k, msg_size + 2, 12); unencrypted_primary: (k, n_rest, 4) — the primary input after the message slots."""
    desc, keep = _vk_desc(vk)
    pr = _a(proofs, 48)
    k = pr.shape[0]
    ct = np.ascontiguousarray(cts, dtype=np.uint64).reshape(k, pubkey.msg_size + 2, 12)
    x = np.ascontiguousarray(unencrypted_primary, dtype=np.uint64).reshape(k, -1, 4)
    ok = np.zeros(k, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_saver_verify_encryption_batch(ctx._h, C.byref(desc), C.byref(pubkey.desc), k, _p(ct), _p(x),
                                                         x.shape[1], _p(pr), _p(ok)))
    return ok.astype(bool)


def verify_encryption_gt_batch(ctx: Context, vk: dict, pubkey: SaverPublicKey, cts, unencrypted_primary, proofs) -> np.ndarray:
    """verify_encryption_batch with the verification key as codec.groth16_vk_from_bytes delivers it (alpha_g1_beta_g2 as a GT
    value): bool per ballot"""
    gt, g2, d2 = _a(vk["alpha_g1_beta_g2"]).reshape(-1), _a(vk["gamma_g2"]).reshape(-1, 24), _a(vk["delta_g2"]).reshape(-1, 24)
    gabc = _a(vk["gamma_ABC_g1"]).reshape(-1, 12)
    if gt.shape[0] != 72:
        raise ValueError("alpha_g1_beta_g2 must be 72 limbs")
    pr = _a(proofs, 48)
    k = pr.shape[0]
    ct = np.ascontiguousarray(cts, dtype=np.uint64).reshape(k, pubkey.msg_size + 2, 12)
    x = np.ascontiguousarray(unencrypted_primary, dtype=np.uint64).reshape(k, -1, 4)
    ok = np.zeros(k, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_saver_verify_encryption_gt_batch(ctx._h, _p(gt), _p(g2), _p(d2), gabc.shape[0], _p(gabc),
                                                            C.byref(pubkey.desc), k, _p(ct), _p(x), x.shape[1], _p(pr), _p(ok)))
    return ok.astype(bool)


def decrypt(ctx: Context, ct, rho: int, svk: dict, gamma_abc_rest, block_bits: int = 4):
    """decrypt_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:401-441): (message ints, verify_c0 point).
    svk: the SAVER verification-key arrays (rho_g2, rho_sv_g2, rho_rhov_g2).  ValueError("Decryption failed.") where the
    reference asserts."""
    c = _a(ct, 12)
    n = c.shape[0] - 2
    sv, rv, g = _a(svk["rho_sv_g2"], 24), _a(svk["rho_rhov_g2"], 24), _a(gamma_abc_rest, 12)
    if n < 0 or sv.shape[0] < n or rv.shape[0] < n or g.shape[0] < n:
        raise ValueError("Cipher text size should match the verification key (exclusive of 2 elements in CT).")
    m = np.zeros((max(n, 1), 4), dtype=np.uint64)
    c0 = np.zeros(12, dtype=np.uint64)
    ctx.check(ctx._lib.vsp_saver_decrypt(ctx._h, n, _p(c), _p(_fr(rho)), _p(sv), _p(rv), _p(g), int(block_bits), _p(m), _p(c0)))
    return [int(x) for x in m[:n, 0]], c0


def verify_decryption(ctx: Context, ct, plain, svk: dict, gamma_abc_rest, proof) -> bool:
    """verify_decryption_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:548-575); proof = decrypt's verify_c0."""
    c = _a(ct, 12)
    n = c.shape[0] - 2
    pl = np.concatenate([_fr(int(x)) for x in plain]).reshape(-1, 4) if len(plain) else np.zeros((1, 4), np.uint64)
    if len(plain) != n:
        raise ValueError("Cipher text size should be equal to the plain text size (exclusive of 2 element in CT).")
    ok = np.zeros(1, dtype=np.uint8)
    ctx.check(ctx._lib.vsp_saver_verify_decryption(ctx._h, n, _p(c), _p(pl), _p(_a(svk["rho_g2"], 24)), _p(_a(svk["rho_sv_g2"], 24)),
                                                   _p(_a(svk["rho_rhov_g2"], 24)), _p(_a(gamma_abc_rest, 12)), _p(_a(proof, 12)),
                                                   _p(ok)))
    return bool(ok[0])
