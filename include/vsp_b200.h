/* vsp_b200.h — C ABI of the B200-native VoteSAVER ballot-proof hot path (libvsp_b200.so).
 *
 * One shared library, plain pointers and sizes, no C++ or torch types.  Every entry point replaces one
 * interface of the reference (paths relative to the reference tree NilFoundation/vote-saver-protocol @ ae4f646c):
 *
 *   vsp_domain_plan / vsp_fr_fft / vsp_fr_multiply_by_coset / vsp_fr_divide_by_z_on_coset
 *       math::make_evaluation_domain<Fr>(m)           libs/math/include/nil/crypto3/math/algorithms/make_evaluation_domain.hpp:51-106
 *       evaluation_domain<Fr>::fft / inverse_fft      libs/math/include/nil/crypto3/math/domains/evaluation_domain.hpp:78,83
 *                                                     (basic_radix2_domain.hpp:69-96, step_radix2_domain.hpp:73-166)
 *       math::multiply_by_coset                       libs/math/include/nil/crypto3/math/coset.hpp:37-43
 *       evaluation_domain<Fr>::divide_by_z_on_coset   basic_radix2_domain.hpp:118-124, step_radix2_domain.hpp:222-246
 *   vsp_g1_msm / vsp_g2_msm / vsp_g*_bases_*
 *       algebra::multiexp<multiexp_method_BDLO12>(bases_begin, bases_end, scalars_begin, scalars_end, chunks)
 *                                                     call site libs/containers/include/nil/crypto3/container/sparse_vector.hpp:258-261
 *   vsp_pk_load / vsp_witness_map / vsp_prove
 *       zk::prove<r1cs_gg_ppzksnark<bls12<381>, ..., proving_mode::encrypted_input>>(pk, pubkey, primary, auxiliary, r)
 *                                                     call site libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:353-354
 *       proving-key field order                       libs/marshalling/zk/include/nil/crypto3/marshalling/zk/types/r1cs_gg_ppzksnark/fast_proving_key.hpp:66-103
 *       R1CS shape                                    .../types/r1cs_gg_ppzksnark/r1cs.hpp:57-113
 *   vsp_saver_encrypt / vsp_saver_rerandomize
 *       encrypt_op<elgamal_verifiable<Curve>>::process       elgamal_verifiable.hpp:322-357
 *       rerandomize_op<elgamal_verifiable<Curve>>::process   elgamal_verifiable.hpp:632-667
 *
 * Conventions
 *   - Every function returns 0 on success and a negative VSP_E_* code on failure; it never aborts and never throws.
 *     vsp_last_error(ctx) returns the text of the last failure on that context.
 *   - There is no CPU fallback: without a usable CUDA device vsp_ctx_create fails with VSP_E_NO_DEVICE.
 *   - A context is bound to one GPU and one CUDA stream and is not thread-safe; distinct contexts are independent.
 *   - Numbers are canonical (non-Montgomery) little-endian 64-bit limbs:  Fr = 4 limbs (< r), Fq = 6 limbs (< p).
 *     G1 affine = x | y (12 limbs); G2 affine = x.c0 | x.c1 | y.c0 | y.c1 (24 limbs); the point at infinity is
 *     all-zero (mirrors fast_curve_element's X | Y | is_infinity, .../marshalling/algebra/types/fast_curve_element.hpp:56-97).
 *     Outputs are canonical affine, so they compare bit for bit with the reference's to_affine() results.
 *   - Pointers are host pointers unless the parameter name starts with d_ (device pointer on the context's GPU).
 *   - Groth16's blinding scalars r, s are explicit inputs (the reference draws them inside zk::prove).
 */
#ifndef VSP_B200_H
#define VSP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* exported from libvsp_b200.so (the library is built with -fvisibility=hidden) */
#if defined(__GNUC__)
#define VSP_API __attribute__((visibility("default")))
#else
#define VSP_API
#endif

#define VSP_OK 0
#define VSP_E_INVALID (-1)    /* bad argument (the reference throws std::invalid_argument / asserts) */
#define VSP_E_NO_DEVICE (-2)  /* no CUDA device / context creation failed */
#define VSP_E_CUDA (-3)       /* CUDA runtime error, text in vsp_last_error */
#define VSP_E_DOMAIN (-4)     /* no radix-2 family evaluation domain of that size */
#define VSP_E_NOMEM (-5)

#define VSP_DOMAIN_BASIC_RADIX2 1
#define VSP_DOMAIN_EXTENDED_RADIX2 2 /* needs log m = 33 for Fr: reported by vsp_domain_plan, not executable */
#define VSP_DOMAIN_STEP_RADIX2 3

typedef struct vsp_ctx vsp_ctx;
typedef struct vsp_g1_bases vsp_g1_bases;
typedef struct vsp_g2_bases vsp_g2_bases;
typedef struct vsp_pk vsp_pk;
typedef struct vsp_saver_pub vsp_saver_pub;

/* ---- context ------------------------------------------------------------------------------ */
/* cuda_stream: an existing CUstream / cudaStream_t to launch on (e.g. torch's current stream), or NULL for a
 * private non-blocking stream. */
VSP_API int vsp_ctx_create(int device, void *cuda_stream, vsp_ctx **out);
VSP_API void vsp_ctx_destroy(vsp_ctx *ctx);
VSP_API const char *vsp_last_error(const vsp_ctx *ctx);
/* text of the last non-fatal degradation on this context ("" if none): today only "a window table did not fit in
 * device memory, multiexps over these bases run table-less" (same results, slower for small queries) */
VSP_API const char *vsp_last_warning(const vsp_ctx *ctx);
/* ballots per chunk of the batched entry points (vsp_prove_batch*, vsp_saver_encrypt_batch*): 0 restores the default
 * (environment VSP_BATCH_CHUNK, else 256).  The library further clamps it by the circuit's domain size m to
 * max(1, 2^23 / m), so a batch of any size over any circuit runs in bounded scratch; inside a chunk a multiexp whose
 * ballots x points x windows would pass 2^32 keys is itself run in sub-batches.  Results do not depend on it. */
VSP_API int vsp_ctx_set_batch_chunk(vsp_ctx *ctx, int ballots);
VSP_API int vsp_ctx_synchronize(vsp_ctx *ctx);
/* Scratch memory.  Every temporary device buffer of a call comes from a cache of device blocks the context owns
 * (stream-ordered like cudaMallocAsync, but dependent only on the order of the host's calls): the first run of a
 * workload cudaMallocs its blocks, every later run of the same workload makes no driver allocation at all (the
 * reference allocates its std::vectors per call, e.g. r1cs_gg_ppzksnark/prover.hpp:117-160; this is their device-side
 * stand-in).  Unused blocks go back to the driver on vsp_ctx_trim, when more than VSP_SCRATCH_CACHE_GB (environment,
 * default 64) sit unused at the moment a new block is needed, when a cudaMalloc fails, and on vsp_ctx_destroy.
 * vsp_ctx_trim synchronises the device.  vsp_ctx_scratch_stats: bytes the cache owns, bytes of them not in use, and
 * the number of cudaMalloc calls it has made so far (any pointer may be NULL). */
VSP_API int vsp_ctx_trim(vsp_ctx *ctx);
VSP_API int vsp_ctx_scratch_stats(const vsp_ctx *ctx, uint64_t *owned_bytes, uint64_t *unused_bytes,
                                  uint64_t *driver_allocations);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
VSP_API uint64_t vsp_ctx_launch_count(const vsp_ctx *ctx);
VSP_API const char *vsp_version(void);
/* live per-kernel-class device timing (CUDA events around every launch, on the launching stream).  Classes:
 * 0 G1 bucket accumulation, 1 G2 bucket accumulation, 2 multiexp sort (digits/scan/scatter), 3 multiexp reduction,
 * 4 NTT, 5 QAP (R1CS evaluation, pointwise), 6 fixed-base multiples + proof assembly, 7 other.
 * enable(on) synchronises, clears the counters and switches recording; read() returns launches, summed device ms and
 * summed algorithmic work units (32-bit multiply-adds for classes 0/1) since enable. */
#define VSP_KCLASS_COUNT 8
VSP_API int vsp_ctx_profile_enable(vsp_ctx *ctx, int on);
VSP_API int vsp_ctx_profile_read(vsp_ctx *ctx, int kclass, uint64_t *launches, double *total_ms, double *units);

/* ---- evaluation domains over Fr ------------------------------------------------------------ */
VSP_API int vsp_domain_plan(uint64_t min_size, int *kind, uint64_t *m);
/* fft / inverse_fft of the domain of size m (m as returned by vsp_domain_plan).  n_in <= m values are read; the
 * basic domain zero-pads short input, the step domain's inverse_fft rejects it (VSP_E_INVALID) as the reference does. */
VSP_API int vsp_fr_fft(vsp_ctx *ctx, uint64_t m, int inverse, const uint64_t *in, uint64_t n_in, uint64_t *out);
VSP_API int vsp_fr_fft_dev(vsp_ctx *ctx, uint64_t m, int inverse, uint64_t *d_inout);
/* vsp_fr_fft_dev for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Basic radix-2 domains only, 2 <= m <= VSP_NOALLOC_FFT_MAX_M = 2^16, m a power of two;
 * anything else is rejected (VSP_E_DOMAIN for a size that is not a power of two, VSP_E_INVALID beyond the maximum,
 * message in vsp_last_error).  Same kernels and bit-identical results as vsp_fr_fft_dev, stream-ordered on the
 * context's stream, but the transform's ping-pong buffer and twiddle table are two file-scope __device__ arrays
 * (2 MiB + 1 MiB, held by every GPU that loads the library until exit), so the call makes no cudaMalloc / cudaFree /
 * cudaMallocAsync and calls no allocating library routine.  The first call on each device resolves the arrays'
 * addresses, fills the twiddle table and synchronises the context's stream once: make it outside graph capture.
 * Every call on a GPU shares the one scratch array, whatever context or stream it comes from: calls that overlap on
 * the GPU race; calls on one stream are ordered.  Nothing in the scratch needs clearing between calls. */
#define VSP_NOALLOC_FFT_MAX_M (1ull << 16)
VSP_API int vsp_fr_fft_noalloc_dev(vsp_ctx *ctx, uint64_t m, int inverse, uint64_t *d_inout);
VSP_API int vsp_fr_multiply_by_coset(vsp_ctx *ctx, uint64_t n, const uint64_t g[4], uint64_t *a_inout);
VSP_API int vsp_fr_multiply_by_coset_dev(vsp_ctx *ctx, uint64_t n, const uint64_t g[4], uint64_t *d_inout);
VSP_API int vsp_fr_divide_by_z_on_coset(vsp_ctx *ctx, uint64_t m, uint64_t *a_inout);
VSP_API int vsp_fr_divide_by_z_on_coset_dev(vsp_ctx *ctx, uint64_t m, uint64_t *d_inout);

/* ---- multiexp ------------------------------------------------------------------------------- */
/* one-shot: sum_i scalars[i] * bases[i] */
VSP_API int vsp_g1_msm(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, const uint64_t *scalars, uint64_t *out);
VSP_API int vsp_g2_msm(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, const uint64_t *scalars, uint64_t *out);
/* resident bases (a proving-key query): upload once, run many scalar vectors */
VSP_API int vsp_g1_bases_create(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, vsp_g1_bases **out);
VSP_API void vsp_g1_bases_destroy(vsp_g1_bases *b);
VSP_API int vsp_g2_bases_create(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, vsp_g2_bases **out);
VSP_API void vsp_g2_bases_destroy(vsp_g2_bases *b);
/* the same with the choice of building the window table (table = 0: the points only, multiexps over the handle take the
 * one-shot path with the bases already resident — what algebra::multiexp costs without the upload) */
VSP_API int vsp_g1_bases_create_ex(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, int table, vsp_g1_bases **out);
VSP_API int vsp_g2_bases_create_ex(vsp_ctx *ctx, uint64_t n, const uint64_t *bases, int table, vsp_g2_bases **out);
/* synthetic bases for the measurement sweeps, built on the device:
 * bases[i] = k_i * G1_generator, k_i = splitmix64(seed + i) | 1.  vsp_g1_bases_read copies a range back as canonical
 * affine points so small cases can be checked against the oracle. */
VSP_API int vsp_g1_bases_generate(vsp_ctx *ctx, uint64_t n, uint64_t seed, vsp_g1_bases **out);
/* the same with the choice of building the window table (table = 0: points only; multiexps over the handle then take the
 * one-shot path of vsp_g1_msm with the bases already resident) */
VSP_API int vsp_g1_bases_generate_ex(vsp_ctx *ctx, uint64_t n, uint64_t seed, int table, vsp_g1_bases **out);
VSP_API int vsp_g1_bases_read(vsp_ctx *ctx, const vsp_g1_bases *b, uint64_t first, uint64_t count, uint64_t *out);
/* the same over G2 (k_i * G2_generator), written straight to a host array of n canonical affine points */
VSP_API int vsp_g2_points_generate(vsp_ctx *ctx, uint64_t n, uint64_t seed, uint64_t *out);
/* scalars: n_scalars <= n canonical Fr (host, or device when scalars_on_device != 0); out: host affine point */
VSP_API int vsp_g1_msm_bases(vsp_ctx *ctx, const vsp_g1_bases *b, uint64_t n_scalars, const uint64_t *scalars,
                     int scalars_on_device, uint64_t *out);
VSP_API int vsp_g2_msm_bases(vsp_ctx *ctx, const vsp_g2_bases *b, uint64_t n_scalars, const uint64_t *scalars,
                     int scalars_on_device, uint64_t *out);
/* same, result left on the device as a canonical affine point (12 / 24 limbs) — no host synchronisation */
VSP_API int vsp_g1_msm_bases_dev(vsp_ctx *ctx, const vsp_g1_bases *b, uint64_t n_scalars, const uint64_t *d_scalars,
                         uint64_t *d_out);
VSP_API int vsp_g2_msm_bases_dev(vsp_ctx *ctx, const vsp_g2_bases *b, uint64_t n_scalars, const uint64_t *d_scalars,
                         uint64_t *d_out);
/* sum of k affine points (the join step of a point-range sharded MSM after the all-gather) */
VSP_API int vsp_g1_sum(vsp_ctx *ctx, uint64_t k, const uint64_t *points, uint64_t *out);
VSP_API int vsp_g2_sum(vsp_ctx *ctx, uint64_t k, const uint64_t *points, uint64_t *out);
/* the same with the k points and the result on the device (e.g. the all_gather output buffer); stream-ordered */
VSP_API int vsp_g1_sum_dev(vsp_ctx *ctx, uint64_t k, const uint64_t *d_points, uint64_t *d_out);
VSP_API int vsp_g2_sum_dev(vsp_ctx *ctx, uint64_t k, const uint64_t *d_points, uint64_t *d_out);

/* ---- R1CS, proving key, prover ------------------------------------------------------------------ */
typedef struct {
    uint64_t num_inputs;      /* primary_input_size */
    uint64_t num_aux;         /* auxiliary_input_size */
    uint64_t num_constraints;
    /* CSR per matrix; column 0 is the constant-one variable */
    const uint64_t *a_row, *b_row, *c_row; /* num_constraints + 1 */
    const uint32_t *a_col, *b_col, *c_col;
    const uint64_t *a_val, *b_val, *c_val; /* canonical Fr */
} vsp_r1cs_t;

typedef struct {
    vsp_r1cs_t cs;
    const uint64_t *alpha_g1, *beta_g1, *beta_g2, *delta_g1, *delta_g2;
    const uint64_t *A_query; /* num_inputs + num_aux + 1 G1 */
    uint64_t B_count;        /* sparse B_query: sorted variable indices and their (G2, G1) pairs */
    const uint32_t *B_idx;
    const uint64_t *B_g2, *B_g1;
    uint64_t H_count; /* >= m - 1 */
    const uint64_t *H_query;
    const uint64_t *L_query; /* num_aux G1 */
} vsp_proving_key_t;

typedef struct {
    uint64_t msg_size;
    const uint64_t *delta_g1;
    const uint64_t *delta_s_g1; /* msg_size G1 */
    const uint64_t *t_g1;       /* msg_size G1 */
    const uint64_t *t_g2;       /* msg_size + 1 G2 (not used by encrypt / rerandomize) */
    const uint64_t *delta_sum_s_g1;
    const uint64_t *gamma_inverse_sum_s_g1;
} vsp_saver_pubkey_t;

VSP_API int vsp_pk_load(vsp_ctx *ctx, const vsp_proving_key_t *desc, vsp_pk **out);
VSP_API void vsp_pk_destroy(vsp_pk *pk);
VSP_API uint64_t vsp_pk_domain_size(const vsp_pk *pk);

/* r1cs_to_qap::witness_map(cs, primary, auxiliary, 0, 0, 0): h_out receives coefficients_for_H[0..m] (m + 1 Fr) */
VSP_API int vsp_witness_map(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *primary, const uint64_t *aux, uint64_t *h_out);

/* proof = g_A (12 limbs) | g_B (24 limbs) | g_C (12 limbs).  r_saver / gamma_inverse_sum_s_g1 both non-NULL selects
 * proving_mode::encrypted_input: g_C += r_saver * gamma_inverse_sum_s_g1. */
VSP_API int vsp_prove(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *primary, const uint64_t *aux, const uint64_t r[4],
              const uint64_t s[4], const uint64_t *r_saver, const uint64_t *gamma_inverse_sum_s_g1,
              uint64_t *out_proof);
/* k independent ballots on one proving key: witnesses laid out ballot-major; rs = (r, s, r_saver) per ballot
 * (r_saver ignored when gamma_inverse_sum_s_g1 is NULL).  out_proofs: k * 48 limbs. */
VSP_API int vsp_prove_batch(vsp_ctx *ctx, const vsp_pk *pk, uint64_t k, const uint64_t *primary, const uint64_t *aux,
                    const uint64_t *rs, const uint64_t *gamma_inverse_sum_s_g1, uint64_t *out_proofs);

/* device-resident variant: d_assignments = k full assignments (1 || primary || auxiliary), num_inputs + num_aux + 1
 * canonical Fr each; d_proofs = k * 48 limbs on the device.  Stream-ordered, no host synchronisation. */
VSP_API int vsp_prove_batch_dev(vsp_ctx *ctx, const vsp_pk *pk, uint64_t k, const uint64_t *d_assignments,
                                const uint64_t *rs, const uint64_t *gamma_inverse_sum_s_g1, uint64_t *d_proofs);

/* Point-range sharded proving (one proof over G GPUs; BASELINE configs[4]).  A sliced key holds, for each of the five
 * queries, only a contiguous range: desc->A_query / B_* / H_query / L_query point at the slices (B_idx: the slice's
 * global variable indices, B_count = B_hi - B_lo; H_count = H_hi - H_lo), the constraint system is replicated.
 * vsp_prove_partial returns this rank's share of the five multiexps as canonical affine points
 *   out_partials = A(w) | B1(w) | H(h) | L(w_aux)  (G1, 12 limbs each) | B2(w) (G2, 24 limbs)  = 72 limbs;
 * the shares are summed per component across ranks (one all_gather + vsp_g1_sum / vsp_g2_sum) and vsp_prove_assemble
 * builds the proof from the sums (it only needs alpha/beta/delta, so any key of the same circuit serves). */
typedef struct {
    uint64_t A_lo, A_hi; /* within [0, num_inputs + num_aux + 1) */
    uint64_t B_lo, B_hi; /* within the sparse B_query's entries */
    uint64_t H_lo, H_hi; /* within [0, m - 1) */
    uint64_t L_lo, L_hi; /* within [0, num_aux) */
} vsp_pk_slice_t;
VSP_API int vsp_pk_load_slice(vsp_ctx *ctx, const vsp_proving_key_t *desc, const vsp_pk_slice_t *slice, vsp_pk **out);
VSP_API int vsp_prove_partial(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *primary, const uint64_t *aux,
                              uint64_t *out_partials);
VSP_API int vsp_prove_assemble(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *sums, const uint64_t r[4],
                               const uint64_t s[4], const uint64_t *r_saver, const uint64_t *gamma_inverse_sum_s_g1,
                               uint64_t *out_proof);
/* Device-resident forms of the two calls above, for one process per GPU under torch.distributed / NCCL:
 * vsp_prove_partial_dev leaves this rank's five shares as 144 limbs of raw Montgomery XYZZ at d_partials
 * (evA | evB1 | evH | evL: 24 limbs each; evB2: 48), the ranks exchange them with ONE all_gather straight out of and into
 * device memory, and vsp_prove_assemble_dev sums the world x 144 gathered limbs on the device and leaves the proof
 * (48 limbs) at d_proof.  Both are stream-ordered on the context's stream and never block the host; d_assignment is
 * (1 || primary || auxiliary) canonical Fr. */
VSP_API int vsp_prove_partial_dev(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *d_assignment, uint64_t *d_partials);
VSP_API int vsp_prove_assemble_dev(vsp_ctx *ctx, const vsp_pk *pk, uint64_t world, const uint64_t *d_gathered,
                                   const uint64_t r[4], const uint64_t s[4], const uint64_t *r_saver,
                                   const uint64_t *gamma_inverse_sum_s_g1, uint64_t *d_proof);
/* vsp_prove_partial_dev in two calls, with room for an exchange between them: the three coset extensions of the witness
 * map (inverse_fft, multiply_by_coset, fft of aA, aB, aC) are independent, so G ranks can each extend the vectors of
 * `vector_mask` (bit 0 aA, bit 1 aB, bit 2 aC) into d_abc (3 m Fr, caller-owned device memory), broadcast them to each
 * other while the A / L / B multiexps already run on the side streams, and finish: H on the coset, its coefficients, the H
 * multiexp.  begin with vector_mask = 7 followed by finish equals vsp_prove_partial_dev. */
VSP_API int vsp_prove_partial_begin_dev(vsp_ctx *ctx, const vsp_pk *pk, const uint64_t *d_assignment, unsigned vector_mask,
                                        uint64_t *d_abc, uint64_t *d_partials);
VSP_API int vsp_prove_partial_finish_dev(vsp_ctx *ctx, const vsp_pk *pk, uint64_t *d_abc, uint64_t *d_partials);

/* SAVER: ciphertext (msg_size + 2 G1) and the encrypted_input proof in one call */
VSP_API int vsp_saver_encrypt(vsp_ctx *ctx, const vsp_pk *pk, const vsp_saver_pubkey_t *pub, const uint64_t *gamma_abc_rest,
                      const uint64_t *msg, const uint64_t *primary, const uint64_t *aux, const uint64_t r_saver[4],
                      const uint64_t r[4], const uint64_t s[4], uint64_t *out_ct, uint64_t *out_proof);
VSP_API int vsp_saver_ciphertext(vsp_ctx *ctx, const vsp_saver_pubkey_t *pub, const uint64_t *gamma_abc_rest,
                         const uint64_t *msg, const uint64_t r_saver[4], uint64_t *out_ct);
/* Batched SAVER (the 4,096-ballot configuration): the public key's 3 msg_size + 2 G1 bases stay on the GPU as 8-bit
 * fixed-base tables; k ballots are encrypted and proved in chunks that share every launch.  The message of ballot b
 * is the first msg_size primary inputs ("plain text should be a prefix of the primary input",
 * elgamal_verifiable.hpp:336-339).  rs = k x (r, s, r_saver).  out_cts: k x (msg_size + 2) G1; out_proofs: k x 48 limbs.
 * The _dev variant takes k device-resident assignments (1 || primary || auxiliary) and leaves both results on the
 * device, stream-ordered. */
VSP_API int vsp_saver_pub_load(vsp_ctx *ctx, const vsp_saver_pubkey_t *pub, const uint64_t *gamma_abc_rest,
                               vsp_saver_pub **out);
VSP_API void vsp_saver_pub_destroy(vsp_saver_pub *sp);
VSP_API int vsp_saver_encrypt_batch(vsp_ctx *ctx, const vsp_pk *pk, const vsp_saver_pub *sp, uint64_t k,
                                    const uint64_t *primary, const uint64_t *aux, const uint64_t *rs, uint64_t *out_cts,
                                    uint64_t *out_proofs);
VSP_API int vsp_saver_encrypt_batch_dev(vsp_ctx *ctx, const vsp_pk *pk, const vsp_saver_pub *sp, uint64_t k,
                                        const uint64_t *d_assignments, const uint64_t *rs, uint64_t *d_cts,
                                        uint64_t *d_proofs);
/* rnd = r | z1 | z2 */
VSP_API int vsp_saver_rerandomize(vsp_ctx *ctx, const vsp_saver_pubkey_t *pub, const uint64_t *delta_g2, const uint64_t *rnd,
                          const uint64_t *ct_in, const uint64_t *proof_in, uint64_t *ct_out, uint64_t *proof_out);
/* rerandomize_op<elgamal_verifiable>::process for k ballots in one call (the second half of the reference's timed vote
 * phase, bin/cli/include/nil/vote_saver/common.hpp:1132-1145): rnd = k x (r | z1 | z2); cts k x (msg_size + 2) G1;
 * proofs k x 48 limbs.  Key material comes from the resident handles (pk: delta_g2 and gamma_inverse_sum_s_g1 tables; sp:
 * delta, delta_s, delta_sum_s tables), so every multiple of it is fixed-base.  The _dev variant takes and leaves the
 * cipher texts and proofs on the device (in and out may alias), e.g. chained after vsp_saver_encrypt_batch_dev. */
VSP_API int vsp_saver_rerandomize_batch(vsp_ctx *ctx, const vsp_pk *pk, const vsp_saver_pub *sp, uint64_t k,
                                        const uint64_t *rnd, const uint64_t *cts_in, const uint64_t *proofs_in,
                                        uint64_t *cts_out, uint64_t *proofs_out);
VSP_API int vsp_saver_rerandomize_batch_dev(vsp_ctx *ctx, const vsp_pk *pk, const vsp_saver_pub *sp, uint64_t k,
                                            const uint64_t *rnd, const uint64_t *d_cts_in, const uint64_t *d_proofs_in,
                                            uint64_t *d_cts_out, uint64_t *d_proofs_out);

/* ---- key generation: the callers that produce the keys the path consumes ------------------ */
/* zk::generate<r1cs_gg_ppzksnark<bls12<381>, ..., proving_mode::encrypted_input>>(constraint_system)
 *     call site bin/cli/include/nil/vote_saver/common.hpp:916-917 (generator source in un-vendored crypto3-zk @ d335bb8b).
 * toxic = t | alpha | beta | gamma | delta (5 canonical Fr; the reference draws them inside generate — explicit here so
 * the key pair is reproducible bit for bit).  Every array of `out` is caller-allocated:
 *   A_query (n_v + 1) G1; B_idx / B_g2 / B_g1 capacity n_v + 1 (B_count is written: the sparse B_query's non-zero
 *   entries); H_query (m - 1) G1, m from vsp_domain_plan(num_constraints + num_inputs + 1); L_query num_aux G1;
 *   gamma_ABC_g1 (num_inputs + 1) G1 = first | rest; the rest single points.  gamma_g1 / delta_g1 are the
 *   encrypted_input-mode verification-key extras SAVER key generation reads (elgamal_verifiable.hpp:233, 250-260).
 * The filled proving-key half maps field for field onto vsp_proving_key_t / vsp_pk_load. */
typedef struct {
    uint64_t *alpha_g1, *beta_g1, *beta_g2, *delta_g1, *delta_g2;
    uint64_t *A_query;
    uint64_t B_count; /* out */
    uint32_t *B_idx;
    uint64_t *B_g2, *B_g1;
    uint64_t H_count; /* out: m - 1 */
    uint64_t *H_query;
    uint64_t *L_query;
    uint64_t *gamma_g2, *gamma_g1, *gamma_ABC_g1;
} vsp_keypair_t;
VSP_API int vsp_groth16_generate(vsp_ctx *ctx, const vsp_r1cs_t *cs, const uint64_t *toxic, vsp_keypair_t *out);

/* generate_keypair_op<elgamal_verifiable<Curve>>::process (elgamal_verifiable.hpp:218-271).
 * rnd = rho | t_0 | (s_i, v_i, t_i) for i < msg_size: 3 msg_size + 2 canonical Fr, consumed in the reference's order.
 * Public key = (delta_g1 [the caller's], delta_s_g1, t_g1, t_g2, delta_sum_s_g1, gamma_inverse_sum_s_g1),
 * secret key = rho, verification key = (rho_g2, rho_sv_g2, rho_rhov_g2); all arrays caller-allocated. */
typedef struct {
    uint64_t *delta_s_g1;             /* msg_size G1 */
    uint64_t *t_g1;                   /* msg_size G1 */
    uint64_t *t_g2;                   /* msg_size + 1 G2 */
    uint64_t *delta_sum_s_g1;         /* 1 G1 */
    uint64_t *gamma_inverse_sum_s_g1; /* 1 G1 */
    uint64_t *rho_g2;                 /* 1 G2 */
    uint64_t *rho_sv_g2;              /* msg_size G2 */
    uint64_t *rho_rhov_g2;            /* msg_size G2 */
    uint64_t *rho;                    /* 4 limbs */
} vsp_saver_keypair_t;
VSP_API int vsp_saver_keygen(vsp_ctx *ctx, uint64_t msg_size, const uint64_t *delta_g1, const uint64_t *gamma_g1,
                             const uint64_t *gamma_abc_rest, const uint64_t *rnd, vsp_saver_keypair_t *out);

/* building blocks: out[i] = scalars[i] * points[i] (variable base), and fixed-base multiples of the group generators
 * (`scalar * g1_type::value_type::one()`, e.g. elgamal_verifiable.hpp:233-249) from an 8-bit table, one thread per scalar */
VSP_API int vsp_g1_mul_batch(vsp_ctx *ctx, uint64_t n, const uint64_t *points, const uint64_t *scalars, uint64_t *out);
VSP_API int vsp_g1_generator_mul_batch(vsp_ctx *ctx, uint64_t n, const uint64_t *scalars, uint64_t *out);
VSP_API int vsp_g2_generator_mul_batch(vsp_ctx *ctx, uint64_t n, const uint64_t *scalars, uint64_t *out);

/* ---- pairing side: the verifiers that consume what the path produces ------------------------- */
/* r1cs_gg_ppzksnark verification key, group members only: the reference stores e(alpha_g1, beta_g2) as the GT element
 * alpha_g1_beta_g2 (libs/marshalling/zk/.../r1cs_gg_ppzksnark/verification_key.hpp:69); here its two preimages are
 * passed and the pairing joins the product.  gamma_ABC_g1 = first | rest, gamma_ABC_count points. */
typedef struct {
    const uint64_t *alpha_g1, *beta_g2, *gamma_g2, *delta_g2;
    uint64_t gamma_ABC_count;
    const uint64_t *gamma_ABC_g1;
} vsp_verification_key_t;
/* algebra::pair_reduced<Curve>(P, Q) (call sites elgamal_verifiable.hpp:415-424, 559-571), batched as n_checks products
 * of `pairs` pairings: g1 / g2 hold n_checks x pairs canonical affine points; gt_out[c] = prod_j e(g1[c][j], g2[c][j])
 * as a canonical Fq12 (72 limbs, tower order c0.c0 | c0.c1 | c0.c2 | c1.c0 | c1.c1 | c1.c2 over Fq2 = c0 | c1,
 * Fq12 = Fq6[w]/(w^2 - v), Fq6 = Fq2[v]/(v^3 - (1 + u))); is_one[c] = 1 iff the product is the identity.  Either
 * output may be NULL.  A pair with a point at infinity contributes 1.  The value is crypto3's pair_reduced, i.e. the
 * Miller value to the power 3 (p^12 - 1) / r (the cube of the textbook reduced pairing; pinned on the alpha_g1_beta_g2
 * stored in the reference's bin/cli/src/data.bin). */
VSP_API int vsp_pairing_products(vsp_ctx *ctx, uint64_t n_checks, uint64_t pairs, const uint64_t *g1, const uint64_t *g2,
                                 uint64_t *gt_out, uint8_t *is_one);
/* zk::verify<r1cs_gg_ppzksnark<...>>(vk, primary_input, proof) for k proofs on one key (basic mode;
 * libs/marshalling/zk/test/r1cs_gg_ppzksnark.cpp:160-163): ok_out[b] = 1 iff
 *   e(A, B) = e(alpha, beta) e(gamma_ABC.first + sum_i x_i gamma_ABC.rest[i], gamma) e(C, delta).
 * primary: k x n_primary canonical Fr; vk->gamma_ABC_count must be n_primary + 1; proofs: k x 48 limbs. */
VSP_API int vsp_groth16_verify_batch(vsp_ctx *ctx, const vsp_verification_key_t *vk, uint64_t k, const uint64_t *primary,
                                     uint64_t n_primary, const uint64_t *proofs, uint8_t *ok_out);
/* The same with the verification key in the reference's own form (libs/marshalling/zk/.../verification_key.hpp:69-75:
 * alpha_g1_beta_g2 = pair_reduced(alpha_g1, beta_g2) as an Fq12 value, not alpha and beta): ok_out[b] = 1 iff
 *   e(A, B) = alpha_g1_beta_g2 * e(gamma_ABC.first + sum_i x_i gamma_ABC.rest[i], gamma) * e(C, delta)
 * compared as canonical Fq12 values.  alpha_g1_beta_g2: 72 limbs in the order vsp_pairing_products writes, which is the
 * order the reference serialises (pinned on bin/cli/src/data.bin, tests/golden/databin_groth16.json). */
VSP_API int vsp_groth16_verify_gt_batch(vsp_ctx *ctx, const uint64_t *alpha_g1_beta_g2, const uint64_t *gamma_g2,
                                        const uint64_t *delta_g2, uint64_t gamma_ABC_count, const uint64_t *gamma_ABC_g1,
                                        uint64_t k, const uint64_t *primary, uint64_t n_primary, const uint64_t *proofs,
                                        uint8_t *ok_out);
/* Both verifiers first check what the prover sent (A, B, C, the cipher text): coordinates reduced mod p, on the curve /
 * twist, in the order-r subgroup (proof.is_well_formed() and the decoders' guarantees in the reference); an ill-formed
 * element gives ok_out[b] = 0.  The point at infinity is well formed. */
/* verify_encryption_op<elgamal_verifiable>::process -> zk::verify<..., encrypted_input>(ct, vk, pubkey,
 * unencrypted_primary_input, proof) (elgamal_verifiable.hpp:486-490) for k ballots: ok_out[b] = 1 iff
 *   prod_{i <= n} e(ct_i, t_g2[i]) = e(ct_{n+1}, G2)   and
 *   e(A, B) = e(alpha, beta) e(gamma_ABC.first + sum_{i<=n} ct_i + sum_j x_j gamma_ABC.rest[n + j], gamma) e(C, delta),
 * n = pub->msg_size.  cts: k x (n + 2) G1; unencrypted_primary: k x n_unencrypted Fr (the primary input after the
 * message slots); vk->gamma_ABC_count must be 1 + n + n_unencrypted. */
VSP_API int vsp_saver_verify_encryption_batch(vsp_ctx *ctx, const vsp_verification_key_t *vk, const vsp_saver_pubkey_t *pub,
                                              uint64_t k, const uint64_t *cts, const uint64_t *unencrypted_primary,
                                              uint64_t n_unencrypted, const uint64_t *proofs, uint8_t *ok_out);

/* The same with the verification key in the reference's own form, as vsp_codec_groth16_vk_read delivers it
 * (alpha_g1_beta_g2: 72 limbs; verification_key.hpp:128-152) — what a verifier that only holds the key blobs calls
 * (verify_encryption in the vote phase's own check, bin/cli/include/nil/vote_saver/common.hpp:1163-1169). */
VSP_API int vsp_saver_verify_encryption_gt_batch(vsp_ctx *ctx, const uint64_t *alpha_g1_beta_g2, const uint64_t *gamma_g2,
                                                 const uint64_t *delta_g2, uint64_t gamma_ABC_count,
                                                 const uint64_t *gamma_ABC_g1, const vsp_saver_pubkey_t *pub, uint64_t k,
                                                 const uint64_t *cts, const uint64_t *unencrypted_primary,
                                                 uint64_t n_unencrypted, const uint64_t *proofs, uint8_t *ok_out);

/* decrypt_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:401-441): ct = msg_size + 2 G1 (e.g. the tally's sum of
 * ballots' ciphertexts), rho the secret key, (rho_sv_g2, rho_rhov_g2) from the SAVER verification key, gamma_abc_rest the
 * first msg_size points of gamma_ABC_g1.rest.  m_out: msg_size Fr, each the discrete log found by the reference's search
 * over exp = 0 .. 2^block_bits (scheme_type::block_bits, default 4); verify_c0_out = rho * ct[0], the decryption proof.
 * VSP_E_INVALID with text "Decryption failed." where the reference asserts. */
VSP_API int vsp_saver_decrypt(vsp_ctx *ctx, uint64_t msg_size, const uint64_t *ct, const uint64_t rho[4],
                              const uint64_t *rho_sv_g2, const uint64_t *rho_rhov_g2, const uint64_t *gamma_abc_rest,
                              unsigned block_bits, uint64_t *m_out, uint64_t *verify_c0_out);
/* verify_decryption_op<elgamal_verifiable>::process (elgamal_verifiable.hpp:548-575): proof = verify_c0 of decrypt */
VSP_API int vsp_saver_verify_decryption(vsp_ctx *ctx, uint64_t msg_size, const uint64_t *ct, const uint64_t *plain,
                                        const uint64_t *rho_g2, const uint64_t *rho_sv_g2, const uint64_t *rho_rhov_g2,
                                        const uint64_t *gamma_abc_rest, const uint64_t *proof, uint8_t *ok_out);

/* ---- witness generation for the ballot circuit --------------------------------------------------
 * Replaces the per-vote generate_r1cs_witness sequence of bin/cli (common.hpp:1110-1128) for k ballots at once.  The
 * circuit (components::encrypted_input_voting over Pedersen-on-Jubjub hashes and an arity-2 Merkle path) lives in the
 * un-vendored crypto3-blueprint; vote-saver-protocol_b200/voting.py restates it and produces both the R1CS and the
 * layout tables this program is created from (parity with crypto3's wire values unpinned; every witness satisfies that
 * R1CS).  vsp_voting_witness_batch_dev fills k assignments (1 || primary || auxiliary, canonical Fr, num_vars + 1 elements
 * each) in device memory, the input format of vsp_saver_encrypt_batch_dev: votes / indices k u32 (the chosen option and the
 * voter's leaf position), eid the election id bits, sks k x 4 limbs (255-bit secrets), paths k x depth x 4 limbs (sibling
 * digests, leaf level first). */
typedef struct vsp_voting_program vsp_voting_program;
VSP_API int vsp_voting_program_create(vsp_ctx *ctx, const uint32_t *scalars, uint64_t n_scalars, const uint32_t *hashes,
                                      uint64_t n_hashes, const uint32_t *in_idx, uint64_t n_in, const uint32_t *levels,
                                      uint64_t depth, const uint64_t *tables, uint64_t n_table_rows,
                                      vsp_voting_program **out);
VSP_API void vsp_voting_program_destroy(vsp_voting_program *p);
VSP_API uint64_t vsp_voting_program_num_vars(const vsp_voting_program *p);
VSP_API int vsp_voting_witness_batch_dev(vsp_ctx *ctx, const vsp_voting_program *p, uint64_t k, const uint32_t *votes,
                                         uint64_t eid, const uint64_t *sks, const uint32_t *indices, const uint64_t *paths,
                                         uint64_t *d_assignments);

/* ---- wire codecs: the reference's byte formats either side of the path (host only, no GPU work) ------------- */
/* Formats (relative to libs/marshalling): big-endian fixed-width integers (Fr 32 B, Fq 48 B, size_t 8 B), sequences
 * with an 8-byte count prefix; compressed points as processing/curve_element.hpp:93-174 writes them (G1 48 B,
 * G2 96 B = x.c1 | x.c0, flag bits 0x80 compressed / 0x40 infinity / 0x20 y > (p-1)/2); "fast" points X | Y |
 * is_infinity (types/fast_curve_element.hpp:56-97).  These are what bin/cli's vote phase reads (common.hpp:1004-1006)
 * and writes (common.hpp:1154-1160). */
VSP_API int vsp_codec_g1_compress(const uint64_t *point, uint8_t *out48);
VSP_API int vsp_codec_g2_compress(const uint64_t *point, uint8_t *out96);
/* proof g_A | g_B | g_C (48 limbs) -> 192 bytes (zk/.../r1cs_gg_ppzksnark/proof.hpp:63-71) */
VSP_API int vsp_codec_proof_write(const uint64_t *proof, uint8_t *out192);
/* the decoders (processing/curve_element.hpp:176-395, proof.hpp:86-102): compressed octets -> canonical affine limbs;
 * VSP_E_INVALID for a malformed encoding (C bit clear, non-canonical infinity, x >= p) or an x with no curve point */
VSP_API int vsp_codec_g1_decompress(const uint8_t *in48, uint64_t *point);
VSP_API int vsp_codec_g2_decompress(const uint8_t *in96, uint64_t *point);
VSP_API int vsp_codec_proof_read(const uint8_t *in192, uint64_t *proof);
/* n cipher-text G1 points -> 8 + 48 n bytes; n Fr -> 8 + 32 n bytes (zk/.../r1cs_gg_ppzksnark/primary_input.hpp:61-97) */
VSP_API int vsp_codec_ciphertext_write(const uint64_t *ct, uint64_t n, uint8_t *out);
VSP_API int vsp_codec_primary_input_write(const uint64_t *values, uint64_t n, uint8_t *out);
VSP_API int vsp_codec_primary_input_read(const uint8_t *in, size_t len, uint64_t *n, uint64_t *values, uint64_t capacity);
/* r1cs_gg_ppzksnark_fast_proving_key (zk/.../r1cs_gg_ppzksnark/fast_proving_key.hpp:66-103): parse in two steps —
 * sizes, then fill caller-allocated arrays addressed by a vsp_proving_key_t (rows: num_constraints + 1; cols / vals:
 * nnz_*; A_query: A_count; B_idx / B_g2 / B_g1: B_count; H_query: H_count; L_query: L_count).  The filled descriptor
 * goes straight into vsp_pk_load.  vsp_codec_fast_pk_write is the inverse (out == NULL: size only). */
typedef struct {
    uint64_t num_inputs, num_aux, num_constraints, nnz_a, nnz_b, nnz_c;
    uint64_t A_count, B_count, B_domain_size, H_count, L_count;
    uint64_t bytes; /* length of the serialised key inside the blob */
} vsp_fast_pk_sizes_t;
VSP_API int vsp_codec_fast_pk_sizes(const uint8_t *blob, size_t len, vsp_fast_pk_sizes_t *out);
VSP_API int vsp_codec_fast_pk_read(const uint8_t *blob, size_t len, vsp_proving_key_t *desc);
VSP_API int vsp_codec_fast_pk_write(const vsp_proving_key_t *desc, uint8_t *out, size_t capacity, size_t *written);

/* The inverse of vsp_codec_ciphertext_write (deserialize_ct, bin/cli/include/nil/vote_saver/common.hpp:776): *n receives
 * the count; ct == NULL queries it; capacity in points.  VSP_E_INVALID for a truncated blob or a malformed point. */
VSP_API int vsp_codec_ciphertext_read(const uint8_t *in, size_t len, uint64_t *n, uint64_t *ct, uint64_t capacity);
/* elgamal_verifiable public key, the vote phase's pk_eid blob (common.hpp:691, 1002;
 * libs/marshalling/zk/.../pubkey/types/elgamal_verifiable.hpp, elgamal_verifiable_public_key / fill_public_key):
 * delta_g1 | delta_sum_s_g1 | gamma_inverse_sum_s_g1 | delta_s_g1[msg_size] | t_g1[msg_size] | t_g2[msg_size + 1], every
 * point compressed, every list count-prefixed.  write: out == NULL gives the size only.  read: pub == NULL gives
 * *msg_size only; otherwise pub->msg_size must equal it and pub's six arrays, caller-allocated for that size, are filled
 * (the filled struct goes straight into vsp_saver_pub_load / vsp_saver_encrypt). */
VSP_API int vsp_codec_saver_pubkey_write(const vsp_saver_pubkey_t *pub, uint8_t *out, size_t capacity, size_t *written);
VSP_API int vsp_codec_saver_pubkey_read(const uint8_t *blob, size_t len, uint64_t *msg_size, vsp_saver_pubkey_t *pub);
/* elgamal_verifiable verification key (same file, elgamal_verifiable_verification_key / fill_verification_key;
 * common.hpp:706): rho_g2 | rho_sv_g2[msg_size] | rho_rhov_g2[msg_size], what vsp_saver_decrypt /
 * vsp_saver_verify_decryption take.  read: rho_g2 == NULL gives *msg_size only; otherwise capacity must equal it. */
VSP_API int vsp_codec_saver_vk_write(uint64_t msg_size, const uint64_t *rho_g2, const uint64_t *rho_sv_g2,
                                     const uint64_t *rho_rhov_g2, uint8_t *out, size_t capacity, size_t *written);
VSP_API int vsp_codec_saver_vk_read(const uint8_t *blob, size_t len, uint64_t *msg_size, uint64_t *rho_g2,
                                    uint64_t *rho_sv_g2, uint64_t *rho_rhov_g2, uint64_t capacity);
/* elgamal_verifiable private key (same file, fill_private_key; common.hpp:722): rho as 32 big-endian bytes; the reader
 * rejects a value that is not reduced mod r */
VSP_API int vsp_codec_saver_sk_write(const uint64_t rho[4], uint8_t *out32);
VSP_API int vsp_codec_saver_sk_read(const uint8_t *in, size_t len, uint64_t rho[4]);
/* r1cs_gg_ppzksnark_extended_verification_key, the vote and tally phases' verification key blob (common.hpp:734, 1005;
 * libs/marshalling/zk/.../r1cs_gg_ppzksnark/verification_key.hpp:128-152): alpha_g1_beta_g2 as 12 Fq coefficients in tower
 * order (72 limbs here, the order vsp_pairing_products writes and vsp_groth16_verify_gt_batch reads) | gamma_g2 | delta_g2 |
 * delta_g1 | gamma_g1 | gamma_ABC_g1 = first | rest, the rest a sparse_vector (indices | values | domain_size;
 * accumulation_vector.hpp:52-57, sparse_vector.hpp:49-58).  gamma_ABC_g1 here is dense, gamma_ABC_count = 1 + domain size
 * points; the writer lists every position of the rest (as the reference's generator builds it), the reader leaves
 * positions the blob does not list at infinity.  write: out == NULL gives the size only.  read: alpha_g1_beta_g2 == NULL
 * gives *gamma_ABC_count only; otherwise capacity must equal it.  A rest whose domain size exceeds 2^24 is refused
 * (VSP_E_INVALID): the dense array it would need is not worth allocating on the word of an untrusted blob. */
VSP_API int vsp_codec_groth16_vk_write(const uint64_t *alpha_g1_beta_g2, const uint64_t *gamma_g2, const uint64_t *delta_g2,
                                       const uint64_t *delta_g1, const uint64_t *gamma_g1, uint64_t gamma_ABC_count,
                                       const uint64_t *gamma_ABC_g1, uint8_t *out, size_t capacity, size_t *written);
VSP_API int vsp_codec_groth16_vk_read(const uint8_t *blob, size_t len, uint64_t *gamma_ABC_count, uint64_t *alpha_g1_beta_g2,
                                      uint64_t *gamma_g2, uint64_t *delta_g2, uint64_t *delta_g1, uint64_t *gamma_g1,
                                      uint64_t *gamma_ABC_g1, uint64_t capacity);

/* The vote phase's remaining inputs (common.hpp:996-1003), host only.
 * serialize_bitarray / deserialize_bitarray (common.hpp:574-612): bits holds n_bits bytes of 0 / 1; bit i travels in octet
 * i / 8, most significant bit first; the blob is exactly ceil(n_bits / 8) octets. */
VSP_API int vsp_codec_bitarray_write(const uint8_t *bits, uint64_t n_bits, uint8_t *out);
VSP_API int vsp_codec_bitarray_read(const uint8_t *in, size_t len, uint64_t n_bits, uint8_t *bits);
/* get_multi_field_element_from_bits (common.hpp:543-570): ceil(n_bits / 254) canonical Fr, bit j of a chunk weighs 2^j */
VSP_API int vsp_codec_bits_to_field_elements(const uint8_t *bits, uint64_t n_bits, uint64_t *out);
/* deserialize_merkle_tree (common.hpp:614-636; libs/containers/include/nil/crypto3/container/merkle/tree.hpp:424-427,
 * 492-499): 2^(depth + 1) - 1 digests of 255 bits as 32-octet bit arrays, leaves first, then row by row, the root last.
 * digests: 4 limbs each, bit i of the integer = digest bit i. */
VSP_API int vsp_codec_merkle_tree_write(const uint64_t *digests, uint64_t depth, uint8_t *out);
VSP_API int vsp_codec_merkle_tree_read(const uint8_t *blob, size_t len, uint64_t depth, uint64_t *digests);
/* containers::merkle_proof<hash, 2>(tree, index) (libs/containers/.../merkle/proof.hpp:65-91): the depth sibling digests
 * from the leaf level up, i.e. one row of vsp_voting_witness_batch_dev's paths */
VSP_API int vsp_codec_merkle_copath(const uint64_t *digests, uint64_t depth, uint64_t index, uint64_t *path);

/* ---- measurement helpers ---------------------------------------------------------------------- */
/* 32x32->64 multiply-adds the G1 bucket-accumulation kernel of this build issues per bucket addition (the reference
 * algorithm's mixed Jacobian addition is 11 x 300 = 3,300): lets a roofline in reference units be read as pipe use */
VSP_API int vsp_accumulate_mads_per_add(vsp_ctx *ctx, double *out);
/* sustained 32-bit multiply-add issue rate: runs a dependent-free IMAD.WIDE carry-chain kernel on all SMs for
 * `iters` rounds and returns multiply-adds per second (the P_int roofline denominator). */
VSP_API int vsp_measure_imad_rate(vsp_ctx *ctx, uint64_t iters, double *mads_per_second);
/* the same for the FP64 pipe: independent chains of fused multiply-adds on every SM (the unit of a 52-bit-limb DFMA
 * Montgomery product, two per partial product).  Reported by bench.py beside the integer peak; DESIGN.md section 4 has the
 * go / no-go it supports. */
VSP_API int vsp_measure_dfma_rate(vsp_ctx *ctx, uint64_t iters, double *fmas_per_second);

#ifdef __cplusplus
}
#endif
#endif /* VSP_B200_H */
