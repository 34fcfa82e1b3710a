// common.cuh — context object, error plumbing and stream-ordered scratch memory shared by the .cu files.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/vsp_b200.h"

namespace vsp {
struct NttTables;

// kernel classes for the live per-class device timing (vsp_ctx_profile_*): bench.py's roofline leg
enum KClass {
    KC_MSM_ACC_G1 = 0,  // bucket accumulation, G1 (the dominant kernel of a proof)
    KC_MSM_ACC_G2 = 1,  // bucket accumulation, G2
    KC_MSM_SORT = 2,    // digits, histogram scans, scatter, task list
    KC_MSM_REDUCE = 3,  // heavy-bucket collapse, bucket running sums, window sums, Horner + to-affine
    KC_NTT = 4,         // Stockham passes and step-domain kernels
    KC_QAP = 5,         // R1CS evaluation, pointwise H, coset tables
    KC_FIXED = 6,       // fixed-base scalar multiples, proof assembly, SAVER ciphertext
    KC_OTHER = 7,
    KC_COUNT = 8
};

struct KProfRec {
    int kc;
    cudaEvent_t a, b;
    int slot;       // index into the pinned counter array, or -1
    double scale;   // units = counter * scale (slot >= 0) or scale alone
};
struct KProf {
    bool on = false;
    std::vector<KProfRec> recs;
    double ms[KC_COUNT] = {0};
    double units[KC_COUNT] = {0};
    uint64_t n[KC_COUNT] = {0};
    uint32_t* pinned = nullptr;  // device-side counters copied back (4 bytes each)
    int pinned_cap = 0, pinned_used = 0;
};
}

struct vsp_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    std::string warn;  // last non-fatal degradation (e.g. a window table that did not fit); vsp_last_warning
    uint64_t launches = 0;
    int sm_count = 148;
    vsp::NttTables* ntt = nullptr;  // lazily built twiddle tables (ntt.cu)
    // side streams + events for overlapping independent multiexps inside vsp_prove
    cudaStream_t side[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t copy = nullptr;  // host -> device witness prefetch of the next chunk (prover.cu), created on first use
    cudaEvent_t ev_fork = nullptr;
    cudaEvent_t ev_join[4] = {nullptr, nullptr, nullptr, nullptr};
    // double-buffered scratch of the proof-assembly tail (prover.cu): the tail of chunk i runs on side[1] while the
    // main stream already works on chunk i + 1
    struct TailSlot {
        void* p = nullptr;
        size_t cap = 0;
        cudaEvent_t done = nullptr;
        bool pending = false;
    } tail[2];
    unsigned tail_next = 0;
    void* partial_bsc = nullptr;  // gathered B-query scalars between vsp_prove_partial_begin_dev and _finish_dev
    size_t partial_bsc_cap = 0;
    int batch_chunk = 0;  // ballots per chunk of the batched prover (0: VSP_BATCH_CHUNK or 64); vsp_ctx_set_batch_chunk
    cudaEvent_t ev_main = nullptr;
    vsp::KProf prof;
    int last_prof = -1;  // profile record of the most recent launch (for kprof_units)
};

namespace vsp {

inline int set_error(vsp_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

inline void set_warning(vsp_ctx* ctx, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->warn = buf;
}

#define VSP_CUDA_CHECK(ctx, expr)                                                                          \
    do {                                                                                                   \
        cudaError_t _e = (expr);                                                                           \
        if (_e != cudaSuccess)                                                                             \
            return ::vsp::set_error((ctx), VSP_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                    __FILE__, __LINE__);                                                   \
    } while (0)

#define VSP_TRY(expr)           \
    do {                        \
        int _rc = (expr);       \
        if (_rc != VSP_OK) return _rc; \
    } while (0)

// live per-class timing: when enabled, every launch is bracketed by two events on its own stream
inline int kprof_begin(vsp_ctx* ctx, int kc, cudaStream_t s) {
    if (!ctx->prof.on) return -1;
    KProfRec r;
    r.kc = kc;
    r.slot = -1;
    r.scale = 0.0;
    if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return -1;
    cudaEventRecord(r.a, s);
    ctx->prof.recs.push_back(r);
    return (int)ctx->prof.recs.size() - 1;
}
inline void kprof_end(vsp_ctx* ctx, int idx, cudaStream_t s) {
    if (idx >= 0) cudaEventRecord(ctx->prof.recs[idx].b, s);
}
// algorithmic work of launch `idx`: `scale` units, times a 32-bit device counter read back after the launch if given
inline void kprof_units(vsp_ctx* ctx, int idx, double scale, const uint32_t* d_counter, cudaStream_t s) {
    if (idx < 0) return;
    KProf& p = ctx->prof;
    p.recs[idx].scale = scale;
    if (d_counter && p.pinned && p.pinned_used < p.pinned_cap) {
        p.recs[idx].slot = p.pinned_used++;
        cudaMemcpyAsync(p.pinned + p.recs[idx].slot, d_counter, 4, cudaMemcpyDeviceToHost, s);
    }
}

// counted launch helpers: every kernel of this library goes through them so gpu_launches is an exact count
#define VSP_LAUNCH_K(ctx, kc, kernel, grid, block, smem, strm, ...)  \
    do {                                                             \
        const int _pi = ::vsp::kprof_begin((ctx), (kc), (strm));     \
        kernel<<<(grid), (block), (smem), (strm)>>>(__VA_ARGS__);    \
        ::vsp::kprof_end((ctx), _pi, (strm));                        \
        (ctx)->last_prof = _pi;                                      \
        ++(ctx)->launches;                                           \
    } while (0)
#define VSP_LAUNCH(ctx, kernel, grid, block, smem, strm, ...) \
    VSP_LAUNCH_K(ctx, ::vsp::KC_OTHER, kernel, grid, block, smem, strm, __VA_ARGS__)

// Stream-ordered device buffer (cudaMallocAsync pool: no device synchronisation, memory reused across calls).
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaStream_t s = nullptr;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    int alloc(vsp_ctx* ctx, size_t n, cudaStream_t stream) {
        release();
        if (n == 0) n = 16;
        s = stream;
        cudaError_t e = cudaMallocAsync(&p, n, stream);
        if (e != cudaSuccess) {
            p = nullptr;
            return set_error(ctx, VSP_E_NOMEM, "cudaMallocAsync(%zu) failed: %s", n, cudaGetErrorString(e));
        }
        bytes = n;
        return VSP_OK;
    }
    void release() {
        if (p) cudaFreeAsync(p, s);
        p = nullptr;
        bytes = 0;
    }
    template <class T>
    T* as() const {
        return reinterpret_cast<T*>(p);
    }
};

inline unsigned ceil_log2_u64(uint64_t n) {
    unsigned l = 0;
    while ((uint64_t(1) << l) < n) ++l;
    return l;
}
inline unsigned floor_log2_u64(uint64_t n) {
    unsigned l = 0;
    while ((n >> (l + 1)) != 0) ++l;
    return l;
}

// internal cross-file entry points ------------------------------------------------------------
// ntt.cu
int ntt_tables_ensure(vsp_ctx* ctx, unsigned log_n);
void ntt_tables_free(vsp_ctx* ctx);
// power-of-two NTT of 2^log_n canonical/any-form Fr values, out of place (d_in may equal d_out only if a scratch
// is supplied internally); inverse scales by 2^-log_n.  pre / post: optional per-index multiplier tables.
// `batch` vectors, vector v at element v * stride (0 = packed); the pre / post tables are shared by the batch.
int ntt_pow2(vsp_ctx* ctx, cudaStream_t s, const uint32_t* d_in, uint32_t* d_out, unsigned log_n, bool inverse,
             const uint32_t* d_pre, const uint32_t* d_post, unsigned batch = 1, uint64_t in_stride = 0,
             uint64_t out_stride = 0);
int domain_fft_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, bool inverse, uint32_t* d_inout, unsigned batch = 1,
                   uint64_t stride = 0);
int multiply_by_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t n, const uint64_t g[4], uint32_t* d_inout);
int divide_by_z_on_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, uint32_t* d_inout);

}  // namespace vsp
