// common.cuh — context object, error plumbing and stream-ordered scratch memory shared by the .cu files.
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "../../include/vsp_b200.h"
#include "scratch_class.h"

namespace vsp {
struct NttTables;
struct ScratchCache;

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
    // pinned host staging of the per-chunk scalar uploads (prover.cu): four slots in turn; a slot is refilled only after
    // the copies that read it have completed, so the uploads are truly asynchronous and the launching thread, which runs
    // at most a step ahead, does not wait in practice
    struct StageSlot {
        void* p = nullptr;
        size_t cap = 0;
        cudaEvent_t done = nullptr;
        bool pending = false;
    } stage[4];
    unsigned stage_next = 0;
    void* partial_bsc = nullptr;  // gathered B-query scalars between vsp_prove_partial_begin_dev and _finish_dev
    size_t partial_bsc_cap = 0;
    int batch_chunk = 0;  // ballots per chunk of the batched prover (0: VSP_BATCH_CHUNK or 64); vsp_ctx_set_batch_chunk
    cudaEvent_t ev_main = nullptr;
    vsp::ScratchCache* scratch = nullptr;  // device blocks behind every DevBuf (below)
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

// VSP_HOST_TRACE=<ms>: every CUDA runtime call / kernel launch of this library that holds the calling thread for longer
// than <ms> is reported on stderr with its source line (diagnosis of host-side stalls; off by default: one branch)
struct HostTrace {
    const char* what;
    const char* file;
    int line;
    std::chrono::steady_clock::time_point t0;
    static double threshold() {
        static const double t = getenv("VSP_HOST_TRACE") ? atof(getenv("VSP_HOST_TRACE")) : -1.0;
        return t;
    }
    HostTrace(const char* w, const char* f, int l) : what(w), file(f), line(l) {
        if (threshold() >= 0) t0 = std::chrono::steady_clock::now();
    }
    ~HostTrace() {
        if (threshold() < 0) return;
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (ms > threshold()) fprintf(stderr, "[vsp host trace] %.2f ms in %s (%s:%d)\n", ms, what, file, line);
    }
};

#define VSP_CUDA_CHECK(ctx, expr)                                                                          \
    do {                                                                                                   \
        cudaError_t _e;                                                                                    \
        {                                                                                                  \
            ::vsp::HostTrace _ht(#expr, __FILE__, __LINE__);                                               \
            _e = (expr);                                                                                   \
        }                                                                                                  \
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
        {                                                            \
            ::vsp::HostTrace _ht(#kernel, __FILE__, __LINE__);       \
            kernel<<<(grid), (block), (smem), (strm)>>>(__VA_ARGS__);\
        }                                                            \
        ::vsp::kprof_end((ctx), _pi, (strm));                        \
        (ctx)->last_prof = _pi;                                      \
        ++(ctx)->launches;                                           \
    } while (0)
#define VSP_LAUNCH(ctx, kernel, grid, block, smem, strm, ...) \
    VSP_LAUNCH_K(ctx, ::vsp::KC_OTHER, kernel, grid, block, smem, strm, __VA_ARGS__)

// ---------------------------------------------------------------------------------------------
// Scratch memory of a context: a cache of device blocks the context owns, in place of the driver's stream-ordered pool.
//
// The pool (cudaMallocAsync) was measured to hold the launching thread for 20-800 ms inside ONE allocation of the first
// step after a device synchronisation, in about one bench run in four (its block assignment depends on which frees
// have completed on the GPU, so the free list it finds after a synchronisation varies from run to run).  The cache
// depends only on the order of the host's calls: a step that repeats the same allocations finds the same blocks, and
// after the first step no driver allocation happens at all.
//
// A block is found by (stream, size class): a block of the request's class last used on the SAME stream, else a new one
// from cudaMalloc.  A block never moves between streams, so stream order alone makes reuse safe:
// no events, and no dependency is ever created between the context's streams (taking a block another stream released
// made that stream's work up to the release a prerequisite: measured 2 % on the ballot step, where the multiexps on the
// side streams are meant to overlap).  As with cudaFreeAsync, work on other streams that reads a buffer must have been
// joined into the buffer's stream before the release.  The price is memory: the cache holds the sum of the streams'
// peaks rather than the peak of their sum.  Blocks are returned to the driver by scratch_trim (vsp_ctx_trim), when more
// than VSP_SCRATCH_CACHE_GB (default 64) sit unused at the moment a new block is needed, when a cudaMalloc fails, and
// when the context is destroyed.
// ---------------------------------------------------------------------------------------------
struct ScratchBlock {
    void* p = nullptr;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;  // the one stream this block is used on
};

struct ScratchCache {
    std::map<cudaStream_t, std::multimap<size_t, void*>> unused;  // per stream, by size
    size_t unused_bytes = 0, owned_bytes = 0;
    size_t limit = size_t(getenv("VSP_SCRATCH_CACHE_GB") ? atof(getenv("VSP_SCRATCH_CACHE_GB")) * 1073741824.0 : 64.0 * 1073741824.0);
    uint64_t driver_allocations = 0;  // cudaMalloc calls so far (a steady loop stops adding to it)
};

// every unused block back to the driver (device-wide synchronisation: their last users must be done)
inline void scratch_trim(vsp_ctx* ctx) {
    ScratchCache* c = ctx ? ctx->scratch : nullptr;
    if (!c || c->unused_bytes == 0) return;
    cudaDeviceSynchronize();
    for (auto& per_stream : c->unused)
        for (auto& kv : per_stream.second) {
            cudaFree(kv.second);
            c->owned_bytes -= kv.first;
        }
    c->unused.clear();
    c->unused_bytes = 0;
}

inline int scratch_acquire(vsp_ctx* ctx, size_t n, cudaStream_t stream, ScratchBlock* out) {
    if (!ctx->scratch) ctx->scratch = new ScratchCache();
    ScratchCache& c = *ctx->scratch;
    n = scratch_size_class(n);
    auto& mine = c.unused[stream];
    auto it = mine.find(n);
    if (it != mine.end()) {
        out->p = it->second;
        out->bytes = it->first;
        out->stream = stream;
        c.unused_bytes -= out->bytes;
        mine.erase(it);
        return VSP_OK;
    }
    if (c.unused_bytes > c.limit) scratch_trim(ctx);
    void* p = nullptr;
    cudaError_t e;
    {
        HostTrace ht("cudaMalloc (scratch cache miss)", __FILE__, __LINE__);
        e = cudaMalloc(&p, n);
        if (e != cudaSuccess && c.unused_bytes) {
            cudaGetLastError();
            scratch_trim(ctx);
            e = cudaMalloc(&p, n);
        }
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_error(ctx, VSP_E_NOMEM, "cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(e));
    }
    out->p = p;
    out->bytes = n;
    out->stream = stream;
    c.owned_bytes += n;
    ++c.driver_allocations;
    return VSP_OK;
}

inline void scratch_release(vsp_ctx* ctx, ScratchBlock& b) {
    ctx->scratch->unused[b.stream].emplace(b.bytes, b.p);
    ctx->scratch->unused_bytes += b.bytes;
}

// Stream-ordered device buffer out of the context's scratch cache: no device synchronisation and, once a workload has
// run through once, no driver call beyond an event record.
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaStream_t s = nullptr;
    vsp_ctx* owner = nullptr;
    ScratchBlock blk;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    int alloc(vsp_ctx* ctx, size_t n, cudaStream_t stream) {
        release();
        if (!ctx) return VSP_E_INVALID;
        if (n == 0) n = 16;
        int rc = scratch_acquire(ctx, n, stream, &blk);
        if (rc != VSP_OK) return rc;
        owner = ctx;
        s = stream;
        p = blk.p;
        bytes = n;
        return VSP_OK;
    }
    void release() {
        if (p) scratch_release(owner, blk);
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
