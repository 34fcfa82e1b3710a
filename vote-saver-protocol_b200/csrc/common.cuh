// common.cuh — context object, error plumbing and stream-ordered scratch memory shared by the .cu files.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/vsp_b200.h"

namespace vsp {
struct NttTables;
}

struct vsp_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    uint64_t launches = 0;
    int sm_count = 148;
    vsp::NttTables* ntt = nullptr;  // lazily built twiddle tables (ntt.cu)
    // side streams + events for overlapping independent multiexps inside vsp_prove
    cudaStream_t side[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr;
    cudaEvent_t ev_join[4] = {nullptr, nullptr, nullptr, nullptr};
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

// counted launch helper: every kernel of this library goes through it so gpu_launches is an exact count
#define VSP_LAUNCH(ctx, kernel, grid, block, smem, strm, ...)        \
    do {                                                             \
        kernel<<<(grid), (block), (smem), (strm)>>>(__VA_ARGS__);    \
        ++(ctx)->launches;                                           \
    } while (0)

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
int ntt_pow2(vsp_ctx* ctx, cudaStream_t s, const uint32_t* d_in, uint32_t* d_out, unsigned log_n, bool inverse,
             const uint32_t* d_pre, const uint32_t* d_post);
int domain_fft_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, bool inverse, uint32_t* d_inout);
int multiply_by_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t n, const uint64_t g[4], uint32_t* d_inout);
int divide_by_z_on_coset_dev(vsp_ctx* ctx, cudaStream_t s, uint64_t m, uint32_t* d_inout);

}  // namespace vsp
