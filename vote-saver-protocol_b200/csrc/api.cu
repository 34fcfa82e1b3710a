// api.cu — the extern "C" boundary of libvsp_b200.so (declared in include/vsp_b200.h): context management and the
// host-buffer / device-buffer wrappers around the NTT, domain and multiexp layers.  The prover entry points are in
// prover.cu.
#include "common.cuh"
#include "fp.cuh"
#include "msm.cuh"

#include <new>

using namespace vsp;

namespace vsp {
int domain_plan(uint64_t min_size, int* kind, uint64_t* m);  // ntt.cu
int bases_generate_g1(vsp_ctx*, cudaStream_t, uint32_t*, uint64_t, uint64_t);  // msm_g1.cu
int bases_read_g1(vsp_ctx*, cudaStream_t, const uint32_t*, uint64_t, uint64_t, uint64_t*);
int points_generate_g2(vsp_ctx*, cudaStream_t, uint64_t*, uint64_t, uint64_t);  // msm_g2.cu
}

struct vsp_g1_bases {
    BasesDev b;
};
struct vsp_g2_bases {
    BasesDev b;
};

// ---- live per-kernel-class timing ------------------------------------------------------------------
static void prof_drain(vsp_ctx* ctx) {
    KProf& p = ctx->prof;
    for (KProfRec& r : p.recs) {
        float ms = 0.f;
        if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
            p.ms[r.kc] += ms;
            p.n[r.kc] += 1;
            p.units[r.kc] += r.slot >= 0 ? r.scale * (double)p.pinned[r.slot] : r.scale;
        }
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    p.recs.clear();
    p.pinned_used = 0;
    cudaGetLastError();
}


extern "C" {

const char* vsp_version(void) { return "vsp_b200 0.1 (sm_100a)"; }

int vsp_ctx_create(int device, void* cuda_stream, vsp_ctx** out) {
    if (!out) return VSP_E_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return VSP_E_NO_DEVICE;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        return VSP_E_NO_DEVICE;
    }
    vsp_ctx* ctx = new (std::nothrow) vsp_ctx();
    if (!ctx) return VSP_E_NOMEM;
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    if (cuda_stream) {
        ctx->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
        ctx->own_stream = false;
    } else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete ctx;
            return VSP_E_NO_DEVICE;
        }
        ctx->own_stream = true;
    }
    for (int i = 0; i < 4; ++i) {
        cudaStreamCreateWithFlags(&ctx->side[i], cudaStreamNonBlocking);
        cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming);
    }
    cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ctx->ev_main, cudaEventDisableTiming);
    for (int i = 0; i < 2; ++i) cudaEventCreateWithFlags(&ctx->tail[i].done, cudaEventDisableTiming);
    // scratch comes from the context's own block cache (common.cuh), created on first use: the process's default
    // memory pool is left as the host application configured it
    if (cudaGetLastError() != cudaSuccess) {
        delete ctx;
        return VSP_E_NO_DEVICE;
    }
    *out = ctx;
    return VSP_OK;
}

void vsp_ctx_destroy(vsp_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 4; ++i)
        if (ctx->side[i]) cudaStreamSynchronize(ctx->side[i]);
    ntt_tables_free(ctx);
    prof_drain(ctx);
    if (ctx->prof.pinned) cudaFreeHost(ctx->prof.pinned);
    for (int i = 0; i < 4; ++i) {
        if (ctx->side[i]) cudaStreamDestroy(ctx->side[i]);
        if (ctx->ev_join[i]) cudaEventDestroy(ctx->ev_join[i]);
    }
    if (ctx->copy) {
        cudaStreamSynchronize(ctx->copy);
        cudaStreamDestroy(ctx->copy);
        ctx->copy = nullptr;
    }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_main) cudaEventDestroy(ctx->ev_main);
    for (int i = 0; i < 2; ++i) {
        if (ctx->tail[i].done) cudaEventDestroy(ctx->tail[i].done);
        if (ctx->tail[i].p) cudaFree(ctx->tail[i].p);
        if (i == 0 && ctx->partial_bsc) cudaFree(ctx->partial_bsc);
    }
    for (int i = 0; i < 4; ++i) {
        if (ctx->stage[i].done) cudaEventDestroy(ctx->stage[i].done);
        if (ctx->stage[i].p) cudaFreeHost(ctx->stage[i].p);
    }
    if (ctx->scratch) {
        scratch_trim(ctx);
        delete ctx->scratch;
        ctx->scratch = nullptr;
    }
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* vsp_last_error(const vsp_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
const char* vsp_last_warning(const vsp_ctx* ctx) { return ctx ? ctx->warn.c_str() : ""; }
int vsp_ctx_set_batch_chunk(vsp_ctx* ctx, int ballots) {
    if (!ctx || ballots < 0 || ballots > 1024) return VSP_E_INVALID;
    ctx->batch_chunk = ballots;
    return VSP_OK;
}

int vsp_ctx_trim(vsp_ctx* ctx) {
    if (!ctx) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    scratch_trim(ctx);
    return VSP_OK;
}

int vsp_ctx_scratch_stats(const vsp_ctx* ctx, uint64_t* owned_bytes, uint64_t* unused_bytes, uint64_t* driver_allocations) {
    if (!ctx) return VSP_E_INVALID;
    const ScratchCache* c = ctx->scratch;
    if (owned_bytes) *owned_bytes = c ? c->owned_bytes : 0;
    if (unused_bytes) *unused_bytes = c ? c->unused_bytes : 0;
    if (driver_allocations) *driver_allocations = c ? c->driver_allocations : 0;
    return VSP_OK;
}

int vsp_ctx_synchronize(vsp_ctx* ctx) {
    if (!ctx) return VSP_E_INVALID;
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

uint64_t vsp_ctx_launch_count(const vsp_ctx* ctx) { return ctx ? ctx->launches : 0; }

int vsp_ctx_profile_enable(vsp_ctx* ctx, int on) {
    if (!ctx) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 4; ++i) cudaStreamSynchronize(ctx->side[i]);
    prof_drain(ctx);
    KProf& p = ctx->prof;
    if (on && !p.pinned) {
        p.pinned_cap = 1 << 16;
        if (cudaMallocHost(&p.pinned, (size_t)p.pinned_cap * 4) != cudaSuccess) {
            p.pinned = nullptr;
            p.pinned_cap = 0;
            cudaGetLastError();
        }
    }
    for (int k = 0; k < KC_COUNT; ++k) {
        p.ms[k] = 0;
        p.units[k] = 0;
        p.n[k] = 0;
    }
    p.on = on != 0;
    return VSP_OK;
}

int vsp_ctx_profile_read(vsp_ctx* ctx, int kclass, uint64_t* launches, double* total_ms, double* units) {
    if (!ctx || kclass < 0 || kclass >= KC_COUNT) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    prof_drain(ctx);
    if (launches) *launches = ctx->prof.n[kclass];
    if (total_ms) *total_ms = ctx->prof.ms[kclass];
    if (units) *units = ctx->prof.units[kclass];
    return VSP_OK;
}

// ---- domains ------------------------------------------------------------------------------------
int vsp_domain_plan(uint64_t min_size, int* kind, uint64_t* m) {
    if (!kind || !m) return VSP_E_INVALID;
    return domain_plan(min_size, kind, m);
}

static int check_domain_size(vsp_ctx* ctx, uint64_t m, int* kind) {
    uint64_t mm = 0;
    if (domain_plan(m, kind, &mm) != VSP_OK || mm != m)
        return set_error(ctx, VSP_E_DOMAIN, "%llu is not the size of a radix-2 family domain", (unsigned long long)m);
    if (*kind == VSP_DOMAIN_EXTENDED_RADIX2)
        return set_error(ctx, VSP_E_DOMAIN, "extended_radix2 domains (log m = 33) are not executable");
    return VSP_OK;
}

int vsp_fr_fft_dev(vsp_ctx* ctx, uint64_t m, int inverse, uint64_t* d_inout) {
    if (!ctx || !d_inout) return VSP_E_INVALID;
    int kind;
    VSP_TRY(check_domain_size(ctx, m, &kind));
    cudaSetDevice(ctx->device);
    return domain_fft_dev(ctx, ctx->stream, m, inverse != 0, reinterpret_cast<uint32_t*>(d_inout));
}

int vsp_fr_fft(vsp_ctx* ctx, uint64_t m, int inverse, const uint64_t* in, uint64_t n_in, uint64_t* out) {
    if (!ctx || !in || !out) return VSP_E_INVALID;
    int kind;
    VSP_TRY(check_domain_size(ctx, m, &kind));
    // basic_radix2_domain zero-pads short input and throws on long input (basic_radix2_domain.hpp:70-77);
    // step_radix2_domain::inverse_fft throws unless a.size() == m (step_radix2_domain.hpp:112-114)
    if (n_in > m) return set_error(ctx, VSP_E_INVALID, "expected a.size() <= m");
    if (inverse && kind == VSP_DOMAIN_STEP_RADIX2 && n_in != m)
        return set_error(ctx, VSP_E_INVALID, "step_radix2: expected a.size() == this->m");
    cudaSetDevice(ctx->device);
    DevBuf d;
    VSP_TRY(d.alloc(ctx, m * 32, ctx->stream));
    if (n_in < m) VSP_CUDA_CHECK(ctx, cudaMemsetAsync(d.p, 0, m * 32, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d.p, in, n_in * 32, cudaMemcpyHostToDevice, ctx->stream));
    VSP_TRY(domain_fft_dev(ctx, ctx->stream, m, inverse != 0, d.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out, d.p, m * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

int vsp_fr_multiply_by_coset_dev(vsp_ctx* ctx, uint64_t n, const uint64_t g[4], uint64_t* d_inout) {
    if (!ctx || !g || !d_inout) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return multiply_by_coset_dev(ctx, ctx->stream, n, g, reinterpret_cast<uint32_t*>(d_inout));
}

int vsp_fr_multiply_by_coset(vsp_ctx* ctx, uint64_t n, const uint64_t g[4], uint64_t* a_inout) {
    if (!ctx || !g || (!a_inout && n)) return VSP_E_INVALID;
    if (n == 0) return VSP_OK;
    cudaSetDevice(ctx->device);
    DevBuf d;
    VSP_TRY(d.alloc(ctx, n * 32, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d.p, a_inout, n * 32, cudaMemcpyHostToDevice, ctx->stream));
    VSP_TRY(multiply_by_coset_dev(ctx, ctx->stream, n, g, d.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(a_inout, d.p, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

int vsp_fr_divide_by_z_on_coset_dev(vsp_ctx* ctx, uint64_t m, uint64_t* d_inout) {
    if (!ctx || !d_inout) return VSP_E_INVALID;
    int kind;
    VSP_TRY(check_domain_size(ctx, m, &kind));
    cudaSetDevice(ctx->device);
    return divide_by_z_on_coset_dev(ctx, ctx->stream, m, reinterpret_cast<uint32_t*>(d_inout));
}

int vsp_fr_divide_by_z_on_coset(vsp_ctx* ctx, uint64_t m, uint64_t* a_inout) {
    if (!ctx || !a_inout) return VSP_E_INVALID;
    int kind;
    VSP_TRY(check_domain_size(ctx, m, &kind));
    cudaSetDevice(ctx->device);
    DevBuf d;
    VSP_TRY(d.alloc(ctx, m * 32, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d.p, a_inout, m * 32, cudaMemcpyHostToDevice, ctx->stream));
    VSP_TRY(divide_by_z_on_coset_dev(ctx, ctx->stream, m, d.as<uint32_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(a_inout, d.p, m * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

// ---- multiexp -------------------------------------------------------------------------------------
static int bases_create(vsp_ctx* ctx, bool g2, uint64_t n, const uint64_t* bases, BasesDev* b, bool table = true) {
    if (!ctx || (!bases && n)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return bases_upload(ctx, g2, n, bases, b, table);
}

int vsp_g1_bases_create_ex(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, int table, vsp_g1_bases** out);
int vsp_g2_bases_create_ex(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, int table, vsp_g2_bases** out);
int vsp_g1_bases_create(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, vsp_g1_bases** out) {
    return vsp_g1_bases_create_ex(ctx, n, bases, 1, out);
}
int vsp_g2_bases_create(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, vsp_g2_bases** out) {
    return vsp_g2_bases_create_ex(ctx, n, bases, 1, out);
}
int vsp_g1_bases_create_ex(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, int table, vsp_g1_bases** out) {
    if (!out) return VSP_E_INVALID;
    vsp_g1_bases* h = new (std::nothrow) vsp_g1_bases();
    if (!h) return VSP_E_NOMEM;
    int rc = bases_create(ctx, false, n, bases, &h->b, table != 0);
    if (rc) {
        bases_free(&h->b);
        delete h;
        return rc;
    }
    *out = h;
    return VSP_OK;
}
void vsp_g1_bases_destroy(vsp_g1_bases* b) {
    if (!b) return;
    bases_free(&b->b);
    delete b;
}
int vsp_g2_bases_create_ex(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, int table, vsp_g2_bases** out) {
    if (!out) return VSP_E_INVALID;
    vsp_g2_bases* h = new (std::nothrow) vsp_g2_bases();
    if (!h) return VSP_E_NOMEM;
    int rc = bases_create(ctx, true, n, bases, &h->b, table != 0);
    if (rc) {
        bases_free(&h->b);
        delete h;
        return rc;
    }
    *out = h;
    return VSP_OK;
}
void vsp_g2_bases_destroy(vsp_g2_bases* b) {
    if (!b) return;
    bases_free(&b->b);
    delete b;
}

// synthetic bases for the measurement sweeps: bases[i] = k_i * G, k_i = splitmix64(seed + i) | 1
int vsp_g1_bases_generate_ex(vsp_ctx* ctx, uint64_t n, uint64_t seed, int table, vsp_g1_bases** out);
int vsp_g1_bases_generate(vsp_ctx* ctx, uint64_t n, uint64_t seed, vsp_g1_bases** out) {
    return vsp_g1_bases_generate_ex(ctx, n, seed, 1, out);
}
// table = 0: the points only (multiexps over the handle take the one-shot, table-less path)
int vsp_g1_bases_generate_ex(vsp_ctx* ctx, uint64_t n, uint64_t seed, int table, vsp_g1_bases** out) {
    if (!ctx || !out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    vsp_g1_bases* h = new (std::nothrow) vsp_g1_bases();
    if (!h) return VSP_E_NOMEM;
    h->b.ctx = ctx;
    h->b.n = n;
    h->b.g2 = false;
    cudaError_t e = cudaMalloc(&h->b.d_pts, n ? n * 96 : 16);
    if (e != cudaSuccess) {
        delete h;
        return set_error(ctx, VSP_E_NOMEM, "cudaMalloc(%llu bases) failed: %s", (unsigned long long)n, cudaGetErrorString(e));
    }
    int rc = n ? bases_generate_g1(ctx, ctx->stream, h->b.d_pts, n, seed) : VSP_OK;
    if (rc == VSP_OK) rc = bases_finalize(ctx, &h->b, table != 0);
    if (rc == VSP_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
        rc = set_error(ctx, VSP_E_CUDA, "bases generation failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc) {
        bases_free(&h->b);
        delete h;
        return rc;
    }
    *out = h;
    return VSP_OK;
}
// canonical affine copy of bases [first, first + count) to the host
int vsp_g1_bases_read(vsp_ctx* ctx, const vsp_g1_bases* b, uint64_t first, uint64_t count, uint64_t* out) {
    if (!ctx || !b || (!out && count)) return VSP_E_INVALID;
    if (first > b->b.n || count > b->b.n - first) return set_error(ctx, VSP_E_INVALID, "bases range out of bounds");
    if (count == 0) return VSP_OK;
    cudaSetDevice(ctx->device);
    DevBuf d;
    VSP_TRY(d.alloc(ctx, count * 96, ctx->stream));
    VSP_TRY(bases_read_g1(ctx, ctx->stream, b->b.d_pts, first, count, d.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out, d.p, count * 96, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

int vsp_g2_points_generate(vsp_ctx* ctx, uint64_t n, uint64_t seed, uint64_t* out) {
    if (!ctx || (!out && n)) return VSP_E_INVALID;
    if (n == 0) return VSP_OK;
    cudaSetDevice(ctx->device);
    DevBuf d;
    VSP_TRY(d.alloc(ctx, n * 192, ctx->stream));
    VSP_TRY(points_generate_g2(ctx, ctx->stream, d.as<uint64_t>(), n, seed));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out, d.p, n * 192, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

static int msm_bases_host_out(vsp_ctx* ctx, const BasesDev& b, uint64_t n, const uint64_t* scalars, int on_device,
                              uint64_t* out) {
    if (!ctx || (!scalars && n) || !out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    const size_t out_bytes = b.g2 ? 192 : 96;
    DevBuf d_sc, d_out;
    const uint32_t* sc = reinterpret_cast<const uint32_t*>(scalars);
    if (!on_device && n) {
        VSP_TRY(d_sc.alloc(ctx, n * 32, ctx->stream));
        VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_sc.p, scalars, n * 32, cudaMemcpyHostToDevice, ctx->stream));
        sc = d_sc.as<uint32_t>();
    }
    VSP_TRY(d_out.alloc(ctx, out_bytes, ctx->stream));
    VSP_TRY(msm_run(ctx, ctx->stream, b, n, sc, d_out.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out, d_out.p, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}

int vsp_g1_msm_bases(vsp_ctx* ctx, const vsp_g1_bases* b, uint64_t n_scalars, const uint64_t* scalars,
                     int scalars_on_device, uint64_t* out) {
    if (!b) return VSP_E_INVALID;
    return msm_bases_host_out(ctx, b->b, n_scalars, scalars, scalars_on_device, out);
}
int vsp_g2_msm_bases(vsp_ctx* ctx, const vsp_g2_bases* b, uint64_t n_scalars, const uint64_t* scalars,
                     int scalars_on_device, uint64_t* out) {
    if (!b) return VSP_E_INVALID;
    return msm_bases_host_out(ctx, b->b, n_scalars, scalars, scalars_on_device, out);
}
int vsp_g1_msm_bases_dev(vsp_ctx* ctx, const vsp_g1_bases* b, uint64_t n_scalars, const uint64_t* d_scalars,
                         uint64_t* d_out) {
    if (!ctx || !b || !d_out || (!d_scalars && n_scalars)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return msm_run(ctx, ctx->stream, b->b, n_scalars, reinterpret_cast<const uint32_t*>(d_scalars), d_out);
}

int vsp_g2_msm_bases_dev(vsp_ctx* ctx, const vsp_g2_bases* b, uint64_t n_scalars, const uint64_t* d_scalars,
                         uint64_t* d_out) {
    if (!ctx || !b || !d_out || (!d_scalars && n_scalars)) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return msm_run(ctx, ctx->stream, b->b, n_scalars, reinterpret_cast<const uint32_t*>(d_scalars), d_out);
}

// one-shot multiexp: the bases are used once, so no window table is built for them
static int msm_one_shot(vsp_ctx* ctx, bool g2, uint64_t n, const uint64_t* bases, const uint64_t* scalars, uint64_t* out) {
    BasesDev b;
    int rc = bases_create(ctx, g2, n, bases, &b, false);
    if (rc == VSP_OK) rc = msm_bases_host_out(ctx, b, n, scalars, 0, out);
    bases_free(&b);
    return rc;
}
int vsp_g1_msm(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, const uint64_t* scalars, uint64_t* out) {
    return msm_one_shot(ctx, false, n, bases, scalars, out);
}
int vsp_g2_msm(vsp_ctx* ctx, uint64_t n, const uint64_t* bases, const uint64_t* scalars, uint64_t* out) {
    return msm_one_shot(ctx, true, n, bases, scalars, out);
}

static int points_sum_host(vsp_ctx* ctx, bool g2, uint64_t k, const uint64_t* points, uint64_t* out) {
    if (!ctx || (!points && k) || !out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    const size_t pb = g2 ? 192 : 96;
    DevBuf d_in, d_out;
    VSP_TRY(d_in.alloc(ctx, k * pb, ctx->stream));
    VSP_TRY(d_out.alloc(ctx, pb, ctx->stream));
    if (k) VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_in.p, points, k * pb, cudaMemcpyHostToDevice, ctx->stream));
    VSP_TRY(points_sum_dev(ctx, ctx->stream, g2, k, d_in.as<uint64_t>(), d_out.as<uint64_t>()));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(out, d_out.p, pb, cudaMemcpyDeviceToHost, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return VSP_OK;
}
int vsp_g1_sum(vsp_ctx* ctx, uint64_t k, const uint64_t* points, uint64_t* out) {
    return points_sum_host(ctx, false, k, points, out);
}
int vsp_g2_sum(vsp_ctx* ctx, uint64_t k, const uint64_t* points, uint64_t* out) {
    return points_sum_host(ctx, true, k, points, out);
}
int vsp_g1_sum_dev(vsp_ctx* ctx, uint64_t k, const uint64_t* d_points, uint64_t* d_out) {
    if (!ctx || (!d_points && k) || !d_out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return points_sum_dev(ctx, ctx->stream, false, k, d_points, d_out);
}
int vsp_g2_sum_dev(vsp_ctx* ctx, uint64_t k, const uint64_t* d_points, uint64_t* d_out) {
    if (!ctx || (!d_points && k) || !d_out) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    return points_sum_dev(ctx, ctx->stream, true, k, d_points, d_out);
}

}  // extern "C"

// ---- integer-pipe peak -------------------------------------------------------------------------------
// Each thread runs 8 independent 64-bit multiply-accumulate chains (mad.lo.cc / madc.hi pairs, which ptxas fuses
// into IMAD.WIDE.U32), so the measurement is issue-bound, not latency-bound.  One "multiply-add" = one 32x32->64
// product accumulated into 64 bits: the unit of SURVEY.md 8d's I_mul(n) = 2n^2 + n.
__global__ void __launch_bounds__(256) k_imad_rate(uint32_t* sink, uint32_t iters, uint32_t seed) {
    uint32_t a = seed + threadIdx.x, b = seed * 3u + blockIdx.x;
    uint32_t lo[8], hi[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        lo[i] = a + i;
        hi[i] = b + i;
    }
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                lo[i] = ptx::mad_lo_cc(a, b, lo[i]);
                hi[i] = ptx::madc_hi(a, b, hi[i]);
            }
            a += 0x9e3779b9u;
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= lo[i] ^ hi[i];
    if (x == 0x12345678u) sink[0] = x;  // keep the chains live
}

// ---- the second arithmetic pipe: FP64 fused multiply-adds (the unit of a 52-bit-limb DFMA Montgomery product) --------
__global__ void __launch_bounds__(256) k_dfma_rate(double* sink, uint32_t iters, double seed) {
    double a = seed + threadIdx.x * 1e-9, b = 1.0 + blockIdx.x * 1e-12;
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = a + i;
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], b, a);
            a += 1e-7;
        }
    }
    double x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x += acc[i];
    if (x == 0.12345) sink[0] = x;
}
extern "C" int vsp_measure_dfma_rate(vsp_ctx* ctx, uint64_t iters, double* fmas_per_second) {
    if (!ctx || !fmas_per_second || iters == 0 || iters > 0xffffffffull) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf sink;
    VSP_TRY(sink.alloc(ctx, 16, ctx->stream));
    const unsigned blocks = ctx->sm_count * 8, threads = 256;
    cudaEvent_t e0, e1;
    VSP_CUDA_CHECK(ctx, cudaEventCreate(&e0));
    VSP_CUDA_CHECK(ctx, cudaEventCreate(&e1));
    VSP_LAUNCH(ctx, k_dfma_rate, blocks, threads, 0, ctx->stream, sink.as<double>(), (uint32_t)(iters / 8 + 1), 1.5);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(e0, ctx->stream));
    VSP_LAUNCH(ctx, k_dfma_rate, blocks, threads, 0, ctx->stream, sink.as<double>(), (uint32_t)iters, 2.5);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(e1, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaEventSynchronize(e1));
    float ms = 0;
    VSP_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *fmas_per_second = (double)blocks * threads * (double)iters * 64.0 / (ms * 1e-3);
    return VSP_OK;
}

extern "C" int vsp_accumulate_mads_per_add(vsp_ctx* ctx, double* out) {
    if (!ctx || !out) return VSP_E_INVALID;
    *out = ACC_G1_MADS_PER_ADD;
    return VSP_OK;
}

extern "C" int vsp_measure_imad_rate(vsp_ctx* ctx, uint64_t iters, double* mads_per_second) {
    if (!ctx || !mads_per_second || iters == 0 || iters > 0xffffffffull) return VSP_E_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf sink;
    VSP_TRY(sink.alloc(ctx, 16, ctx->stream));
    const unsigned blocks = ctx->sm_count * 8, threads = 256;
    cudaEvent_t e0, e1;
    VSP_CUDA_CHECK(ctx, cudaEventCreate(&e0));
    VSP_CUDA_CHECK(ctx, cudaEventCreate(&e1));
    VSP_LAUNCH(ctx, k_imad_rate, blocks, threads, 0, ctx->stream, sink.as<uint32_t>(), (uint32_t)(iters / 8 + 1), 1u);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(e0, ctx->stream));
    VSP_LAUNCH(ctx, k_imad_rate, blocks, threads, 0, ctx->stream, sink.as<uint32_t>(), (uint32_t)iters, 7u);
    VSP_CUDA_CHECK(ctx, cudaEventRecord(e1, ctx->stream));
    VSP_CUDA_CHECK(ctx, cudaEventSynchronize(e1));
    float ms = 0;
    VSP_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double mads = (double)blocks * threads * (double)iters * 64.0;
    *mads_per_second = mads / (ms * 1e-3);
    return VSP_OK;
}
