// msm_impl.cuh — signed-digit bucket multiexp (Pippenger) over BLS12-381 G1 / G2 on sm_100a, batched over ballots.
//
// Replaces algebra::multiexp<policies::multiexp_method_BDLO12> and multiexp_with_mixed_addition (un-vendored
// crypto3-algebra @ e4fa2c5e; signature from the call site
// /root/reference/libs/containers/include/nil/crypto3/container/sparse_vector.hpp:258-261) and the prover's
// multiexps over A/B/H/L_query.  Any correct algorithm yields the same group element; results
// leave as canonical affine points, which is what the reference's encoders emit
// (libs/marshalling/algebra/.../processing/curve_element.hpp:112-113).
//
// Resident bases (a proving-key query) carry a table of window multiples  T[w][i] = 2^(c w) * P_i  built once at
// load time, so every window of every scalar drops into ONE bucket set per ballot and no doubling happens at
// multiexp time.  One-shot bases (vsp_g*_msm) have no table: one bucket set per window, joined by a Horner pass.
//
// Pipeline for `batch` scalar vectors over the same bases (all stream-ordered, no host round trip):
//   1 digits      scalar -> W signed c-bit digits; key = (ballot, set, |d| - 1); histogram by key
//   2 scan        exclusive prefix sums -> bucket offsets; buckets split into tasks of <= TASK_POINTS points
//   3 scatter     counting sort of (table index | sign) by key
//   4 accumulate  one thread per task: XYZZ mixed additions of the gathered affine points      <- the hot kernel
//   5 collapse    buckets split over several tasks are summed (thread per bucket; block per very heavy bucket)
//   6 reduce      per bucket set: sum_v v * B_v = sum_j 2^j * (sum of buckets whose value has bit j) — c parallel
//                 tree sums instead of a serial running sum — then a c-step Horner per set
//   7 final       Horner over the sets of a ballot (one-shot bases only), to affine, out of Montgomery form
// 0 and 1 scalars need no special path: a zero contributes no digit, a one lands in bucket (set 0, value 1), which
// step 5 reduces in parallel — the reference's `_with_mixed_addition` pre-pass (skip 0, add 1) is subsumed.
#pragma once
#include <algorithm>
#include <type_traits>

#include "ec.cuh"
#include "msm.cuh"

namespace vsp {
namespace {

constexpr unsigned TASK_POINTS = 64;  // max points one accumulate thread sums
constexpr unsigned KEY_NONE = 0xffffffffu;
constexpr unsigned HEAVY_TASKS = 32;  // buckets with more partials than this are collapsed by a whole block

struct MsmPlan {
    unsigned c;        // window bits
    unsigned windows;  // 255 / c + 1
    unsigned nb_log;   // c - 1: buckets per set = 2^(c-1), bucket index v - 1 for digit magnitude v in [1, 2^(c-1)]
    unsigned sets;     // bucket sets per ballot: 1 with a window table, `windows` without
    unsigned batch;    // ballots
    uint64_t stride;   // table stride between windows (bases.n), 0 without a table
    uint64_t buckets;  // batch * sets << nb_log
};

// ---------------------------------------------------------------------------------------------
// 1. digits + histogram
// ---------------------------------------------------------------------------------------------
constexpr uint32_t AGG_BUCKETS = 32;
__global__ void k_msm_digits(const uint32_t* __restrict__ scalars, uint64_t n, uint64_t scalar_stride, MsmPlan p,
                             uint32_t* __restrict__ keys, uint32_t* __restrict__ counts, int count_here) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const unsigned b = blockIdx.y;
    if (i >= n) return;
    const uint32_t* sp = scalars + 8 * (uint64_t(b) * scalar_stride + i);
    uint32_t s[9];
    const uint4 lo = reinterpret_cast<const uint4*>(sp)[0];
    const uint4 hi = reinterpret_cast<const uint4*>(sp)[1];
    s[0] = lo.x; s[1] = lo.y; s[2] = lo.z; s[3] = lo.w;
    s[4] = hi.x; s[5] = hi.y; s[6] = hi.z; s[7] = hi.w;
    s[8] = 0;
    const uint32_t mask = (1u << p.c) - 1u, half = 1u << (p.c - 1);
    uint32_t carry = 0;
    for (unsigned w = 0; w < p.windows; ++w) {
        const unsigned bit = w * p.c;
        uint32_t d = 0;
        if (bit < 256) {
            const unsigned limb = bit >> 5, sh = bit & 31;
            uint64_t two = (uint64_t(s[limb + 1]) << 32) | s[limb];
            d = (uint32_t)(two >> sh) & mask;
        }
        d += carry;
        const uint32_t set = b * p.sets + (p.sets > 1 ? w : 0u);
        uint32_t key = KEY_NONE;
        if (d > half) {
            const uint32_t mag = (1u << p.c) - d;  // negative digit; d == 2^c (all-ones + carry) is digit 0, carry 1
            carry = 1;
            if (mag) key = ((set << p.nb_log) + mag - 1u) | 0x80000000u;
        } else {
            carry = 0;
            if (d) key = (set << p.nb_log) + d - 1u;
        }
        keys[(uint64_t(b) * p.windows + w) * n + i] = key;
        // a 0/1 (or small-valued) witness sends every lane to the same few counters: for the lowest AGG_BUCKETS
        // buckets of a set, one atomic per distinct bucket in the warp; uniform digits take the plain atomic
        if (count_here && key != KEY_NONE) {
            const uint32_t bk = key & 0x7fffffffu;
            if ((bk & ((1u << p.nb_log) - 1u)) < AGG_BUCKETS) {
                const unsigned grp = __match_any_sync(__activemask(), bk);
                if ((threadIdx.x & 31u) == unsigned(__ffs(grp) - 1)) atomicAdd(&counts[bk], (uint32_t)__popc(grp));
            } else {
                atomicAdd(&counts[bk], 1u);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// 2. scans
// ---------------------------------------------------------------------------------------------
constexpr unsigned SCAN_THREADS = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// MODE 0: values = in[i];  MODE 1: values = ceil(in[i] / TASK_POINTS)
template <int MODE>
__device__ __forceinline__ uint32_t scan_value(const uint32_t* in, uint64_t i, uint64_t n) {
    if (i >= n) return 0;
    const uint32_t v = in[i];
    return MODE == 0 ? v : (v + TASK_POINTS - 1) / TASK_POINTS;
}

template <int MODE>
__global__ void k_scan_tile_sums(const uint32_t* in, uint64_t n, uint32_t* tile_sums) {
    __shared__ uint32_t red[SCAN_THREADS];
    const uint64_t base = uint64_t(blockIdx.x) * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (unsigned k = 0; k < SCAN_ITEMS; ++k) s += scan_value<MODE>(in, base + k, n);
    red[threadIdx.x] = s;
    __syncthreads();
    for (unsigned st = SCAN_THREADS / 2; st > 0; st >>= 1) {
        if (threadIdx.x < st) red[threadIdx.x] += red[threadIdx.x + st];
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = red[0];
}

// single block: exclusive scan of tile sums in place, total to tile_sums[tiles]
__global__ void k_scan_tiles(uint32_t* tile_sums, unsigned tiles) {
    __shared__ uint32_t buf[1024];
    __shared__ uint32_t running;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    for (unsigned base = 0; base < tiles; base += 1024) {
        const unsigned i = base + threadIdx.x;
        const uint32_t v = i < tiles ? tile_sums[i] : 0;
        buf[threadIdx.x] = v;
        __syncthreads();
        for (unsigned off = 1; off < 1024; off <<= 1) {
            uint32_t t = threadIdx.x >= off ? buf[threadIdx.x - off] : 0;
            __syncthreads();
            buf[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < tiles) tile_sums[i] = running + buf[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) running += buf[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_sums[tiles] = running;
}

template <int MODE>
__global__ void k_scan_apply(const uint32_t* in, uint64_t n, const uint32_t* tile_sums, uint32_t* out) {
    __shared__ uint32_t buf[SCAN_THREADS];
    const uint64_t base = uint64_t(blockIdx.x) * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (unsigned k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = scan_value<MODE>(in, base + k, n);
        s += v[k];
    }
    buf[threadIdx.x] = s;
    __syncthreads();
    for (unsigned off = 1; off < SCAN_THREADS; off <<= 1) {
        uint32_t t = threadIdx.x >= off ? buf[threadIdx.x - off] : 0;
        __syncthreads();
        buf[threadIdx.x] += t;
        __syncthreads();
    }
    uint32_t run = tile_sums[blockIdx.x] + buf[threadIdx.x] - s;
#pragma unroll
    for (unsigned k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
    // one past the end: grand total
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == SCAN_THREADS - 1) out[n] = tile_sums[gridDim.x];
}

template <int MODE>
static int exclusive_scan(vsp_ctx* ctx, cudaStream_t s, const uint32_t* d_in, uint64_t n, uint32_t* d_out,
                          uint32_t* d_tile_scratch) {
    const unsigned tiles = (unsigned)((n + SCAN_TILE - 1) / SCAN_TILE);
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, (k_scan_tile_sums<MODE>), tiles, SCAN_THREADS, 0, s, d_in, n, d_tile_scratch);
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_scan_tiles, 1, 1024, 0, s, d_tile_scratch, tiles);
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, (k_scan_apply<MODE>), tiles, SCAN_THREADS, 0, s, d_in, n, d_tile_scratch, d_out);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// 3. scatter (counting sort) and task list
// ---------------------------------------------------------------------------------------------
__global__ void k_msm_scatter(const uint32_t* __restrict__ keys, uint64_t total, uint64_t n, MsmPlan p,
                              const uint32_t* __restrict__ offsets, uint32_t* __restrict__ cursors,
                              uint32_t* __restrict__ sorted) {
    const uint64_t q = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= total) return;
    const uint32_t key = keys[q];
    if (key == KEY_NONE) return;
    const uint32_t bk = key & 0x7fffffffu;
    uint32_t pos;
    if ((bk & ((1u << p.nb_log) - 1u)) < AGG_BUCKETS) {
        // warp-aggregated cursor (see k_msm_digits): lanes holding the same bucket take consecutive slots from one atomic
        const unsigned grp = __match_any_sync(__activemask(), bk);
        const unsigned lane = threadIdx.x & 31u, leader = unsigned(__ffs(grp) - 1);
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&cursors[bk], (uint32_t)__popc(grp));
        base = __shfl_sync(grp, base, leader);
        pos = offsets[bk] + base + (uint32_t)__popc(grp & ((1u << lane) - 1u));
    } else {
        pos = offsets[bk] + atomicAdd(&cursors[bk], 1u);
    }
    const uint64_t row = q / n;  // ballot * windows + window
    const uint32_t i = (uint32_t)(q - row * n);
    const uint32_t w = (uint32_t)(row % p.windows);
    sorted[pos] = (uint32_t)(i + w * p.stride) | (key & 0x80000000u);  // index into the window table
}

// Counting sort with the counters in shared memory.  With a window table a ballot owns ONE bucket set of 2^(c-1) <=
// 32,768 buckets, i.e. <= 128 KB of 32-bit counters: a CTA keeps them on chip, so the two passes over the keys
// (histogram, then scatter with running cursors) cost shared-memory atomics instead of one L2 atomic round trip per
// key (k_msm_scatter: 173 cycles of long-scoreboard stall per issue in the round-1 ncu capture).  The keys of ballot b
// are the contiguous range [b W n, (b + 1) W n); `split` CTAs share a ballot, each taking a contiguous slice, and
// leave their slice's histogram in hist_split so the scatter pass can start each slice's cursors after the slices
// before it.  Order inside a bucket is arbitrary, as it already was: bucket sums do not depend on it.
constexpr unsigned SET_SORT_THREADS = 1024, SET_SORT_UNROLL = 8;
__global__ void __launch_bounds__(SET_SORT_THREADS) k_msm_hist_set(const uint32_t* __restrict__ keys, uint64_t per_set,
                                                                   unsigned nb_log, unsigned split,
                                                                   uint32_t* __restrict__ hist_split,
                                                                   uint32_t* __restrict__ counts) {
    extern __shared__ uint32_t sh_cnt[];
    const unsigned nb = 1u << nb_log, set = blockIdx.y, sp = blockIdx.x;
    for (unsigned j = threadIdx.x; j < nb; j += blockDim.x) sh_cnt[j] = 0;
    __syncthreads();
    const uint64_t slice = (per_set + split - 1) / split;
    const uint64_t lo = uint64_t(sp) * slice, hi = (lo + slice < per_set) ? lo + slice : per_set;
    const uint32_t* kp = keys + uint64_t(set) * per_set;
    // SET_SORT_UNROLL independent loads in flight per thread: the loop is latency-bound otherwise (one CTA per SM)
    for (uint64_t q0 = lo + threadIdx.x; q0 < hi; q0 += uint64_t(SET_SORT_UNROLL) * blockDim.x) {
        uint32_t k[SET_SORT_UNROLL];
#pragma unroll
        for (unsigned u = 0; u < SET_SORT_UNROLL; ++u) {
            const uint64_t q = q0 + uint64_t(u) * blockDim.x;
            k[u] = q < hi ? kp[q] : KEY_NONE;
        }
#pragma unroll
        for (unsigned u = 0; u < SET_SORT_UNROLL; ++u)
            if (k[u] != KEY_NONE) atomicAdd(&sh_cnt[k[u] & (nb - 1u)], 1u);
    }
    __syncthreads();
    uint32_t* hs = hist_split + (uint64_t(set) * split + sp) * nb;
    for (unsigned j = threadIdx.x; j < nb; j += blockDim.x) hs[j] = sh_cnt[j];
    (void)counts;
}
// counts[set][j] = sum over the set's slices of hist_split[set][slice][j]
__global__ void k_msm_sum_split(const uint32_t* __restrict__ hist_split, unsigned nb_log, unsigned split, uint64_t buckets,
                                uint32_t* __restrict__ counts) {
    const uint64_t g = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (g >= buckets) return;
    const uint64_t set = g >> nb_log, j = g & ((uint64_t(1) << nb_log) - 1);
    uint32_t c = 0;
    for (unsigned q = 0; q < split; ++q) c += hist_split[((set * split + q) << nb_log) + j];
    counts[g] = c;
}
__global__ void __launch_bounds__(SET_SORT_THREADS) k_msm_scatter_set(const uint32_t* __restrict__ keys, uint64_t per_set,
                                                                      uint64_t n, MsmPlan p, unsigned split,
                                                                      const uint32_t* __restrict__ hist_split,
                                                                      const uint32_t* __restrict__ offsets,
                                                                      uint32_t* __restrict__ sorted) {
    extern __shared__ uint32_t sh_cur[];
    const unsigned nb = 1u << p.nb_log, set = blockIdx.y, sp = blockIdx.x;
    const uint32_t* og = offsets + (uint64_t(set) << p.nb_log);
    for (unsigned j = threadIdx.x; j < nb; j += blockDim.x) {
        uint32_t c = og[j];
        for (unsigned q = 0; q < sp; ++q) c += hist_split[(uint64_t(set) * split + q) * nb + j];
        sh_cur[j] = c;
    }
    __syncthreads();
    const uint64_t slice = (per_set + split - 1) / split;
    const uint64_t lo = uint64_t(sp) * slice, hi = (lo + slice < per_set) ? lo + slice : per_set;
    const uint32_t* kp = keys + uint64_t(set) * per_set;
    // per_set < 2^32 (checked by the caller): 32-bit index arithmetic, one 32-bit division per key
    const uint32_t lo32 = (uint32_t)lo, hi32 = (uint32_t)hi, n32 = (uint32_t)n, stride32 = (uint32_t)p.stride;
    for (uint32_t q0 = lo32 + threadIdx.x; q0 < hi32; q0 += SET_SORT_UNROLL * SET_SORT_THREADS) {
        uint32_t k[SET_SORT_UNROLL];
#pragma unroll
        for (unsigned u = 0; u < SET_SORT_UNROLL; ++u) {
            const uint32_t q = q0 + u * SET_SORT_THREADS;
            k[u] = (q < hi32 && q >= q0) ? kp[q] : KEY_NONE;
        }
#pragma unroll
        for (unsigned u = 0; u < SET_SORT_UNROLL; ++u) {
            if (k[u] == KEY_NONE) continue;
            const uint32_t q = q0 + u * SET_SORT_THREADS;
            const uint32_t pos = atomicAdd(&sh_cur[k[u] & (nb - 1u)], 1u);
            const uint32_t w = q / n32;
            sorted[pos] = (q - w * n32 + w * stride32) | (k[u] & 0x80000000u);
        }
        if (q0 + SET_SORT_UNROLL * SET_SORT_THREADS < q0) break;  // 32-bit wrap
    }
}


struct Task {
    uint32_t begin, end;  // positions in `sorted`
};

// one thread per bucket: emit its tasks
__global__ void k_msm_tasks(const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ task_off, uint64_t buckets,
                            Task* __restrict__ tasks, uint32_t* __restrict__ heavy, uint32_t* __restrict__ heavy_count) {
    const uint64_t b = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= buckets) return;
    const uint32_t lo = offsets[b], hi = offsets[b + 1];
    const uint32_t t0 = task_off[b], t1 = task_off[b + 1];
    for (uint32_t t = t0; t < t1; ++t) {
        Task k;
        k.begin = lo + (t - t0) * TASK_POINTS;
        k.end = min(hi, k.begin + TASK_POINTS);
        tasks[t] = k;
    }
    if (t1 - t0 > HEAVY_TASKS) heavy[atomicAdd(heavy_count, 1u)] = (uint32_t)b;
}

// Tasks are handed to threads in order of decreasing length (a counting sort on the 1..TASK_POINTS lengths), so the
// 32 lanes of a warp run the same number of additions: with ~30 points per bucket the lengths are Poisson-spread
// and bucket order would leave a third of the lanes idle.  Only the assignment of tasks to threads changes; every
// task still sums its points in the same order into partial[task], so results are unaffected.
constexpr unsigned LEN_BINS = TASK_POINTS + 1;
__global__ void __launch_bounds__(256) k_task_hist(const Task* __restrict__ tasks, const uint32_t* __restrict__ n_tasks_ptr,
                                                   uint32_t* __restrict__ bins) {
    __shared__ uint32_t h[LEN_BINS];
    for (unsigned i = threadIdx.x; i < LEN_BINS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const uint32_t n = *n_tasks_ptr;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x)
        atomicAdd(&h[tasks[t].end - tasks[t].begin], 1u);
    __syncthreads();
    for (unsigned i = threadIdx.x; i < LEN_BINS; i += blockDim.x)
        if (h[i]) atomicAdd(&bins[i], h[i]);
}
// bins[len] -> start position of that length class, longest first; cursors follow at bins[LEN_BINS + len]
__global__ void k_task_bin_offsets(uint32_t* bins) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t run = 0;
    for (int len = (int)LEN_BINS - 1; len >= 0; --len) {
        const uint32_t c = bins[len];
        bins[LEN_BINS + len] = run;
        run += c;
    }
}
__global__ void __launch_bounds__(256) k_task_order(const Task* __restrict__ tasks, const uint32_t* __restrict__ n_tasks_ptr,
                                                    uint32_t* __restrict__ bins, uint32_t* __restrict__ order) {
    __shared__ uint32_t h[LEN_BINS], base[LEN_BINS];
    const uint32_t n = *n_tasks_ptr;
    const uint32_t per_block = (n + gridDim.x - 1) / gridDim.x;
    const uint32_t lo = blockIdx.x * per_block, hi = min(n, lo + per_block);
    for (unsigned i = threadIdx.x; i < LEN_BINS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (uint32_t t = lo + threadIdx.x; t < hi; t += blockDim.x) atomicAdd(&h[tasks[t].end - tasks[t].begin], 1u);
    __syncthreads();
    for (unsigned i = threadIdx.x; i < LEN_BINS; i += blockDim.x) {
        base[i] = h[i] ? atomicAdd(&bins[LEN_BINS + i], h[i]) : 0u;
        h[i] = 0;
    }
    __syncthreads();
    for (uint32_t t = lo + threadIdx.x; t < hi; t += blockDim.x) {
        const uint32_t len = tasks[t].end - tasks[t].begin;
        order[base[len] + atomicAdd(&h[len], 1u)] = t;
    }
}

// ---------------------------------------------------------------------------------------------
// 4. accumulate
// ---------------------------------------------------------------------------------------------
// lanes per thread of the batched-affine kernel (each lane keeps an affine accumulator and a prefix product in local
// memory: 144 B in G1, 288 B in G2)
template <class F>
struct AffineLanes {
    static constexpr int K = std::is_same<F, FqOps>::value ? 16 : 8;
};

// resident blocks per SM the accumulation kernel is compiled for: three in G1 (<= 168 registers per thread; the lazy
// reductions had pushed it to 172, i.e. two blocks), two in G2 (255 registers either way)
template <class F>
struct AccBlocks {
    static constexpr int MIN_BLOCKS = std::is_same<F, FqOps>::value ? 3 : 2;
};

template <class F>
__global__ void __launch_bounds__(128, AccBlocks<F>::MIN_BLOCKS) k_msm_accumulate(const uint32_t* __restrict__ table, const uint32_t* __restrict__ sorted,
                                                        const Task* __restrict__ tasks,
                                                        const uint32_t* __restrict__ order,
                                                        const uint32_t* __restrict__ n_tasks_ptr, uint32_t max_tasks,
                                                        uint32_t* __restrict__ partial) {
    // exact task count is device-resident (task_off[buckets]): no host round trip.  From max_tasks up the launch belongs
    // to k_msm_accumulate_affine.
    const uint32_t n_tasks = *n_tasks_ptr;
    if (n_tasks >= max_tasks) return;
    constexpr int PT_LIMBS = Affine<F>::BYTES / 4;
    // Grid-stride over the length-ordered tasks (a thread's successive tasks are `gridDim.x * blockDim.x` apart in the
    // order, so a warp still holds equal lengths).  By default the grid covers every task with its own thread;
    // VSP_MSM_ACC_BLOCKS = k caps it at k blocks per SM.  Measured on B200 (ballot circuit, 256-ballot steps): two blocks
    // per SM keep the integer pipe as busy as three (accumulation alone 118 -> 124 ms per step), but leaving that room to
    // the other streams' kernels does not shorten the step (1,161 against 1,210 ballots/s; one block per SM 1,062), and
    // neither do two or four independent stream sets proving disjoint halves of a batch concurrently
    // (tools/dual_pipeline_probe.py: 1,315 / 1,319 / 1,325 ballots/s): the step is the sum of its kernels' work.
    for (uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x; tid < n_tasks; tid += gridDim.x * blockDim.x) {
        const uint32_t t = order[tid];
        const Task k = tasks[t];
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t pos = k.begin; pos < k.end; ++pos) {
            const uint32_t v = sorted[pos];
            Affine<F> pt = Affine<F>::load(table + uint64_t(v & 0x7fffffffu) * PT_LIMBS);
            if (v & 0x80000000u) pt.y = F::neg(pt.y);
            ec_madd<F>(acc, pt);
        }
        acc.store(partial + uint64_t(t) * (XYZZ<F>::BYTES / 4));
    }
}

// The same sums in AFFINE coordinates on shared inversions (large launches: the H_query multiexp of a chunk of ballots,
// any multiexp over uniform scalars).  A thread owns K lanes; a lane sums one task at a time, point after point, and
// moves on to its next task when one is done.  In every step the thread performs one addition in each of its K lanes
// and all K share ONE field inversion (Montgomery's trick), which the constant-time division-step inverse makes cheap
// enough (ec.cuh: 5M + 1S + 1/K inversion per addition against the XYZZ mixed addition's 8M + 2S).
//   lane g = thread * K + k takes tasks order[g], order[g + lanes], order[g + 2 lanes], ...: `order` lists the tasks by
//   decreasing length, so the K lanes of a thread and the 32 threads of a warp always hold tasks of (nearly) the same
//   length, and every lane's total is one task from each stratum of the length distribution — the lanes finish together.
// Results land in partial[task] as XYZZ with ZZ = ZZZ = 1, so the collapse / reduction kernels do not care which kernel
// produced them.  Exceptional additions (P + P, P - P, an infinite base) take a divergent slow path; they do not occur
// for independent bases and only cost time in crafted inputs.
// one shared copy of the product / square for the affine kernel: its loop bodies are then a few hundred instructions of
// glue around calls instead of five inlined 6 KB multiplications, which is what the instruction cache wants
template <class F>
struct Shared {
    using E = typename F::T;
    static __device__ __noinline__ E mul(const E& a, const E& b) { return F::mul(a, b); }
    static __device__ __noinline__ E sqr(const E& a) { return F::sqr(a); }
};
template <class F>
VSP_DEV void affine_add_shared(typename F::T& x1, typename F::T& y1, const typename F::T& x2, const typename F::T& y2,
                               const typename F::T& dinv) {
    using E = typename F::T;
    const E lam = Shared<F>::mul(F::sub(y2, y1), dinv);
    const E x3 = F::sub(F::sub(Shared<F>::sqr(lam), x1), x2);
    y1 = F::sub(Shared<F>::mul(lam, F::sub(x1, x3)), y1);
    x1 = x3;
}

template <class F, int K>
__global__ void __launch_bounds__(128, 4) k_msm_accumulate_affine(const uint32_t* __restrict__ table,
                                                               const uint32_t* __restrict__ sorted,
                                                               const Task* __restrict__ tasks,
                                                               const uint32_t* __restrict__ order,
                                                               const uint32_t* __restrict__ n_tasks_ptr, uint32_t min_tasks,
                                                               uint32_t* __restrict__ partial) {
    using E = typename F::T;
    constexpr int PT_LIMBS = Affine<F>::BYTES / 4, XL = XYZZ<F>::BYTES / 4;
    const uint32_t n_tasks = *n_tasks_ptr;
    if (n_tasks < min_tasks) return;  // a small launch: k_msm_accumulate does it
    const uint32_t lanes = gridDim.x * blockDim.x * K;
    const uint32_t lane0 = (blockIdx.x * blockDim.x + threadIdx.x) * K;
    if (lane0 >= n_tasks) return;
    E ax[K], ay[K], pref[K];
    uint32_t t_next[K], t_cur[K], pos[K], end[K];
    uint32_t live = 0;  // bit k: lane k's accumulator holds a point (clear: the point at infinity)
    uint32_t busy = 0;  // bit k: lane k has a current task
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
        t_next[k] = lane0 + k;
        pos[k] = end[k] = 0;
        t_cur[k] = 0;
    }
    const E one = F::one();
    for (;;) {
        // ---- lanes whose task is done store it and open their next one
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            const uint32_t bit = 1u << k;
            if (pos[k] != end[k]) continue;
            if (busy & bit) {
                XYZZ<F> r = (live & bit) ? XYZZ<F>{ax[k], ay[k], one, one} : XYZZ<F>::inf();
                r.store(partial + uint64_t(t_cur[k]) * XL);
                busy &= ~bit;
            }
            if (t_next[k] < n_tasks) {
                const uint32_t t = order[t_next[k]];
                t_next[k] += lanes;
                const Task tk = tasks[t];
                t_cur[k] = t;
                pos[k] = tk.begin;
                end[k] = tk.end;
                busy |= bit;
                live &= ~bit;
                if (pos[k] < end[k]) {  // the task's first point opens the accumulator
                    const uint32_t v = sorted[pos[k]++];
                    Affine<F> pt = Affine<F>::load(table + uint64_t(v & 0x7fffffffu) * PT_LIMBS);
                    if (!pt.is_inf()) {
                        if (v & 0x80000000u) pt.y = F::neg(pt.y);
                        ax[k] = pt.x;
                        ay[k] = pt.y;
                        live |= bit;
                    }
                }
            }
        }
        if (!busy) break;  // every lane has stored its last task
        // ---- phase A: every lane's denominator x2 - x1 (1 where the lane has no ordinary addition this step) and the
        // running products; the gather of lane k + 1 is issued before the product of lane k
        E run = one;
        uint32_t adds = 0;
        auto denominator = [&](int k) -> E {
            E d = one;
            if (pos[k] < end[k] && (live & (1u << k))) {
                const uint32_t v = sorted[pos[k]];
                const E px = F::load(table + uint64_t(v & 0x7fffffffu) * PT_LIMBS);
                d = F::sub(px, ax[k]);
                // x2 == x1 (P + P or P - P), or x2 = 0 (an all-zero base is the point at infinity): slow path in phase B
                if (F::is_zero(d) || F::is_zero(px)) d = one;
            }
            return d;
        };
        E d_next = denominator(0);
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            const E d = d_next;
            if (k + 1 < K) d_next = denominator(k + 1);
            adds |= (pos[k] < end[k]) ? 1u : 0u;
            pref[k] = run;
            run = Shared<F>::mul(run, d);
        }
        if (!adds) continue;  // only single-point tasks were opened: nothing to add, go store them
        E inv = F::inv_fast(run);
        // ---- phase B, lanes in reverse: 1 / d_k = inv * pref_k, then inv *= d_k
#pragma unroll 1
        for (int k = K - 1; k >= 0; --k) {
            const uint32_t bit = 1u << k;
            if (pos[k] >= end[k]) continue;
            const uint32_t v = sorted[pos[k]++];
            Affine<F> pt = Affine<F>::load(table + uint64_t(v & 0x7fffffffu) * PT_LIMBS);
            if (v & 0x80000000u) pt.y = F::neg(pt.y);
            if (!(live & bit)) {  // accumulator at infinity: the point replaces it (its d was 1)
                if (!pt.is_inf()) {
                    ax[k] = pt.x;
                    ay[k] = pt.y;
                    live |= bit;
                }
                continue;
            }
            const E d = F::sub(pt.x, ax[k]);
            if (F::is_zero(d) || F::is_zero(pt.x)) {  // d was 1 in phase A
                if (pt.is_inf()) continue;
                if (F::is_zero(d)) {
                    if (F::eq(pt.y, ay[k])) {
                        if (!ec_affine_dbl<F>(ax[k], ay[k])) live &= ~bit;
                    } else {
                        live &= ~bit;  // P - P
                    }
                } else {  // a finite point with x = 0: one addition on its own inversion
                    const E dinv = F::inv_fast(d);
                    ec_affine_add_with_inverse<F>(ax[k], ay[k], pt.x, pt.y, dinv);
                }
                continue;
            }
            const E dinv = Shared<F>::mul(inv, pref[k]);
            inv = Shared<F>::mul(inv, d);
            affine_add_shared<F>(ax[k], ay[k], pt.x, pt.y, dinv);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// 5. collapse multi-task buckets: partial[task_off[b]] = sum of the bucket's partials
// ---------------------------------------------------------------------------------------------
// light buckets (2 .. HEAVY_TASKS partials): one thread each
template <class F>
__global__ void __launch_bounds__(64) k_msm_collapse_light(const uint32_t* __restrict__ task_off, uint64_t buckets,
                                                           uint32_t* __restrict__ partial) {
    const uint64_t b = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= buckets) return;
    const uint32_t t0 = task_off[b], t1 = task_off[b + 1];
    if (t1 - t0 < 2 || t1 - t0 > HEAVY_TASKS) return;
    constexpr int L = XYZZ<F>::BYTES / 4;
    XYZZ<F> acc = XYZZ<F>::load(partial + uint64_t(t0) * L);
    for (uint32_t t = t0 + 1; t < t1; ++t) ec_add<F>(acc, XYZZ<F>::load(partial + uint64_t(t) * L));
    acc.store(partial + uint64_t(t0) * L);
}

// heavy buckets (e.g. the bucket of all scalars == 1), two launches: the bucket's T tasks are cut into slices of
// `per` = max(COLLAPSE_SLICE_MIN, ceil(T / COLLAPSE_SLICES)) tasks; block (h, slice) sums its slice (strided sums then a
// shared-memory tree) into the slice's first slot, then (second launch, slices == 0) one block per bucket sums the
// slice heads into partial[t0].  A 2^22-point all-ones bucket is 65,536 tasks: 16 + 6 + 6 additions deep instead of 1,030.
constexpr unsigned COLLAPSE_THREADS = 64;
constexpr unsigned COLLAPSE_SLICES = 64;
constexpr unsigned COLLAPSE_SLICE_MIN = 256;
template <class F>
__global__ void __launch_bounds__(COLLAPSE_THREADS) k_msm_collapse_heavy(const uint32_t* __restrict__ heavy,
                                                                         const uint32_t* __restrict__ heavy_count,
                                                                         const uint32_t* __restrict__ task_off,
                                                                         uint32_t* __restrict__ partial, int heads) {
    extern __shared__ uint4 sm4[];
    uint32_t* sm = reinterpret_cast<uint32_t*>(sm4);
    constexpr int L = XYZZ<F>::BYTES / 4;
    const uint32_t nh = *heavy_count;
    for (uint32_t h = blockIdx.x; h < nh; h += gridDim.x) {
        const uint32_t b = heavy[h];
        const uint32_t t0 = task_off[b], t1 = task_off[b + 1];
        const uint32_t per = max(COLLAPSE_SLICE_MIN, (t1 - t0 + COLLAPSE_SLICES - 1) / COLLAPSE_SLICES);
        uint32_t lo, hi, step;
        if (heads) {  // the slice heads t0, t0 + per, ...
            lo = t0;
            hi = t1;
            step = per;
        } else {
            lo = t0 + blockIdx.y * per;
            hi = min(t1, lo + per);
            step = 1;
            if (lo >= t1) continue;  // block-uniform
        }
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint64_t t = lo + uint64_t(threadIdx.x) * step; t < hi; t += uint64_t(COLLAPSE_THREADS) * step) {
            XYZZ<F> q = XYZZ<F>::load(partial + t * L);
            ec_add<F>(acc, q);
        }
        acc.store(sm + threadIdx.x * L);
        __syncthreads();
        for (unsigned st = COLLAPSE_THREADS / 2; st > 0; st >>= 1) {
            if (threadIdx.x < st) {
                XYZZ<F> q = XYZZ<F>::load(sm + (threadIdx.x + st) * L);
                ec_add<F>(acc, q);
                acc.store(sm + threadIdx.x * L);
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) acc.store(partial + uint64_t(lo) * L);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// 6. bucket reduction:  S = sum_v v B_v  over the 2^(c-1) buckets of a set (bucket index i holds value v = i + 1)
//    a) segments of SEG = 2^seg_log buckets, one thread each, top-down running sums (2 additions per bucket):
//           run_t = sum_{i in seg t} B_i        acc_t = sum_{i in seg t} (i - lo_t + 1) B_i
//       so  S = sum_t acc_t + SEG * sum_t t run_t
//    b) the (SEG times smaller) weighted sum over segments by bit planes of t:  sum_t t run_t = sum_j 2^j P_j,
//       P_j = sum of run_t over t with bit j set — parallel tree sums; plane index nseg_log sums the acc_t
//    c) per set: Horner over the planes, log2(SEG) doublings, add sum acc_t
// ---------------------------------------------------------------------------------------------
constexpr unsigned RED_THREADS = 64;
constexpr unsigned SEG_LOG_MAX = 6;

template <class F>
__global__ void __launch_bounds__(64) k_msm_reduce_segments(const uint32_t* __restrict__ partial,
                                                            const uint32_t* __restrict__ task_off, MsmPlan p,
                                                            unsigned seg_log, uint32_t* __restrict__ seg_out) {
    constexpr int L = XYZZ<F>::BYTES / 4;
    const unsigned nseg_log = p.nb_log - seg_log;
    const uint64_t q = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;  // (set, segment)
    const uint64_t n_sets = uint64_t(p.batch) * p.sets;
    if (q >= (n_sets << nseg_log)) return;
    const uint64_t lo = q << seg_log;  // global bucket index of the segment's first bucket (sets are contiguous)
    XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
    for (int i = (1 << seg_log) - 1; i >= 0; --i) {
        const uint32_t t0 = task_off[lo + i], t1 = task_off[lo + i + 1];
        if (t1 > t0) ec_add<F>(run, XYZZ<F>::load(partial + uint64_t(t0) * L));
        ec_add<F>(acc, run);
    }
    acc.store(seg_out + (2 * q) * L);
    run.store(seg_out + (2 * q + 1) * L);
}

// block = (chunk, plane j, set): over the chunk's RED_CHUNK segments, j < nseg_log sums run_t over t with bit j set;
// j == nseg_log sums every acc_t
constexpr unsigned RED_CHUNK = 1024;
template <class F>
__global__ void __launch_bounds__(RED_THREADS) k_msm_reduce_planes(const uint32_t* __restrict__ seg, unsigned nseg_log,
                                                                   unsigned chunks, uint32_t* __restrict__ plane_out) {
    extern __shared__ uint4 sm4[];
    uint32_t* sm = reinterpret_cast<uint32_t*>(sm4);
    constexpr int L = XYZZ<F>::BYTES / 4;
    const unsigned chunk = blockIdx.x, j = blockIdx.y, set = blockIdx.z;
    const uint32_t nseg = 1u << nseg_log;
    const uint32_t lo = chunk * RED_CHUNK, hi = min(nseg, lo + RED_CHUNK);
    const uint32_t* base = seg + (uint64_t(set) << nseg_log) * 2 * L;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t t = lo + threadIdx.x; t < hi; t += RED_THREADS) {
        if (j < nseg_log) {
            if ((t >> j) & 1u) ec_add<F>(acc, XYZZ<F>::load(base + (2 * uint64_t(t) + 1) * L));
        } else {
            ec_add<F>(acc, XYZZ<F>::load(base + (2 * uint64_t(t)) * L));
        }
    }
    acc.store(sm + threadIdx.x * L);
    __syncthreads();
    for (unsigned st = RED_THREADS / 2; st > 0; st >>= 1) {
        if (threadIdx.x < st) {
            ec_add<F>(acc, XYZZ<F>::load(sm + (threadIdx.x + st) * L));
            acc.store(sm + threadIdx.x * L);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) acc.store(plane_out + ((uint64_t(set) * (nseg_log + 1) + j) * chunks + chunk) * L);
}

// block = set: LP lanes per plane sum the plane's chunks (strided, then a shared-memory tree), the planes are combined
// as sum_j 2^j P_j by a pairing tree (level d: P_j += 2^d P_(j+d), so nseg_log - 1 doublings and log2 additions deep
// instead of a serial Horner), then times SEG plus the acc plane.  Everything here is latency of one thread's
// additions (~10 us each), which is what a single large multiexp waits on.
template <class F>
struct RedSets {
    static constexpr unsigned LP = 1536 / XYZZ<F>::BYTES;  // 8 lanes per plane in G1, 4 in G2 (48 KB of shared memory)
    static_assert((LP & (LP - 1)) == 0, "the lane tree needs a power of two");
    static constexpr unsigned THREADS = 32 * LP;
};
template <class F>
__global__ void __launch_bounds__(256) k_msm_reduce_sets(const uint32_t* __restrict__ plane_in, unsigned nseg_log,
                                                         unsigned seg_log, unsigned chunks, uint32_t* __restrict__ set_out) {
    extern __shared__ uint4 sm4[];
    uint32_t* sm = reinterpret_cast<uint32_t*>(sm4);
    constexpr int L = XYZZ<F>::BYTES / 4;
    constexpr unsigned LP = RedSets<F>::LP;
    const unsigned set = blockIdx.x, tid = threadIdx.x, j = tid / LP, l = tid % LP;
    XYZZ<F> acc = XYZZ<F>::inf();
    if (j <= nseg_log)
        for (unsigned ch = l; ch < chunks; ch += LP)
            ec_add<F>(acc, XYZZ<F>::load(plane_in + ((uint64_t(set) * (nseg_log + 1) + j) * chunks + ch) * L));
    acc.store(sm + tid * L);
    __syncthreads();
    for (unsigned st = LP / 2; st > 0; st >>= 1) {
        if (l < st && j <= nseg_log) {
            ec_add<F>(acc, XYZZ<F>::load(sm + (tid + st) * L));
            acc.store(sm + tid * L);
        }
        __syncthreads();
    }
    // plane j now sits at sm[j * LP]; pair them up
    for (unsigned d = 1; d < nseg_log; d <<= 1) {
        if (l == 0 && (j & (2 * d - 1)) == 0 && j + d < nseg_log) {
            XYZZ<F> q = XYZZ<F>::load(sm + (j + d) * LP * L);
            for (unsigned k = 0; k < d; ++k) q = ec_dbl<F>(q);
            ec_add<F>(acc, q);
            acc.store(sm + tid * L);
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (nseg_log == 0) acc = XYZZ<F>::inf();  // no run planes: only the acc plane below
        for (unsigned k = 0; k < seg_log; ++k) acc = ec_dbl<F>(acc);
        ec_add<F>(acc, XYZZ<F>::load(sm + nseg_log * LP * L));
        acc.store(set_out + uint64_t(set) * L);
    }
}

// ---------------------------------------------------------------------------------------------
// 7. final: Horner over a ballot's sets (one-shot bases), affine, canonical
// ---------------------------------------------------------------------------------------------
template <class F>
VSP_DEV void store_canonical_affine(uint64_t* out, const XYZZ<F>& p, bool lowlat = false) {
    Affine<F> a = ec_to_affine<F>(p, lowlat);
    a.x = F::from_mont(a.x);
    a.y = F::from_mont(a.y);
    a.store(out);
}

// block = ballot (lane 0 works): out + ballot * out_stride (u64 limbs)
template <class F>
__global__ void k_msm_final(const uint32_t* __restrict__ set_in, MsmPlan p, uint64_t* __restrict__ out,
                            uint64_t out_stride, int raw) {
    if (threadIdx.x != 0) return;
    constexpr int L = XYZZ<F>::BYTES / 4;
    const unsigned b = blockIdx.x;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = (int)p.sets - 1; w >= 0; --w) {
        if (w != (int)p.sets - 1)
            for (unsigned k = 0; k < p.c; ++k) acc = ec_dbl<F>(acc);
        ec_add<F>(acc, XYZZ<F>::load(set_in + (uint64_t(b) * p.sets + w) * L));
    }
    uint64_t* o = out + uint64_t(b) * out_stride;
    if (raw)
        acc.store(o);  // Montgomery-form XYZZ for the prover's assembly kernel: skips the inversion
    else
        store_canonical_affine<F>(o, acc, true);  // lane 0 of the block works alone
}

template <class F>
__global__ void k_write_infinity(uint64_t* out, uint64_t out_stride, int raw) {
    const int n = (raw ? XYZZ<F>::BYTES : Affine<F>::BYTES) / 8;
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[uint64_t(blockIdx.x) * out_stride + i] = 0;
}

// ---------------------------------------------------------------------------------------------
// bases: Montgomery conversion and the window table  T[w][i] = 2^(c w) P_i
// ---------------------------------------------------------------------------------------------
template <class F>
__global__ void k_bases_to_mont(uint32_t* pts, uint64_t n) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    constexpr int L = Affine<F>::BYTES / 4;
    Affine<F> a = Affine<F>::load(pts + i * L);
    a.x = F::to_mont(a.x);
    a.y = F::to_mont(a.y);
    a.store(pts + i * L);
}

// level w from level w - 1: c doublings per point, then back to affine with one inversion per TABLE_CH points
// (Montgomery's trick).  Runs once per proving-key query.
constexpr int TABLE_CH = 4;
template <class F>
__global__ void __launch_bounds__(64) k_table_level(uint32_t* table, uint64_t n, unsigned level, unsigned c) {
    using E = typename F::T;
    constexpr int L = Affine<F>::BYTES / 4;
    const uint64_t base = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) * TABLE_CH;
    if (base >= n) return;
    const uint32_t* src = table + uint64_t(level - 1) * n * L;
    uint32_t* dst = table + uint64_t(level) * n * L;
    XYZZ<F> r[TABLE_CH];
    E den[TABLE_CH], pre[TABLE_CH];
    for (int q = 0; q < TABLE_CH; ++q) {
        r[q] = XYZZ<F>::inf();
        if (base + q < n) {
            Affine<F> pt = Affine<F>::load(src + (base + q) * L);
            if (!pt.is_inf()) {
                r[q] = ec_dbl_affine<F>(pt);
                for (unsigned k = 1; k < c; ++k) r[q] = ec_dbl<F>(r[q]);
            }
        }
        den[q] = r[q].is_inf() ? F::one() : F::mul(r[q].ZZ, r[q].ZZZ);
        pre[q] = q ? F::mul(pre[q - 1], den[q]) : den[q];
    }
    E inv = F::inv(pre[TABLE_CH - 1]);
    for (int q = TABLE_CH - 1; q >= 0; --q) {
        const E dinv = q ? F::mul(inv, pre[q - 1]) : inv;  // 1 / (ZZ ZZZ) of point q
        inv = F::mul(inv, den[q]);
        if (base + q >= n) continue;
        Affine<F> a = Affine<F>::inf();
        if (!r[q].is_inf()) {
            a.x = F::mul(r[q].X, F::mul(dinv, r[q].ZZZ));
            a.y = F::mul(r[q].Y, F::mul(dinv, r[q].ZZ));
        }
        a.store(dst + (base + q) * L);
    }
}

// window width minimising  (bucket additions) + (reduction: 2 full additions per bucket, ~1.4x a mixed one each).
// With a window table the reduction runs once per (ballot, query) over buckets that a realistic witness fills (a third of
// the ballot circuit's wires are full-width field elements), and its kernels — serial running sums — reach well under
// half the accumulate kernel's pipe utilisation, so a bucket there is weighted by what it measured to cost on the ballot
// circuit (B200, 64-ballot chunks: c = 13 beats 14, 15, 16 at 32k points: 1,139 / 1,101 / 1,092 / 965 proofs/s), not by
// its multiplication count.
static unsigned choose_window(uint64_t n, bool table) {
    if (const char* e = getenv("VSP_MSM_C")) {  // tuning override
        const int c = atoi(e);
        if (c >= 2 && c <= 24) return (unsigned)c;
    }
    double best = 1e300;
    unsigned best_c = 4;
    for (unsigned c = 4; c <= 22; ++c) {
        const double w = 255 / c + 1, red = (table ? 16.0 : 2.8) * double(1u << (c - 1));
        const double cost = table ? w * double(n) + red : w * (double(n) + red);
        if (cost < best) {
            best = cost;
            best_c = c;
        }
    }
    return best_c;
}

// ---------------------------------------------------------------------------------------------
// driver
// ---------------------------------------------------------------------------------------------
template <class F>
static int msm_run_t(vsp_ctx* ctx, cudaStream_t s, const BasesDev& bases, uint64_t n, const uint32_t* d_scalars,
                     uint64_t scalar_stride, unsigned batch, uint64_t* d_out, uint64_t out_stride, bool raw) {
    constexpr int KC_ACC = std::is_same<F, FqOps>::value ? KC_MSM_ACC_G1 : KC_MSM_ACC_G2;
    if (n > bases.n) return set_error(ctx, VSP_E_INVALID, "multiexp: %llu scalars for %llu bases", (unsigned long long)n,
                                      (unsigned long long)bases.n);
    if (batch == 0) return VSP_OK;
    if (n == 0) {
        VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_write_infinity<F>, batch, 32, 0, s, d_out, out_stride, raw ? 1 : 0);
        VSP_CUDA_CHECK(ctx, cudaGetLastError());
        return VSP_OK;
    }
    MsmPlan p;
    p.c = bases.table ? bases.c : choose_window(n, false);
    p.windows = 255 / p.c + 1;
    p.nb_log = p.c - 1;
    p.sets = bases.table ? 1 : p.windows;
    p.batch = batch;
    p.stride = bases.table ? bases.n : 0;
    p.buckets = (uint64_t(batch) * p.sets) << p.nb_log;
    const uint64_t total = n * p.windows * batch;
    if (total >= (uint64_t(1) << 32) || p.buckets >= (uint64_t(1) << 31) || uint64_t(batch) * p.sets > 65535) {
        // keys and sorted positions are 32-bit: a batch too large for one pass runs as consecutive sub-batches on the same
        // stream (same results; the scratch of one sub-batch is released, stream-ordered, before the next allocates)
        const uint64_t per = n * p.windows, sets_buckets = uint64_t(p.sets) << p.nb_log;
        uint64_t sub = std::min<uint64_t>({((uint64_t(1) << 32) - 1) / per, ((uint64_t(1) << 31) - 1) / sets_buckets,
                                           65535 / p.sets});
        if (sub == 0 || batch == 1)
            return set_error(ctx, VSP_E_INVALID, "multiexp: n * windows must be < 2^32 (n = %llu, %u windows)",
                             (unsigned long long)n, p.windows);
        for (unsigned b0 = 0; b0 < batch; b0 += (unsigned)sub) {
            const unsigned bn = (unsigned)std::min<uint64_t>(sub, batch - b0);
            VSP_TRY(msm_run_t<F>(ctx, s, bases, n, d_scalars + 8 * uint64_t(b0) * scalar_stride, scalar_stride, bn,
                                 d_out + uint64_t(b0) * out_stride, out_stride, raw));
        }
        return VSP_OK;
    }
    const unsigned tiles = (unsigned)((p.buckets + SCAN_TILE - 1) / SCAN_TILE);

    DevBuf keys, counts, offsets, cursors, task_off, tile_scratch, sorted, heavy, heavy_count;
    VSP_TRY(keys.alloc(ctx, total * 4, s));
    VSP_TRY(counts.alloc(ctx, (p.buckets + 1) * 4, s));
    VSP_TRY(offsets.alloc(ctx, (p.buckets + 1) * 4, s));
    VSP_TRY(cursors.alloc(ctx, (p.buckets + 1) * 4, s));
    VSP_TRY(task_off.alloc(ctx, (p.buckets + 1) * 4, s));
    VSP_TRY(tile_scratch.alloc(ctx, (tiles + 2) * 4, s));
    VSP_TRY(sorted.alloc(ctx, total * 4, s));
    VSP_TRY(heavy.alloc(ctx, p.buckets * 4, s));
    VSP_TRY(heavy_count.alloc(ctx, 4, s));
    VSP_CUDA_CHECK(ctx, cudaMemsetAsync(counts.p, 0, (p.buckets + 1) * 4, s));
    VSP_CUDA_CHECK(ctx, cudaMemsetAsync(cursors.p, 0, (p.buckets + 1) * 4, s));
    VSP_CUDA_CHECK(ctx, cudaMemsetAsync(heavy_count.p, 0, 4, s));

    // counters in shared memory when a ballot's bucket set fits one CTA (window tables, c <= 16) and the batch fills the
    // SMs; otherwise global atomics
    const bool set_sort = p.sets == 1 && p.nb_log <= 15 && batch >= 8 && !getenv("VSP_MSM_GLOBAL_SORT");
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_digits, dim3((unsigned)((n + 127) / 128), batch), 128, 0, s, d_scalars, n,
                 scalar_stride, p, keys.as<uint32_t>(), counts.as<uint32_t>(), set_sort ? 0 : 1);
    DevBuf hist_split;
    unsigned split = 1;
    const size_t set_smem = (size_t)4 << p.nb_log;
    if (set_sort) {
        VSP_CUDA_CHECK(ctx, cudaFuncSetAttribute(k_msm_hist_set, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
        VSP_CUDA_CHECK(ctx, cudaFuncSetAttribute(k_msm_scatter_set, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
        split = (2u * (unsigned)ctx->sm_count + batch - 1) / batch;
        if (split < 1) split = 1;
        if (split > 8) split = 8;
        VSP_TRY(hist_split.alloc(ctx, ((size_t)batch * split << p.nb_log) * 4, s));
        VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_hist_set, dim3(split, batch), SET_SORT_THREADS, set_smem, s, keys.as<uint32_t>(),
                     n * p.windows, p.nb_log, split, hist_split.as<uint32_t>(), counts.as<uint32_t>());
        VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_sum_split, (unsigned)((p.buckets + 255) / 256), 256, 0, s,
                     hist_split.as<uint32_t>(), p.nb_log, split, p.buckets, counts.as<uint32_t>());
    }
    VSP_TRY(exclusive_scan<0>(ctx, s, counts.as<uint32_t>(), p.buckets, offsets.as<uint32_t>(),
                              tile_scratch.as<uint32_t>()));
    VSP_TRY(exclusive_scan<1>(ctx, s, counts.as<uint32_t>(), p.buckets, task_off.as<uint32_t>(),
                              tile_scratch.as<uint32_t>()));
    if (set_sort)
        VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_scatter_set, dim3(split, batch), SET_SORT_THREADS, set_smem, s,
                     keys.as<uint32_t>(), n * p.windows, n, p, split, hist_split.as<uint32_t>(), offsets.as<uint32_t>(),
                     sorted.as<uint32_t>());
    else
        VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_scatter, (unsigned)((total + 255) / 256), 256, 0, s, keys.as<uint32_t>(), total,
                     n, p, offsets.as<uint32_t>(), cursors.as<uint32_t>(), sorted.as<uint32_t>());
    // upper bound on the number of tasks, known without a device read-back: ceil(count / T) <= count / T + 1
    const uint64_t max_tasks = total / TASK_POINTS + p.buckets;
    DevBuf tasks, partial;
    VSP_TRY(tasks.alloc(ctx, max_tasks * sizeof(Task), s));
    VSP_TRY(partial.alloc(ctx, max_tasks * XYZZ<F>::BYTES, s));
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_msm_tasks, (unsigned)((p.buckets + 127) / 128), 128, 0, s, offsets.as<uint32_t>(),
                 task_off.as<uint32_t>(), p.buckets, tasks.as<Task>(), heavy.as<uint32_t>(), heavy_count.as<uint32_t>());
    // the exact task count lives on the device (task_off[buckets]); the grid covers the bound and extra threads exit
    DevBuf bins, order;
    VSP_TRY(bins.alloc(ctx, 2 * LEN_BINS * 4, s));
    VSP_TRY(order.alloc(ctx, max_tasks * 4, s));
    VSP_CUDA_CHECK(ctx, cudaMemsetAsync(bins.p, 0, 2 * LEN_BINS * 4, s));
    const uint32_t* d_n_tasks = task_off.as<uint32_t>() + p.buckets;
    const unsigned order_blocks = (unsigned)std::min<uint64_t>((max_tasks + 2047) / 2048, 1184);
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_task_hist, order_blocks, 256, 0, s, tasks.as<Task>(), d_n_tasks, bins.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_task_bin_offsets, 1, 32, 0, s, bins.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_SORT, k_task_order, order_blocks, 256, 0, s, tasks.as<Task>(), d_n_tasks, bins.as<uint32_t>(),
                 order.as<uint32_t>());
    // Two kernels can do the accumulation.  The XYZZ one (a thread per task) is the default.  VSP_MSM_AFFINE=1 hands every
    // launch to the batched-affine kernel instead, VSP_MSM_AFFINE=auto only launches with enough tasks to fill the GPU
    // (the device-resident task count decides; the other kernel's threads exit at once).  Measured on B200 the affine
    // kernel issues 29 % fewer heavy-pipe multiplies per addition but runs 1.6x SLOWER (a 256-ballot step's G1 accumulation
    // 175 ms against 110 ms): its per-lane state lives in local memory, misses L1 (8.7 % hit rate, 600 B of DRAM traffic
    // per addition) and leaves the warps waiting on loads and on the serial division-step chain.  Kept for that evidence
    // and as a starting point (profiles/r02_ncu_affine_accumulate.txt); results are bit-identical either way.
    constexpr int AK = AffineLanes<F>::K;
    uint32_t affine_min = 0xffffffffu;
    unsigned affine_blocks = 0;
    if (const char* e = getenv("VSP_MSM_AFFINE")) {
        if (e[0] == '1' || e[0] == 'a') {
            int per_sm = 0;
            VSP_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_msm_accumulate_affine<F, AK>, 128, 0));
            const uint64_t resident_threads = uint64_t(ctx->sm_count) * std::max(per_sm, 1) * 128;
            affine_min = e[0] == '1' ? 1u : (uint32_t)std::min<uint64_t>(resident_threads * AK / 2, 0xfffffffeull);
            affine_blocks = (unsigned)std::min<uint64_t>((max_tasks + uint64_t(AK) * 128 - 1) / (uint64_t(AK) * 128),
                                                         resident_threads / 128);
        }
    }
    // VSP_MSM_ACC_SMEM (bytes of unused dynamic shared memory per block) caps the kernel's resident blocks per SM from
    // the outside: an occupancy experiment (does leaving register-file room for the other streams' kernels pay?)
    size_t acc_smem = 0;
    if (const char* e = getenv("VSP_MSM_ACC_SMEM")) {
        acc_smem = (size_t)atol(e);
        if (acc_smem > 48 * 1024)
            VSP_CUDA_CHECK(ctx, cudaFuncSetAttribute(k_msm_accumulate<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)acc_smem));
    }
    unsigned acc_per_sm = 0;
    if (const char* e = getenv("VSP_MSM_ACC_BLOCKS")) acc_per_sm = (unsigned)atoi(e);  // 0: one block per 128 tasks (uncapped)
    uint64_t acc_blocks = (max_tasks + 127) / 128;
    if (acc_per_sm) acc_blocks = std::min<uint64_t>(acc_blocks, uint64_t(ctx->sm_count) * acc_per_sm);
    VSP_LAUNCH_K(ctx, KC_ACC, k_msm_accumulate<F>, (unsigned)acc_blocks, 128, acc_smem, s, bases.d_pts,
                 sorted.as<uint32_t>(), tasks.as<Task>(), order.as<uint32_t>(), d_n_tasks, affine_min, partial.as<uint32_t>());
    // algorithmic work of this launch for the live roofline: bucket additions scheduled (= non-zero digits,
    // offsets[buckets]) x 11 field multiplications (the reference's mixed Jacobian add) x I_mul(limbs)
    kprof_units(ctx, ctx->last_prof, 11.0 * (std::is_same<F, FqOps>::value ? 300.0 : 3.0 * 300.0),
                offsets.as<uint32_t>() + p.buckets, s);
    if (affine_blocks)
        VSP_LAUNCH_K(ctx, KC_ACC, (k_msm_accumulate_affine<F, AK>), affine_blocks, 128, 0, s, bases.d_pts, sorted.as<uint32_t>(),
                     tasks.as<Task>(), order.as<uint32_t>(), d_n_tasks, affine_min, partial.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_collapse_light<F>, (unsigned)((p.buckets + 63) / 64), 64, 0, s,
                 task_off.as<uint32_t>(), p.buckets, partial.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_collapse_heavy<F>, dim3(256, COLLAPSE_SLICES), COLLAPSE_THREADS,
                 COLLAPSE_THREADS * XYZZ<F>::BYTES, s, heavy.as<uint32_t>(), heavy_count.as<uint32_t>(),
                 task_off.as<uint32_t>(), partial.as<uint32_t>(), 0);
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_collapse_heavy<F>, 256, COLLAPSE_THREADS, COLLAPSE_THREADS * XYZZ<F>::BYTES, s,
                 heavy.as<uint32_t>(), heavy_count.as<uint32_t>(), task_off.as<uint32_t>(), partial.as<uint32_t>(), 1);
    const unsigned n_sets = batch * p.sets;
    // segment length: keep ~2^14 segment threads in flight (each runs a serial chain of 2 * SEG additions); measured on the
    // ballot circuit (c = 13, 64-ballot chunks): SEG = 16 beats 8 and 32 (reduction 37 / 42 / 51 ms per step)
    unsigned seg_log = 1;
    {
        const unsigned total_log = p.nb_log + floor_log2_u64(n_sets);
        if (total_log > 14 + 1) seg_log = total_log - 14;
        if (p.sets == 1 && seg_log > 4) seg_log = 4;  // batched ballots: SEG = 16 measured best at 64, 128 and 256 per chunk
        if (const char* e = getenv("VSP_MSM_SEG_LOG")) seg_log = (unsigned)atoi(e);  // tuning override
        if (seg_log < 1) seg_log = 1;
        if (seg_log > SEG_LOG_MAX) seg_log = SEG_LOG_MAX;
        if (seg_log > p.nb_log) seg_log = p.nb_log;
    }
    const unsigned nseg_log = p.nb_log - seg_log;
    const uint64_t n_seg = uint64_t(n_sets) << nseg_log;
    const unsigned chunks = (unsigned)(((uint64_t(1) << nseg_log) + RED_CHUNK - 1) / RED_CHUNK);
    DevBuf segs, planes, sets;
    VSP_TRY(segs.alloc(ctx, (size_t)n_seg * 2 * XYZZ<F>::BYTES, s));
    VSP_TRY(planes.alloc(ctx, (size_t)n_sets * (nseg_log + 1) * chunks * XYZZ<F>::BYTES, s));
    VSP_TRY(sets.alloc(ctx, (size_t)n_sets * XYZZ<F>::BYTES, s));
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_reduce_segments<F>, (unsigned)((n_seg + 63) / 64), 64, 0, s,
                 partial.as<uint32_t>(), task_off.as<uint32_t>(), p, seg_log, segs.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_reduce_planes<F>, dim3(chunks, nseg_log + 1, n_sets), RED_THREADS,
                 RED_THREADS * XYZZ<F>::BYTES, s, segs.as<uint32_t>(), nseg_log, chunks, planes.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_reduce_sets<F>, n_sets, RedSets<F>::THREADS,
                 RedSets<F>::THREADS * XYZZ<F>::BYTES, s, planes.as<uint32_t>(), nseg_log, seg_log, chunks,
                 sets.as<uint32_t>());
    VSP_LAUNCH_K(ctx, KC_MSM_REDUCE, k_msm_final<F>, batch, 32, 0, s, sets.as<uint32_t>(), p, d_out, out_stride,
                 raw ? 1 : 0);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    return VSP_OK;
}

// after d_pts[0, n) holds Montgomery affine points: fix the window width and, for resident bases, build the table
template <class F>
static int bases_finalize_t(vsp_ctx* ctx, cudaStream_t s, BasesDev* b, bool table) {
    constexpr int L = Affine<F>::BYTES / 4;
    b->c = b->force_c ? b->force_c : choose_window(b->n ? b->n : 1, table);
    b->windows = 255 / b->c + 1;
    b->table = false;
    if (!table || b->n == 0) return VSP_OK;
    if (b->n * b->windows >= (uint64_t(1) << 31)) return VSP_OK;  // too large to index: stay table-less
    uint32_t* t = nullptr;
    const size_t bytes = (size_t)b->n * b->windows * Affine<F>::BYTES;
    if (cudaMalloc(&t, bytes) != cudaSuccess) {
        cudaGetLastError();
        // not enough memory for the table: the table-less path still works (same results), but a small-n multiexp is an
        // order of magnitude slower without it, so the caller is told (vsp_last_warning)
        set_warning(ctx, "window table of %llu points x %u windows (%.1f GB) did not fit: multiexps over these bases run table-less",
                    (unsigned long long)b->n, b->windows, bytes / 1e9);
        return VSP_OK;
    }
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(t, b->d_pts, (size_t)b->n * Affine<F>::BYTES, cudaMemcpyDeviceToDevice, s));
    const uint64_t threads = (b->n + TABLE_CH - 1) / TABLE_CH;
    for (unsigned w = 1; w < b->windows; ++w)
        VSP_LAUNCH_K(ctx, KC_OTHER, k_table_level<F>, (unsigned)((threads + 63) / 64), 64, 0, s, t, b->n, w, b->c);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));
    cudaFree(b->d_pts);
    b->d_pts = t;
    b->table = true;
    (void)L;
    return VSP_OK;
}

// ---------------------------------------------------------------------------------------------
// sum of k affine points (join step of a sharded multiexp)
// ---------------------------------------------------------------------------------------------
template <class F>
__global__ void k_points_sum(const uint64_t* pts, uint64_t k, uint64_t* out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    constexpr int L8 = Affine<F>::BYTES / 8;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint64_t i = 0; i < k; ++i) {
        Affine<F> a = Affine<F>::load(pts + i * L8);
        if (a.is_inf()) continue;
        a.x = F::to_mont(a.x);
        a.y = F::to_mont(a.y);
        ec_madd<F>(acc, a);
    }
    store_canonical_affine<F>(out, acc, true);
}

}  // namespace

// one translation unit per curve keeps ptxas time down (msm_g1.cu / msm_g2.cu)
#define VSP_MSM_INSTANTIATE(SUFFIX, F)                                                                              \
    int msm_run_##SUFFIX(vsp_ctx* ctx, cudaStream_t s, const BasesDev& bases, uint64_t n, const uint32_t* d_scalars, \
                         uint64_t scalar_stride, unsigned batch, uint64_t* d_out, uint64_t out_stride, bool raw) {  \
        return msm_run_t<F>(ctx, s, bases, n, d_scalars, scalar_stride, batch, d_out, out_stride, raw);             \
    }                                                                                                               \
    int bases_finalize_##SUFFIX(vsp_ctx* ctx, cudaStream_t s, BasesDev* b, bool table) {                            \
        return bases_finalize_t<F>(ctx, s, b, table);                                                               \
    }                                                                                                               \
    int points_sum_##SUFFIX(vsp_ctx* ctx, cudaStream_t s, uint64_t k, const uint64_t* d_points, uint64_t* d_out) {  \
        VSP_LAUNCH_K(ctx, KC_FIXED, k_points_sum<F>, 1, 32, 0, s, d_points, k, d_out);                              \
        VSP_CUDA_CHECK(ctx, cudaGetLastError());                                                                    \
        return VSP_OK;                                                                                              \
    }                                                                                                               \
    int bases_to_mont_##SUFFIX(vsp_ctx* ctx, cudaStream_t s, uint32_t* d_pts, uint64_t n) {                         \
        VSP_LAUNCH_K(ctx, KC_OTHER, k_bases_to_mont<F>, (unsigned)((n + 127) / 128), 128, 0, s, d_pts, n);          \
        VSP_CUDA_CHECK(ctx, cudaGetLastError());                                                                    \
        return VSP_OK;                                                                                              \
    }

}  // namespace vsp
