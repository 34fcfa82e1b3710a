// voting.cu — batched witness generation for the encrypted_input ballot circuit on the GPU.
//
// Replaces the nine generate_r1cs_witness steps bin/cli runs per vote on the host
// (/root/reference/bin/cli/include/nil/vote_saver/common.hpp:1110-1128: merkle path, address bits, m / eid / sk blocks,
// the voting component (Pedersen hashes over Jubjub: pk = H(sk), sn = H(eid || sk), the arity-2 Merkle path of
// libs/containers/include/nil/crypto3/container/merkle/proof.hpp:66-92), the three multipacking components).  The
// components themselves live in the un-vendored crypto3-blueprint / crypto3-hash submodules, so the circuit is this
// repository's restatement of the construction they port (Zcash Sapling's Pedersen hash as ethsnarks arithmetises it);
// its layout is defined once, in vote-saver-protocol_b200/voting.py, which hands this file the tables it walks.  Parity
// with crypto3's wire values is therefore unpinned; every witness made here must satisfy that module's R1CS and equal
// its host restatement bit for bit (tests/test_gpu_voting.py).
//
// One call fills k full assignments (1 || primary || auxiliary, canonical Fr) in device memory, ready for
// vsp_saver_encrypt_batch_dev.  A Pedersen hash is a chain of Montgomery-form additions with one field inversion each, so
// the parallelism is across ballots, the hashes of a stage and the (up to three) segments of a hash: one thread per
// (ballot, hash, segment), the constant-time division-step inverse keeping a warp's lanes in lock step; the Merkle levels
// are stages of a stream-ordered launch sequence.
#include "common.cuh"
#include "fp.cuh"

#include <cstring>
#include <new>
#include <vector>

using namespace vsp;

namespace {
constexpr unsigned WPS = 63;        // windows per segment
constexpr unsigned DIGEST = 255;    // digest bits
constexpr unsigned CHUNK = 254;     // bits per packed field element (common.hpp:861)

struct HashDesc {
    uint32_t in_off, n_bits, var_base, out_base, windows, segments;
};
struct LevelDesc {
    uint32_t sib, left, right, hash;
};
struct Layout {
    uint32_t depth, eid_bits, msg_size, num_vars, num_inputs, m, eid_packed, sn_packed, rt_packed, eid, sk, address, h_pk, h_sn,
        max_seg;
};

VSP_DEV uint32_t window_offset(uint32_t w) { return 5 * w - 3 * ((w + WPS - 1) / WPS); }

VSP_DEV Fr fr_small(uint32_t v) {
    Fr r = Fr::zero();
    r.l[0] = v;
    return r;
}
VSP_DEV void store_canonical(uint32_t* w, uint32_t var, const Fr& mont) { fp_store(w + 8 * uint64_t(var), fp_from_mont(mont)); }
VSP_DEV void store_small(uint32_t* w, uint32_t var, uint32_t v) { fp_store(w + 8 * uint64_t(var), fr_small(v)); }
VSP_DEV uint32_t load_bit(const uint32_t* w, uint32_t var) { return w[8 * uint64_t(var)] & 1u; }

struct Consts {
    Fr a, d, scale;  // Montgomery form: Montgomery-curve A, Edwards d, sqrt(-(A + 2))
};

// the plain inputs of every ballot: constant 1, the one-hot vote, eid / sk / address bits, the sibling digests of every level
__global__ void k_vote_inputs(Layout L, const LevelDesc* __restrict__ levels, const uint32_t* __restrict__ votes, uint64_t eid,
                              const uint64_t* __restrict__ sks, const uint32_t* __restrict__ indices,
                              const uint64_t* __restrict__ paths, unsigned k, uint32_t* __restrict__ assign) {
    const unsigned per = 1 + 1 + L.eid_bits + DIGEST + L.depth + L.depth * DIGEST;
    const uint64_t q = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= uint64_t(k) * per) return;
    const unsigned b = (unsigned)(q / per);
    unsigned i = (unsigned)(q % per);
    uint32_t* w = assign + 8 * uint64_t(b) * (L.num_vars + 1);
    if (i == 0) return store_small(w, 0, 1);
    if (--i == 0) return store_small(w, L.m + votes[b], 1);
    if (--i < L.eid_bits) return store_small(w, L.eid + i, (uint32_t)((eid >> i) & 1));
    i -= L.eid_bits;
    if (i < DIGEST) return store_small(w, L.sk + i, (uint32_t)((sks[4 * uint64_t(b) + i / 64] >> (i % 64)) & 1));
    i -= DIGEST;
    if (i < L.depth) return store_small(w, L.address + i, (indices[b] >> i) & 1u);
    i -= L.depth;
    const unsigned lvl = i / DIGEST, bit = i % DIGEST;
    store_small(w, levels[lvl].sib + bit, (uint32_t)((paths[(uint64_t(b) * L.depth + lvl) * 4 + bit / 64] >> (bit % 64)) & 1));
}

// thread = (ballot, hash of this stage, segment): the segment's windows — lookup, conditional negation, Montgomery sum —
// then its Edwards form
__global__ void __launch_bounds__(64) k_pedersen_segment(Layout L, const HashDesc* __restrict__ hashes, const uint32_t* __restrict__ in_idx,
                                                         const uint32_t* __restrict__ tables, Consts cst, uint32_t h0, unsigned k,
                                                         uint32_t* __restrict__ assign) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= k) return;
    const HashDesc h = hashes[h0 + blockIdx.y / L.max_seg];
    const unsigned seg = blockIdx.y % L.max_seg;
    if (seg >= h.segments) return;
    uint32_t* w = assign + 8 * uint64_t(b) * (L.num_vars + 1);
    const unsigned w0 = seg * WPS, w1 = min(h.windows, w0 + WPS);
    Fr ax = Fr::zero(), ay = Fr::zero();
    for (unsigned win = w0; win < w1; ++win) {
        const unsigned j = win - w0;
        uint32_t bits[3];
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const unsigned i = 3 * win + t;
            bits[t] = i < h.n_bits ? load_bit(w, in_idx[h.in_off + i]) : 0u;  // padding bits are 0
        }
        const uint32_t* tp = tables + 8 * (uint64_t((seg * WPS + j) * 4 + bits[0] + 2 * bits[1]) * 2);
        const Fr px = fp_load<FrCfg>(tp);
        Fr py = fp_load<FrCfg>(tp + 8);
        if (bits[2]) py = fp_neg(py);
        const uint32_t o = h.var_base + window_offset(win);
        store_small(w, o, bits[0] & bits[1]);
        store_canonical(w, o + 1, py);
        if (j == 0) {
            ax = px;
            ay = py;
        } else {
            const Fr lam = fp_mul(fp_sub(py, ay), fp_inv_safegcd(fp_sub(px, ax)));
            const Fr xs = fp_sub(fp_sub(fp_sub(fp_sqr(lam), cst.a), ax), px);
            const Fr ys = fp_sub(fp_mul(lam, fp_sub(ax, xs)), ay);
            store_canonical(w, o + 2, lam);
            store_canonical(w, o + 3, xs);
            store_canonical(w, o + 4, ys);
            ax = xs;
            ay = ys;
        }
    }
    // (x, y) -> u = scale x / y, v = (x - 1) / (x + 1) on one inversion
    const Fr one = Fr::one();
    const Fr xp1 = fp_add(ax, one);
    const Fr inv = fp_inv_safegcd(fp_mul(ay, xp1));
    const Fr u = fp_mul(fp_mul(cst.scale, ax), fp_mul(inv, xp1));
    const Fr v = fp_mul(fp_sub(ax, one), fp_mul(inv, ay));
    const uint32_t o2 = h.var_base + window_offset(h.windows) + 2 * seg;
    store_canonical(w, o2, u);
    store_canonical(w, o2 + 1, v);
}

// thread = (ballot, hash of this stage): the Edwards sum of the segments and the digest bits
__global__ void __launch_bounds__(64) k_pedersen_finish(Layout L, const HashDesc* __restrict__ hashes, Consts cst, uint32_t h0, unsigned k,
                                                        uint32_t* __restrict__ assign) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= k) return;
    const HashDesc h = hashes[h0 + blockIdx.y];
    uint32_t* w = assign + 8 * uint64_t(b) * (L.num_vars + 1);
    const uint32_t base2 = h.var_base + window_offset(h.windows);
    Fr cu = fp_to_mont(fp_load<FrCfg>(w + 8 * uint64_t(base2))), cv = fp_to_mont(fp_load<FrCfg>(w + 8 * uint64_t(base2 + 1)));
    const Fr one = Fr::one();
    for (unsigned s = 1; s < h.segments; ++s) {
        const Fr su = fp_to_mont(fp_load<FrCfg>(w + 8 * uint64_t(base2 + 2 * s)));
        const Fr sv = fp_to_mont(fp_load<FrCfg>(w + 8 * uint64_t(base2 + 2 * s + 1)));
        const Fr U = fp_mul(fp_add(cu, cv), fp_add(su, sv));
        const Fr A = fp_mul(sv, cu), B = fp_mul(su, cv);
        const Fr Cc = fp_mul(fp_mul(cst.d, A), B);
        const Fr p1 = fp_add(one, Cc), m1 = fp_sub(one, Cc);
        const Fr inv = fp_inv_safegcd(fp_mul(p1, m1));
        const Fr u3 = fp_mul(fp_add(A, B), fp_mul(inv, m1));
        const Fr v3 = fp_mul(fp_sub(fp_sub(U, A), B), fp_mul(inv, p1));
        const uint32_t o = base2 + 2 * h.segments + 6 * (s - 1);
        store_canonical(w, o, U);
        store_canonical(w, o + 1, A);
        store_canonical(w, o + 2, B);
        store_canonical(w, o + 3, Cc);
        store_canonical(w, o + 4, u3);
        store_canonical(w, o + 5, v3);
        cu = u3;
        cv = v3;
    }
    const Fr c = fp_from_mont(cu);
    for (unsigned i = 0; i < DIGEST; ++i) store_small(w, h.out_base + i, (c.l[i >> 5] >> (i & 31)) & 1u);
}

// thread = (ballot, bit): left / right children of a Merkle level from the running digest, the sibling and the address bit
__global__ void k_merkle_select(Layout L, LevelDesc lv, uint32_t cur, unsigned lvl, unsigned k, uint32_t* __restrict__ assign) {
    const uint64_t q = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= uint64_t(k) * DIGEST) return;
    const unsigned b = (unsigned)(q / DIGEST), i = (unsigned)(q % DIGEST);
    uint32_t* w = assign + 8 * uint64_t(b) * (L.num_vars + 1);
    const uint32_t a = load_bit(w, L.address + lvl), s = load_bit(w, lv.sib + i), c = load_bit(w, cur + i);
    store_small(w, lv.left + i, a ? s : c);   // address bit 1: the running digest is the right child
    store_small(w, lv.right + i, a ? c : s);
}

// thread = ballot: the packed primary inputs (multipacking, 254 bits per element)
__global__ void k_vote_pack(Layout L, uint32_t sn_bits, uint32_t rt_bits, unsigned k, uint32_t* __restrict__ assign) {
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= k) return;
    uint32_t* w = assign + 8 * uint64_t(b) * (L.num_vars + 1);
    auto pack = [&](uint32_t first_bit, unsigned nbits, uint32_t dst) {
        for (unsigned c0 = 0; c0 < nbits; c0 += CHUNK) {
            Fr v = Fr::zero();
            const unsigned n = min(CHUNK, nbits - c0);
            for (unsigned i = 0; i < n; ++i) v.l[i >> 5] |= load_bit(w, first_bit + c0 + i) << (i & 31);
            fp_store(w + 8 * uint64_t(dst + c0 / CHUNK), v);
        }
    };
    pack(L.eid, L.eid_bits, L.eid_packed);
    pack(sn_bits, DIGEST, L.sn_packed);
    pack(rt_bits, DIGEST, L.rt_packed);
}

__global__ void k_tables_to_mont(uint32_t* t, uint64_t n) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp_store(t + 8 * i, fp_to_mont(fp_load<FrCfg>(t + 8 * i)));
}
}  // namespace

struct vsp_voting_program {
    vsp_ctx* ctx = nullptr;
    Layout L;
    std::vector<HashDesc> hashes;
    std::vector<LevelDesc> levels;
    HashDesc* d_hashes = nullptr;
    LevelDesc* d_levels = nullptr;
    uint32_t* d_in_idx = nullptr;
    uint32_t* d_tables = nullptr;  // Montgomery form; the last three rows are A, d, scale
    Consts cst;
};

extern "C" void vsp_voting_program_destroy(vsp_voting_program* p) {
    if (!p) return;
    cudaSetDevice(p->ctx->device);
    cudaDeviceSynchronize();
    if (p->d_hashes) cudaFree(p->d_hashes);
    if (p->d_levels) cudaFree(p->d_levels);
    if (p->d_in_idx) cudaFree(p->d_in_idx);
    if (p->d_tables) cudaFree(p->d_tables);
    delete p;
}

// scalars: the 15 layout fields of voting.py's device_program(); hashes: n_hashes x 6 u32; in_idx: the input-bit variable of
// every hash; levels: depth x 4 u32; tables: canonical Fr rows, (segment, window, k) -> x | y, followed by A, d, scale
extern "C" int vsp_voting_program_create(vsp_ctx* ctx, const uint32_t* scalars, uint64_t n_scalars, const uint32_t* hashes,
                                         uint64_t n_hashes, const uint32_t* in_idx, uint64_t n_in, const uint32_t* levels,
                                         uint64_t depth, const uint64_t* tables, uint64_t n_table_rows,
                                         vsp_voting_program** out) {
    if (!ctx || !scalars || !hashes || !in_idx || !tables || !out || n_scalars != 15 || n_hashes == 0 || (depth && !levels))
        return VSP_E_INVALID;
    *out = nullptr;
    cudaSetDevice(ctx->device);
    vsp_voting_program* p = new (std::nothrow) vsp_voting_program();
    if (!p) return VSP_E_NOMEM;
    p->ctx = ctx;
    memcpy(&p->L, scalars, sizeof(Layout));
    const Layout& L = p->L;
    bool ok = L.depth == depth && L.max_seg >= 1 && L.max_seg <= 8 && n_table_rows == uint64_t(L.max_seg) * WPS * 4 * 2 + 3 &&
              n_hashes == 2 + depth && L.h_pk < n_hashes && L.h_sn < n_hashes && L.msg_size >= 1 && L.eid_bits <= 64;
    p->hashes.resize(n_hashes);
    memcpy(p->hashes.data(), hashes, n_hashes * sizeof(HashDesc));
    p->levels.resize(depth);
    if (depth) memcpy(p->levels.data(), levels, depth * sizeof(LevelDesc));
    for (const HashDesc& h : p->hashes)
        ok = ok && h.segments >= 1 && h.segments <= L.max_seg && h.windows == (h.n_bits + 2) / 3 && uint64_t(h.in_off) + h.n_bits <= n_in &&
             h.out_base + DIGEST <= L.num_vars + 1;
    for (uint64_t i = 0; ok && i < n_in; ++i) ok = in_idx[i] <= L.num_vars;
    for (const LevelDesc& lv : p->levels) ok = ok && lv.hash < n_hashes && lv.right + DIGEST <= L.num_vars + 1;
    if (!ok) {
        delete p;
        return set_error(ctx, VSP_E_INVALID, "voting program: inconsistent layout tables");
    }
    cudaStream_t s = ctx->stream;
    int rc = VSP_OK;
    auto up = [&](void** d, const void* h, size_t bytes) {
        if (rc != VSP_OK) return;
        if (cudaMalloc(d, bytes ? bytes : 16) != cudaSuccess || (bytes && cudaMemcpyAsync(*d, h, bytes, cudaMemcpyHostToDevice, s) != cudaSuccess))
            rc = set_error(ctx, VSP_E_CUDA, "voting program upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    };
    up((void**)&p->d_hashes, hashes, n_hashes * sizeof(HashDesc));
    up((void**)&p->d_levels, levels, depth * sizeof(LevelDesc));
    up((void**)&p->d_in_idx, in_idx, n_in * 4);
    up((void**)&p->d_tables, tables, n_table_rows * 32);
    if (rc == VSP_OK) {
        VSP_LAUNCH_K(ctx, KC_OTHER, k_tables_to_mont, (unsigned)((n_table_rows + 127) / 128), 128, 0, s, p->d_tables, n_table_rows);
        uint32_t c[24];
        if (cudaMemcpyAsync(c, p->d_tables + 8 * (n_table_rows - 3), 96, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess)
            rc = set_error(ctx, VSP_E_CUDA, "voting program upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        memcpy(p->cst.a.l, c, 32);
        memcpy(p->cst.d.l, c + 8, 32);
        memcpy(p->cst.scale.l, c + 16, 32);
    }
    if (rc != VSP_OK) {
        vsp_voting_program_destroy(p);
        return rc;
    }
    *out = p;
    return VSP_OK;
}

extern "C" uint64_t vsp_voting_program_num_vars(const vsp_voting_program* p) { return p ? p->L.num_vars : 0; }

// votes / indices: k u32; sks: k x 4 u64 (bit i of the secret = bit i of the integer); paths: k x depth x 4 u64 sibling digests,
// leaf level first; d_assignments: k x (num_vars + 1) x 4 u64 on the device.  The small per-ballot inputs are host arrays.
extern "C" int vsp_voting_witness_batch_dev(vsp_ctx* ctx, const vsp_voting_program* p, uint64_t k, const uint32_t* votes,
                                            uint64_t eid, const uint64_t* sks, const uint32_t* indices, const uint64_t* paths,
                                            uint64_t* d_assignments) {
    if (!ctx || !p || !d_assignments || (k && (!votes || !sks || !indices || (p->L.depth && !paths)))) return VSP_E_INVALID;
    if (k == 0) return VSP_OK;
    if (k > (1u << 22)) return set_error(ctx, VSP_E_INVALID, "voting witness: batch too large");
    cudaSetDevice(ctx->device);
    const Layout& L = p->L;
    for (uint64_t b = 0; b < k; ++b) {
        if (votes[b] >= L.msg_size) return set_error(ctx, VSP_E_INVALID, "vote %u out of range (msg_size %u)", votes[b], L.msg_size);
        if (L.depth < 32 && (indices[b] >> L.depth)) return set_error(ctx, VSP_E_INVALID, "voter index %u outside a depth-%u tree", indices[b], L.depth);
        if (sks[4 * b + 3] >> 63) return set_error(ctx, VSP_E_INVALID, "secret key wider than 255 bits");
    }
    if (L.eid_bits < 64 && (eid >> L.eid_bits)) return set_error(ctx, VSP_E_INVALID, "eid wider than eid_bits");
    cudaStream_t s = ctx->stream;
    const unsigned K = (unsigned)k;
    DevBuf d_votes, d_sks, d_idx, d_paths;
    VSP_TRY(d_votes.alloc(ctx, k * 4, s));
    VSP_TRY(d_sks.alloc(ctx, k * 32, s));
    VSP_TRY(d_idx.alloc(ctx, k * 4, s));
    VSP_TRY(d_paths.alloc(ctx, k * (L.depth ? L.depth : 1) * 32, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_votes.p, votes, k * 4, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_sks.p, sks, k * 32, cudaMemcpyHostToDevice, s));
    VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_idx.p, indices, k * 4, cudaMemcpyHostToDevice, s));
    if (L.depth) VSP_CUDA_CHECK(ctx, cudaMemcpyAsync(d_paths.p, paths, k * L.depth * 32, cudaMemcpyHostToDevice, s));
    uint32_t* a = reinterpret_cast<uint32_t*>(d_assignments);
    VSP_CUDA_CHECK(ctx, cudaMemsetAsync(a, 0, k * (uint64_t(L.num_vars) + 1) * 32, s));
    const uint64_t per = 2 + L.eid_bits + DIGEST + L.depth + uint64_t(L.depth) * DIGEST;
    VSP_LAUNCH_K(ctx, KC_QAP, k_vote_inputs, (unsigned)((k * per + 127) / 128), 128, 0, s, L, p->d_levels, d_votes.as<uint32_t>(), eid,
                 d_sks.as<uint64_t>(), d_idx.as<uint32_t>(), d_paths.as<uint64_t>(), K, a);
    auto run_hashes = [&](uint32_t h0, unsigned n) {
        VSP_LAUNCH_K(ctx, KC_QAP, k_pedersen_segment, dim3((K + 63) / 64, n * L.max_seg), 64, 0, s, L, p->d_hashes, p->d_in_idx,
                     p->d_tables, p->cst, h0, K, a);
        VSP_LAUNCH_K(ctx, KC_QAP, k_pedersen_finish, dim3((K + 63) / 64, n), 64, 0, s, L, p->d_hashes, p->cst, h0, K, a);
    };
    // stage 0: pk = H(sk) and sn = H(eid || sk); they are hashes 0 and 1 of the program
    if (L.h_pk + 1 == L.h_sn) {
        run_hashes(L.h_pk, 2);
    } else {
        run_hashes(L.h_pk, 1);
        run_hashes(L.h_sn, 1);
    }
    uint32_t cur = p->hashes[L.h_pk].out_base;
    for (unsigned lvl = 0; lvl < L.depth; ++lvl) {
        const LevelDesc lv = p->levels[lvl];
        VSP_LAUNCH_K(ctx, KC_QAP, k_merkle_select, (unsigned)((k * DIGEST + 127) / 128), 128, 0, s, L, lv, cur, lvl, K, a);
        run_hashes(lv.hash, 1);
        cur = p->hashes[lv.hash].out_base;
    }
    VSP_LAUNCH_K(ctx, KC_QAP, k_vote_pack, (K + 63) / 64, 64, 0, s, L, p->hashes[L.h_sn].out_base, cur, K, a);
    VSP_CUDA_CHECK(ctx, cudaGetLastError());
    VSP_CUDA_CHECK(ctx, cudaStreamSynchronize(s));  // the per-ballot inputs are host arrays of the caller
    return VSP_OK;
}
