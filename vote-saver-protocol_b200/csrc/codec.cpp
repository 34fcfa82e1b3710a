// codec.cpp — the reference's wire formats either side of the hot path (host only, no GPU work): what
// process_encrypted_input_mode_vote_phase reads (the "fast" uncompressed proving key) and writes (compressed proof,
// cipher text, primary-input slices).  SURVEY.md 8f rank 1.
//
// Formats restated from (paths relative to /root/reference/libs/marshalling):
//   integers            fixed width, big-endian (bin/cli/include/nil/vote_saver/common.hpp:180), width = ceil(bits / 8):
//                       Fr 32 B, Fq 48 B  (multiprecision/.../detail/integral/basic_fixed_precision_type.hpp:84-95);
//                       std::size_t 8 B
//   sequences           8-byte count prefix, then the elements (option::sequence_size_field_prefix<integral<size_t>>)
//   field_element       Fq2 = c0 | c1   (algebra/.../types/field_element.hpp:343-390: data[0], data[1])
//   fast_curve_element  X | Y | is_infinity:u8, rebuilt as (X, Y, 1)   (algebra/.../types/fast_curve_element.hpp:56-97)
//   curve_element       compressed: G1 48 B big-endian x; G2 96 B x.c1 | x.c0; top bits of byte 0: C = 0x80, I = 0x40,
//                       S = 0x20 set iff y > (p - 1) / 2, for G2 decided on y.c1 unless it is zero, then y.c0
//                       (algebra/.../processing/curve_element.hpp:93-174, processing/detail/curve_element.hpp:46-92)
//   proof               g_A | g_B | g_C compressed = 192 B   (zk/.../r1cs_gg_ppzksnark/proof.hpp:63-71)
//   primary input       count | Fr...;  encrypted primary input (cipher text) count | compressed G1...
//                       (zk/.../r1cs_gg_ppzksnark/primary_input.hpp:61-97)
//   fast proving key    alpha_g1 beta_g1 beta_g2 delta_g1 delta_g2 | A_query | B_query (indices, (g: G2, h: G1) pairs,
//                       domain_size) | H_query | L_query | constraint system
//                       (zk/.../r1cs_gg_ppzksnark/fast_proving_key.hpp:66-103, sparse_vector.hpp:103-115,
//                        fast_knowledge_commitment.hpp:62-68, r1cs.hpp:57-113: linear term = index:size_t | coeff:Fr)
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/vsp_b200.h"

namespace {

constexpr size_t FR_BYTES = 32, FQ_BYTES = 48, G1_FAST = 2 * FQ_BYTES + 1, G2_FAST = 4 * FQ_BYTES + 1;

// (p - 1) / 2, little-endian 64-bit limbs
const uint64_t HALF_P[6] = {0xdcff7fffffffd555ull, 0x0f55ffff58a9ffffull, 0xb39869507b587b12ull,
                            0xb23ba5c279c2895full, 0x258dd3db21a5d66bull, 0x0d0088f51cbff34dull};

// big-endian bytes <-> little-endian limbs
void put_be(uint8_t* out, const uint64_t* limbs, size_t bytes) {
    for (size_t k = 0; k < bytes; ++k) {
        const size_t bit = 8 * (bytes - 1 - k);
        out[k] = (uint8_t)(limbs[bit / 64] >> (bit % 64));
    }
}
void get_be(const uint8_t* in, uint64_t* limbs, size_t bytes) {
    for (size_t i = 0; i < bytes / 8; ++i) limbs[i] = 0;
    for (size_t k = 0; k < bytes; ++k) {
        const size_t bit = 8 * (bytes - 1 - k);
        limbs[bit / 64] |= (uint64_t)in[k] << (bit % 64);
    }
}
void put_u64(uint8_t* out, uint64_t v) { put_be(out, &v, 8); }
uint64_t get_u64(const uint8_t* in) {
    uint64_t v;
    get_be(in, &v, 8);
    return v;
}
bool is_zero(const uint64_t* x, int n) {
    uint64_t a = 0;
    for (int i = 0; i < n; ++i) a |= x[i];
    return a == 0;
}
// y > (p - 1) / 2
bool fq_is_high(const uint64_t* y) {
    for (int i = 5; i >= 0; --i) {
        if (y[i] > HALF_P[i]) return true;
        if (y[i] < HALF_P[i]) return false;
    }
    return false;
}

void g1_compress(const uint64_t* p, uint8_t* out) {
    if (is_zero(p, 12)) {
        memset(out, 0, FQ_BYTES);
        out[0] = 0x80 | 0x40;
        return;
    }
    put_be(out, p, FQ_BYTES);
    out[0] |= 0x80 | (fq_is_high(p + 6) ? 0x20 : 0);
}
void g2_compress(const uint64_t* p, uint8_t* out) {
    if (is_zero(p, 24)) {
        memset(out, 0, 2 * FQ_BYTES);
        out[0] = 0x80 | 0x40;
        return;
    }
    put_be(out, p + 6, FQ_BYTES);        // x.c1
    put_be(out + FQ_BYTES, p, FQ_BYTES);  // x.c0
    const uint64_t *y0 = p + 12, *y1 = p + 18;
    const bool high = is_zero(y1, 6) ? fq_is_high(y0) : fq_is_high(y1);
    out[0] |= 0x80 | (high ? 0x20 : 0);
}

// fast_curve_element: X | Y | is_infinity
void fast_write(uint8_t* out, const uint64_t* p, int fq_count) {  // fq_count = 2 (G1) or 4 (G2)
    const bool inf = is_zero(p, 6 * fq_count);
    if (inf) {
        // the zero point's affine form is (0, 1) in Jacobian conventions; readers ignore X, Y when the flag is set
        memset(out, 0, FQ_BYTES * fq_count);
        out[FQ_BYTES * (fq_count / 2) + FQ_BYTES - 1] = 1;
    } else {
        for (int i = 0; i < fq_count; ++i) put_be(out + FQ_BYTES * i, p + 6 * i, FQ_BYTES);
    }
    out[FQ_BYTES * fq_count] = inf ? 1 : 0;
}
void fast_read(const uint8_t* in, uint64_t* p, int fq_count) {
    if (in[FQ_BYTES * fq_count]) {
        memset(p, 0, 48 * fq_count);
        return;
    }
    for (int i = 0; i < fq_count; ++i) get_be(in + FQ_BYTES * i, p + 6 * i, FQ_BYTES);
}

// ---- host Fq / Fq2 for the decoders: y from x (processing/curve_element.hpp:176-395 take the square root of x^3 + b
// and pick the root whose sign matches the S bit).  Plain 6 x 64-bit limbs, schoolbook product + shift-subtract-free
// Montgomery reduction; decoding is a few points per ballot, nowhere near the hot path.
typedef unsigned __int128 u128;
const uint64_t P64[6] = {0xb9feffffffffaaabull, 0x1eabfffeb153ffffull, 0x6730d2a0f6b0f624ull,
                         0x64774b84f38512bfull, 0x4b1ba7b6434bacd7ull, 0x1a0111ea397fe69aull};
const uint64_t P_INV64 = 0x89f3fffcfffcfffdull;  // -p^-1 mod 2^64
struct HFq {
    uint64_t l[6];
};
bool hfq_geq_p(const uint64_t* x) {
    for (int i = 5; i >= 0; --i) {
        if (x[i] > P64[i]) return true;
        if (x[i] < P64[i]) return false;
    }
    return true;
}
bool hfq_is_zero(const HFq& a) { return is_zero(a.l, 6); }
bool hfq_eq(const HFq& a, const HFq& b) { return memcmp(a.l, b.l, 48) == 0; }
HFq hfq_add(const HFq& a, const HFq& b) {
    HFq r;
    u128 c = 0;
    for (int i = 0; i < 6; ++i) {
        c += (u128)a.l[i] + b.l[i];
        r.l[i] = (uint64_t)c;
        c >>= 64;
    }
    if (hfq_geq_p(r.l)) {
        u128 br = 0;
        for (int i = 0; i < 6; ++i) {
            const u128 d = (u128)r.l[i] - P64[i] - (uint64_t)br;
            r.l[i] = (uint64_t)d;
            br = (d >> 64) & 1;
        }
    }
    return r;
}
HFq hfq_neg(const HFq& a) {
    if (hfq_is_zero(a)) return a;
    HFq r;
    u128 br = 0;
    for (int i = 0; i < 6; ++i) {
        const u128 d = (u128)P64[i] - a.l[i] - (uint64_t)br;
        r.l[i] = (uint64_t)d;
        br = (d >> 64) & 1;
    }
    return r;
}
HFq hfq_sub(const HFq& a, const HFq& b) { return hfq_add(a, hfq_neg(b)); }
// Montgomery product a b / 2^384 mod p (CIOS)
HFq hfq_mmul(const HFq& a, const HFq& b) {
    uint64_t t[8] = {0};
    for (int i = 0; i < 6; ++i) {
        u128 c = 0;
        for (int j = 0; j < 6; ++j) {
            c += (u128)a.l[j] * b.l[i] + t[j];
            t[j] = (uint64_t)c;
            c >>= 64;
        }
        c += t[6];
        t[6] = (uint64_t)c;
        t[7] = (uint64_t)(c >> 64);
        const uint64_t m = t[0] * P_INV64;
        c = (u128)m * P64[0] + t[0];
        c >>= 64;
        for (int j = 1; j < 6; ++j) {
            c += (u128)m * P64[j] + t[j];
            t[j - 1] = (uint64_t)c;
            c >>= 64;
        }
        c += t[6];
        t[5] = (uint64_t)c;
        t[6] = t[7] + (uint64_t)(c >> 64);
    }
    HFq r;
    memcpy(r.l, t, 48);
    if (t[6] || hfq_geq_p(r.l)) {
        u128 br = 0;
        for (int i = 0; i < 6; ++i) {
            const u128 d = (u128)r.l[i] - P64[i] - (uint64_t)br;
            r.l[i] = (uint64_t)d;
            br = (d >> 64) & 1;
        }
    }
    return r;
}
// R^2 mod p by doubling 1 -> 2^768 mod p
HFq hfq_r2() {
    static HFq r2;
    static bool have = false;
    if (!have) {
        HFq x = {{1, 0, 0, 0, 0, 0}};
        for (int i = 0; i < 768; ++i) x = hfq_add(x, x);
        r2 = x;
        have = true;
    }
    return r2;
}
HFq hfq_to_mont(const HFq& a) { return hfq_mmul(a, hfq_r2()); }
HFq hfq_from_mont(const HFq& a) {
    const HFq one = {{1, 0, 0, 0, 0, 0}};
    return hfq_mmul(a, one);
}
// a^e (Montgomery form in and out), e little-endian limbs
HFq hfq_pow(const HFq& a, const uint64_t* e, int limbs) {
    HFq acc = hfq_to_mont(HFq{{1, 0, 0, 0, 0, 0}});
    for (int bit = 64 * limbs - 1; bit >= 0; --bit) {
        acc = hfq_mmul(acc, acc);
        if ((e[bit / 64] >> (bit % 64)) & 1) acc = hfq_mmul(acc, a);
    }
    return acc;
}
// square root (p = 3 mod 4: a^((p + 1) / 4)); false if a is not a square
bool hfq_sqrt(const HFq& a, HFq* out) {
    uint64_t e[6];
    u128 c = 1;
    for (int i = 0; i < 6; ++i) {  // p + 1
        c += P64[i];
        e[i] = (uint64_t)c;
        c >>= 64;
    }
    for (int i = 0; i < 6; ++i) e[i] = (e[i] >> 2) | (i < 5 ? e[i + 1] << 62 : 0);
    const HFq r = hfq_pow(a, e, 6);
    if (!hfq_eq(hfq_mmul(r, r), a)) return false;
    *out = r;
    return true;
}
HFq hfq_inv(const HFq& a) {
    uint64_t e[6];
    memcpy(e, P64, 48);
    e[0] -= 2;  // p - 2 (no borrow: the low limb ends in ...aaab)
    return hfq_pow(a, e, 6);
}
struct HFq2 {
    HFq c0, c1;
};
HFq2 hfq2_mul(const HFq2& a, const HFq2& b) {
    const HFq t0 = hfq_mmul(a.c0, b.c0), t1 = hfq_mmul(a.c1, b.c1);
    return HFq2{hfq_sub(t0, t1), hfq_add(hfq_mmul(a.c0, b.c1), hfq_mmul(a.c1, b.c0))};
}
// a square root in Fq2 = Fq[u] / (u^2 + 1) by the norm: x0^2 = (a0 +- sqrt(a0^2 + a1^2)) / 2, x1 = a1 / (2 x0)
bool hfq2_sqrt(const HFq2& a, HFq2* out) {
    const HFq zero = {{0, 0, 0, 0, 0, 0}};
    if (hfq_is_zero(a.c1)) {
        HFq r;
        if (hfq_sqrt(a.c0, &r)) {
            *out = HFq2{r, zero};
            return true;
        }
        if (hfq_sqrt(hfq_neg(a.c0), &r)) {
            *out = HFq2{zero, r};
            return true;
        }
        return false;
    }
    HFq n;
    if (!hfq_sqrt(hfq_add(hfq_mmul(a.c0, a.c0), hfq_mmul(a.c1, a.c1)), &n)) return false;
    const HFq two = hfq_to_mont(HFq{{2, 0, 0, 0, 0, 0}}), inv2 = hfq_inv(two);
    for (int sgn = 0; sgn < 2; ++sgn) {
        const HFq nn = sgn ? hfq_neg(n) : n;
        HFq x0;
        if (!hfq_sqrt(hfq_mmul(hfq_add(a.c0, nn), inv2), &x0) || hfq_is_zero(x0)) continue;
        const HFq x1 = hfq_mmul(a.c1, hfq_inv(hfq_add(x0, x0)));
        const HFq2 r = {x0, x1}, rr = hfq2_mul(r, r);
        if (hfq_eq(rr.c0, a.c0) && hfq_eq(rr.c1, a.c1)) {
            *out = r;
            return true;
        }
    }
    return false;
}

// compressed octets -> canonical affine limbs.  VSP_E_INVALID where the reference's decoder fails: C bit clear, a
// non-canonical infinity, x >= p, x not on the curve.
int g1_decompress(const uint8_t* in, uint64_t* p) {
    if (!(in[0] & 0x80)) return VSP_E_INVALID;
    if (in[0] & 0x40) {
        if (in[0] & 0x3f) return VSP_E_INVALID;
        for (size_t i = 1; i < FQ_BYTES; ++i)
            if (in[i]) return VSP_E_INVALID;
        memset(p, 0, 96);
        return VSP_OK;
    }
    uint8_t xb[FQ_BYTES];
    memcpy(xb, in, FQ_BYTES);
    xb[0] &= 0x1f;
    HFq x;
    get_be(xb, x.l, FQ_BYTES);
    if (hfq_geq_p(x.l)) return VSP_E_INVALID;
    const HFq xm = hfq_to_mont(x), b = hfq_to_mont(HFq{{4, 0, 0, 0, 0, 0}});
    HFq ym;
    if (!hfq_sqrt(hfq_add(hfq_mmul(hfq_mmul(xm, xm), xm), b), &ym)) return VSP_E_INVALID;
    HFq y = hfq_from_mont(ym);
    if (fq_is_high(y.l) != bool(in[0] & 0x20)) y = hfq_neg(y);
    memcpy(p, x.l, 48);
    memcpy(p + 6, y.l, 48);
    return VSP_OK;
}
int g2_decompress(const uint8_t* in, uint64_t* p) {
    if (!(in[0] & 0x80)) return VSP_E_INVALID;
    if (in[0] & 0x40) {
        if (in[0] & 0x3f) return VSP_E_INVALID;
        for (size_t i = 1; i < 2 * FQ_BYTES; ++i)
            if (in[i]) return VSP_E_INVALID;
        memset(p, 0, 192);
        return VSP_OK;
    }
    uint8_t xb[FQ_BYTES];
    memcpy(xb, in, FQ_BYTES);
    xb[0] &= 0x1f;
    HFq x1, x0;
    get_be(xb, x1.l, FQ_BYTES);
    get_be(in + FQ_BYTES, x0.l, FQ_BYTES);
    if (hfq_geq_p(x0.l) || hfq_geq_p(x1.l)) return VSP_E_INVALID;
    const HFq2 xm = {hfq_to_mont(x0), hfq_to_mont(x1)};
    const HFq four = hfq_to_mont(HFq{{4, 0, 0, 0, 0, 0}});
    HFq2 rhs = hfq2_mul(hfq2_mul(xm, xm), xm);
    rhs.c0 = hfq_add(rhs.c0, four);
    rhs.c1 = hfq_add(rhs.c1, four);
    HFq2 ym;
    if (!hfq2_sqrt(rhs, &ym)) return VSP_E_INVALID;
    HFq y0 = hfq_from_mont(ym.c0), y1 = hfq_from_mont(ym.c1);
    const bool high = hfq_is_zero(y1) ? fq_is_high(y0.l) : fq_is_high(y1.l);
    if (high != bool(in[0] & 0x20)) {
        y0 = hfq_neg(y0);
        y1 = hfq_neg(y1);
    }
    memcpy(p, x0.l, 48);
    memcpy(p + 6, x1.l, 48);
    memcpy(p + 12, y0.l, 48);
    memcpy(p + 18, y1.l, 48);
    return VSP_OK;
}

struct Reader {
    const uint8_t* p;
    size_t left;
    bool ok = true;
    const uint8_t* take(size_t n) {
        if (!ok || left < n) {
            ok = false;
            return nullptr;
        }
        const uint8_t* r = p;
        p += n;
        left -= n;
        return r;
    }
    uint64_t u64() {
        const uint8_t* q = take(8);
        return q ? get_u64(q) : 0;
    }
};

}  // namespace

extern "C" {

int vsp_codec_g1_compress(const uint64_t* point, uint8_t* out48) {
    if (!point || !out48) return VSP_E_INVALID;
    g1_compress(point, out48);
    return VSP_OK;
}
int vsp_codec_g2_compress(const uint64_t* point, uint8_t* out96) {
    if (!point || !out96) return VSP_E_INVALID;
    g2_compress(point, out96);
    return VSP_OK;
}
// compressed octets -> canonical affine limbs (processing/curve_element.hpp:176-395); VSP_E_INVALID for a malformed
// encoding or an x with no point on the curve.  Subgroup membership is the verifier's check, as in the reference.
int vsp_codec_g1_decompress(const uint8_t* in48, uint64_t* point) {
    if (!in48 || !point) return VSP_E_INVALID;
    return g1_decompress(in48, point);
}
int vsp_codec_g2_decompress(const uint8_t* in96, uint64_t* point) {
    if (!in96 || !point) return VSP_E_INVALID;
    return g2_decompress(in96, point);
}
// 192 bytes -> proof (g_A | g_B | g_C as 48 limbs)   (proof.hpp:86-102)
int vsp_codec_proof_read(const uint8_t* in192, uint64_t* proof) {
    if (!in192 || !proof) return VSP_E_INVALID;
    int rc = g1_decompress(in192, proof);
    if (rc == VSP_OK) rc = g2_decompress(in192 + 48, proof + 12);
    if (rc == VSP_OK) rc = g1_decompress(in192 + 144, proof + 36);
    return rc;
}
// proof (g_A | g_B | g_C as 48 limbs) -> 192 bytes
int vsp_codec_proof_write(const uint64_t* proof, uint8_t* out192) {
    if (!proof || !out192) return VSP_E_INVALID;
    g1_compress(proof, out192);
    g2_compress(proof + 12, out192 + 48);
    g1_compress(proof + 36, out192 + 144);
    return VSP_OK;
}
// cipher text of n G1 points -> 8 + 48 n bytes
int vsp_codec_ciphertext_write(const uint64_t* ct, uint64_t n, uint8_t* out) {
    if ((!ct && n) || !out) return VSP_E_INVALID;
    put_u64(out, n);
    for (uint64_t i = 0; i < n; ++i) g1_compress(ct + 12 * i, out + 8 + 48 * i);
    return VSP_OK;
}
// n Fr -> 8 + 32 n bytes
int vsp_codec_primary_input_write(const uint64_t* values, uint64_t n, uint8_t* out) {
    if ((!values && n) || !out) return VSP_E_INVALID;
    put_u64(out, n);
    for (uint64_t i = 0; i < n; ++i) put_be(out + 8 + FR_BYTES * i, values + 4 * i, FR_BYTES);
    return VSP_OK;
}
// 8 + 32 n bytes -> n Fr; *n receives the count; values may be NULL to query it; capacity in elements
int vsp_codec_primary_input_read(const uint8_t* in, size_t len, uint64_t* n, uint64_t* values, uint64_t capacity) {
    if (!in || !n || len < 8) return VSP_E_INVALID;
    const uint64_t cnt = get_u64(in);
    if (cnt > (len - 8) / FR_BYTES) return VSP_E_INVALID;
    *n = cnt;
    if (!values) return VSP_OK;
    if (capacity < cnt) return VSP_E_INVALID;
    for (uint64_t i = 0; i < cnt; ++i) get_be(in + 8 + FR_BYTES * i, values + 4 * i, FR_BYTES);
    return VSP_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// fast proving key  (r1cs_proving_key.bin as bin/cli reads it: common.hpp:1004-1006)
// ---------------------------------------------------------------------------------------------------------------------
namespace {

// walks one linear combination; with cols/vals non-null also stores it at position *pos
bool walk_lc(Reader& r, uint64_t* nnz, uint32_t* cols, uint64_t* vals, uint64_t n_vars) {
    const uint64_t terms = r.u64();
    if (!r.ok || terms > r.left / (8 + FR_BYTES)) return r.ok = false;
    for (uint64_t t = 0; t < terms; ++t) {
        const uint64_t idx = r.u64();
        const uint8_t* c = r.take(FR_BYTES);
        if (!c) return false;
        if (cols) {
            if (idx > n_vars) return r.ok = false;
            cols[*nnz] = (uint32_t)idx;
            get_be(c, vals + 4 * *nnz, FR_BYTES);
        }
        ++*nnz;
    }
    return true;
}

// one pass over the blob: sizes only (desc == nullptr) or fill the caller's arrays
int fast_pk_walk(const uint8_t* blob, size_t len, vsp_fast_pk_sizes_t* sz, vsp_proving_key_t* d) {
    Reader r {blob, len};
    auto g1 = [&](const uint64_t* dst) {
        const uint8_t* q = r.take(G1_FAST);
        if (q && dst) fast_read(q, const_cast<uint64_t*>(dst), 2);
    };
    auto g2 = [&](const uint64_t* dst) {
        const uint8_t* q = r.take(G2_FAST);
        if (q && dst) fast_read(q, const_cast<uint64_t*>(dst), 4);
    };
    auto g1_list = [&](const uint64_t* dst, uint64_t* count) {
        const uint64_t n = r.u64();
        if (!r.ok || n > r.left / G1_FAST) {
            r.ok = false;
            return;
        }
        *count = n;
        for (uint64_t i = 0; i < n; ++i) g1(dst ? dst + 12 * i : nullptr);
    };
    g1(d ? d->alpha_g1 : nullptr);
    g1(d ? d->beta_g1 : nullptr);
    g2(d ? d->beta_g2 : nullptr);
    g1(d ? d->delta_g1 : nullptr);
    g2(d ? d->delta_g2 : nullptr);
    vsp_fast_pk_sizes_t s;
    memset(&s, 0, sizeof(s));
    g1_list(d ? d->A_query : nullptr, &s.A_count);
    // B_query: indices | (g: G2, h: G1) pairs | domain_size
    const uint64_t n_idx = r.u64();
    if (!r.ok || n_idx > r.left / 8) return VSP_E_INVALID;
    for (uint64_t i = 0; i < n_idx; ++i) {
        const uint64_t v = r.u64();
        if (d) {
            if (v > 0xffffffffull) return VSP_E_INVALID;
            const_cast<uint32_t*>(d->B_idx)[i] = (uint32_t)v;
        }
    }
    const uint64_t n_val = r.u64();
    if (!r.ok || n_val != n_idx || n_val > r.left / (G1_FAST + G2_FAST)) return VSP_E_INVALID;
    for (uint64_t i = 0; i < n_val; ++i) {
        g2(d ? d->B_g2 + 24 * i : nullptr);
        g1(d ? d->B_g1 + 12 * i : nullptr);
    }
    s.B_count = n_val;
    s.B_domain_size = r.u64();
    g1_list(d ? d->H_query : nullptr, &s.H_count);
    g1_list(d ? d->L_query : nullptr, &s.L_count);
    // constraint system: primary_input_size | auxiliary_input_size | constraints
    s.num_inputs = r.u64();
    s.num_aux = r.u64();
    s.num_constraints = r.u64();
    if (!r.ok || s.num_constraints > r.left / 24) return VSP_E_INVALID;
    const uint64_t n_vars = s.num_inputs + s.num_aux;
    uint64_t nnz[3] = {0, 0, 0};
    for (uint64_t i = 0; i < s.num_constraints; ++i) {
        for (int k = 0; k < 3; ++k) {
            uint64_t* row = nullptr;
            uint32_t* col = nullptr;
            uint64_t* val = nullptr;
            if (d) {
                row = const_cast<uint64_t*>(k == 0 ? d->cs.a_row : k == 1 ? d->cs.b_row : d->cs.c_row);
                col = const_cast<uint32_t*>(k == 0 ? d->cs.a_col : k == 1 ? d->cs.b_col : d->cs.c_col);
                val = const_cast<uint64_t*>(k == 0 ? d->cs.a_val : k == 1 ? d->cs.b_val : d->cs.c_val);
                row[i] = nnz[k];
            }
            if (!walk_lc(r, &nnz[k], col, val, n_vars)) return VSP_E_INVALID;
            if (d) row[i + 1] = nnz[k];
        }
    }
    if (!r.ok) return VSP_E_INVALID;
    s.nnz_a = nnz[0];
    s.nnz_b = nnz[1];
    s.nnz_c = nnz[2];
    s.bytes = len - r.left;
    if (d) {
        if (s.num_constraints == 0) {
            const_cast<uint64_t*>(d->cs.a_row)[0] = const_cast<uint64_t*>(d->cs.b_row)[0] = 0;
            const_cast<uint64_t*>(d->cs.c_row)[0] = 0;
        }
        d->cs.num_inputs = s.num_inputs;
        d->cs.num_aux = s.num_aux;
        d->cs.num_constraints = s.num_constraints;
        d->B_count = s.B_count;
        d->H_count = s.H_count;
    }
    if (sz) *sz = s;
    return VSP_OK;
}

}  // namespace

extern "C" {

int vsp_codec_fast_pk_sizes(const uint8_t* blob, size_t len, vsp_fast_pk_sizes_t* out) {
    if (!blob || !out) return VSP_E_INVALID;
    return fast_pk_walk(blob, len, out, nullptr);
}

// desc's array pointers must address caller-allocated writable memory of the sizes vsp_codec_fast_pk_sizes reported
// (rows: num_constraints + 1; cols / vals: nnz; A_query: A_count G1; B_*: B_count; H_query: H_count; L_query: L_count);
// the scalar fields of desc are filled in
int vsp_codec_fast_pk_read(const uint8_t* blob, size_t len, vsp_proving_key_t* desc) {
    if (!blob || !desc) return VSP_E_INVALID;
    return fast_pk_walk(blob, len, nullptr, desc);
}

// serialised size of a key / write it (the inverse of vsp_codec_fast_pk_read); A_query holds num_inputs + num_aux + 1
// points, L_query num_aux, H_query H_count; B's domain_size is written as num_inputs + num_aux + 1
int vsp_codec_fast_pk_write(const vsp_proving_key_t* d, uint8_t* out, size_t capacity, size_t* written) {
    if (!d || !written) return VSP_E_INVALID;
    const uint64_t nv1 = d->cs.num_inputs + d->cs.num_aux + 1, nc = d->cs.num_constraints;
    const uint64_t nnz = d->cs.a_row[nc] + d->cs.b_row[nc] + d->cs.c_row[nc];
    const size_t need = 3 * G1_FAST + 2 * G2_FAST + 8 + nv1 * G1_FAST + 8 + d->B_count * 8 + 8 +
                        d->B_count * (G1_FAST + G2_FAST) + 8 + 8 + d->H_count * G1_FAST + 8 + d->cs.num_aux * G1_FAST +
                        24 + nc * 24 + nnz * (8 + FR_BYTES);
    *written = need;
    if (!out) return VSP_OK;
    if (capacity < need) return VSP_E_INVALID;
    uint8_t* p = out;
    auto g1 = [&](const uint64_t* src) { fast_write(p, src, 2); p += G1_FAST; };
    auto g2 = [&](const uint64_t* src) { fast_write(p, src, 4); p += G2_FAST; };
    auto u64 = [&](uint64_t v) { put_u64(p, v); p += 8; };
    g1(d->alpha_g1);
    g1(d->beta_g1);
    g2(d->beta_g2);
    g1(d->delta_g1);
    g2(d->delta_g2);
    u64(nv1);
    for (uint64_t i = 0; i < nv1; ++i) g1(d->A_query + 12 * i);
    u64(d->B_count);
    for (uint64_t i = 0; i < d->B_count; ++i) u64(d->B_idx[i]);
    u64(d->B_count);
    for (uint64_t i = 0; i < d->B_count; ++i) {
        g2(d->B_g2 + 24 * i);
        g1(d->B_g1 + 12 * i);
    }
    u64(nv1);
    u64(d->H_count);
    for (uint64_t i = 0; i < d->H_count; ++i) g1(d->H_query + 12 * i);
    u64(d->cs.num_aux);
    for (uint64_t i = 0; i < d->cs.num_aux; ++i) g1(d->L_query + 12 * i);
    u64(d->cs.num_inputs);
    u64(d->cs.num_aux);
    u64(nc);
    const uint64_t* rows[3] = {d->cs.a_row, d->cs.b_row, d->cs.c_row};
    const uint32_t* cols[3] = {d->cs.a_col, d->cs.b_col, d->cs.c_col};
    const uint64_t* vals[3] = {d->cs.a_val, d->cs.b_val, d->cs.c_val};
    for (uint64_t i = 0; i < nc; ++i)
        for (int k = 0; k < 3; ++k) {
            u64(rows[k][i + 1] - rows[k][i]);
            for (uint64_t q = rows[k][i]; q < rows[k][i + 1]; ++q) {
                u64(cols[k][q]);
                put_be(p, vals[k] + 4 * q, FR_BYTES);
                p += FR_BYTES;
            }
        }
    return (size_t)(p - out) == need ? VSP_OK : VSP_E_INVALID;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// The other key and cipher-text blobs of the vote and tally phases, all in compressed curve_element form:
//   cipher text (reader)          count | G1...                                 deserialize_ct, common.hpp:776
//   SAVER public key  pk_eid      delta_g1 | delta_sum_s_g1 | gamma_inverse_sum_s_g1 | delta_s_g1[] | t_g1[] | t_g2[]
//                                 (zk/.../pubkey/types/elgamal_verifiable.hpp: elgamal_verifiable_public_key, fill_public_key;
//                                  read by the vote phase, common.hpp:691, 1002)
//   SAVER verification key        rho_g2 | rho_sv_g2[] | rho_rhov_g2[]           (same file: fill_verification_key)
//   SAVER secret key              rho: Fr                                        (same file: fill_private_key)
//   Groth16 extended verif. key   alpha_g1_beta_g2 (12 Fq, tower order) | gamma_g2 | delta_g2 | delta_g1 | gamma_g1 |
//                                 gamma_ABC_g1 = first | rest as sparse_vector (indices[] | values[] | domain_size)
//                                 (zk/.../r1cs_gg_ppzksnark/verification_key.hpp:128-152, accumulation_vector.hpp:52-57,
//                                  sparse_vector.hpp:49-58; read by the vote phase, common.hpp:734, 1005)
// ---------------------------------------------------------------------------------------------------------------------
namespace {

constexpr size_t G1_C = FQ_BYTES, G2_C = 2 * FQ_BYTES;

// count-prefixed list of compressed points of `width` bytes; dst == nullptr: walk only.  `expect` != ~0: the count the
// caller's arrays were sized for
bool point_list(Reader& r, size_t width, uint64_t expect, uint64_t* dst, uint64_t* count) {
    const uint64_t n = r.u64();
    if (!r.ok || n > r.left / width) return r.ok = false;
    if (count) *count = n;
    if (dst && n != expect) return r.ok = false;
    const size_t limbs = width == G1_C ? 12 : 24;
    for (uint64_t i = 0; i < n; ++i) {
        const uint8_t* q = r.take(width);
        if (!q) return false;
        if (dst && (width == G1_C ? g1_decompress(q, dst + limbs * i) : g2_decompress(q, dst + limbs * i)) != VSP_OK)
            return r.ok = false;
    }
    return true;
}
bool point_one(Reader& r, size_t width, uint64_t* dst) {
    const uint8_t* q = r.take(width);
    if (!q) return false;
    if (dst && (width == G1_C ? g1_decompress(q, dst) : g2_decompress(q, dst)) != VSP_OK) return r.ok = false;
    return true;
}

struct Writer {
    uint8_t* p;
    void g1(const uint64_t* src) { g1_compress(src, p); p += G1_C; }
    void g2(const uint64_t* src) { g2_compress(src, p); p += G2_C; }
    void u64(uint64_t v) { put_u64(p, v); p += 8; }
    void g1_list(const uint64_t* src, uint64_t n) { u64(n); for (uint64_t i = 0; i < n; ++i) g1(src + 12 * i); }
    void g2_list(const uint64_t* src, uint64_t n) { u64(n); for (uint64_t i = 0; i < n; ++i) g2(src + 24 * i); }
};

}  // namespace

extern "C" {

// 8 + 48 n bytes -> n cipher-text G1 points; *n receives the count; ct may be NULL to query it; capacity in points
int vsp_codec_ciphertext_read(const uint8_t* in, size_t len, uint64_t* n, uint64_t* ct, uint64_t capacity) {
    if (!in || !n) return VSP_E_INVALID;
    Reader r {in, len};
    uint64_t cnt = 0;
    if (!point_list(r, G1_C, 0, nullptr, &cnt)) return VSP_E_INVALID;
    *n = cnt;
    if (!ct) return VSP_OK;
    if (capacity < cnt) return VSP_E_INVALID;
    Reader r2 {in, len};
    return point_list(r2, G1_C, cnt, ct, nullptr) ? VSP_OK : VSP_E_INVALID;
}

// size of / the serialised public key: msg_size points in delta_s_g1 and t_g1, msg_size + 1 in t_g2 (what
// generate_keypair_op builds, elgamal_verifiable.hpp:218-271)
int vsp_codec_saver_pubkey_write(const vsp_saver_pubkey_t* pub, uint8_t* out, size_t capacity, size_t* written) {
    if (!pub || !written) return VSP_E_INVALID;
    const uint64_t n = pub->msg_size;
    const size_t need = 3 * G1_C + 3 * 8 + 2 * n * G1_C + (n + 1) * G2_C;
    *written = need;
    if (!out) return VSP_OK;
    if (capacity < need) return VSP_E_INVALID;
    Writer w {out};
    w.g1(pub->delta_g1);
    w.g1(pub->delta_sum_s_g1);
    w.g1(pub->gamma_inverse_sum_s_g1);
    w.g1_list(pub->delta_s_g1, n);
    w.g1_list(pub->t_g1, n);
    w.g2_list(pub->t_g2, n + 1);
    return (size_t)(w.p - out) == need ? VSP_OK : VSP_E_INVALID;
}

// pub == NULL: *msg_size receives the blob's message size (the length of its delta_s_g1 list).  Otherwise pub->msg_size
// must equal it, pub's six arrays are caller-allocated for that size and are filled.  VSP_E_INVALID for a truncated blob,
// a malformed point, or lists that are not msg_size | msg_size | msg_size + 1 long.
int vsp_codec_saver_pubkey_read(const uint8_t* blob, size_t len, uint64_t* msg_size, vsp_saver_pubkey_t* pub) {
    if (!blob || (!msg_size && !pub)) return VSP_E_INVALID;
    Reader r {blob, len};
    auto w = [](const uint64_t* q) { return const_cast<uint64_t*>(q); };
    if (pub && (!pub->delta_g1 || !pub->delta_sum_s_g1 || !pub->gamma_inverse_sum_s_g1 || !pub->delta_s_g1 || !pub->t_g1 ||
                !pub->t_g2))
        return VSP_E_INVALID;
    point_one(r, G1_C, pub ? w(pub->delta_g1) : nullptr);
    point_one(r, G1_C, pub ? w(pub->delta_sum_s_g1) : nullptr);
    point_one(r, G1_C, pub ? w(pub->gamma_inverse_sum_s_g1) : nullptr);
    uint64_t n0 = 0, n1 = 0, n2 = 0;
    const uint64_t n = pub ? pub->msg_size : 0;
    if (!r.ok || !point_list(r, G1_C, n, pub ? w(pub->delta_s_g1) : nullptr, &n0)) return VSP_E_INVALID;
    if (!point_list(r, G1_C, n, pub ? w(pub->t_g1) : nullptr, &n1)) return VSP_E_INVALID;
    if (!point_list(r, G2_C, n + 1, pub ? w(pub->t_g2) : nullptr, &n2)) return VSP_E_INVALID;
    if (n1 != n0 || n2 != n0 + 1) return VSP_E_INVALID;
    if (msg_size) *msg_size = n0;
    return VSP_OK;
}

int vsp_codec_saver_vk_write(uint64_t msg_size, const uint64_t* rho_g2, const uint64_t* rho_sv_g2,
                             const uint64_t* rho_rhov_g2, uint8_t* out, size_t capacity, size_t* written) {
    if (!written || !rho_g2 || (msg_size && (!rho_sv_g2 || !rho_rhov_g2))) return VSP_E_INVALID;
    const size_t need = G2_C + 2 * 8 + 2 * msg_size * G2_C;
    *written = need;
    if (!out) return VSP_OK;
    if (capacity < need) return VSP_E_INVALID;
    Writer w {out};
    w.g2(rho_g2);
    w.g2_list(rho_sv_g2, msg_size);
    w.g2_list(rho_rhov_g2, msg_size);
    return VSP_OK;
}

// rho_g2 == NULL: *msg_size receives the length of the two lists.  Otherwise the three arrays hold 1 | capacity |
// capacity G2 points and capacity must equal that length.
int vsp_codec_saver_vk_read(const uint8_t* blob, size_t len, uint64_t* msg_size, uint64_t* rho_g2, uint64_t* rho_sv_g2,
                            uint64_t* rho_rhov_g2, uint64_t capacity) {
    if (!blob || (!msg_size && !rho_g2)) return VSP_E_INVALID;
    if (rho_g2 && (!rho_sv_g2 || !rho_rhov_g2)) return VSP_E_INVALID;
    Reader r {blob, len};
    uint64_t n0 = 0, n1 = 0;
    if (!point_one(r, G2_C, rho_g2)) return VSP_E_INVALID;
    if (!point_list(r, G2_C, capacity, rho_g2 ? rho_sv_g2 : nullptr, &n0)) return VSP_E_INVALID;
    if (!point_list(r, G2_C, capacity, rho_g2 ? rho_rhov_g2 : nullptr, &n1)) return VSP_E_INVALID;
    if (n0 != n1) return VSP_E_INVALID;
    if (msg_size) *msg_size = n0;
    return VSP_OK;
}

// rho <-> 32 big-endian bytes; the reader rejects a value that is not reduced mod r
int vsp_codec_saver_sk_write(const uint64_t rho[4], uint8_t* out32) {
    if (!rho || !out32) return VSP_E_INVALID;
    put_be(out32, rho, FR_BYTES);
    return VSP_OK;
}
int vsp_codec_saver_sk_read(const uint8_t* in, size_t len, uint64_t rho[4]) {
    static const uint64_t R[4] = {0xffffffff00000001ull, 0x53bda402fffe5bfeull, 0x3339d80809a1d805ull, 0x73eda753299d7d48ull};
    if (!in || !rho || len < FR_BYTES) return VSP_E_INVALID;
    get_be(in, rho, FR_BYTES);
    for (int i = 3; i >= 0; --i) {
        if (rho[i] < R[i]) return VSP_OK;
        if (rho[i] > R[i]) return VSP_E_INVALID;
    }
    return VSP_E_INVALID;  // rho == r
}

// gamma_ABC_g1 = first | rest, gamma_ABC_count = 1 + the rest's domain size; the rest is written dense (indices
// 0 .. count - 2), the way sparse_vector(std::vector&&) builds it (libs/containers/.../sparse_vector.hpp:81-84)
int vsp_codec_groth16_vk_write(const uint64_t* alpha_g1_beta_g2, const uint64_t* gamma_g2, const uint64_t* delta_g2,
                               const uint64_t* delta_g1, const uint64_t* gamma_g1, uint64_t gamma_ABC_count,
                               const uint64_t* gamma_ABC_g1, uint8_t* out, size_t capacity, size_t* written) {
    if (!written || !alpha_g1_beta_g2 || !gamma_g2 || !delta_g2 || !delta_g1 || !gamma_g1 || !gamma_ABC_g1 ||
        gamma_ABC_count == 0)
        return VSP_E_INVALID;
    const uint64_t n = gamma_ABC_count - 1;
    const size_t need = 12 * FQ_BYTES + 2 * G2_C + 3 * G1_C + 8 + 8 * n + 8 + n * G1_C + 8;
    *written = need;
    if (!out) return VSP_OK;
    if (capacity < need) return VSP_E_INVALID;
    Writer w {out};
    for (int i = 0; i < 12; ++i) {
        put_be(w.p, alpha_g1_beta_g2 + 6 * i, FQ_BYTES);
        w.p += FQ_BYTES;
    }
    w.g2(gamma_g2);
    w.g2(delta_g2);
    w.g1(delta_g1);
    w.g1(gamma_g1);
    w.g1(gamma_ABC_g1);
    w.u64(n);
    for (uint64_t i = 0; i < n; ++i) w.u64(i);
    w.g1_list(gamma_ABC_g1 + 12, n);
    w.u64(n);
    return (size_t)(w.p - out) == need ? VSP_OK : VSP_E_INVALID;
}

// alpha_g1_beta_g2 == NULL: *gamma_ABC_count receives 1 + the rest's domain size.  Otherwise every array is filled
// (alpha_g1_beta_g2: 72 limbs in vsp_pairing_products' order; gamma_ABC_g1: capacity points, capacity == that count;
// positions the sparse rest does not list are the point at infinity).  VSP_E_INVALID for a truncated blob, a malformed
// point, an Fq coefficient >= p, indices that do not increase or reach the domain size.
int vsp_codec_groth16_vk_read(const uint8_t* blob, size_t len, uint64_t* gamma_ABC_count, uint64_t* alpha_g1_beta_g2,
                              uint64_t* gamma_g2, uint64_t* delta_g2, uint64_t* delta_g1, uint64_t* gamma_g1,
                              uint64_t* gamma_ABC_g1, uint64_t capacity) {
    if (!blob || (!gamma_ABC_count && !alpha_g1_beta_g2)) return VSP_E_INVALID;
    const bool fill = alpha_g1_beta_g2 != nullptr;
    if (fill && (!gamma_g2 || !delta_g2 || !delta_g1 || !gamma_g1 || !gamma_ABC_g1)) return VSP_E_INVALID;
    Reader r {blob, len};
    for (int i = 0; i < 12; ++i) {
        const uint8_t* q = r.take(FQ_BYTES);
        if (!q) return VSP_E_INVALID;
        uint64_t c[6];
        get_be(q, c, FQ_BYTES);
        if (hfq_geq_p(c)) return VSP_E_INVALID;
        if (fill) memcpy(alpha_g1_beta_g2 + 6 * i, c, 48);
    }
    if (!point_one(r, G2_C, fill ? gamma_g2 : nullptr) || !point_one(r, G2_C, fill ? delta_g2 : nullptr) ||
        !point_one(r, G1_C, fill ? delta_g1 : nullptr) || !point_one(r, G1_C, fill ? gamma_g1 : nullptr) ||
        !point_one(r, G1_C, fill ? gamma_ABC_g1 : nullptr))
        return VSP_E_INVALID;
    const uint64_t n_idx = r.u64();
    if (!r.ok || n_idx > r.left / 8) return VSP_E_INVALID;
    const uint8_t* idx = r.take(8 * n_idx);
    const uint64_t n_val = r.u64();
    if (!idx || !r.ok || n_val != n_idx || n_val > r.left / G1_C) return VSP_E_INVALID;
    const uint8_t* val = r.take(G1_C * n_val);
    const uint64_t domain = r.u64();
    // gamma_ABC_g1 is handed out dense: a domain past 2^24 positions (1.6 GB of points) is refused rather than allocated
    if (!val || !r.ok || domain > (uint64_t(1) << 24) || n_idx > domain) return VSP_E_INVALID;
    if (gamma_ABC_count) *gamma_ABC_count = domain + 1;
    uint64_t prev = 0;
    for (uint64_t i = 0; i < n_idx; ++i) {
        const uint64_t k = get_u64(idx + 8 * i);
        if (k >= domain || (i && k <= prev)) return VSP_E_INVALID;
        prev = k;
    }
    if (!fill) return VSP_OK;
    if (capacity != domain + 1) return VSP_E_INVALID;
    memset(gamma_ABC_g1 + 12, 0, (size_t)domain * 96);
    for (uint64_t i = 0; i < n_idx; ++i)
        if (g1_decompress(val + G1_C * i, gamma_ABC_g1 + 12 * (1 + get_u64(idx + 8 * i))) != VSP_OK) return VSP_E_INVALID;
    return VSP_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// The vote phase's remaining inputs (bin/cli/include/nil/vote_saver/common.hpp:996-1003): bit arrays (the voter's secret
// key, the voters' public keys), the voters' Merkle tree and the copath taken from it, and the packing of bit strings
// into primary-input field elements.
//   bit array        n bits -> ceil(n / 8) octets, bit i in octet i / 8, most significant bit first
//                    (serialize_bitarray / deserialize_bitarray, common.hpp:574-612: stream_endian::big_octet_big_bit)
//   Merkle tree      2^(depth + 1) - 1 digests of 255 bits, each a 32-octet bit array; leaves first, then row by row, the
//                    root last (deserialize_merkle_tree, common.hpp:614-636; libs/containers/.../merkle/tree.hpp:424-427,
//                    492-499).  Here a digest is the 4-limb integer whose bit i is digest bit i.
//   copath           arity 2: the sibling of the node on the way up, leaf level first
//                    (libs/containers/.../merkle/proof.hpp:65-91)
//   packing          get_multi_field_element_from_bits (common.hpp:557-570): chunks of 254 bits, bit j of a chunk has
//                    weight 2^j (get_field_element_from_bits, common.hpp:543-555)
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr uint64_t DIGEST_BITS = 255, DIGEST_OCTETS = 32, PACK_CHUNK = 254, MAX_TREE_DEPTH = 40;
}

extern "C" {

int vsp_codec_bitarray_write(const uint8_t* bits, uint64_t n_bits, uint8_t* out) {
    if ((!bits && n_bits) || !out) return VSP_E_INVALID;
    memset(out, 0, (size_t)((n_bits + 7) / 8));
    for (uint64_t i = 0; i < n_bits; ++i)
        if (bits[i]) out[i / 8] |= (uint8_t)(0x80u >> (i % 8));
    return VSP_OK;
}

// len must be exactly ceil(n_bits / 8) (the reference asserts it); padding bits are ignored, as the reference's copy_n does
int vsp_codec_bitarray_read(const uint8_t* in, size_t len, uint64_t n_bits, uint8_t* bits) {
    if (!in || (!bits && n_bits) || len != (size_t)((n_bits + 7) / 8)) return VSP_E_INVALID;
    for (uint64_t i = 0; i < n_bits; ++i) bits[i] = (in[i / 8] >> (7 - i % 8)) & 1;
    return VSP_OK;
}

// out: ceil(n_bits / 254) canonical Fr
int vsp_codec_bits_to_field_elements(const uint8_t* bits, uint64_t n_bits, uint64_t* out) {
    if ((!bits && n_bits) || (!out && n_bits)) return VSP_E_INVALID;
    for (uint64_t c = 0; c * PACK_CHUNK < n_bits; ++c) {
        uint64_t* v = out + 4 * c;
        v[0] = v[1] = v[2] = v[3] = 0;
        const uint64_t lo = c * PACK_CHUNK, hi = lo + PACK_CHUNK < n_bits ? lo + PACK_CHUNK : n_bits;
        for (uint64_t i = lo; i < hi; ++i)
            if (bits[i]) v[(i - lo) / 64] |= uint64_t(1) << ((i - lo) % 64);
    }
    return VSP_OK;
}

int vsp_codec_merkle_tree_write(const uint64_t* digests, uint64_t depth, uint8_t* out) {
    if (!digests || !out || depth > MAX_TREE_DEPTH) return VSP_E_INVALID;
    const uint64_t nodes = (uint64_t(2) << depth) - 1;
    for (uint64_t k = 0; k < nodes; ++k) {
        const uint64_t* d = digests + 4 * k;
        if (d[3] >> 63) return VSP_E_INVALID;  // a digest has 255 bits
        uint8_t* o = out + DIGEST_OCTETS * k;
        memset(o, 0, DIGEST_OCTETS);
        for (uint64_t i = 0; i < DIGEST_BITS; ++i)
            if ((d[i / 64] >> (i % 64)) & 1) o[i / 8] |= (uint8_t)(0x80u >> (i % 8));
    }
    return VSP_OK;
}

// digests: 2^(depth + 1) - 1 entries of 4 limbs.  VSP_E_INVALID unless len is exactly that many 32-octet digests.
int vsp_codec_merkle_tree_read(const uint8_t* blob, size_t len, uint64_t depth, uint64_t* digests) {
    if (!blob || !digests || depth > MAX_TREE_DEPTH) return VSP_E_INVALID;
    const uint64_t nodes = (uint64_t(2) << depth) - 1;
    if (len / DIGEST_OCTETS != nodes || len % DIGEST_OCTETS) return VSP_E_INVALID;
    for (uint64_t k = 0; k < nodes; ++k) {
        uint64_t* d = digests + 4 * k;
        d[0] = d[1] = d[2] = d[3] = 0;
        const uint8_t* o = blob + DIGEST_OCTETS * k;
        for (uint64_t i = 0; i < DIGEST_BITS; ++i)
            if ((o[i / 8] >> (7 - i % 8)) & 1) d[i / 64] |= uint64_t(1) << (i % 64);
    }
    return VSP_OK;
}

// path: depth digests, leaf level first — one row of vsp_voting_witness_batch_dev's `paths`
int vsp_codec_merkle_copath(const uint64_t* digests, uint64_t depth, uint64_t index, uint64_t* path) {
    if (!digests || (!path && depth) || depth > MAX_TREE_DEPTH || index >= (uint64_t(1) << depth)) return VSP_E_INVALID;
    uint64_t cur = index, row_len = uint64_t(1) << depth, row_begin = 0;
    for (uint64_t level = 0; level < depth; ++level) {
        const uint64_t sibling = (cur - row_begin) % 2 ? cur - 1 : cur + 1;
        memcpy(path + 4 * level, digests + 4 * sibling, 32);
        cur = row_len + row_begin + (cur - row_begin) / 2;
        row_begin += row_len;
        row_len /= 2;
    }
    return VSP_OK;
}

}  // extern "C"
