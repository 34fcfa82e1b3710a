// host_fr.h — a few dozen host-side Fr operations for *setup scalars only*: the handful of constants the
// reference also computes once per domain on the CPU (omega.pow(k), Z.inversed(), value_type(n).inversed();
// basic_radix2_domain.hpp:118-124, step_radix2_domain.hpp:222-246).  O(log n) work per call; no vector data ever
// goes through this file.
#pragma once
#include <cstdint>
#include <cstring>

namespace vsp {
namespace hostfr {

typedef unsigned __int128 u128;

struct Fr4 {
    uint64_t v[4];
};

static const uint64_t MOD[4] = {0xffffffff00000001ULL, 0x53bda402fffe5bfeULL, 0x3339d80809a1d805ULL,
                                0x73eda753299d7d48ULL};

inline bool geq(const uint64_t* a, const uint64_t* b) {
    for (int i = 3; i >= 0; --i)
        if (a[i] != b[i]) return a[i] > b[i];
    return true;
}
inline void sub_n(uint64_t* r, const uint64_t* a, const uint64_t* b) {
    uint64_t br = 0;
    for (int i = 0; i < 4; ++i) {
        u128 d = (u128)a[i] - b[i] - br;
        r[i] = (uint64_t)d;
        br = (uint64_t)(d >> 64) & 1;
    }
}
inline Fr4 add(const Fr4& a, const Fr4& b) {
    Fr4 r;
    u128 c = 0;
    for (int i = 0; i < 4; ++i) {
        c += (u128)a.v[i] + b.v[i];
        r.v[i] = (uint64_t)c;
        c >>= 64;
    }
    if (geq(r.v, MOD)) sub_n(r.v, r.v, MOD);
    return r;
}
inline Fr4 sub(const Fr4& a, const Fr4& b) {
    Fr4 r;
    if (geq(a.v, b.v)) {
        sub_n(r.v, a.v, b.v);
    } else {
        uint64_t t[4];
        sub_n(t, b.v, a.v);
        sub_n(r.v, MOD, t);
    }
    return r;
}
// plain (non-Montgomery) modular product by shift-and-add on canonical values
inline Fr4 mul(const Fr4& a, const Fr4& b) {
    Fr4 acc = {{0, 0, 0, 0}};
    for (int i = 255; i >= 0; --i) {
        acc = add(acc, acc);
        if ((b.v[i >> 6] >> (i & 63)) & 1) acc = add(acc, a);
    }
    return acc;
}
inline Fr4 from_u64(uint64_t x) {
    Fr4 r = {{x, 0, 0, 0}};
    return r;
}
inline Fr4 pow_u64(Fr4 base, uint64_t e) {
    Fr4 acc = from_u64(1);
    while (e) {
        if (e & 1) acc = mul(acc, base);
        base = mul(base, base);
        e >>= 1;
    }
    return acc;
}
inline Fr4 inv(const Fr4& a) {
    uint64_t e[4], two[4] = {2, 0, 0, 0};
    sub_n(e, MOD, two);
    Fr4 acc = from_u64(1), base = a;
    for (int i = 0; i < 256; ++i) {
        if ((e[i >> 6] >> (i & 63)) & 1) acc = mul(acc, base);
        base = mul(base, base);
    }
    return acc;
}
// 2^256 mod r, to hand Montgomery-form constants to kernels
inline Fr4 to_mont(const Fr4& a) {
    Fr4 r = a;
    for (int i = 0; i < 256; ++i) r = add(r, r);
    return r;
}
// unity_root(2^log_n) = root_of_unity^(2^(32 - log_n)), root_of_unity = 7^((r-1)/2^32)
inline Fr4 unity_root_pow2(unsigned log_n) {
    static const Fr4 ROOT = {{0x3829971f439f0d2bULL, 0xb63683508c2280b9ULL, 0xd09b681922c813b4ULL, 0x16a2a19edfe81f20ULL}};
    Fr4 w = ROOT;
    for (unsigned i = 32; i > log_n; --i) w = mul(w, w);
    return w;
}
inline bool is_zero(const Fr4& a) { return (a.v[0] | a.v[1] | a.v[2] | a.v[3]) == 0; }

}  // namespace hostfr
}  // namespace vsp
