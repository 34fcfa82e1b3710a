/* ORACLE (test infrastructure) — short-Weierstrass a = 0 Jacobian group law, instantiated twice
 * (G1 over Fq, G2 over Fq2) by vsp_oracle.c.  Not part of the product path.
 *
 * Restates what crypto3-algebra's `curves::detail::bls12_381 g1/g2, coordinates::jacobian_with_a4_0`
 * computes (type named at /root/reference/bin/cli/include/nil/vote_saver/common.hpp:119; the source is in the
 * un-vendored submodule crypto3-algebra @ e4fa2c5e, /root/reference/.SUBMODULES.json:17).  Formulas are the
 * standard EFD dbl-2009-l / add-2007-bl / madd-2007-bl; any correct law yields the same canonical affine
 * result, which is what parity is compared on.
 *
 * Required macros: CV(name), F_T, F_ADD, F_SUB, F_MUL, F_SQR, F_NEG, F_INV, F_ISZERO, F_EQ, F_ONE(dst), F_ZERO(dst)
 */

typedef struct { F_T X, Y, Z; } CV(jac);
typedef struct { F_T x, y; } CV(aff); /* infinity = (0, 0): never on y^2 = x^3 + b with b != 0 */

static inline int CV(is_inf)(const CV(jac) *p) { return F_ISZERO(&p->Z); }
static inline int CV(aff_is_inf)(const CV(aff) *p) { return F_ISZERO(&p->x) && F_ISZERO(&p->y); }

static inline void CV(set_inf)(CV(jac) *p) {
    F_ONE(&p->X);
    F_ONE(&p->Y);
    F_ZERO(&p->Z);
}

static inline void CV(from_aff)(CV(jac) *r, const CV(aff) *a) {
    if (CV(aff_is_inf)(a)) {
        CV(set_inf)(r);
        return;
    }
    r->X = a->x;
    r->Y = a->y;
    F_ONE(&r->Z);
}

static void CV(to_aff)(CV(aff) *r, const CV(jac) *p) {
    if (CV(is_inf)(p)) {
        F_ZERO(&r->x);
        F_ZERO(&r->y);
        return;
    }
    F_T zi, zi2, zi3;
    F_INV(&zi, &p->Z);
    F_SQR(&zi2, &zi);
    F_MUL(&zi3, &zi2, &zi);
    F_MUL(&r->x, &p->X, &zi2);
    F_MUL(&r->y, &p->Y, &zi3);
}

static void CV(dbl)(CV(jac) *r, const CV(jac) *p) {
    if (CV(is_inf)(p)) {
        *r = *p;
        return;
    }
    F_T A, B, C, D, E, Fv, t, X3, Y3, Z3;
    F_SQR(&A, &p->X);
    F_SQR(&B, &p->Y);
    F_SQR(&C, &B);
    F_ADD(&t, &p->X, &B);
    F_SQR(&t, &t);
    F_SUB(&t, &t, &A);
    F_SUB(&t, &t, &C);
    F_ADD(&D, &t, &t);
    F_ADD(&E, &A, &A);
    F_ADD(&E, &E, &A);
    F_SQR(&Fv, &E);
    F_ADD(&t, &D, &D);
    F_SUB(&X3, &Fv, &t);
    F_MUL(&Z3, &p->Y, &p->Z);
    F_ADD(&Z3, &Z3, &Z3);
    F_ADD(&C, &C, &C);
    F_ADD(&C, &C, &C);
    F_ADD(&C, &C, &C);
    F_SUB(&t, &D, &X3);
    F_MUL(&Y3, &E, &t);
    F_SUB(&Y3, &Y3, &C);
    r->X = X3;
    r->Y = Y3;
    r->Z = Z3;
}

static void CV(add)(CV(jac) *r, const CV(jac) *p, const CV(jac) *q) {
    if (CV(is_inf)(p)) {
        *r = *q;
        return;
    }
    if (CV(is_inf)(q)) {
        *r = *p;
        return;
    }
    F_T Z1Z1, Z2Z2, U1, U2, S1, S2, H, Rr, HH, HHH, V, t, X3, Y3, Z3;
    F_SQR(&Z1Z1, &p->Z);
    F_SQR(&Z2Z2, &q->Z);
    F_MUL(&U1, &p->X, &Z2Z2);
    F_MUL(&U2, &q->X, &Z1Z1);
    F_MUL(&t, &q->Z, &Z2Z2);
    F_MUL(&S1, &p->Y, &t);
    F_MUL(&t, &p->Z, &Z1Z1);
    F_MUL(&S2, &q->Y, &t);
    if (F_EQ(&U1, &U2)) {
        if (F_EQ(&S1, &S2)) {
            CV(dbl)(r, p);
        } else {
            CV(set_inf)(r);
        }
        return;
    }
    F_SUB(&H, &U2, &U1);
    F_SUB(&Rr, &S2, &S1);
    F_SQR(&HH, &H);
    F_MUL(&HHH, &H, &HH);
    F_MUL(&V, &U1, &HH);
    F_SQR(&X3, &Rr);
    F_SUB(&X3, &X3, &HHH);
    F_SUB(&X3, &X3, &V);
    F_SUB(&X3, &X3, &V);
    F_SUB(&t, &V, &X3);
    F_MUL(&Y3, &Rr, &t);
    F_MUL(&t, &S1, &HHH);
    F_SUB(&Y3, &Y3, &t);
    F_MUL(&Z3, &p->Z, &q->Z);
    F_MUL(&Z3, &Z3, &H);
    r->X = X3;
    r->Y = Y3;
    r->Z = Z3;
}

/* mixed addition: q affine (Z2 = 1) */
static void CV(madd)(CV(jac) *r, const CV(jac) *p, const CV(aff) *q) {
    if (CV(aff_is_inf)(q)) {
        *r = *p;
        return;
    }
    if (CV(is_inf)(p)) {
        CV(from_aff)(r, q);
        return;
    }
    F_T Z1Z1, U2, S2, H, Rr, HH, HHH, V, t, X3, Y3, Z3;
    F_SQR(&Z1Z1, &p->Z);
    F_MUL(&U2, &q->x, &Z1Z1);
    F_MUL(&t, &p->Z, &Z1Z1);
    F_MUL(&S2, &q->y, &t);
    if (F_EQ(&p->X, &U2)) {
        if (F_EQ(&p->Y, &S2)) {
            CV(dbl)(r, p);
        } else {
            CV(set_inf)(r);
        }
        return;
    }
    F_SUB(&H, &U2, &p->X);
    F_SUB(&Rr, &S2, &p->Y);
    F_SQR(&HH, &H);
    F_MUL(&HHH, &H, &HH);
    F_MUL(&V, &p->X, &HH);
    F_SQR(&X3, &Rr);
    F_SUB(&X3, &X3, &HHH);
    F_SUB(&X3, &X3, &V);
    F_SUB(&X3, &X3, &V);
    F_SUB(&t, &V, &X3);
    F_MUL(&Y3, &Rr, &t);
    F_MUL(&t, &p->Y, &HHH);
    F_SUB(&Y3, &Y3, &t);
    F_MUL(&Z3, &p->Z, &H);
    r->X = X3;
    r->Y = Y3;
    r->Z = Z3;
}

static inline void CV(neg)(CV(jac) *r, const CV(jac) *p) {
    r->X = p->X;
    F_NEG(&r->Y, &p->Y);
    r->Z = p->Z;
}

/* k: canonical 4x64 little-endian scalar (< r).  Plain MSB-first double-and-add. */
static void CV(mul)(CV(jac) *r, const u64 k[4], const CV(jac) *p) {
    CV(jac) acc;
    CV(set_inf)(&acc);
    int started = 0;
    for (int i = 255; i >= 0; --i) {
        if (started) CV(dbl)(&acc, &acc);
        if ((k[i >> 6] >> (i & 63)) & 1) {
            CV(add)(&acc, &acc, p);
            started = 1;
        }
    }
    *r = acc;
}

/* libff/crypto3 multiexp_inner<BDLO12>  [UPSTREAM-RECALL]: window c = log2(n) - (log2(n)/3 - 2), groups from the
 * top, c doublings between groups, running-sum bucket reduction.  Bases affine (as after fast deserialisation,
 * /root/reference/libs/marshalling/algebra/include/nil/crypto3/marshalling/algebra/types/fast_curve_element.hpp:94-96). */
static void CV(bdlo12_inner)(CV(jac) *out, const CV(aff) *bases, const u64 *scalars, size_t n) {
    CV(set_inf)(out);
    if (n == 0) return;
    size_t log2_length = 0;
    while ((n >> (log2_length + 1)) != 0) ++log2_length;
    long cc = (long)log2_length - ((long)(log2_length / 3) - 2);
    size_t c = cc < 1 ? 1 : (size_t)cc;
    size_t num_bits = 0;
    for (size_t i = 0; i < n; ++i) {
        const u64 *s = scalars + 4 * i;
        for (int l = 3; l >= 0; --l) {
            if (s[l]) {
                size_t b = 64 * l + (64 - __builtin_clzll(s[l]));
                if (b > num_bits) num_bits = b;
                break;
            }
        }
    }
    size_t num_groups = (num_bits + c - 1) / c;
    size_t nb = (size_t)1 << c;
    CV(jac) *buckets = (CV(jac) *)malloc(nb * sizeof(CV(jac)));
    unsigned char *nz = (unsigned char *)malloc(nb);
    int result_nonzero = 0;
    for (size_t k = num_groups; k-- > 0;) {
        if (result_nonzero)
            for (size_t i = 0; i < c; ++i) CV(dbl)(out, out);
        memset(nz, 0, nb);
        for (size_t i = 0; i < n; ++i) {
            const u64 *s = scalars + 4 * i;
            size_t bit = k * c, id = 0;
            for (size_t j = 0; j < c && bit + j < 256; ++j) id |= (size_t)((s[(bit + j) >> 6] >> ((bit + j) & 63)) & 1) << j;
            if (id == 0) continue;
            if (nz[id]) {
                CV(madd)(&buckets[id], &buckets[id], &bases[i]);
            } else {
                CV(from_aff)(&buckets[id], &bases[i]);
                nz[id] = 1;
            }
        }
        CV(jac) running;
        int running_nonzero = 0;
        for (size_t i = nb - 1; i > 0; --i) {
            if (nz[i]) {
                if (running_nonzero)
                    CV(add)(&running, &running, &buckets[i]);
                else
                    running = buckets[i];
                running_nonzero = 1;
            }
            if (running_nonzero) {
                if (result_nonzero)
                    CV(add)(out, out, &running);
                else
                    *out = running;
                result_nonzero = 1;
            }
        }
    }
    free(buckets);
    free(nz);
}

/* multiexp<BDLO12>(bases, scalars, chunks): chunked split as libff multi_exp does; one OpenMP task per chunk
 * (the #ifdef MULTICORE hook at /root/reference/libs/containers/include/nil/crypto3/container/sparse_vector.hpp:224-229). */
static void CV(multiexp)(CV(jac) *out, const CV(aff) *bases, const u64 *scalars, size_t n, size_t chunks) {
    if (chunks <= 1 || n < chunks) {
        CV(bdlo12_inner)(out, bases, scalars, n);
        return;
    }
    size_t one = n / chunks;
    CV(jac) *part = (CV(jac) *)malloc(chunks * sizeof(CV(jac)));
#pragma omp parallel for schedule(static, 1)
    for (long i = 0; i < (long)chunks; ++i) {
        size_t lo = (size_t)i * one, hi = ((size_t)i == chunks - 1) ? n : lo + one;
        CV(bdlo12_inner)(&part[i], bases + lo, scalars + 4 * lo, hi - lo);
    }
    CV(set_inf)(out);
    for (size_t i = 0; i < chunks; ++i) CV(add)(out, out, &part[i]);
    free(part);
}

/* multiexp_with_mixed_addition: skip 0, mixed-add 1, BDLO12 on the rest  [UPSTREAM-RECALL] */
static void CV(multiexp_mixed)(CV(jac) *out, const CV(aff) *bases, const u64 *scalars, size_t n, size_t chunks) {
    CV(jac) acc;
    CV(set_inf)(&acc);
    CV(aff) *g = (CV(aff) *)malloc((n ? n : 1) * sizeof(CV(aff)));
    u64 *p = (u64 *)malloc((n ? n : 1) * 4 * sizeof(u64));
    size_t cnt = 0;
    for (size_t i = 0; i < n; ++i) {
        const u64 *s = scalars + 4 * i;
        if ((s[0] | s[1] | s[2] | s[3]) == 0) continue;
        if (s[0] == 1 && (s[1] | s[2] | s[3]) == 0) {
            CV(madd)(&acc, &acc, &bases[i]);
            continue;
        }
        g[cnt] = bases[i];
        memcpy(p + 4 * cnt, s, 32);
        ++cnt;
    }
    CV(jac) rest;
    CV(multiexp)(&rest, g, p, cnt, chunks);
    CV(add)(out, &acc, &rest);
    free(g);
    free(p);
}
