/*
 * oracle/ckks_fast.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, fast path).
 *
 * Same primitives as oracle/ckks_prims.c with identical (canonical) outputs, written the way a tuned CPU
 * library does them -- Shoup twiddles with Harvey lazy butterflies for the NTT (what OpenFHE's NativeInteger
 * ModMulFastConst precomputation amounts to), Barrett products, Montgomery-form basis-conversion tables --
 * and OpenMP across limbs / coefficient blocks (OpenFHE parallelises across RNS towers the same way).
 * tests/test_oracle.py checks every function here against its slow twin in ckks_prims.c.
 * This is what bench.py times as the CPU baseline ("port"); the product (CUDA) path never links it.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>

typedef unsigned __int128 u128;
typedef uint64_t u64;

static inline u64 mulhi(u64 a, u64 b) { return (u64)(((u128)a * b) >> 64); }
static inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
static inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
static inline int bitlen(u64 x) { return x ? 64 - __builtin_clzll(x) : 0; }

typedef struct { u64 q, mu; int bits; } barrett_t;
static inline barrett_t barrett_init(u64 q) {
    barrett_t b; b.q = q; b.bits = bitlen(q);
    b.mu = (u64)((((u128)1) << (2 * b.bits)) / q);
    return b;
}
/* a*b mod q for a,b < q < 2^62 (HAC 14.42, base 2) */
static inline u64 barrett_mul(u64 a, u64 b, const barrett_t *m) {
    u128 z = (u128)a * b;
    u64 q1 = (u64)(z >> (m->bits - 1));
    u64 q3 = (u64)(((u128)q1 * m->mu) >> (m->bits + 1));
    u64 r = (u64)z - q3 * m->q;
    if (r >= m->q) r -= m->q;
    if (r >= m->q) r -= m->q;
    return r;
}
static inline u64 shoup_lazy(u64 x, u64 w, u64 wsh, u64 q) { return x * w - mulhi(x, wsh) * q; }   /* in [0,2q) */

/* wsh[i] = floor(w[i] * 2^64 / q) */
void orf_shoup_table(u64 *wsh, const u64 *w, size_t n, u64 q) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) wsh[i] = (u64)((((u128)w[i]) << 64) / q);
}

static void ntt_fwd_one(u64 *a, int logn, u64 q, const u64 *W, const u64 *Wsh) {
    const uint32_t n = 1u << logn; uint32_t t = n; const u64 q2 = q << 1;
    for (uint32_t m = 1; m < n; m <<= 1) {
        t >>= 1;
        for (uint32_t i = 0; i < m; i++) {
            const u64 w = W[m + i], wsh = Wsh[m + i];
            u64 *x = a + 2 * (size_t)i * t, *y = x + t;
            for (uint32_t j = 0; j < t; j++) {
                u64 X = x[j]; if (X >= q2) X -= q2;
                u64 T = shoup_lazy(y[j], w, wsh, q);
                x[j] = X + T; y[j] = X - T + q2;
            }
        }
    }
    for (uint32_t j = 0; j < n; j++) { u64 v = a[j]; if (v >= q2) v -= q2; if (v >= q) v -= q; a[j] = v; }
}
static void ntt_inv_one(u64 *a, int logn, u64 q, const u64 *W, const u64 *Wsh, u64 ninv) {
    const uint32_t n = 1u << logn; uint32_t t = 1; const u64 q2 = q << 1;
    for (uint32_t m = n >> 1; m >= 1; m >>= 1) {
        for (uint32_t i = 0; i < m; i++) {
            const u64 w = W[m + i], wsh = Wsh[m + i];
            u64 *x = a + 2 * (size_t)i * t, *y = x + t;
            for (uint32_t j = 0; j < t; j++) {
                u64 X = x[j], Y = y[j], s = X + Y;
                if (s >= q2) s -= q2;
                x[j] = s; y[j] = shoup_lazy(X - Y + q2, w, wsh, q);
            }
        }
        t <<= 1;
    }
    const u64 ninv_sh = (u64)((((u128)ninv) << 64) / q);
    for (uint32_t j = 0; j < n; j++) { u64 v = shoup_lazy(a[j], ninv, ninv_sh, q); if (v >= q) v -= q; a[j] = v; }
}

/* rows limbs a[r][n]; row r uses modulus id ids[r]; tables tab[M][n], tab_sh[M][n] (forward: psi powers
 * bit-reversed; inverse: psi^-1 powers); ninv[M] = n^-1 mod q (inverse only) */
void orf_ntt_batch(u64 *a, int rows, int logn, const int32_t *ids, const u64 *moduli, const u64 *tab, const u64 *tab_sh,
                   const u64 *ninv, int inverse) {
    const size_t n = (size_t)1 << logn;
#pragma omp parallel for schedule(dynamic, 1)
    for (int r = 0; r < rows; r++) {
        const int id = ids[r];
        if (inverse) ntt_inv_one(a + r * n, logn, moduli[id], tab + id * n, tab_sh + id * n, ninv[id]);
        else ntt_fwd_one(a + r * n, logn, moduli[id], tab + id * n, tab_sh + id * n);
    }
}

/* ---- pointwise ops over one limb (OpenMP over coefficients) ---- */
void orf_vadd(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = addmod(a[i], b[i], q);
}
void orf_vsub(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = submod(a[i], b[i], q);
}
void orf_vneg(u64 *o, const u64 *a, size_t n, u64 q) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = a[i] ? q - a[i] : 0;
}
void orf_vmul(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) {
    const barrett_t m = barrett_init(q);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = barrett_mul(a[i], b[i], &m);
}
void orf_vmuladd(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) {
    const barrett_t m = barrett_init(q);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = addmod(o[i], barrett_mul(a[i], b[i], &m), q);
}
void orf_vmuls(u64 *o, const u64 *a, u64 s, size_t n, u64 q) {
    s %= q;
    const u64 ssh = (u64)((((u128)s) << 64) / q);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) { u64 v = shoup_lazy(a[i], s, ssh, q); o[i] = v >= q ? v - q : v; }
}
void orf_vadds(u64 *o, const u64 *a, u64 s, size_t n, u64 q) {
    s %= q;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) o[i] = addmod(a[i], s, q);
}

static inline uint32_t bitrev32(uint32_t x, int bits) {
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
    x = __builtin_bswap32(x);
    return bits ? x >> (32 - bits) : 0;
}
void orf_automorph_eval(u64 *o, const u64 *a, int logn, u64 k) {
    const uint32_t n = 1u << logn; const u64 mask = ((u64)n << 1) - 1;
#pragma omp parallel for schedule(static)
    for (long j = 0; j < (long)n; j++) {
        uint32_t idx = (uint32_t)((((2 * (u64)j + 1) * k) & mask) >> 1);
        o[bitrev32((uint32_t)j, logn)] = a[bitrev32(idx, logn)];
    }
}

/* ApproxSwitchCRTBasis with Montgomery-form tables: out_j = REDC(sum_i y_i * (hat_ij * 2^64 mod t_j)) */
void orf_approx_switch_crt_basis(u64 *out, const u64 *in, size_t n, int nin, int nout, const u64 *qin, const u64 *qout,
                                 const u64 *qhatinv, const u64 *qhat_mod_out) {
    u64 *hm = (u64 *)malloc(sizeof(u64) * (size_t)nin * nout), *qinv = (u64 *)malloc(sizeof(u64) * nout);
    u64 hsh[128];
    for (int i = 0; i < nin; i++) hsh[i] = (u64)((((u128)qhatinv[i]) << 64) / qin[i]);
    for (int j = 0; j < nout; j++) {
        u64 inv = 1; for (int k = 0; k < 6; k++) inv *= 2 - qout[j] * inv;
        qinv[j] = inv;
        for (int i = 0; i < nin; i++) hm[(size_t)i * nout + j] = (u64)((((u128)qhat_mod_out[(size_t)i * nout + j]) << 64) % qout[j]);
    }
#pragma omp parallel for schedule(static)
    for (long c = 0; c < (long)n; c++) {
        u64 y[128];
        for (int i = 0; i < nin; i++) { u64 v = shoup_lazy(in[(size_t)i * n + c], qhatinv[i], hsh[i], qin[i]); y[i] = v >= qin[i] ? v - qin[i] : v; }
        for (int j = 0; j < nout; j++) {
            u128 acc = 0;
            for (int i = 0; i < nin; i++) acc += (u128)y[i] * hm[(size_t)i * nout + j];
            u64 lo = (u64)acc, hi = (u64)(acc >> 64), t = mulhi(lo * qinv[j], qout[j]);
            out[(size_t)j * n + c] = hi >= t ? hi - t : hi + qout[j] - t;
        }
    }
    free(hm); free(qinv);
}

void orf_switch_modulus(u64 *o, const u64 *a, size_t n, u64 qold, u64 qnew) {
    const u64 half = qold >> 1;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) {
        u64 v = a[i];
        if (v > half) { u64 neg = (qold - v) % qnew; o[i] = neg ? qnew - neg : 0; } else o[i] = v % qnew;
    }
}
