/*
 * oracle/ckks_prims.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, plain C + __int128).
 *
 * Slow, obviously-correct restatement of the RNS-CKKS array primitives that the
 * reference (oksuman/sorting-fhe) reaches through OpenFHE 1.1.4 (`cc->Eval*`,
 * src/sort_algo.h, src/sign.cpp, src/comparison.cpp, src/rotation.h).  OpenFHE is an
 * un-vendored dependency that is absent from /root/reference, so every function
 * restates OpenFHE's *published algorithm*; bit-parity with OpenFHE itself is
 * "parity unpinned" (no reference test pins a limb).  The product (CUDA) path must
 * never call into this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do.
 *
 * All vectors are uint64 residues in [0,q).  Layout: limb-major [limb][n].
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

typedef unsigned __int128 u128;
typedef uint64_t u64;

static inline u64 mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }
static inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
static inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }

u64 orc_powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q; a %= q;
    while (e) { if (e & 1) r = mulmod(r, a, q); a = mulmod(a, a, q); e >>= 1; }
    return r;
}
u64 orc_invmod(u64 a, u64 q) { return orc_powmod(a, q - 2, q); }

/* deterministic Miller-Rabin for 64-bit */
int orc_is_prime(u64 n) {
    if (n < 2) return 0;
    static const u64 sp[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (int i = 0; i < 12; i++) { if (n % sp[i] == 0) return n == sp[i]; }
    u64 d = n - 1; int s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (int i = 0; i < 12; i++) {
        u64 x = orc_powmod(sp[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) { x = mulmod(x, x, n); if (x == n - 1) { comp = 0; break; } }
        if (comp) return 0;
    }
    return 1;
}

/* OpenFHE FirstPrime/PreviousPrime/NextPrime (nbtheory) restated: primes == 1 mod m */
u64 orc_first_prime(int nbits, u64 m) {
    u64 q = (u64)1 << nbits, r = q % m;
    u64 qn = q + 1 - r;
    if (r > 0) qn += m;
    while (!orc_is_prime(qn)) qn += m;
    return qn;
}
u64 orc_prev_prime(u64 q, u64 m) { u64 p = q - m; while (!orc_is_prime(p)) p -= m; return p; }
u64 orc_next_prime(u64 q, u64 m) { u64 p = q + m; while (!orc_is_prime(p)) p += m; return p; }

/* minimal primitive m-th root of unity mod q (OpenFHE RootOfUnity picks the minimum
 * over all primitive roots so that contexts agree); m is a power of two. */
u64 orc_min_root_of_unity(u64 m, u64 q) {
    u64 g = 0;
    for (u64 c = 2;; c++) { /* any element whose (q-1)/m power has order exactly m */
        u64 r = orc_powmod(c, (q - 1) / m, q);
        if (orc_powmod(r, m / 2, q) == q - 1) { g = r; break; }
    }
    u64 g2 = mulmod(g, g, q), x = g, mn = g;
    for (u64 i = 1; i < m / 2; i++) { x = mulmod(x, g2, q); if (x < mn) mn = x; }
    return mn;
}

static inline uint32_t bitrev(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

/* table[bitrev(i)] = psi^i, i<n  (OpenFHE rootOfUnityReverse tables) */
void orc_make_psi_table(u64 *tab, int logn, u64 psi, u64 q) {
    uint32_t n = 1u << logn; u64 x = 1;
    for (uint32_t i = 0; i < n; i++) { tab[bitrev(i, logn)] = x; x = mulmod(x, psi, q); }
}

/* Forward negacyclic NTT, Cooley-Tukey, natural in -> bit-reversed out
 * (OpenFHE ForwardTransformToBitReverse). */
void orc_ntt_fwd(u64 *a, int logn, u64 q, const u64 *psi_br) {
    uint32_t n = 1u << logn, t = n;
    for (uint32_t m = 1; m < n; m <<= 1) {
        t >>= 1;
        for (uint32_t i = 0; i < m; i++) {
            u64 w = psi_br[m + i];
            uint32_t j1 = 2 * i * t;
            for (uint32_t j = j1; j < j1 + t; j++) {
                u64 u = a[j], v = mulmod(a[j + t], w, q);
                a[j] = addmod(u, v, q); a[j + t] = submod(u, v, q);
            }
        }
    }
}

/* Inverse: Gentleman-Sande, bit-reversed in -> natural out, scaled by n^-1
 * (OpenFHE InverseTransformFromBitReverse). */
void orc_ntt_inv(u64 *a, int logn, u64 q, const u64 *psi_inv_br) {
    uint32_t n = 1u << logn, t = 1;
    for (uint32_t m = n >> 1; m >= 1; m >>= 1) {
        for (uint32_t i = 0; i < m; i++) {
            u64 w = psi_inv_br[m + i];
            uint32_t j1 = 2 * i * t;
            for (uint32_t j = j1; j < j1 + t; j++) {
                u64 u = a[j], v = a[j + t];
                a[j] = addmod(u, v, q); a[j + t] = mulmod(submod(u, v, q), w, q);
            }
        }
        t <<= 1;
    }
    u64 ninv = orc_invmod(n % q, q);
    for (uint32_t j = 0; j < n; j++) a[j] = mulmod(a[j], ninv, q);
}

/* ---- pointwise vector ops over one limb ---- */
void orc_vadd(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = addmod(a[i], b[i], q); }
void orc_vsub(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = submod(a[i], b[i], q); }
void orc_vmul(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = mulmod(a[i], b[i], q); }
void orc_vmuladd(u64 *o, const u64 *a, const u64 *b, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = addmod(o[i], mulmod(a[i], b[i], q), q); }
void orc_vmuls(u64 *o, const u64 *a, u64 s, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = mulmod(a[i], s, q); }
void orc_vadds(u64 *o, const u64 *a, u64 s, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = addmod(a[i], s, q); }
void orc_vneg(u64 *o, const u64 *a, size_t n, u64 q) { for (size_t i = 0; i < n; i++) o[i] = a[i] ? q - a[i] : 0; }

/* Automorphism X -> X^k on an EVALUATION-format (bit-reversed NTT) limb:
 * OpenFHE PrecomputeAutoMap + AutomorphismTransform: out[rev(j)] = in[rev(idx)],
 * idx = (((2j+1)*k) mod 2n) >> 1. */
void orc_automorph_eval(u64 *o, const u64 *a, int logn, u64 k) {
    uint32_t n = 1u << logn; u64 m = (u64)n << 1;
    for (uint32_t j = 0; j < n; j++) {
        u64 jt = 2 * (u64)j + 1;
        uint32_t idx = (uint32_t)(((jt * k) % m) >> 1);
        o[bitrev(j, logn)] = a[bitrev(idx, logn)];
    }
}

/* ApproxSwitchCRTBasis (OpenFHE DCRTPoly): in[nin][n] coefficient-format limbs mod
 * qin[i]; y_i = x_i * qhatinv[i] mod qin[i]; out_j = sum_i y_i * qhat_mod_out[i*nout+j]
 * mod qout[j] (128-bit accumulate, one reduction, no correction term). */
void orc_approx_switch_crt_basis(u64 *out, const u64 *in, size_t n, int nin, int nout,
                                 const u64 *qin, const u64 *qout, const u64 *qhatinv,
                                 const u64 *qhat_mod_out) {
    for (size_t c = 0; c < n; c++) {
        u64 y[128];
        for (int i = 0; i < nin; i++) y[i] = mulmod(in[(size_t)i * n + c], qhatinv[i], qin[i]);
        for (int j = 0; j < nout; j++) {
            u128 acc = 0;
            for (int i = 0; i < nin; i++) acc += (u128)y[i] * qhat_mod_out[(size_t)i * nout + j];
            out[(size_t)j * n + c] = (u64)(acc % qout[j]);
        }
    }
}

/* centred modulus switch of one coefficient-format limb from qold to qnew
 * (OpenFHE NativeVector::SwitchModulus) */
void orc_switch_modulus(u64 *o, const u64 *a, size_t n, u64 qold, u64 qnew) {
    u64 half = qold >> 1;
    for (size_t i = 0; i < n; i++) {
        u64 v = a[i];
        if (v > half) { /* negative: v - qold */
            u64 neg = (qold - v) % qnew;
            o[i] = neg ? qnew - neg : 0;
        } else o[i] = v % qnew;
    }
}

/* ---- counter-based sampling spec shared with the product ---- */
static inline u64 mix64(u64 x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31;
    return x;
}
u64 orc_hash(u64 seed, u64 stream, u64 idx) {
    return mix64(mix64(seed + 0x9E3779B97F4A7C15ULL * (stream + 1)) ^ mix64(idx + 0xD1B54A32D192ED03ULL));
}
/* kind 0: ternary {-1,0,1}; kind 1: centred binomial(20) error (sigma ~3.16); both as
 * residues mod q.  kind 2: uniform mod q. */
void orc_sample(u64 *o, size_t n, u64 q, u64 seed, u64 stream, int kind) {
    for (size_t i = 0; i < n; i++) {
        u64 h = orc_hash(seed, stream, i);
        if (kind == 0) { u64 t = h % 3; o[i] = t == 0 ? 0 : (t == 1 ? 1 : q - 1); }
        else if (kind == 1) {
            int e = __builtin_popcountll(h & 0xFFFFF) - __builtin_popcountll((h >> 20) & 0xFFFFF);
            o[i] = e >= 0 ? (u64)e : q - (u64)(-e);
        } else o[i] = h % q;
    }
}

/* ---- CKKS special FFT (HEAAN/OpenFHE DiscreteFourierTransform::FFTSpecial{,Inv}) ----
 * vals: interleaved re,im; size = slots; ksi: table of M+1 complex roots exp(2*pi*i*k/M)
 * interleaved; rot[j] = 5^j mod M.  Separate mul/add (compile with -ffp-contract=off) so
 * that a GPU port using __dmul_rn/__dadd_rn is bit-identical. */
static void bitrev_cplx(double *v, uint32_t size) {
    int bits = 0; while ((1u << bits) < size) bits++;
    for (uint32_t i = 0; i < size; i++) {
        uint32_t j = bitrev(i, bits);
        if (i < j) { double r = v[2*i], m = v[2*i+1]; v[2*i] = v[2*j]; v[2*i+1] = v[2*j+1]; v[2*j] = r; v[2*j+1] = m; }
    }
}
void orc_fft_special_inv(double *v, uint32_t size, const double *ksi, const uint32_t *rot, uint32_t M) {
    for (uint32_t len = size; len >= 2; len >>= 1) {
        uint32_t lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (uint32_t i = 0; i < size; i += len)
            for (uint32_t j = 0; j < lenh; j++) {
                uint32_t idx = (lenq - (rot[j] % lenq)) * gap;
                double ar = v[2*(i+j)], ai = v[2*(i+j)+1], br = v[2*(i+j+lenh)], bi = v[2*(i+j+lenh)+1];
                double ur = ar + br, ui = ai + bi, vr = ar - br, vi = ai - bi;
                double wr = ksi[2*idx], wi = ksi[2*idx+1];
                double p1 = vr * wr, p2 = vi * wi, p3 = vr * wi, p4 = vi * wr;
                v[2*(i+j)] = ur; v[2*(i+j)+1] = ui;
                v[2*(i+j+lenh)] = p1 - p2; v[2*(i+j+lenh)+1] = p3 + p4;
            }
    }
    bitrev_cplx(v, size);
    double inv = 1.0 / (double)size;
    for (uint32_t i = 0; i < 2 * size; i++) v[i] = v[i] * inv;
}
void orc_fft_special(double *v, uint32_t size, const double *ksi, const uint32_t *rot, uint32_t M) {
    bitrev_cplx(v, size);
    for (uint32_t len = 2; len <= size; len <<= 1) {
        uint32_t lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (uint32_t i = 0; i < size; i += len)
            for (uint32_t j = 0; j < lenh; j++) {
                uint32_t idx = (rot[j] % lenq) * gap;
                double ur = v[2*(i+j)], ui = v[2*(i+j)+1], br = v[2*(i+j+lenh)], bi = v[2*(i+j+lenh)+1];
                double wr = ksi[2*idx], wi = ksi[2*idx+1];
                double p1 = br * wr, p2 = bi * wi, p3 = br * wi, p4 = bi * wr;
                double vr = p1 - p2, vi = p3 + p4;
                v[2*(i+j)] = ur + vr; v[2*(i+j)+1] = ui + vi;
                v[2*(i+j+lenh)] = ur - vr; v[2*(i+j+lenh)+1] = ui - vi;
            }
    }
}

/* round-half-even double -> residues (|x| < 2^63).  x = value * scale computed by caller. */
void orc_round_to_rns(u64 *o, const double *x, size_t n, u64 q) {
    for (size_t i = 0; i < n; i++) {
        double r = nearbyint(x[i]);
        int64_t v = (int64_t)r;
        if (v >= 0) o[i] = (u64)v % q; else { u64 m = ((u64)(-v)) % q; o[i] = m ? q - m : 0; }
    }
}

/* two-limb centred CRT to double: v mod (q0*q1) centred, as double (decode step) */
void orc_crt2_to_double(double *o, const u64 *a0, const u64 *a1, size_t n, u64 q0, u64 q1) {
    u64 q0inv_q1 = orc_invmod(q0 % q1, q1);
    u128 Q = (u128)q0 * q1, half = Q >> 1;
    for (size_t i = 0; i < n; i++) {
        /* v = a0 + q0 * ((a1 - a0) * q0^-1 mod q1) */
        u64 d = submod(a1[i] % q1, a0[i] % q1, q1);
        u64 k = mulmod(d, q0inv_q1, q1);
        u128 v = (u128)a0[i] + (u128)q0 * k;
        if (v > half) { u128 neg = Q - v; o[i] = -((double)(u64)(neg >> 64) * 18446744073709551616.0 + (double)(u64)neg); }
        else o[i] = (double)(u64)(v >> 64) * 18446744073709551616.0 + (double)(u64)v;
    }
}
void orc_crt1_to_double(double *o, const u64 *a0, size_t n, u64 q0) {
    u64 half = q0 >> 1;
    for (size_t i = 0; i < n; i++) o[i] = a0[i] > half ? -(double)(q0 - a0[i]) : (double)a0[i];
}
