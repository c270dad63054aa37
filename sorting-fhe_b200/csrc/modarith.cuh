// modarith.cuh -- 64-bit modular arithmetic for sm_100a integer pipes.
//
// Moduli are NTT-friendly primes q < 2^61 (OpenFHE 1.1.4 chain: 60-bit first/special primes,
// 40..59-bit scaling primes, 20-bit FLEXIBLEAUTOEXT extra prime; SURVEY.md 8(c')).
// Three reduction flavours, all returning canonical residues so results are bit-identical to
// the oracle's `unsigned __int128 %`:
//   * Shoup   : x * w mod q for a long-lived constant w with w' = floor(w * 2^64 / q)
//   * Barrett : a * b mod q for two data operands (a,b < q)
//   * Montgomery reduce of a 128-bit accumulator against constants stored as c * 2^64 mod q
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

struct ModInfo {   // one entry per modulus (Q primes then P primes), lives in global memory
    u64 q;
    u64 mu;        // floor(2^(2*bits) / q), bits = bit length of q
    u64 qinv;      // q^-1 mod 2^64 (Montgomery)
    u64 r2;        // 2^128 mod q (to enter Montgomery form)
    u64 ninv;      // n^-1 mod q
    u64 ninv_sh;   // Shoup companion of ninv
    u64 qrecip;    // floor(2^64 / q): x - mulhi(x, qrecip) * q lies in [0, 2q) for any x < 2^64
    int bits;
    int pad;
};

__device__ __forceinline__ u64 add_mod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
__device__ __forceinline__ u64 sub_mod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
__device__ __forceinline__ u64 neg_mod(u64 a, u64 q) { return a ? q - a : 0; }

// x * w mod q, result in [0, 2q): valid for any x < 2^64
__device__ __forceinline__ u64 shoup_lazy(u64 x, u64 w, u64 wsh, u64 q) {
    u64 hi = __umul64hi(x, wsh);
    return x * w - hi * q;
}
__device__ __forceinline__ u64 shoup_mul(u64 x, u64 w, u64 wsh, u64 q) {
    u64 r = shoup_lazy(x, w, wsh, q);
    return r >= q ? r - q : r;
}

// a * b mod q, a,b < q (classical Barrett, HAC 14.42 with base 2)
__device__ __forceinline__ u64 barrett_mul(u64 a, u64 b, const ModInfo& m) {
    u64 lo = a * b, hi = __umul64hi(a, b);
    int s1 = m.bits - 1;                       // z >> (bits-1)  (fits 64 bits)
    u64 q1 = (s1 == 0) ? lo : ((lo >> s1) | (hi << (64 - s1)));
    u64 plo = q1 * m.mu, phi = __umul64hi(q1, m.mu);
    int s2 = m.bits + 1;                       // >> (bits+1), bits+1 <= 62
    u64 q3 = (plo >> s2) | (phi << (64 - s2));
    u64 r = lo - q3 * m.q;
    if (r >= m.q) r -= m.q;
    if (r >= m.q) r -= m.q;
    return r;
}

// (hi:lo) * 2^-64 mod q, requires (hi:lo) < q * 2^64; canonical result
__device__ __forceinline__ u64 mont_reduce(u64 lo, u64 hi, u64 q, u64 qinv) {
    u64 m = lo * qinv;
    u64 t = __umul64hi(m, q);
    return hi >= t ? hi - t : hi + q - t;
}

// 128-bit accumulate acc += a * b
__device__ __forceinline__ void mac128(u64& lo, u64& hi, u64 a, u64 b) {
    u64 pl = a * b, ph = __umul64hi(a, b);
    lo += pl;
    hi += ph + (lo < pl);
}
