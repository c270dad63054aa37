// keyswitch.cu -- hybrid key switching kernels K2-K4: ApproxSwitchCRTBasis
// (used by ModUp and ModDown), the evaluation-key inner product, and the ModDown epilogue.
//
// Replaces OpenFHE KeySwitchHYBRID::EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCoreExt /
// DCRTPoly::ApproxModDown, reached from every cc->EvalMult(ct,ct) relinearisation and
// cc->EvalRotate of the reference (src/sign.cpp:23-56, src/rotation.h:224,231,
// src/sort_algo.h:361,496,580,746).  Semantics pinned by oracle/ckks_ops.py key_switch_core.
//
// The evaluation key stream (2 * digits * (l+K) limbs) is the irreducible HBM term; keys are
// kept in Montgomery form so that sum_j d_j * key_j is one 128-bit accumulation + one reduction.
#include "engine.h"
#include <algorithm>
#include <cstdlib>

#define KL(c) (c)->stats.kernel_launches++; if ((c)->dry) return

// Batched ApproxSwitchCRTBasis on PRE-SCALED inputs: the inverse NTT that produced `in` already multiplied limb i by
// n^-1 * Qhat_i^-1 (one fused Shoup factor), so y_i = in_i and
//   out[j][x] = sum_i y_i[x] * (Qhat_i mod t_j)   mod t_j            (no correction term, 128-bit accumulate)
// Up to 4 groups per launch: the digits of a ModUp, or the two polynomials of a ModDown.
__global__ void __launch_bounds__(256)
k_conv_batch(u64* __restrict__ out, const u64* __restrict__ in, const ConvBatch B, const ModInfo* __restrict__ mi, int n, int ysplit) {
    pdl_sync();
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    const int gi = blockIdx.y / ysplit, part = blockIdx.y % ysplit;
    const ConvGroup g = B.g[gi];
    u64 y[MAX_CONV_IN];
#pragma unroll 4
    for (int i = 0; i < g.nin; i++) y[i] = in[(size_t)(g.in_row0 + i) * n + x];
    for (int j = part; j < g.nout; j += ysplit) {
        const ModInfo m = mi[B.out_ids[gi][j]];
        u64 lo = 0, hi = 0;
        for (int i = 0; i < g.nin; i++) mac128(lo, hi, y[i], __ldg(g.hatmod + (size_t)i * g.nout + j));
        out[(size_t)(g.out_row0 + j) * n + x] = mont_reduce(lo, hi, m.q, m.qinv);
    }
}
// The same conversion arranged for the fma pipe, for groups of at most 12 inputs (every key switch of the sorters'
// parameter sets: alpha = 11 limbs per digit, 8 special primes).  The constants are cut into 20-bit digits (two for a
// 40-bit target modulus, three for a 60-bit one) and staged in shared memory; the residues of two coefficients stay in
// registers as 32-bit halves (y0, y1 < 2^28).  Every partial product digit x half is below 2^52, so twelve of them add
// up in a 64-bit column sum with no carry: a term costs four (six) IMAD.WIDE.U32 and nothing else, instead of a full
// 64 x 64 -> 128 multiply-accumulate with its carry chain.  The columns are recombined once per output,
//   total = c00 + c10 2^20 + c01 2^32 + c20 2^40 + c11 2^52 + c21 2^72,
// which is the integer k_conv_batch accumulates; same Montgomery reduction, same result.
__device__ __forceinline__ void add_shifted(u64& lo, u64& hi, u64 s, int sh) {   // (hi:lo) += s << sh, 0 < sh < 64
    const u64 t = s << sh;
    lo += t;
    hi += (s >> (64 - sh)) + (lo < t);
}
__device__ __forceinline__ void madw(u64& acc, u32 a, u32 b) { asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(a), "r"(b)); }
template <int NI>
__global__ void __launch_bounds__(256)
k_conv_split(u64* __restrict__ out, const u64* __restrict__ in, const __grid_constant__ ConvBatch B, const ModInfo* __restrict__ mi, int n,
             int ysplit) {
    pdl_sync();
    extern __shared__ uint4 sh_h[];                       // [jc][NI]: digits (h & 0xFFFFF, (h >> 20) & 0xFFFFF, h >> 40, -)
    const int gi = blockIdx.y / ysplit, part = blockIdx.y - gi * ysplit;
    const int nin = B.g[gi].nin, nout = B.g[gi].nout;
    const int per = (nout + ysplit - 1) / ysplit, jlo = part * per, jc = min(per, nout - jlo);
    if (jc <= 0) return;
    const u64* __restrict__ hat = B.g[gi].hatmod;
    for (int t = threadIdx.x; t < NI * jc; t += blockDim.x) {
        const int jj = t / NI, i = t - jj * NI;
        const u64 h = i < nin ? hat[(size_t)i * nout + jlo + jj] : 0;
        sh_h[t] = make_uint4((u32)h & 0xFFFFFu, (u32)(h >> 20) & 0xFFFFFu, (u32)(h >> 40), 0u);
    }
    __syncthreads();
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (x >= n) return;
    u32 a0[NI], a1[NI], b0[NI], b1[NI];
    const u64* src = in + (size_t)B.g[gi].in_row0 * n + x;
#pragma unroll
    for (int i = 0; i < NI; i++) {
        ulonglong2 v = make_ulonglong2(0, 0);
        if (i < nin) v = *reinterpret_cast<const ulonglong2*>(src + (size_t)i * n);
        a0[i] = (u32)v.x; a1[i] = (u32)(v.x >> 32); b0[i] = (u32)v.y; b1[i] = (u32)(v.y >> 32);
    }
    u64* dst = out + (size_t)(B.g[gi].out_row0 + jlo) * n + x;
    for (int jj = 0; jj < jc; jj++) {
        const uint4* hp = sh_h + jj * NI;
        const ModInfo m = mi[B.out_ids[gi][jlo + jj]];
        u64 a00 = 0, a10 = 0, a01 = 0, a11 = 0, a20 = 0, a21 = 0, b00 = 0, b10 = 0, b01 = 0, b11 = 0, b20 = 0, b21 = 0;
        if (m.bits <= 40) {
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const uint4 h = hp[i];
                madw(a00, h.x, a0[i]); madw(a10, h.y, a0[i]); madw(a01, h.x, a1[i]); madw(a11, h.y, a1[i]);
                madw(b00, h.x, b0[i]); madw(b10, h.y, b0[i]); madw(b01, h.x, b1[i]); madw(b11, h.y, b1[i]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const uint4 h = hp[i];
                madw(a00, h.x, a0[i]); madw(a10, h.y, a0[i]); madw(a20, h.z, a0[i]);
                madw(a01, h.x, a1[i]); madw(a11, h.y, a1[i]); madw(a21, h.z, a1[i]);
                madw(b00, h.x, b0[i]); madw(b10, h.y, b0[i]); madw(b20, h.z, b0[i]);
                madw(b01, h.x, b1[i]); madw(b11, h.y, b1[i]); madw(b21, h.z, b1[i]);
            }
        }
        u64 al = a00, ah = a21 << 8, bl = b00, bh = b21 << 8;
        add_shifted(al, ah, a10, 20); add_shifted(al, ah, a01, 32); add_shifted(al, ah, a20, 40); add_shifted(al, ah, a11, 52);
        add_shifted(bl, bh, b10, 20); add_shifted(bl, bh, b01, 32); add_shifted(bl, bh, b20, 40); add_shifted(bl, bh, b11, 52);
        *reinterpret_cast<ulonglong2*>(dst + (size_t)jj * n) = make_ulonglong2(mont_reduce(al, ah, m.q, m.qinv), mont_reduce(bl, bh, m.q, m.qinv));
    }
}
template <int NI>
static void launch_conv_split(Context* c, dim3 grid, size_t smem, u64* out, const u64* in, const ConvBatch& B, int ysplit) {
    launch_pdl(c, k_conv_split<NI>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
}
// The conversion on the FP64 pipe.  B200 issues DFMA at 64 lanes/clk/SM, twice the rate ncu shows for IMAD.WIDE (fma pipe
// 36 % busy at 4 cycles per warp instruction in k_conv_split), and a double holds a 53-bit integer exactly: residues and
// constants are cut into 21-bit pieces (two for a modulus of at most 42 bits -- the 40-bit scaling primes straddle 2^40 --,
// three up to 60 bits), every product is below 2^42 and a column (all products of equal weight, at most 3 x 12 of them)
// stays below 2^48.  A term costs 4 (narrow
// residue x narrow constant) to 9 DFMA and nothing else; the columns are turned back into integers once per output,
//   total = s0 + s1 2^21 + s2 2^42 + s3 2^63 + s4 2^84,
// the same integer k_conv_batch / k_conv_split accumulate, followed by the same Montgomery reduction.
// WMODE: 0 = every source modulus has at most 42 bits, 1 = only the first one is wider (the digit that holds q0),
// 2 = any (third piece kept for every source limb: ModDown from the 60-bit special primes).
__device__ __forceinline__ double piece_to_double(u32 v) { return __dadd_rn(__hiloint2double(0x43300000, (int)v), -4503599627370496.0); }
__device__ __forceinline__ u64 column_to_u64(double s) {   // integer s in [0, 2^52)
    return (u64)__double_as_longlong(__dadd_rn(s, 4503599627370496.0)) & 0x000FFFFFFFFFFFFFULL;
}
template <int NI, int WMODE>
__global__ void __launch_bounds__(256)
k_conv_fp(u64* __restrict__ out, const u64* __restrict__ in, const __grid_constant__ ConvBatch B, const ModInfo* __restrict__ mi, int n,
          int ysplit) {
    pdl_sync();
    extern __shared__ double2 sh_c[];                     // [jc][NI][2]: (c0, c1), (c2, 0) -- 21-bit digits of the constants
    const int gi = blockIdx.y / ysplit, part = blockIdx.y - gi * ysplit;
    const int nin = B.g[gi].nin, nout = B.g[gi].nout;
    const int per = (nout + ysplit - 1) / ysplit, jlo = part * per, jc = min(per, nout - jlo);
    if (jc <= 0) return;
    const u64* __restrict__ hat = B.g[gi].hatmod;
    for (int t = threadIdx.x; t < NI * jc; t += blockDim.x) {
        const int jj = t / NI, i = t - jj * NI;
        const u64 h = i < nin ? hat[(size_t)i * nout + jlo + jj] : 0;
        sh_c[2 * t] = make_double2((double)((u32)h & 0x1FFFFFu), (double)((u32)(h >> 21) & 0x1FFFFFu));
        sh_c[2 * t + 1] = make_double2((double)(u32)(h >> 42), 0.0);
    }
    __syncthreads();
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    constexpr int NW = WMODE == 2 ? NI : 1;
    double y0[NI], y1[NI], y2[NW];
    const u64* src = in + (size_t)B.g[gi].in_row0 * n + x;
#pragma unroll
    for (int i = 0; i < NI; i++) {
        const u64 v = i < nin ? src[(size_t)i * n] : 0;
        y0[i] = piece_to_double((u32)v & 0x1FFFFFu);
        y1[i] = piece_to_double((u32)(v >> 21) & 0x1FFFFFu);
        if (i < NW) y2[i] = piece_to_double((u32)(v >> 42));
    }
    u64* dst = out + (size_t)(B.g[gi].out_row0 + jlo) * n + x;
    for (int jj = 0; jj < jc; jj++) {
        const double2* hp = sh_c + 2 * jj * NI;
        const ModInfo m = mi[B.out_ids[gi][jlo + jj]];
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
        if (m.bits <= 42) {
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const double2 c = hp[2 * i];
                s0 = __fma_rn(y0[i], c.x, s0); s1 = __fma_rn(y0[i], c.y, s1);
                s1 = __fma_rn(y1[i], c.x, s1); s2 = __fma_rn(y1[i], c.y, s2);
                if (WMODE == 2 || (WMODE == 1 && i == 0)) { s2 = __fma_rn(y2[i], c.x, s2); s3 = __fma_rn(y2[i], c.y, s3); }
            }
        } else {
#pragma unroll
            for (int i = 0; i < NI; i++) {
                const double2 c = hp[2 * i];
                const double c2 = hp[2 * i + 1].x;
                s0 = __fma_rn(y0[i], c.x, s0); s1 = __fma_rn(y0[i], c.y, s1); s2 = __fma_rn(y0[i], c2, s2);
                s1 = __fma_rn(y1[i], c.x, s1); s2 = __fma_rn(y1[i], c.y, s2); s3 = __fma_rn(y1[i], c2, s3);
                if (WMODE == 2 || (WMODE == 1 && i == 0)) {
                    s2 = __fma_rn(y2[i], c.x, s2); s3 = __fma_rn(y2[i], c.y, s3); s4 = __fma_rn(y2[i], c2, s4);
                }
            }
        }
        u64 lo = column_to_u64(s0), hi = column_to_u64(s4) << 20;
        add_shifted(lo, hi, column_to_u64(s1), 21); add_shifted(lo, hi, column_to_u64(s2), 42); add_shifted(lo, hi, column_to_u64(s3), 63);
        dst[(size_t)jj * n] = mont_reduce(lo, hi, m.q, m.qinv);
    }
}
// k_conv_fp2: the same contraction with (1) no global ModInfo load per output -- the per-output scalars are staged in
// shared memory next to the constants -- and (2) for target moduli of at most 42 bits (30 of the 32 Q limbs) an FP64
// epilogue: the constants are the PLAIN residues Qhat_i mod t_j, the exact column sums s_k < 2^48 are folded with
//   out = s0 + s1 * (2^21 mod t) + s2 * (2^42 mod t) [+ s3 * (2^63 mod t)]   mod t
// by exact floating-point modular products (fp_mulmod of ntt.cu: every intermediate an integer below 2^53), instead of
// rebuilding a 128-bit integer and Montgomery-reducing it on the integer pipes (~45 instructions per output, as many as
// the contraction itself in k_conv_fp).  60-bit targets keep the Montgomery-form constants and the integer epilogue.
// Both give the canonical residue of the same integer sum, so the output equals k_conv_batch's bit for bit.
#define CV_MAGIC 6755399441055744.0
__device__ __forceinline__ double cv_mulmod(double a, double w, double q, double qinv) {
    const double h = __dmul_rn(a, w);
    const double l = __fma_rn(a, w, -h);
    const double qh = __dadd_rn(__fma_rn(h, qinv, CV_MAGIC), -CV_MAGIC);
    return __dadd_rn(__fma_rn(-qh, q, h), l);
}
struct ConvOut { double q, qinv, w2, w3; u64 qi, qinv64; int narrow, pad; };   // 56 bytes -> padded to 64
#ifndef CONV_FOLD
// 1: two-column epilogue for narrow targets (constants C_k = Qhat_i 2^(21k) mod t per input piece, one modular product per output
// instead of three; see narrow_out).  Measured on B200 and left OFF: key switch 507.8 vs 499.3 us at l = 32, base-conversion class
// 9.88 vs 9.36 ms and 55.2-56.2 vs 52.6 ms per DirectSort N=128 sort (tools/build_variant.sh fold "-DCONV_FOLD=1" keyswitch) -- the
// per-CTA staging of the piece constants (two modular products per table entry ahead of the barrier) costs more than the 14 FP64
// instructions per output it saves.
#define CONV_FOLD 0
#endif
#ifndef CONV_UNROLL
#define CONV_UNROLL 1   /* 2 (two outputs in flight) measured 1 % slower at l = 32: 511.8 vs 506.0 us per key switch */
#endif
static constexpr int kConvUnroll = CONV_UNROLL;
template <int NI, int WMODE>
__global__ void __launch_bounds__(256)
k_conv_fp2(u64* __restrict__ out, const u64* __restrict__ in, const __grid_constant__ ConvBatch B, const ModInfo* __restrict__ mi, int n,
           int ysplit) {
    pdl_trigger();   // the constants below come from context tables: staged while the producer of `in` is still running
    extern __shared__ double4 sh_k[];                     // [jc][NI] (c0, c1, c2, -) then [jc] ConvOut
    const int gi = blockIdx.y / ysplit, part = blockIdx.y - gi * ysplit;
    const int nin = B.g[gi].nin, nout = B.g[gi].nout;
    const int per = (nout + ysplit - 1) / ysplit, jlo = part * per, jc = min(per, nout - jlo);
    if (jc <= 0) return;
#if CONV_FOLD
    double2* sh_x = reinterpret_cast<double2*>(sh_k + (size_t)per * NI);   // [jc][NI] third-piece constants of narrow targets
    ConvOut* sh_o = reinterpret_cast<ConvOut*>(sh_x + (size_t)per * NI);
#else
    ConvOut* sh_o = reinterpret_cast<ConvOut*>(sh_k + (size_t)per * NI);
#endif
    for (int t = threadIdx.x; t < jc; t += blockDim.x) {
        const ModInfo m = mi[B.out_ids[gi][jlo + t]];
        ConvOut o;
        o.qi = m.q; o.qinv64 = m.qinv; o.q = (double)m.q; o.qinv = __drcp_rn(o.q); o.narrow = m.bits <= 42 ? 1 : 0; o.pad = 0;
        o.w2 = o.narrow ? (double)((1ull << 42) % m.q) : 0.0;
        o.w3 = o.narrow ? (double)((1ull << 63) % m.q) : 0.0;
        sh_o[t] = o;
    }
    for (int t = threadIdx.x; t < NI * jc; t += blockDim.x) {
        const int jj = t / NI, i = t - jj * NI;
        const ModInfo mo = mi[B.out_ids[gi][jlo + jj]];
        const bool narrow = mo.bits <= 42;
        const u64* tab = narrow ? B.g[gi].hatplain : B.g[gi].hatmod;
        const u64 h = i < nin ? tab[(size_t)i * nout + jlo + jj] : 0;
#if CONV_FOLD
        if (narrow) {
            // C_k = Qhat_i 2^(21 k) mod t for the three input pieces k, each cut into its low and high 21 bits (see narrow_out)
            const double q = (double)mo.q, qinv = __drcp_rn(q);
            double h1 = cv_mulmod((double)h, 2097152.0, q, qinv); if (h1 < 0.0) h1 = __dadd_rn(h1, q);
            double h2 = cv_mulmod(h1, 2097152.0, q, qinv); if (h2 < 0.0) h2 = __dadd_rn(h2, q);
            const u64 u1 = column_to_u64(h1), u2 = column_to_u64(h2);
            sh_k[t] = make_double4((double)((u32)h & 0x1FFFFFu), (double)(u32)(h >> 21), (double)((u32)u1 & 0x1FFFFFu), (double)(u32)(u1 >> 21));
            sh_x[t] = make_double2((double)((u32)u2 & 0x1FFFFFu), (double)(u32)(u2 >> 21));
        } else {
            sh_k[t] = make_double4((double)((u32)h & 0x1FFFFFu), (double)((u32)(h >> 21) & 0x1FFFFFu), (double)(u32)(h >> 42), 0.0);
            sh_x[t] = make_double2(0.0, 0.0);
        }
#else
        sh_k[t] = make_double4((double)((u32)h & 0x1FFFFFu), (double)((u32)(h >> 21) & 0x1FFFFFu), (double)(u32)(h >> 42), 0.0);
#endif
    }
    __syncthreads();
    pdl_wait();
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    constexpr int NW = WMODE == 2 ? NI : 1;
    double y0[NI], y1[NI], y2[NW];
    const u64* src = in + (size_t)B.g[gi].in_row0 * n + x;
#pragma unroll
    for (int i = 0; i < NI; i++) {
        const u64 v = i < nin ? src[(size_t)i * n] : 0;
        y0[i] = piece_to_double((u32)v & 0x1FFFFFu);
        y1[i] = piece_to_double((u32)(v >> 21) & 0x1FFFFFu);
        if (i < NW) y2[i] = piece_to_double((u32)(v >> 42));
    }
    u64* dst = out + (size_t)(B.g[gi].out_row0 + jlo) * n + x;
    // Outputs come in at most three runs of one kind (q0, the other Q limbs, the P limbs); each run is a branch-free loop.
    // The sums are exact integers below 2^53: splitting the middle columns into two accumulators changes nothing but the
    // dependency chains (the ncu capture of round 2 had fixed-latency dependency waits as the top stall of this kernel).
#if CONV_FOLD
    // Narrow targets (t <= 42 bits).  Writing x_i = sum_k y_k[i] 2^(21 k), the output is sum_i sum_k y_k[i] C_k[i] mod t with
    // C_k[i] = Qhat_i 2^(21 k) mod t = C_k0 + C_k1 2^21: TWO exact column sums S0 = sum y_k C_k0, S1 = sum y_k C_k1 (at most
    // 3 x 12 terms below 2^42 each) and out = S0 + S1 2^21 mod t -- one floating-point modular product in the epilogue instead of
    // the three that the weight-ordered columns s0..s3 of k_conv_fp need (27 -> 13 FP64 instructions per output, the contraction's
    // 2 FMAs per piece unchanged).  Same residue, bit for bit.
    auto narrow_out = [&](const int jj) {
        const double4* hp = sh_k + jj * NI;
        const double2* hx = sh_x + jj * NI;
        const ConvOut o = sh_o[jj];
        double s0 = 0.0, s0b = 0.0, s1 = 0.0, s1b = 0.0;
#pragma unroll
        for (int i = 0; i < NI; i++) {
            const double4 c = hp[i];
            s0 = __fma_rn(y0[i], c.x, s0); s1 = __fma_rn(y0[i], c.y, s1);
            s0b = __fma_rn(y1[i], c.z, s0b); s1b = __fma_rn(y1[i], c.w, s1b);
            if (WMODE == 2 || (WMODE == 1 && i == 0)) { const double2 e = hx[i]; s0 = __fma_rn(y2[i], e.x, s0); s1 = __fma_rn(y2[i], e.y, s1); }
        }
        s0 = __dadd_rn(s0, s0b); s1 = __dadd_rn(s1, s1b);
        double r = __dadd_rn(s0, cv_mulmod(s1, 2097152.0, o.q, o.qinv));
        const double t = __dadd_rn(__fma_rn(r, o.qinv, CV_MAGIC), -CV_MAGIC);
        r = __fma_rn(-t, o.q, r);
        if (r < 0.0) r = __dadd_rn(r, o.q);
        dst[(size_t)jj * n] = column_to_u64(r);
    };
#else
    auto narrow_out = [&](const int jj) {
        const double4* hp = sh_k + jj * NI;
        const ConvOut o = sh_o[jj];
        double s0 = 0.0, s1 = 0.0, s1b = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int i = 0; i < NI; i++) {
            const double4 c = hp[i];
            s0 = __fma_rn(y0[i], c.x, s0); s1 = __fma_rn(y0[i], c.y, s1);
            s1b = __fma_rn(y1[i], c.x, s1b); s2 = __fma_rn(y1[i], c.y, s2);
            if (WMODE == 2 || (WMODE == 1 && i == 0)) { s2 = __fma_rn(y2[i], c.x, s2); s3 = __fma_rn(y2[i], c.y, s3); }
        }
        s1 = __dadd_rn(s1, s1b);
        double r = __dadd_rn(s0, cv_mulmod(s1, 2097152.0, o.q, o.qinv));
        r = __dadd_rn(r, cv_mulmod(s2, o.w2, o.q, o.qinv));
        if (WMODE != 0) r = __dadd_rn(r, cv_mulmod(s3, o.w3, o.q, o.qinv));
        const double t = __dadd_rn(__fma_rn(r, o.qinv, CV_MAGIC), -CV_MAGIC);
        r = __fma_rn(-t, o.q, r);
        if (r < 0.0) r = __dadd_rn(r, o.q);
        dst[(size_t)jj * n] = column_to_u64(r);
    };
#endif
    auto wide_out = [&](const int jj) {
        const double4* hp = sh_k + jj * NI;
        const ConvOut o = sh_o[jj];
        double s0 = 0.0, s1 = 0.0, s1b = 0.0, s2 = 0.0, s2b = 0.0, s3 = 0.0, s4 = 0.0;
#pragma unroll
        for (int i = 0; i < NI; i++) {
            const double4 c = hp[i];
            s0 = __fma_rn(y0[i], c.x, s0); s1 = __fma_rn(y0[i], c.y, s1); s2 = __fma_rn(y0[i], c.z, s2);
            s1b = __fma_rn(y1[i], c.x, s1b); s2b = __fma_rn(y1[i], c.y, s2b); s3 = __fma_rn(y1[i], c.z, s3);
            if (WMODE == 2 || (WMODE == 1 && i == 0)) {
                s2 = __fma_rn(y2[i], c.x, s2); s3 = __fma_rn(y2[i], c.y, s3); s4 = __fma_rn(y2[i], c.z, s4);
            }
        }
        s1 = __dadd_rn(s1, s1b); s2 = __dadd_rn(s2, s2b);
        u64 lo = column_to_u64(s0), hi = column_to_u64(s4) << 20;
        add_shifted(lo, hi, column_to_u64(s1), 21); add_shifted(lo, hi, column_to_u64(s2), 42); add_shifted(lo, hi, column_to_u64(s3), 63);
        dst[(size_t)jj * n] = mont_reduce(lo, hi, o.qi, o.qinv64);
    };
    int jj = 0;
    while (jj < jc) {
        const int kind = sh_o[jj].narrow;
        int end = jj + 1;
        while (end < jc && sh_o[end].narrow == kind) end++;
        constexpr int kU = WMODE == 2 ? 1 : kConvUnroll;   // all-wide inputs (ModDown): two outputs in flight need 116 registers
        if (kind) {
#pragma unroll kU
            for (; jj < end; jj++) narrow_out(jj);
        } else {
#pragma unroll kU
            for (; jj < end; jj++) wide_out(jj);
        }
    }
}
template <int NI>
static void launch_conv_fp2(Context* c, dim3 grid, int per, int wmode, u64* out, const u64* in, const ConvBatch& B, int ysplit) {
    const size_t smem = (size_t)NI * per * (sizeof(double4) + (CONV_FOLD ? sizeof(double2) : 0)) + (size_t)per * sizeof(ConvOut);
    if (wmode == 0) launch_pdl(c, k_conv_fp2<NI, 0>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
    else if (wmode == 1) launch_pdl(c, k_conv_fp2<NI, 1>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
    else launch_pdl(c, k_conv_fp2<NI, 2>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
}
template <int NI>
static void launch_conv_fp(Context* c, dim3 grid, size_t smem, int wmode, u64* out, const u64* in, const ConvBatch& B, int ysplit) {
    if (wmode == 0) launch_pdl(c, k_conv_fp<NI, 0>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
    else if (wmode == 1) launch_pdl(c, k_conv_fp<NI, 1>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
    else launch_pdl(c, k_conv_fp<NI, 2>, grid, 256, smem, out, in, B, c->d_mod, c->n, ysplit);
}
void launch_conv_batch(Context* c, u64* out, const u64* in, const ConvBatch& B) {
    KL(c);
    double rows = 0; int max_in = 0, max_out = 0;
    for (int i = 0; i < B.ng; i++) { rows += B.g[i].nin + B.g[i].nout; max_in = std::max(max_in, B.g[i].nin); max_out = std::max(max_out, B.g[i].nout); }
    ProfScope ps(c, KC_BASECONV, 8.0 * c->n * rows);
    // Outputs of a group can be split over `ysplit` CTAs per coefficient block (CKKS_B200_CONV_YSPLIT); every CTA converts the group's
    // inputs to pieces itself, and one CTA per block measured best at every level: inside DirectSort N=128 ysplit 1 / 2 / 3 = 44.48 /
    // 45.12 / 45.95 ms per sort, key switch at l = 32 466.1 vs 471.2 us (2 only from 17 / 24 / 30 outputs on: 44.81 / 44.71 / 44.70 ms).
    static const int ysplit = getenv("CKKS_B200_CONV_YSPLIT") ? std::max(1, atoi(getenv("CKKS_B200_CONV_YSPLIT"))) : 1;
    static const bool generic = getenv("CKKS_B200_GENERIC_CONV") != nullptr;
    static const bool int_conv = getenv("CKKS_B200_INT_CONV") != nullptr;   // A/B switch: the integer-pipe kernel
    static const bool old_fp = getenv("CKKS_B200_OLD_CONV") != nullptr;                 // A/B switch: round 1's k_conv_fp
    if (B.split_ok && !generic && !int_conv && !old_fp && c->n >= 512) {
        int wmode = 0;
        for (int i = 0; i < B.ng; i++) wmode = std::max(wmode, B.g[i].wide == 0 ? 0 : B.g[i].wide == 1 ? 1 : 2);
        dim3 grid((c->n + 255) / 256, B.ng * ysplit);
        const int ni = max_in <= 2 ? 2 : max_in <= 4 ? 4 : max_in <= 6 ? 6 : max_in <= 8 ? 8 : max_in <= 10 ? 10 : max_in <= 11 ? 11 : 12;
        const int per = (max_out + ysplit - 1) / ysplit;
        switch (ni) {
            case 2: launch_conv_fp2<2>(c, grid, per, wmode, out, in, B, ysplit); break;
            case 4: launch_conv_fp2<4>(c, grid, per, wmode, out, in, B, ysplit); break;
            case 6: launch_conv_fp2<6>(c, grid, per, wmode, out, in, B, ysplit); break;
            case 8: launch_conv_fp2<8>(c, grid, per, wmode, out, in, B, ysplit); break;
            case 10: launch_conv_fp2<10>(c, grid, per, wmode, out, in, B, ysplit); break;
            case 11: launch_conv_fp2<11>(c, grid, per, wmode, out, in, B, ysplit); break;
            default: launch_conv_fp2<12>(c, grid, per, wmode, out, in, B, ysplit); break;
        }
    } else if (B.split_ok && !generic && !int_conv && c->n >= 512) {
        int wmode = 0;
        for (int i = 0; i < B.ng; i++) wmode = std::max(wmode, B.g[i].wide == 0 ? 0 : B.g[i].wide == 1 ? 1 : 2);
        dim3 grid((c->n + 255) / 256, B.ng * ysplit);
        const int ni = max_in <= 2 ? 2 : max_in <= 4 ? 4 : max_in <= 6 ? 6 : max_in <= 8 ? 8 : max_in <= 10 ? 10 : max_in <= 11 ? 11 : 12;
        const size_t smem = (size_t)ni * ((max_out + ysplit - 1) / ysplit) * 2 * sizeof(double2);
        switch (ni) {
            case 2: launch_conv_fp<2>(c, grid, smem, wmode, out, in, B, ysplit); break;
            case 4: launch_conv_fp<4>(c, grid, smem, wmode, out, in, B, ysplit); break;
            case 6: launch_conv_fp<6>(c, grid, smem, wmode, out, in, B, ysplit); break;
            case 8: launch_conv_fp<8>(c, grid, smem, wmode, out, in, B, ysplit); break;
            case 10: launch_conv_fp<10>(c, grid, smem, wmode, out, in, B, ysplit); break;
            case 11: launch_conv_fp<11>(c, grid, smem, wmode, out, in, B, ysplit); break;
            default: launch_conv_fp<12>(c, grid, smem, wmode, out, in, B, ysplit); break;
        }
    } else if (B.split_ok && !generic && c->n >= 512) {
        dim3 grid((c->n / 2 + 255) / 256, B.ng * ysplit);
        const int ni = max_in <= 2 ? 2 : max_in <= 4 ? 4 : max_in <= 6 ? 6 : max_in <= 8 ? 8 : max_in <= 10 ? 10 : max_in <= 11 ? 11 : 12;
        const size_t smem = (size_t)ni * ((max_out + ysplit - 1) / ysplit) * sizeof(uint4);
        switch (ni) {
            case 2: launch_conv_split<2>(c, grid, smem, out, in, B, ysplit); break;
            case 4: launch_conv_split<4>(c, grid, smem, out, in, B, ysplit); break;
            case 6: launch_conv_split<6>(c, grid, smem, out, in, B, ysplit); break;
            case 8: launch_conv_split<8>(c, grid, smem, out, in, B, ysplit); break;
            case 10: launch_conv_split<10>(c, grid, smem, out, in, B, ysplit); break;
            case 11: launch_conv_split<11>(c, grid, smem, out, in, B, ysplit); break;
            default: launch_conv_split<12>(c, grid, smem, out, in, B, ysplit); break;
        }
    } else {
        dim3 grid((c->n + 255) / 256, B.ng * ysplit);
        launch_pdl(c, k_conv_batch, grid, 256, 0, out, in, B, c->d_mod, c->n, ysplit);
    }
}

// acc[h][e][x] = sum_j dig_j[e][x] * key[j][h][mod(e)][x]; dig_j[e] is the source polynomial row e
// when e is one of digit j's own limbs, else row of the converted complement (compact layout).
struct DigitMap { int lo[4], hi[4], base[4]; };   // base: first row of digit j's converted complement in conv
// NP = number of digits (compile time, 0 = runtime loop): with the loop unrolled every digit's three 16-byte loads are issued before
// the first multiply-accumulate, so a thread keeps NP x 48 bytes in flight instead of 48 -- the kernel streams the evaluation key
// (two thirds of its traffic) straight from HBM and was at 65 % of the copy peak inside a sort.
template <int NP>
__global__ void __launch_bounds__(256)
k_ks_inner(u64* __restrict__ acc, const u64* __restrict__ src, const u64* __restrict__ conv, const u64* __restrict__ key,
           DigitMap dm, int nparts, int nl, int size_q, int size_p, int M, const ModInfo* __restrict__ mi, int n) {
    pdl_trigger();
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);       // two coefficients per thread, 16-byte accesses
    if (x >= n) return;
    const int e = blockIdx.y, next = nl + size_p;
    const int mid = e < nl ? e : size_q + (e - nl);
    const ModInfo m = mi[mid];
    // (measured and dropped: prefetching this CTA's evaluation-key lines into L2 here, ahead of the wait -- 500.8 vs 498.9 us per key
    // switch at l = 32, the inner-product class 5.81 vs 5.64 ms per sort)
    pdl_wait();
    u64 l0 = 0, h0 = 0, l1 = 0, h1 = 0, l2 = 0, h2 = 0, l3 = 0, h3 = 0;
    auto digit_ptr = [&](int j) -> const u64* {
        if (e >= dm.lo[j] && e < dm.hi[j]) return src + (size_t)e * n + x;
        const int r = e < dm.lo[j] ? e : e - (dm.hi[j] - dm.lo[j]);
        return conv + ((size_t)dm.base[j] + r) * n + x;
    };
    if (NP > 0) {
        ulonglong2 d[NP > 0 ? NP : 1], k0[NP > 0 ? NP : 1], k1[NP > 0 ? NP : 1];
#pragma unroll
        for (int j = 0; j < NP; j++) {
            d[j] = *reinterpret_cast<const ulonglong2*>(digit_ptr(j));
            const u64* kb = key + ((size_t)(j * 2 + 0) * M + mid) * n + x;
            k0[j] = *reinterpret_cast<const ulonglong2*>(kb);
            k1[j] = *reinterpret_cast<const ulonglong2*>(kb + (size_t)M * n);
        }
#pragma unroll
        for (int j = 0; j < NP; j++) {
            mac128(l0, h0, d[j].x, k0[j].x); mac128(l1, h1, d[j].x, k1[j].x);
            mac128(l2, h2, d[j].y, k0[j].y); mac128(l3, h3, d[j].y, k1[j].y);
        }
    } else {
        for (int j = 0; j < nparts; j++) {
            const ulonglong2 d = *reinterpret_cast<const ulonglong2*>(digit_ptr(j));
            const u64* kb = key + ((size_t)(j * 2 + 0) * M + mid) * n + x;
            const ulonglong2 k0 = *reinterpret_cast<const ulonglong2*>(kb), k1 = *reinterpret_cast<const ulonglong2*>(kb + (size_t)M * n);
            mac128(l0, h0, d.x, k0.x); mac128(l1, h1, d.x, k1.x);
            mac128(l2, h2, d.y, k0.y); mac128(l3, h3, d.y, k1.y);
        }
    }
    *reinterpret_cast<ulonglong2*>(acc + (size_t)e * n + x) = make_ulonglong2(mont_reduce(l0, h0, m.q, m.qinv), mont_reduce(l2, h2, m.q, m.qinv));
    *reinterpret_cast<ulonglong2*>(acc + ((size_t)next + e) * n + x) = make_ulonglong2(mont_reduce(l1, h1, m.q, m.qinv), mont_reduce(l3, h3, m.q, m.qinv));
}
void launch_ks_inner2(Context* c, u64* acc, const u64* src, const u64* conv, const EvalKey& k, const KsTables& t) {
    KL(c);
    ProfScope ps(c, KC_KSINNER, 8.0 * c->n * (t.nl + c->size_p) * (3.0 * t.nparts + 2));
    DigitMap dm;
    for (int j = 0, b = 0; j < 4; j++) {
        dm.lo[j] = j < t.nparts ? t.dig_lo[j] : 0; dm.hi[j] = j < t.nparts ? t.dig_hi[j] : 0; dm.base[j] = b;
        if (j < t.nparts) b += t.modup[j].nout;
    }
    dim3 grid((c->n / 2 + 255) / 256, t.nl + c->size_p);
#define KS_INNER_LAUNCH(NPV) launch_pdl(c, k_ks_inner<NPV>, grid, 256, 0, acc, src, conv, k.d, dm, t.nparts, t.nl, c->size_q, c->size_p, c->M(), c->d_mod, c->n)
    switch (t.nparts) {
        case 1: KS_INNER_LAUNCH(1); break;
        case 2: KS_INNER_LAUNCH(2); break;
        case 3: KS_INNER_LAUNCH(3); break;
        case 4: KS_INNER_LAUNCH(4); break;
        default: KS_INNER_LAUNCH(0); break;
    }
#undef KS_INNER_LAUNCH
}

// ApproxModDown epilogue for both polynomials h = 0,1 of a key switch:
//   out[h][i] = addend_h[i] + (acc[h][i] - conv[h][i]) * P^-1 mod q_i          (addends optional)
// With auto_k != 0 the result is written through the automorphism X -> X^k (EVALUATION-format index permutation
// of OpenFHE's PrecomputeAutoMap): out[p] = value[src(p)].
__global__ void k_moddown_final(u64* out, const u64* acc, const u64* conv, const u64* add0, const u64* add1, const u64* pinv,
                                const u64* pinv_sh, const ModInfo* mi, int n, int logn, int nl, int next, u32 auto_k) {
    pdl_sync();
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int h = blockIdx.y / nl, i = blockIdx.y % nl;
    const u64 q = mi[i].q;
    int x = p;
    if (auto_k) {
        u32 j = __brev((u32)p) >> (32 - logn);
        u32 idx = (u32)((((2ull * j + 1) * auto_k) & (2ull * n - 1)) >> 1);
        x = (int)(__brev(idx) >> (32 - logn));
    }
    const size_t o = (size_t)i * n + x;
    u64 r = shoup_mul(sub_mod(acc[(size_t)h * next * n + o], conv[(size_t)h * nl * n + o], q), pinv[i], pinv_sh[i], q);
    const u64* addend = h ? add1 : add0;
    if (addend) r = add_mod(r, addend[o], q);
    out[((size_t)h * nl + i) * n + p] = r;
}
// the same epilogue without an automorphism (relinearisation), two coefficients per thread with 16-byte accesses
__global__ void k_moddown_final2(u64* out, const u64* acc, const u64* conv, const u64* add0, const u64* add1, const u64* pinv,
                                 const u64* pinv_sh, const ModInfo* mi, int n, int nl, int next) {
    pdl_sync();
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (x >= n) return;
    const int h = blockIdx.y / nl, i = blockIdx.y % nl;
    const u64 q = mi[i].q, w = pinv[i], ws = pinv_sh[i];
    const size_t o = (size_t)i * n + x;
    const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(acc + (size_t)h * next * n + o);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(conv + (size_t)h * nl * n + o);
    u64 r0 = shoup_mul(sub_mod(a.x, b.x, q), w, ws, q), r1 = shoup_mul(sub_mod(a.y, b.y, q), w, ws, q);
    const u64* addend = h ? add1 : add0;
    if (addend) { const ulonglong2 d = *reinterpret_cast<const ulonglong2*>(addend + o); r0 = add_mod(r0, d.x, q); r1 = add_mod(r1, d.y, q); }
    *reinterpret_cast<ulonglong2*>(out + ((size_t)h * nl + i) * n + x) = make_ulonglong2(r0, r1);
}
void launch_moddown_final(Context* c, u64* out, const u64* acc, const u64* conv, const u64* add0, const u64* add1, int nl,
                          uint32_t auto_k, const KsTables& t) {
    KL(c);
    ProfScope ps(c, KC_MODDOWN, 8.0 * c->n * nl * (6 + (add0 ? 1 : 0) + (add1 ? 1 : 0)));
    if (!auto_k) {
        launch_pdl(c, k_moddown_final2, dim3((c->n / 2 + 255) / 256, 2 * nl), 256, 0, out, acc, conv, add0, add1, t.d_pinv, t.d_pinv_sh, c->d_mod,
                   c->n, nl, nl + c->size_p);
        return;
    }
    dim3 grid((c->n + 255) / 256, 2 * nl);
    launch_pdl(c, k_moddown_final, grid, 256, 0, out, acc, conv, add0, add1, t.d_pinv, t.d_pinv_sh, c->d_mod, c->n, c->log_n, nl,
                                                 nl + c->size_p, auto_k);
}
