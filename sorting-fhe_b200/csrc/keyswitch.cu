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

#define KL(c) (c)->stats.kernel_launches++; if ((c)->dry) return
#define MAX_CONV_IN 24

struct ConvIds { short in_ids[MAX_CONV_IN]; short out_ids[MAX_LIMBS]; };

// out[j][x] = sum_i ([in[i][x] * hatinv_i]_{q_i}) * hatmod[i][j]  mod t_j     (no correction term)
__global__ void __launch_bounds__(256)
k_base_conv(u64* __restrict__ out, const u64* __restrict__ in, ConvIds ids, int nin, int nout,
            const u64* __restrict__ hatinv, const u64* __restrict__ hatinv_sh, const u64* __restrict__ hatmod,
            const ModInfo* __restrict__ mi, int n) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    u64 y[MAX_CONV_IN];
#pragma unroll 4
    for (int i = 0; i < nin; i++) {
        u64 q = mi[ids.in_ids[i]].q;
        y[i] = shoup_mul(in[(size_t)i * n + x], hatinv[i], hatinv_sh[i], q);
    }
    // targets are split over grid.y to expose more parallelism
    for (int j = blockIdx.y; j < nout; j += gridDim.y) {
        const ModInfo m = mi[ids.out_ids[j]];
        u64 lo = 0, hi = 0;
        for (int i = 0; i < nin; i++) mac128(lo, hi, y[i], hatmod[(size_t)i * nout + j]);
        out[(size_t)j * n + x] = mont_reduce(lo, hi, m.q, m.qinv);
    }
}
void launch_base_conv(Context* c, u64* out, const u64* in, const ConvTables& t) {
    KL(c);
    ProfScope ps(c, KC_BASECONV, 8.0 * c->n * (t.nin + t.nout));
    ConvIds ids;
    for (int i = 0; i < t.nin; i++) ids.in_ids[i] = t.in_ids[i];
    for (int j = 0; j < t.nout; j++) ids.out_ids[j] = t.out_ids[j];
    int ysplit = t.nout < 4 ? t.nout : 4;
    dim3 grid((c->n + 255) / 256, ysplit);
    k_base_conv<<<grid, 256, 0, c->stream>>>(out, in, ids, t.nin, t.nout, t.d_hatinv, t.d_hatinv_sh, t.d_hatmod, c->d_mod, c->n);
}

// acc[h][e][x] = sum_j dig_j[e][x] * key[j][h][mod(e)][x]; dig_j[e] is the source polynomial row e
// when e is one of digit j's own limbs, else row of the converted complement (compact layout).
struct DigitMap { int lo[4], hi[4]; };
__global__ void __launch_bounds__(256)
k_ks_inner(u64* __restrict__ acc, const u64* __restrict__ src, const u64* __restrict__ conv, const u64* __restrict__ key,
           DigitMap dm, int nparts, int nl, int size_q, int size_p, int M, const ModInfo* __restrict__ mi, int n) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    const int e = blockIdx.y, next = nl + size_p;
    const int mid = e < nl ? e : size_q + (e - nl);
    const ModInfo m = mi[mid];
    u64 l0 = 0, h0 = 0, l1 = 0, h1 = 0;
    for (int j = 0; j < nparts; j++) {
        u64 d;
        if (e >= dm.lo[j] && e < dm.hi[j]) d = src[(size_t)e * n + x];
        else {
            int r = e < dm.lo[j] ? e : e - (dm.hi[j] - dm.lo[j]);
            d = conv[((size_t)j * next + r) * n + x];
        }
        const u64* kb = key + ((size_t)(j * 2 + 0) * M + mid) * n + x;
        mac128(l0, h0, d, kb[0]);
        mac128(l1, h1, d, kb[(size_t)M * n]);
    }
    acc[(size_t)e * n + x] = mont_reduce(l0, h0, m.q, m.qinv);
    acc[((size_t)next + e) * n + x] = mont_reduce(l1, h1, m.q, m.qinv);
}
void launch_ks_inner2(Context* c, u64* acc, const u64* src, const u64* conv, const EvalKey& k, const KsTables& t) {
    KL(c);
    ProfScope ps(c, KC_KSINNER, 8.0 * c->n * (t.nl + c->size_p) * (3.0 * t.nparts + 2));
    DigitMap dm;
    for (int j = 0; j < 4; j++) { dm.lo[j] = j < t.nparts ? t.dig_lo[j] : 0; dm.hi[j] = j < t.nparts ? t.dig_hi[j] : 0; }
    dim3 grid((c->n + 255) / 256, t.nl + c->size_p);
    k_ks_inner<<<grid, 256, 0, c->stream>>>(acc, src, conv, k.d, dm, t.nparts, t.nl, c->size_q, c->size_p, c->M(), c->d_mod, c->n);
}

// out[i] = addend[i] + (acc[i] - conv[i]) * P^-1 mod q_i   (ApproxModDown epilogue; addend optional).
// With auto_k != 0 the result is written through the automorphism X -> X^k (EVALUATION-format index
// permutation of OpenFHE's PrecomputeAutoMap): out[p] = value[src(p)].
__global__ void k_moddown_final(u64* out, const u64* acc, const u64* conv, const u64* addend, const u64* pinv,
                                const u64* pinv_sh, const ModInfo* mi, int n, int logn, u32 auto_k) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int i = blockIdx.y; u64 q = mi[i].q;
    int x = p;
    if (auto_k) {
        u32 j = __brev((u32)p) >> (32 - logn);
        u32 idx = (u32)((((2ull * j + 1) * auto_k) & (2ull * n - 1)) >> 1);
        x = (int)(__brev(idx) >> (32 - logn));
    }
    size_t o = (size_t)i * n + x;
    u64 r = shoup_mul(sub_mod(acc[o], conv[o], q), pinv[i], pinv_sh[i], q);
    if (addend) r = add_mod(r, addend[o], q);
    out[(size_t)i * n + p] = r;
}
void launch_moddown_final(Context* c, u64* out, const u64* acc, const u64* conv, const u64* addend, int nl,
                          uint32_t auto_k, const KsTables& t) {
    KL(c);
    ProfScope ps(c, KC_MODDOWN, 8.0 * c->n * nl * (addend ? 4 : 3));
    dim3 grid((c->n + 255) / 256, nl);
    k_moddown_final<<<grid, 256, 0, c->stream>>>(out, acc, conv, addend, t.d_pinv, t.d_pinv_sh, c->d_mod, c->n, c->log_n, auto_k);
}
