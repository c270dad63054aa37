// ntt.cu -- batched negacyclic NTT / inverse NTT over all RNS limbs (kernel K1).
//
// Replaces OpenFHE ChineseRemainderTransformFTT::ForwardTransformToBitReverse /
// InverseTransformFromBitReverse, which every cc->EvalMult / EvalRotate / rescale of the
// reference reaches (src/sign.cpp:23-56, src/rotation.h:224, src/sort_algo.h:358).
// Semantics (pinned by oracle/ckks_prims.c orc_ntt_fwd/orc_ntt_inv): Cooley-Tukey, natural in ->
// bit-reversed out with table W[m+i] = psi^bitrev(m+i); inverse Gentleman-Sande with n^-1.
//
// Design: a limb of n = 2^logn words is processed in <= 2 HBM passes.  A CTA stages a tile of
// 2^tb (tb = min(11, logn)) words in shared memory; every thread owns 8 words in registers and
// performs up to 3 butterfly stages per shared-memory round trip (radix-8), Harvey lazy
// butterflies with Shoup twiddles on the integer pipes.  The "strided" phase handles the first
// s1 = logn - tb stages on tiles of 2^s1 rows x 2^(tb-s1) contiguous columns (coalesced 256 B+
// runs); the "local" phase handles the last tb stages on contiguous 2^tb-word chunks.
// Index math was validated against the oracle by tools/emulate_ntt_indexing.py.
#include "engine.h"

static constexpr int TB_MAX = 11;

__device__ __forceinline__ int padi(int e) { return e + (e >> 3); }

template <bool INV, int S>
__device__ __forceinline__ void ntt_pass(u64* sm, int tb, int sl_lo, int s0, u32 pre, const u64* __restrict__ W,
                                         const u64* __restrict__ Wsh, u64 q) {
    constexpr int G = 8 >> S;
    constexpr int E = 1 << S;
    const int lowbits = tb - sl_lo - S;
    const u64 q2 = q << 1;
#pragma unroll
    for (int g = 0; g < G; g++) {
        const int u = threadIdx.x * G + g;
        const int e0 = ((u >> lowbits) << (lowbits + S)) | (u & ((1 << lowbits) - 1));
        const u32 top = (u32)e0 >> (tb - sl_lo);
        u64 r[E];
#pragma unroll
        for (int k = 0; k < E; k++) r[k] = sm[padi(e0 + (k << lowbits))];
        if (!INV) {
#pragma unroll
            for (int j = 0; j < S; j++) {
                const int sl = sl_lo + j, s = s0 + sl, bit = 1 << (S - 1 - j);
#pragma unroll
                for (int k = 0; k < E; k++) {
                    if (k & bit) continue;
                    const u32 ti = (1u << s) + ((pre << sl) | (top << j) | (u32)(k >> (S - j)));
                    const u64 w = W[ti], wsh = Wsh[ti];
                    u64 X = r[k];
                    if (X >= q2) X -= q2;
                    const u64 T = shoup_lazy(r[k | bit], w, wsh, q);
                    r[k] = X + T;
                    r[k | bit] = X - T + q2;
                }
            }
        } else {
#pragma unroll
            for (int j = S - 1; j >= 0; j--) {
                const int sl = sl_lo + j, s = s0 + sl, bit = 1 << (S - 1 - j);
#pragma unroll
                for (int k = 0; k < E; k++) {
                    if (k & bit) continue;
                    const u32 ti = (1u << s) + ((pre << sl) | (top << j) | (u32)(k >> (S - j)));
                    const u64 w = W[ti], wsh = Wsh[ti];
                    const u64 X = r[k], Y = r[k | bit];
                    u64 sum = X + Y;
                    if (sum >= q2) sum -= q2;
                    r[k] = sum;
                    r[k | bit] = shoup_lazy(X - Y + q2, w, wsh, q);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < E; k++) sm[padi(e0 + (k << lowbits))] = r[k];
    }
}

template <bool INV>
__global__ void __launch_bounds__(256)
ntt_kernel(u64* __restrict__ data, LimbMods mods, const ModInfo* __restrict__ modinfo, const u64* __restrict__ tw,
           const u64* __restrict__ tw_sh, int logn, int tb, int s0, int ns, int strided, int s1, int final_phase) {
    extern __shared__ u64 sm[];
    const int row = blockIdx.y;
    const int mid = mods.m[row % mods.n];
    const ModInfo mi = modinfo[mid];
    const u64 q = mi.q;
    const u64* W = tw + ((size_t)mid << logn);
    const u64* Wsh = tw_sh + ((size_t)mid << logn);
    u64* base = data + ((size_t)row << logn);
    const int tile = blockIdx.x, tsize = 1 << tb, cb = tb - s1;
    const u32 pre = strided ? 0u : (u32)tile;

    for (int e = threadIdx.x; e < tsize; e += blockDim.x) {
        const size_t g = strided ? (((size_t)(e >> cb) << (logn - s1)) | (size_t)((tile << cb) + (e & ((1 << cb) - 1))))
                                 : (((size_t)tile << tb) + e);
        sm[padi(e)] = base[g];
    }
    __syncthreads();
    if (!INV) {
        for (int sl = 0; sl < ns;) {
            const int S = min(3, ns - sl);
            if (S == 3) ntt_pass<false, 3>(sm, tb, sl, s0, pre, W, Wsh, q);
            else if (S == 2) ntt_pass<false, 2>(sm, tb, sl, s0, pre, W, Wsh, q);
            else ntt_pass<false, 1>(sm, tb, sl, s0, pre, W, Wsh, q);
            sl += S;
            __syncthreads();
        }
    } else {
        // groups are the forward groups taken in reverse order
        const int rem = ns % 3;  // forward grouping: 3,3,...,rem  -> last group has size rem (if rem != 0)
        int sl_hi = ns;
        bool first = true;
        while (sl_hi > 0) {
            const int S = (first && rem) ? rem : 3;
            first = false;
            const int sl = sl_hi - S;
            if (S == 3) ntt_pass<true, 3>(sm, tb, sl, s0, pre, W, Wsh, q);
            else if (S == 2) ntt_pass<true, 2>(sm, tb, sl, s0, pre, W, Wsh, q);
            else ntt_pass<true, 1>(sm, tb, sl, s0, pre, W, Wsh, q);
            sl_hi = sl;
            __syncthreads();
        }
    }
    const u64 q2 = q << 1;
    for (int e = threadIdx.x; e < tsize; e += blockDim.x) {
        const size_t g = strided ? (((size_t)(e >> cb) << (logn - s1)) | (size_t)((tile << cb) + (e & ((1 << cb) - 1))))
                                 : (((size_t)tile << tb) + e);
        u64 v = sm[padi(e)];
        if (!INV) { if (v >= q2) v -= q2; }
        else if (final_phase) v = shoup_lazy(v, mi.ninv, mi.ninv_sh, q);
        if (v >= q) v -= q;
        base[g] = v;
    }
}

void launch_ntt(Context* c, u64* data, int rows, const LimbMods& m, bool inverse) {
    c->stats.kernel_launches += (c->log_n > TB_MAX) ? 2 : 1;
    if (c->dry || rows == 0) return;
    ProfScope ps(c, KC_NTT, 16.0 * rows * c->n, (c->log_n > TB_MAX) ? 2 : 1);
    const int logn = c->log_n, tb = logn < TB_MAX ? logn : TB_MAX, s1 = logn - tb;
    const int threads = (1 << tb) / 8;
    const size_t smem = (size_t)((1 << tb) + ((1 << tb) >> 3) + 8) * sizeof(u64);
    dim3 grid(c->n >> tb, rows);
    if (!inverse) {
        if (s1 > 0)
            ntt_kernel<false><<<grid, threads, smem, c->stream>>>(data, m, c->d_mod, c->d_psi, c->d_psi_sh, logn, tb, 0, s1, 1, s1, 0);
        ntt_kernel<false><<<grid, threads, smem, c->stream>>>(data, m, c->d_mod, c->d_psi, c->d_psi_sh, logn, tb, s1, tb, 0, s1, 1);
    } else {
        ntt_kernel<true><<<grid, threads, smem, c->stream>>>(data, m, c->d_mod, c->d_ipsi, c->d_ipsi_sh, logn, tb, s1, tb, 0, s1,
                                                             s1 == 0);
        if (s1 > 0)
            ntt_kernel<true><<<grid, threads, smem, c->stream>>>(data, m, c->d_mod, c->d_ipsi, c->d_ipsi_sh, logn, tb, 0, s1, 1, s1, 1);
    }
}
