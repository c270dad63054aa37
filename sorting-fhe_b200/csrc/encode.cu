// encode.cu -- CKKS canonical-embedding encode / decode on the GPU (kernel K9).
//
// Replaces OpenFHE MakeCKKSPackedPlaintext -> CKKSPackedEncoding::Encode (DiscreteFourierTransform::
// FFTSpecialInv, scale, round, RNS expand, NTT) and Decode; the reference executes 258 encodes per
// DirectSort N=128 (src/sort_algo.h:341,356,573,706,715).  Floating-point operation order is the
// oracle's (oracle/ckks_prims.c orc_fft_special{,_inv}); every product/sum uses the explicit
// round-to-nearest intrinsics (no FMA contraction) so coefficients are bit-identical to the oracle.
#include "engine.h"
#include <cstring>

#define KL(c) (c)->stats.kernel_launches++; if ((c)->dry) return

__global__ void k_fft_inv_stage(double2* v, int size, int len, const double2* ksi, const uint32_t* rot, int M) {
    pdl_sync();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= size / 2) return;
    int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
    int i = (t / lenh) * len, j = t % lenh;
    int idx = (lenq - (int)(rot[j] % (uint32_t)lenq)) * gap;
    double2 a = v[i + j], b = v[i + j + lenh], w = ksi[idx];
    double ur = __dadd_rn(a.x, b.x), ui = __dadd_rn(a.y, b.y), vr = __dsub_rn(a.x, b.x), vi = __dsub_rn(a.y, b.y);
    double p1 = __dmul_rn(vr, w.x), p2 = __dmul_rn(vi, w.y), p3 = __dmul_rn(vr, w.y), p4 = __dmul_rn(vi, w.x);
    v[i + j] = make_double2(ur, ui);
    v[i + j + lenh] = make_double2(__dsub_rn(p1, p2), __dadd_rn(p3, p4));
}
__global__ void k_fft_fwd_stage(double2* v, int size, int len, const double2* ksi, const uint32_t* rot, int M) {
    pdl_sync();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= size / 2) return;
    int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
    int i = (t / lenh) * len, j = t % lenh;
    int idx = (int)(rot[j] % (uint32_t)lenq) * gap;
    double2 u = v[i + j], b = v[i + j + lenh], w = ksi[idx];
    double p1 = __dmul_rn(b.x, w.x), p2 = __dmul_rn(b.y, w.y), p3 = __dmul_rn(b.x, w.y), p4 = __dmul_rn(b.y, w.x);
    double vr = __dsub_rn(p1, p2), vi = __dadd_rn(p3, p4);
    v[i + j] = make_double2(__dadd_rn(u.x, vr), __dadd_rn(u.y, vi));
    v[i + j + lenh] = make_double2(__dsub_rn(u.x, vr), __dsub_rn(u.y, vi));
}
// out[i] = in[bitrev(i)] * scale (* sf, tracking max |.| as raw double bits when maxbits != nullptr)
__global__ void k_bitrev_scale(double2* out, const double2* in, int size, int bits, double scale, int do_scale, double sf,
                               u64* maxbits) {
    pdl_sync();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= size) return;
    int j = bits ? (int)(__brev((u32)i) >> (32 - bits)) : 0;
    double2 a = in[j];
    if (do_scale) { a.x = __dmul_rn(a.x, scale); a.y = __dmul_rn(a.y, scale); }
    if (maxbits) {
        a.x = __dmul_rn(a.x, sf); a.y = __dmul_rn(a.y, sf);
        double m = fmax(fabs(a.x), fabs(a.y));
        if (m > 0.0) atomicMax(maxbits, (u64)__double_as_longlong(m));   // positive doubles order like their bit patterns
    }
    out[i] = a;
}
// coefficient x of limb row l: slot k = x / gap (x < nh: real part, else imaginary part)
// v holds the scaled slot values.  Values beyond 61 bits are divided by 2^logApprox before rounding and the
// residues multiplied back by 2^logApprox mod q (OpenFHE CKKSPackedEncoding::Encode, 64-bit native path).
__global__ void k_encode_expand(u64* out, const double2* v, const ModInfo* mi, int n, int slots, const u64* maxbits) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n) return;
    int l = blockIdx.y, nh = n >> 1, gap = nh / slots;
    const ModInfo m = mi[l];
    u64 q = m.q, r = 0;
    int xx = x < nh ? x : x - nh;
    if (xx % gap == 0) {
        int log_approx = 0;
        u64 mb = *maxbits;
        if (mb) {   // ceil(log2(max)) from the exponent: frexp gives max = f * 2^e, f in [0.5, 1)
            int e; double f = frexp(__longlong_as_double((long long)mb), &e);
            int logc = (f == 0.5) ? e - 1 : e;
            if (logc > 61) log_approx = logc - 61;
        }
        double2 z = v[xx / gap];
        double val = rint(ldexp(x < nh ? z.x : z.y, -log_approx));
        long long iv = (long long)val;
        {   // |iv| mod q without a 64-bit division: x - mulhi(x, floor(2^64 / q)) q lies in [0, 2q)
            u64 a = iv >= 0 ? (u64)iv : (u64)(-iv);
            a -= __umul64hi(a, m.qrecip) * q;
            if (a >= q) a -= q;
            r = iv >= 0 ? a : (a ? q - a : 0);
        }
        if (log_approx) r = barrett_mul(r, (1ull << log_approx) % q, m);
    }
    out[(size_t)l * n + x] = r;
}

// stages of FFTSpecialInv with len <= 2048 on one contiguous block of min(size, 2048) points held in shared memory,
// followed by the bit-reversal permutation, the 1/size and scaling-factor multiplications and the max-|.| reduction.
// from_real: the block is built from the real input vector (imaginary parts and entries >= nvals are zero).
__global__ void __launch_bounds__(1024)
k_fft_inv_tail(double2* __restrict__ out, const double2* __restrict__ v, const double* __restrict__ real_in, int nvals, int from_real,
               int size, int bits, const double2* __restrict__ ksi, const uint32_t* __restrict__ rot, int M, double inv_size,
               double sf, u64* maxbits) {
    pdl_sync();
    extern __shared__ double2 blk[];
    const int bsz = size < 2048 ? size : 2048, base = blockIdx.x * bsz;
    for (int i = threadIdx.x; i < bsz; i += blockDim.x) {
        const int g = base + i;
        blk[i] = from_real ? make_double2(g < nvals ? real_in[g] : 0.0, 0.0) : v[g];
    }
    __syncthreads();
    for (int len = bsz; len >= 2; len >>= 1) {
        // len, lenq = 4 len and M are powers of two: shifts and masks instead of integer division
        const int lenh = len >> 1, lenq = len << 2, lg = 31 - __clz(lenh), gshift = (31 - __clz(M)) - (31 - __clz(lenq));
        for (int t = threadIdx.x; t < bsz / 2; t += blockDim.x) {
            const int i = (t >> lg) << (lg + 1), j = t & (lenh - 1);
            const int idx = (lenq - (int)(rot[j] & (uint32_t)(lenq - 1))) << gshift;
            const double2 a = blk[i + j], b = blk[i + j + lenh], w = ksi[idx];
            const double ur = __dadd_rn(a.x, b.x), ui = __dadd_rn(a.y, b.y), vr = __dsub_rn(a.x, b.x), vi = __dsub_rn(a.y, b.y);
            const double p1 = __dmul_rn(vr, w.x), p2 = __dmul_rn(vi, w.y), p3 = __dmul_rn(vr, w.y), p4 = __dmul_rn(vi, w.x);
            blk[i + j] = make_double2(ur, ui);
            blk[i + j + lenh] = make_double2(__dsub_rn(p1, p2), __dadd_rn(p3, p4));
        }
        __syncthreads();
    }
    double mx = 0.0;
    for (int i = threadIdx.x; i < bsz; i += blockDim.x) {
        const int g = base + i;
        const int o = bits ? (int)(__brev((u32)g) >> (32 - bits)) : 0;
        double2 a = blk[i];
        a.x = __dmul_rn(__dmul_rn(a.x, inv_size), sf); a.y = __dmul_rn(__dmul_rn(a.y, inv_size), sf);
        mx = fmax(mx, fmax(fabs(a.x), fabs(a.y)));
        out[o] = a;
    }
    if (mx > 0.0) atomicMax(maxbits, (u64)__double_as_longlong(mx));
}
// one stage with len > 2048 (pairs further apart than a block); from_real as above
__global__ void k_fft_inv_stage_big(double2* v, const double* __restrict__ real_in, int nvals, int from_real, int size, int len,
                                    const double2* ksi, const uint32_t* rot, int M) {
    pdl_sync();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= size / 2) return;
    const int lenh = len >> 1, lenq = len << 2, lg = 31 - __clz(lenh), gshift = (31 - __clz(M)) - (31 - __clz(lenq));
    const int i = (t >> lg) << (lg + 1), j = t & (lenh - 1);
    const int idx = (lenq - (int)(rot[j] & (uint32_t)(lenq - 1))) << gshift;
    double2 a, b;
    if (from_real) {
        a = make_double2(i + j < nvals ? real_in[i + j] : 0.0, 0.0);
        b = make_double2(i + j + lenh < nvals ? real_in[i + j + lenh] : 0.0, 0.0);
    } else { a = v[i + j]; b = v[i + j + lenh]; }
    const double2 w = ksi[idx];
    double ur = __dadd_rn(a.x, b.x), ui = __dadd_rn(a.y, b.y), vr = __dsub_rn(a.x, b.x), vi = __dsub_rn(a.y, b.y);
    double p1 = __dmul_rn(vr, w.x), p2 = __dmul_rn(vi, w.y), p3 = __dmul_rn(vr, w.y), p4 = __dmul_rn(vi, w.x);
    v[i + j] = make_double2(ur, ui);
    v[i + j + lenh] = make_double2(__dsub_rn(p1, p2), __dadd_rn(p3, p4));
}

// host staging: a ring of pinned chunks so that MakeCKKSPackedPlaintext never synchronises the stream
static double* stage_values(Context* c, const double* h_values, int nvals) {
    constexpr int RING = 32;
    if (!c->enc_pinned) {
        cudaHostAlloc((void**)&c->enc_pinned, (size_t)RING * (c->n / 2) * sizeof(double), cudaHostAllocDefault);
        c->enc_events.resize(RING, nullptr);
    }
    const int slot = (int)(c->enc_ring_pos++ % RING);
    if (c->enc_events[slot]) cudaEventSynchronize(c->enc_events[slot]);
    else cudaEventCreateWithFlags(&c->enc_events[slot], cudaEventDisableTiming);
    double* p = c->enc_pinned + (size_t)slot * (c->n / 2);
    memcpy(p, h_values, (size_t)nvals * sizeof(double));
    return p;
}

// is_complex: h_values holds nvals (re, im) pairs (bootstrapping's linear-transform diagonals); nvals <= ring / 4
void launch_encode(Context* c, u64* out, const double* h_values, int nvals, int slots, double sf, int nl, int lrep, int is_complex) {
    const int n_c = c->n >> lrep;   // compact row length (engine.h Ct::lrep)
    int bits = 0; while ((1 << bits) < slots) bits++;
    const int big_stages = bits > 11 ? bits - 11 : 0;
    c->stats.kernel_launches += big_stages + 2;
    if (c->dry) return;
    {
        ProfScope ps(c, KC_ENCODE, 8.0 * n_c * nl, big_stages + 2);
        double* d_in = (double*)dev_alloc(c, (size_t)(nvals > 0 ? nvals : 1));
        double2* a = (double2*)dev_alloc(c, 2 * (size_t)slots);
        double2* b = (double2*)dev_alloc(c, 2 * (size_t)slots);
        u64* mx = dev_alloc(c, 1);
        int from_real = 1;
        if (is_complex) {
            if (2 * nvals > c->n / 2) throw std::runtime_error("complex encode: at most ring / 4 values");
            double* staged = stage_values(c, h_values, 2 * nvals);
            cudaMemsetAsync(a, 0, (size_t)slots * sizeof(double2), c->stream);
            cudaMemcpyAsync(a, staged, (size_t)2 * nvals * sizeof(double), cudaMemcpyHostToDevice, c->stream);
            from_real = 0;
        } else {
            double* staged = stage_values(c, h_values, nvals);
            cudaMemcpyAsync(d_in, staged, (size_t)nvals * sizeof(double), cudaMemcpyHostToDevice, c->stream);
        }
        cudaEventRecord(c->enc_events[(c->enc_ring_pos - 1) % c->enc_events.size()], c->stream);
        cudaMemsetAsync(mx, 0, sizeof(u64), c->stream);
        const int M = 2 * c->n;
        for (int len = slots; len > 2048; len >>= 1) {
            launch_pdl(c, k_fft_inv_stage_big, (slots / 2 + 255) / 256, 256, 0, a, d_in, nvals, from_real, slots, len,
                                                                                  (const double2*)c->d_ksi, c->d_rot, M);
            from_real = 0;
        }
        const int bsz = slots < 2048 ? slots : 2048, threads = bsz / 2 < 32 ? 32 : (bsz / 2 > 1024 ? 1024 : bsz / 2);
        launch_pdl(c, k_fft_inv_tail, slots / bsz, threads, (size_t)bsz * sizeof(double2), 
            b, a, d_in, nvals, from_real, slots, bits, (const double2*)c->d_ksi, c->d_rot, M, 1.0 / (double)slots, sf, mx);
        dim3 grid((n_c + 255) / 256, nl);
        launch_pdl(c, k_encode_expand, grid, 256, 0, out, b, c->d_mod, n_c, slots, mx);
        dev_free(c, d_in); dev_free(c, a); dev_free(c, b); dev_free(c, mx);
    }
    launch_ntt_ex(c, out, out, nl, mods_q(nl), false, nullptr, RowMap{}, RowMap{}, nullptr, lrep);
}

// centred CRT of limbs (0,1) (or limb 0 alone) -> double, then gather slots with 1/sf
__global__ void k_crt_gather(double2* v, const u64* m, int nlimbs, const ModInfo* mi, int n, int slots, double sf, u64 q0inv_q1) {
    pdl_sync();
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= slots) return;
    int nh = n >> 1, gap = nh / slots;
    double o[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        int x = k * gap + h * nh;
        u64 a0 = m[x], q0 = mi[0].q;
        double d;
        if (nlimbs >= 2) {
            const ModInfo m1 = mi[1];
            u64 q1 = m1.q, a1 = m[(size_t)n + x];
            u64 dd = sub_mod(a1 % q1, a0 % q1, q1);
            u64 kk = barrett_mul(dd, q0inv_q1, m1);
            unsigned __int128 val = (unsigned __int128)a0 + (unsigned __int128)q0 * kk;
            unsigned __int128 Q = (unsigned __int128)q0 * q1, half = Q >> 1;
            if (val > half) { unsigned __int128 neg = Q - val;
                d = -__dadd_rn(__dmul_rn(__ull2double_rn((u64)(neg >> 64)), 18446744073709551616.0), __ull2double_rn((u64)neg)); }
            else d = __dadd_rn(__dmul_rn(__ull2double_rn((u64)(val >> 64)), 18446744073709551616.0), __ull2double_rn((u64)val));
        } else {
            d = a0 > (q0 >> 1) ? -__ull2double_rn(q0 - a0) : __ull2double_rn(a0);
        }
        o[h] = __ddiv_rn(d, sf);
    }
    v[k] = make_double2(o[0], o[1]);
}

void launch_decode(Context* c, double* h_out, const u64* m, int nlimbs, double sf, int slots) {
    int bits = 0; while ((1 << bits) < slots) bits++;
    c->stats.kernel_launches += bits + 2;
    if (c->dry) { for (int i = 0; i < slots; i++) h_out[i] = 0.0; return; }
    double2* a = (double2*)dev_alloc(c, 2 * (size_t)slots);
    double2* b = (double2*)dev_alloc(c, 2 * (size_t)slots);
    u64 q0inv = 0;
    if (nlimbs >= 2) {
        u64 q0 = c->moduli[0], q1 = c->moduli[1];
        unsigned __int128 r = 1, bb = q0 % q1; u64 e = q1 - 2;
        while (e) { if (e & 1) r = r * bb % q1; bb = bb * bb % q1; e >>= 1; }
        q0inv = (u64)r;
    }
    launch_pdl(c, k_crt_gather, (slots + 255) / 256, 256, 0, a, m, nlimbs, c->d_mod, c->n, slots, sf, q0inv);
    launch_pdl(c, k_bitrev_scale, (slots + 255) / 256, 256, 0, b, a, slots, bits, 1.0, 0, 1.0, nullptr);
    const int M = 2 * c->n;
    for (int len = 2; len <= slots; len <<= 1)
        launch_pdl(c, k_fft_fwd_stage, (slots / 2 + 255) / 256, 256, 0, b, slots, len, (const double2*)c->d_ksi, c->d_rot, M);
    std::vector<double> h(2 * (size_t)slots);
    cudaMemcpyAsync(h.data(), b, h.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
    cudaStreamSynchronize(c->stream);
    for (int i = 0; i < slots; i++) h_out[i] = h[2 * i];
    dev_free(c, a);
    dev_free(c, b);
}
