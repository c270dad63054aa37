// kernels.cu -- limb-wise pointwise kernels (K5-K8 of SURVEY.md 2.2), rescale, sampling.
//
// All are HBM-streaming integer kernels: one thread per coefficient, rows = parts x limbs on
// grid.y, coalesced 8-byte accesses, per-limb modulus looked up from a by-value LimbMods.
// They replace the DCRTPoly operator+ / operator* / AutomorphismTransform /
// DropLastElementAndScale calls OpenFHE executes under the reference's cc->EvalAdd / EvalMult /
// EvalRotate / implicit rescale; semantics pinned by
// oracle/ckks_prims.c (orc_v*, orc_automorph_eval, orc_switch_modulus).
#include "engine.h"
#include <map>

struct LimbScalars { u64 v[MAX_LIMBS]; u64 sh[MAX_LIMBS]; };

static inline dim3 grid_for(const Context* c, int rows) { return dim3((c->n + 255) / 256, rows); }
// streaming kernels below move two coefficients (16 bytes) per thread per operand
static inline dim3 grid_for2(const Context* c, int rows) { return dim3((c->n / 2 + 255) / 256, rows); }
typedef ulonglong2 u64x2;
__device__ __forceinline__ u64x2 ld2(const u64* p, size_t o) { return *reinterpret_cast<const u64x2*>(p + o); }
__device__ __forceinline__ void st2(u64* p, size_t o, u64 a, u64 b) { *reinterpret_cast<u64x2*>(p + o) = make_ulonglong2(a, b); }
#define X2_PROLOGUE int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x); if (x >= n) return
#define KL(c) (c)->stats.kernel_launches++; if ((c)->dry) return

LimbMods mods_q(int nl) { LimbMods m; m.n = nl; for (int i = 0; i < nl; i++) m.m[i] = (short)i; return m; }
LimbMods mods_qp(const Context* c, int nl) {
    LimbMods m; m.n = nl + c->size_p;
    for (int i = 0; i < nl; i++) m.m[i] = (short)i;
    for (int k = 0; k < c->size_p; k++) m.m[nl + k] = (short)(c->size_q + k);
    return m;
}
LimbMods mods_range(int lo, int hi) { LimbMods m; m.n = hi - lo; for (int i = lo; i < hi; i++) m.m[i - lo] = (short)i; return m; }

static u64 host_shoup(u64 w, u64 q) { return (u64)((((unsigned __int128)w) << 64) / q); }

__global__ void k_addsub(u64* out, const u64* a, const u64* b, LimbMods mods, const ModInfo* mi, int n, int op) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y; u64 q = mi[mods.m[row % mods.n]].q;
    size_t o = (size_t)row * n + x;
    out[o] = op == 0 ? add_mod(a[o], b[o], q) : sub_mod(a[o], b[o], q);
}
void launch_add(Context* c, u64* out, const u64* a, const u64* b, int rows, const LimbMods& m, int op) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 24.0 * rows * c->n); launch_pdl(c, k_addsub, grid_for(c, rows), 256, 0, out, a, b, m, c->d_mod, c->n, op);
}

__global__ void k_neg(u64* out, const u64* a, LimbMods mods, const ModInfo* mi, int n) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y; u64 q = mi[mods.m[row % mods.n]].q;
    size_t o = (size_t)row * n + x; out[o] = neg_mod(a[o], q);
}
void launch_neg(Context* c, u64* out, const u64* a, int rows, const LimbMods& m) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 16.0 * rows * c->n); launch_pdl(c, k_neg, grid_for(c, rows), 256, 0, out, a, m, c->d_mod, c->n);
}

// ---- FP64 modular products for moduli of at most 44 bits (the 40-bit scaling primes: 30 of the 32 Q limbs).  Same arithmetic as
// ntt.cu's FP64 butterflies: h = a*b (rounded), l = fma(a, b, -h) (exact low part), qh = rint(h / q), r = fma(-qh, q, h) + l, every
// intermediate an exactly representable integer, |r| <= q/2 (1 + 2^-8) + 2^28 for operands below q.  Six FP64-pipe instructions
// instead of the ~45 integer instructions of the 128-bit Barrett product; the canonical residue is the same, bit for bit.
#define PW_FP_BITS 44
#define PW_MAGIC 6755399441055744.0   // 1.5 * 2^52
#define PW_2P52 4503599627370496.0
__device__ __forceinline__ double pw_from_u64(u64 v) { return __dadd_rn(__longlong_as_double((long long)(v | 0x4330000000000000ULL)), -PW_2P52); }   // v < 2^52
__device__ __forceinline__ u64 pw_to_u64(double y) { return (u64)__double_as_longlong(__dadd_rn(y, PW_2P52)) & 0x000FFFFFFFFFFFFFULL; }                 // integer y in [0, 2^52)
__device__ __forceinline__ double pw_mulmod(double a, double b, double q, double qinv) {
    const double h = __dmul_rn(a, b);
    const double l = __fma_rn(a, b, -h);
    const double qh = __dadd_rn(__fma_rn(h, qinv, PW_MAGIC), -PW_MAGIC);
    return __dadd_rn(__fma_rn(-qh, q, h), l);
}
__device__ __forceinline__ u64 pw_canon_reduced(double y, double q) { if (y < 0.0) y = __dadd_rn(y, q); return pw_to_u64(y); }   // |y| < q
__device__ __forceinline__ u64 pw_canon(double x, double q, double qinv) {                                                        // |x| < 2^51
    const double t = __dadd_rn(__fma_rn(x, qinv, PW_MAGIC), -PW_MAGIC);
    return pw_canon_reduced(__fma_rn(-t, q, x), q);
}

__global__ void k_mul(u64* out, const u64* a, const u64* b, LimbMods mods, const ModInfo* mi, int n, int rows_b, int lrep_b) {
    pdl_trigger();
    X2_PROLOGUE;
    int row = blockIdx.y; const ModInfo m = mi[mods.m[row % mods.n]];
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    size_t o = (size_t)row * n + x;
    const u64x2 va = ld2(a, o);
    u64x2 vb;
    if (lrep_b == 0) vb = ld2(b, (size_t)(row % rows_b) * n + x);
    else { const u64 t = b[(size_t)(row % rows_b) * (n >> lrep_b) + (x >> lrep_b)]; vb = make_ulonglong2(t, t); }   // x even: x, x+1 share a group
    if (m.bits <= PW_FP_BITS) {
        const double q = (double)m.q, qinv = __drcp_rn(q);
        st2(out, o, pw_canon_reduced(pw_mulmod(pw_from_u64(va.x), pw_from_u64(vb.x), q, qinv), q),
            pw_canon_reduced(pw_mulmod(pw_from_u64(va.y), pw_from_u64(vb.y), q, qinv), q));
        return;
    }
    st2(out, o, barrett_mul(va.x, vb.x, m), barrett_mul(va.y, vb.y, m));
}
void launch_mul(Context* c, u64* out, const u64* a, const u64* b, int rows_out, int rows_b, const LimbMods& m, int lrep_b) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * (2.0 * rows_out + (double)rows_out / (1 << lrep_b)));
    launch_pdl(c, k_mul, grid_for2(c, rows_out), 256, 0, out, a, b, m, c->d_mod, c->n, rows_b, lrep_b);
}

__global__ void k_mul_scalar(u64* out, const u64* d, LimbMods mods, const ModInfo* mi, int n, LimbScalars s) {
    pdl_trigger();
    X2_PROLOGUE;
    int row = blockIdx.y, l = row % mods.n; u64 q = mi[mods.m[l]].q;
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    size_t o = (size_t)row * n + x;
    const u64x2 v = ld2(d, o);
    st2(out, o, shoup_mul(v.x, s.v[l], s.sh[l], q), shoup_mul(v.y, s.v[l], s.sh[l], q));
}
void launch_mul_scalar(Context* c, u64* out, const u64* data, int rows, const LimbMods& m, const u64* hs, int lrep) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 16.0 * rows * (c->n >> lrep));
    LimbScalars s;
    for (int i = 0; i < m.n; i++) { s.v[i] = hs[i]; s.sh[i] = host_shoup(hs[i], c->moduli[m.m[i]]); }
    const int n = c->n >> lrep;
    launch_pdl(c, k_mul_scalar, dim3((n / 2 + 255) / 256, rows), 256, 0, out, data, m, c->d_mod, n, s);
}

__global__ void k_add_scalar(u64* out, const u64* d, const ModInfo* mi, int n, int nl, LimbScalars s, int negate) {
    pdl_trigger();
    X2_PROLOGUE;
    int row = blockIdx.y, l = row % nl; u64 q = mi[l].q;
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    size_t o = (size_t)row * n + x;
    u64x2 v = ld2(d, o);
    if (negate == 1) { v.x = neg_mod(v.x, q); v.y = neg_mod(v.y, q); }
    else if (negate == 2) { v.x = add_mod(v.x, v.x, q); v.y = add_mod(v.y, v.y, q); }   // 2 * ct + const (Chebyshev recurrences)
    if (row < nl) { v.x = add_mod(v.x, s.v[l], q); v.y = add_mod(v.y, s.v[l], q); }
    st2(out, o, v.x, v.y);
}
void launch_add_scalar(Context* c, u64* out, const u64* in, int parts, int nl, const u64* hs, int negate) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 16.0 * parts * nl * c->n);
    LimbScalars s;
    for (int i = 0; i < nl; i++) { s.v[i] = hs[i]; s.sh[i] = 0; }
    launch_pdl(c, k_add_scalar, grid_for2(c, parts * nl), 256, 0, out, in, c->d_mod, c->n, nl, s, negate);
}

__global__ void k_addsub_parts(u64* out, const u64* xs, int xparts, const u64* ys, int yparts, const ModInfo* mi, int n, int nl,
                               int sub, int lrx, int lry) {
    pdl_trigger();
    X2_PROLOGUE;
    int row = blockIdx.y, p = row / nl; u64 q = mi[row % nl].q;
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    size_t o = (size_t)row * n + x;
    const u64x2 z = make_ulonglong2(0, 0);
    u64x2 a = z, b = z;
    if (p < xparts) { if (lrx == 0) a = ld2(xs, o); else { const u64 t = xs[(size_t)row * (n >> lrx) + (x >> lrx)]; a = make_ulonglong2(t, t); } }
    if (p < yparts) { if (lry == 0) b = ld2(ys, o); else { const u64 t = ys[(size_t)row * (n >> lry) + (x >> lry)]; b = make_ulonglong2(t, t); } }
    if (sub) st2(out, o, sub_mod(a.x, b.x, q), sub_mod(a.y, b.y, q));
    else st2(out, o, add_mod(a.x, b.x, q), add_mod(a.y, b.y, q));
}
void launch_addsub_parts(Context* c, u64* out, const u64* x, int xparts, const u64* y, int yparts, int nl, bool sub, int lrx, int lry) {
    int parts = xparts > yparts ? xparts : yparts;
    KL(c); ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * nl * ((double)xparts / (1 << lrx) + (double)yparts / (1 << lry) + parts));
    launch_pdl(c, k_addsub_parts, grid_for2(c, parts * nl), 256, 0, out, x, xparts, y, yparts, c->d_mod, c->n, nl, sub ? 1 : 0, lrx, lry);
}

// compact plaintext rows [rows][n >> lrep] -> full rows [rows][n]
__global__ void k_expand_rep(u64* out, const u64* in, int n, int lrep) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    out[(size_t)blockIdx.y * n + x] = in[(size_t)blockIdx.y * (n >> lrep) + (x >> lrep)];
}
void launch_expand_rep(Context* c, u64* out, const u64* in, int rows, int lrep) {
    KL(c); ProfScope ps(c, KC_COPY, 8.0 * rows * c->n); launch_pdl(c, k_expand_rep, grid_for(c, rows), 256, 0, out, in, c->n, lrep);
}

// out[row] = sum_i in_i[row] * s_i[limb]  (EvalLinearWSum: one pass instead of cnt scalar products + cnt-1 additions)
#define WSUM_MAX_IN 16
#define WSUM_MAX_PAIRS 640
struct WsumArgs { int cnt, acc; const u64* in[WSUM_MAX_IN]; u64 pairs[WSUM_MAX_PAIRS][2]; };   // passed by value (<= 32 KB of parameters)
__global__ void k_wsum(u64* out, const __grid_constant__ WsumArgs A, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y, l = row % nl; const u64 q = mi[l].q;
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    const size_t o = (size_t)row * n + x;
    u64 r0 = 0, r1 = 0;
    if (A.acc) { const u64x2 v = ld2(out, o); r0 = v.x; r1 = v.y; }
    for (int i = 0; i < A.cnt; i++) {
        const u64x2 v = ld2(A.in[i], o);
        const u64 w = A.pairs[i * nl + l][0], ws = A.pairs[i * nl + l][1];
        r0 = add_mod(r0, shoup_mul(v.x, w, ws, q), q);
        r1 = add_mod(r1, shoup_mul(v.y, w, ws, q), q);
    }
    st2(out, o, r0, r1);
}
void launch_wsum(Context* c, u64* out, const u64* const* ins, const u64* hs, int cnt, int parts, int nl) {
    int per = WSUM_MAX_PAIRS / nl; if (per > WSUM_MAX_IN) per = WSUM_MAX_IN;
    for (int i0 = 0; i0 < cnt; i0 += per) {
        const int k = cnt - i0 < per ? cnt - i0 : per;
        c->stats.kernel_launches++;
        if (c->dry) continue;
        ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * parts * nl * (k + 1 + (i0 ? 1 : 0)));
        WsumArgs A; A.cnt = k; A.acc = i0 ? 1 : 0;
        for (int i = 0; i < k; i++) {
            A.in[i] = ins[i0 + i];
            for (int l = 0; l < nl; l++) {
                const u64 w = hs[(size_t)(i0 + i) * nl + l];
                A.pairs[i * nl + l][0] = w; A.pairs[i * nl + l][1] = host_shoup(w, c->moduli[l]);
            }
        }
        launch_pdl(c, k_wsum, grid_for2(c, parts * nl), 256, 0, out, A, c->d_mod, c->n, nl);
    }
}

// G weighted sums over the SAME inputs (the leaves of one Paterson-Stockmeyer evaluation all combine the baby-step basis T_1..T_k-1,
// cheby.cu wsum_precompute): out_g[row] = sum_i in_i[row] * s_{g,i}[limb].  Every input word is loaded once per G outputs; on limbs
// of at most 44 bits the products are reduced on the FP64 pipe and the centred residues summed exactly in a double, the 60-bit rows
// keep Shoup products.  W = device table [G][cnt][nl] of (scalar, Shoup companion).  Same canonical residues as G k_wsum launches.
template <int G>
__global__ void __launch_bounds__(256) k_wsum_multi(const __grid_constant__ WsumMultiArgs A, const u64* __restrict__ W, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y, l = row % nl; const ModInfo m = mi[l];
    pdl_wait();
    const size_t o = (size_t)row * n + x;
    const int cnt = A.cnt;
    if (m.bits <= PW_FP_BITS) {
        const double q = (double)m.q, qinv = __drcp_rn(q);
        double s0[G], s1[G];
#pragma unroll
        for (int g = 0; g < G; g++) { s0[g] = 0.0; s1[g] = 0.0; }
#pragma unroll 4
        for (int i = 0; i < cnt; i++) {
            const u64x2 v = ld2(A.in[i], o);
            const double a0 = pw_from_u64(v.x), a1 = pw_from_u64(v.y);
#pragma unroll
            for (int g = 0; g < G; g++) {
                const double w = pw_from_u64(__ldg(W + 2 * ((size_t)(g * cnt + i) * nl + l)));
                s0[g] = __dadd_rn(s0[g], pw_mulmod(a0, w, q, qinv));
                s1[g] = __dadd_rn(s1[g], pw_mulmod(a1, w, q, qinv));
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) st2(A.out[g], o, pw_canon(s0[g], q, qinv), pw_canon(s1[g], q, qinv));
        return;
    }
    u64 r0[G], r1[G];
#pragma unroll
    for (int g = 0; g < G; g++) { r0[g] = 0; r1[g] = 0; }
    for (int i = 0; i < cnt; i++) {
        const u64x2 v = ld2(A.in[i], o);
#pragma unroll
        for (int g = 0; g < G; g++) {
            const ulonglong2 w = __ldg(reinterpret_cast<const ulonglong2*>(W) + ((size_t)(g * cnt + i) * nl + l));
            r0[g] = add_mod(r0[g], shoup_mul(v.x, w.x, w.y, m.q), m.q);
            r1[g] = add_mod(r1[g], shoup_mul(v.y, w.x, w.y, m.q), m.q);
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) st2(A.out[g], o, r0[g], r1[g]);
}
// hs: host scalars [groups][cnt][nl] (0 for the inputs a group does not use)
void launch_wsum_multi(Context* c, u64* const* outs, int groups, const u64* const* ins, const u64* hs, int cnt, int parts, int nl) {
    KL(c);
    ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * parts * nl * (cnt + groups));
    const size_t nw = (size_t)groups * cnt * nl;
    std::vector<u64> pairs(2 * nw);
    for (int g = 0; g < groups; g++)
        for (int i = 0; i < cnt; i++)
            for (int l = 0; l < nl; l++) {
                const size_t k = (size_t)(g * cnt + i) * nl + l;
                pairs[2 * k] = hs[k]; pairs[2 * k + 1] = host_shoup(hs[k], c->moduli[l]);
            }
    u64* W = dev_alloc(c, 2 * nw);
    cudaMemcpyAsync(W, pairs.data(), pairs.size() * sizeof(u64), cudaMemcpyHostToDevice, c->stream);   // pageable source: staged before the call returns
    WsumMultiArgs A; A.cnt = cnt;
    for (int i = 0; i < cnt; i++) A.in[i] = ins[i];
    for (int g = 0; g < WSUM_GROUPS; g++) A.out[g] = g < groups ? outs[g] : nullptr;
    const dim3 grid = grid_for2(c, parts * nl);
    switch (groups) {
        case 1: launch_pdl(c, k_wsum_multi<1>, grid, 256, 0, A, W, c->d_mod, c->n, nl); break;
        case 2: launch_pdl(c, k_wsum_multi<2>, grid, 256, 0, A, W, c->d_mod, c->n, nl); break;
        case 3: launch_pdl(c, k_wsum_multi<3>, grid, 256, 0, A, W, c->d_mod, c->n, nl); break;
        default: launch_pdl(c, k_wsum_multi<4>, grid, 256, 0, A, W, c->d_mod, c->n, nl); break;
    }
    dev_free(c, W);
}

// out[p][l] = init[p][l] + sum_i ct_i[p][l] (.) pt_i[l]   -- the  T = EvalAdd(T, EvalMult(ct_i, pt_i))  chains of the
// sorters' masked baby-step sums (src/sort_algo.h:336-345, 567-577) in one pass: every product goes into a 128-bit
// accumulator (<= 16 terms of q^2 < q * 2^64 for q < 2^60), reduced once.  Modular addition and multiplication are
// exact, so the result equals the chain's bit for bit.
__global__ void k_ip_plain(u64* out, const __grid_constant__ IpArgs A, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y, l = row % nl; const ModInfo m = mi[l];
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    const size_t o = (size_t)row * n + x;
    const u64x2 v0 = ld2(A.init, o);
    u64 l0 = v0.x, h0 = 0, l1 = v0.y, h1 = 0;
    for (int i = 0; i < A.cnt; i++) {
        const u64x2 a = ld2(A.ct[i], o);
        u64x2 b;
        if (A.lrep[i] == 0) b = ld2(A.pt[i], (size_t)l * n + x);
        else { const u64 t = A.pt[i][(size_t)l * (n >> A.lrep[i]) + (x >> A.lrep[i])]; b = make_ulonglong2(t, t); }
        mac128(l0, h0, a.x, b.x); mac128(l1, h1, a.y, b.y);
    }
    const u64 t0 = mont_reduce(l0, h0, m.q, m.qinv), t1 = mont_reduce(l1, h1, m.q, m.qinv);   // T * 2^-64
    st2(out, o, mont_reduce(t0 * m.r2, __umul64hi(t0, m.r2), m.q, m.qinv), mont_reduce(t1 * m.r2, __umul64hi(t1, m.r2), m.q, m.qinv));
}
void launch_ip_plain(Context* c, u64* out, const IpArgs& A, int nl) {
    KL(c);
    double pt = 0; for (int i = 0; i < A.cnt; i++) pt += 1.0 / (1 << A.lrep[i]);
    ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * nl * (2.0 * (A.cnt + 2) + pt));
    launch_pdl(c, k_ip_plain, grid_for2(c, 2 * nl), 256, 0, out, A, c->d_mod, c->n, nl);
}

// G inner products over the same ciphertexts (evaluator.cu eval_inner_product_plain_multi): every ciphertext word is loaded once
// and multiplied into G 128-bit accumulators, each reduced exactly as k_ip_plain reduces its own.
template <int G>
__global__ void __launch_bounds__(256) k_ip_plain_multi(const __grid_constant__ IpMultiArgs A, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y, l = row % nl; const ModInfo m = mi[l];
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    const size_t o = (size_t)row * n + x;
    const u64x2 v0 = ld2(A.init, o);
    if (m.bits <= PW_FP_BITS) {
        // 40-bit limbs: every product is reduced on the FP64 pipe and the (at most 16) centred residues are summed exactly in a
        // double -- 15 FP64 instructions per term against ~30 integer ones for the 128-bit multiply-accumulate, which is what lets
        // four accumulators per pass stay ahead of the memory system
        const double q = (double)m.q, qinv = __drcp_rn(q);
        double s0[G], s1[G];
#pragma unroll
        for (int g = 0; g < G; g++) { s0[g] = pw_from_u64(v0.x); s1[g] = pw_from_u64(v0.y); }
#pragma unroll 4
        for (int i = 0; i < A.cnt; i++) {       // unrolled: the loads of four terms are in flight together
            const u64x2 a = ld2(A.ct[i], o);
            const double a0 = pw_from_u64(a.x), a1 = pw_from_u64(a.y);
#pragma unroll
            for (int g = 0; g < G; g++) {
                const int lr = A.lrep[g][i];
                double b0, b1;
                if (lr == 0) { const u64x2 b = ld2(A.pt[g][i], (size_t)l * n + x); b0 = pw_from_u64(b.x); b1 = pw_from_u64(b.y); }
                else { b0 = b1 = pw_from_u64(A.pt[g][i][(size_t)l * (n >> lr) + (x >> lr)]); }
                s0[g] = __dadd_rn(s0[g], pw_mulmod(a0, b0, q, qinv));
                s1[g] = __dadd_rn(s1[g], pw_mulmod(a1, b1, q, qinv));
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) st2(A.out[g], o, pw_canon(s0[g], q, qinv), pw_canon(s1[g], q, qinv));
        return;
    }
    u64 l0[G], h0[G], l1[G], h1[G];
#pragma unroll
    for (int g = 0; g < G; g++) { l0[g] = v0.x; h0[g] = 0; l1[g] = v0.y; h1[g] = 0; }
    for (int i = 0; i < A.cnt; i++) {
        const u64x2 a = ld2(A.ct[i], o);
#pragma unroll
        for (int g = 0; g < G; g++) {
            u64x2 b;
            const int lr = A.lrep[g][i];
            if (lr == 0) b = ld2(A.pt[g][i], (size_t)l * n + x);
            else { const u64 t = A.pt[g][i][(size_t)l * (n >> lr) + (x >> lr)]; b = make_ulonglong2(t, t); }
            mac128(l0[g], h0[g], a.x, b.x); mac128(l1[g], h1[g], a.y, b.y);
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) {
        const u64 t0 = mont_reduce(l0[g], h0[g], m.q, m.qinv), t1 = mont_reduce(l1[g], h1[g], m.q, m.qinv);
        st2(A.out[g], o, mont_reduce(t0 * m.r2, __umul64hi(t0, m.r2), m.q, m.qinv), mont_reduce(t1 * m.r2, __umul64hi(t1, m.r2), m.q, m.qinv));
    }
}
void launch_ip_plain_multi(Context* c, const IpMultiArgs& A, int nl) {
    KL(c);
    double pt = 0;
    for (int g = 0; g < A.groups; g++) for (int i = 0; i < A.cnt; i++) pt += 1.0 / (1 << A.lrep[g][i]);
    ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * nl * (2.0 * (A.cnt + 1 + A.groups) + pt));
    const dim3 grid = grid_for2(c, 2 * nl);
    switch (A.groups) {
        case 1: launch_pdl(c, k_ip_plain_multi<1>, grid, 256, 0, A, c->d_mod, c->n, nl); break;
        case 2: launch_pdl(c, k_ip_plain_multi<2>, grid, 256, 0, A, c->d_mod, c->n, nl); break;
        case 3: launch_pdl(c, k_ip_plain_multi<3>, grid, 256, 0, A, c->d_mod, c->n, nl); break;
        default: launch_pdl(c, k_ip_plain_multi<4>, grid, 256, 0, A, c->d_mod, c->n, nl); break;
    }
}

// d[row][x] <- d[row][x] mod q_limb for words < 2^64: folds a multi-GPU integer sum of residues (NCCL all-reduce of the raw
// limbs of partial ciphertexts, at most 8 addends < 2^61) back to canonical residues
__global__ void k_mod_reduce(u64* d, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y; const ModInfo m = mi[row % nl];
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    const size_t o = (size_t)row * n + x;
    u64x2 v = ld2(d, o);
    v.x -= __umul64hi(v.x, m.qrecip) * m.q; if (v.x >= m.q) v.x -= m.q;
    v.y -= __umul64hi(v.y, m.qrecip) * m.q; if (v.y >= m.q) v.y -= m.q;
    st2(d, o, v.x, v.y);
}
void launch_mod_reduce(Context* c, u64* d, int rows, int nl) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 16.0 * c->n * rows);
    launch_pdl(c, k_mod_reduce, grid_for2(c, rows), 256, 0, d, c->d_mod, c->n, nl);
}

// out[row] = sum_i in_i[row]  (EvalAddMany)
struct ManyArgs { int cnt; const u64* in[32]; };
__global__ void k_add_many(u64* out, const ManyArgs A, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    const int row = blockIdx.y; const u64 q = mi[row % nl].q;
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    const size_t o = (size_t)row * n + x;
    u64x2 r = ld2(A.in[0], o);
    for (int i = 1; i < A.cnt; i++) { const u64x2 v = ld2(A.in[i], o); r.x = add_mod(r.x, v.x, q); r.y = add_mod(r.y, v.y, q); }
    st2(out, o, r.x, r.y);
}
void launch_add_many(Context* c, u64* out, const u64* const* ins, int cnt, int rows, int nl) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 8.0 * c->n * rows * (cnt + 1));
    ManyArgs A; A.cnt = cnt;
    for (int i = 0; i < cnt; i++) A.in[i] = ins[i];
    launch_pdl(c, k_add_many, grid_for2(c, rows), 256, 0, out, A, c->d_mod, c->n, nl);
}

// (a0,a1) x (b0,b1) -> (a0b0, a0b1 + a1b0, a1b1), one pass over the four inputs
__global__ void k_tensor(u64* d3, const u64* a, const u64* b, const ModInfo* mi, int n, int nl) {
    pdl_trigger();
    X2_PROLOGUE;
    int l = blockIdx.y; const ModInfo m = mi[l];
    pdl_wait();   // the modulus entry above comes from a context table: fetched while the predecessor drains
    size_t o = (size_t)l * n + x, P = (size_t)nl * n;
    const u64x2 a0 = ld2(a, o), a1 = ld2(a, o + P), b0 = ld2(b, o), b1 = ld2(b, o + P);
    if (m.bits <= PW_FP_BITS) {
        const double q = (double)m.q, qinv = __drcp_rn(q);
        const double x0 = pw_from_u64(a0.x), x1 = pw_from_u64(a1.x), y0 = pw_from_u64(b0.x), y1 = pw_from_u64(b1.x);
        const double u0 = pw_from_u64(a0.y), u1 = pw_from_u64(a1.y), v0 = pw_from_u64(b0.y), v1 = pw_from_u64(b1.y);
        st2(d3, o, pw_canon_reduced(pw_mulmod(x0, y0, q, qinv), q), pw_canon_reduced(pw_mulmod(u0, v0, q, qinv), q));
        st2(d3, o + P, pw_canon(__dadd_rn(pw_mulmod(x0, y1, q, qinv), pw_mulmod(x1, y0, q, qinv)), q, qinv),
            pw_canon(__dadd_rn(pw_mulmod(u0, v1, q, qinv), pw_mulmod(u1, v0, q, qinv)), q, qinv));
        st2(d3, o + 2 * P, pw_canon_reduced(pw_mulmod(x1, y1, q, qinv), q), pw_canon_reduced(pw_mulmod(u1, v1, q, qinv), q));
        return;
    }
    st2(d3, o, barrett_mul(a0.x, b0.x, m), barrett_mul(a0.y, b0.y, m));
    st2(d3, o + P, add_mod(barrett_mul(a0.x, b1.x, m), barrett_mul(a1.x, b0.x, m), m.q),
        add_mod(barrett_mul(a0.y, b1.y, m), barrett_mul(a1.y, b0.y, m), m.q));
    st2(d3, o + 2 * P, barrett_mul(a1.x, b1.x, m), barrett_mul(a1.y, b1.y, m));
}
void launch_tensor(Context* c, u64* d3, const u64* a, const u64* b, int nl) {
    KL(c); ProfScope ps(c, KC_POINTWISE, 56.0 * nl * c->n); launch_pdl(c, k_tensor, grid_for2(c, nl), 256, 0, d3, a, b, c->d_mod, c->n, nl);
}

// out[p] = in[brev(idx)], j = brev(p), idx = ((2j+1)k mod 2n) >> 1  (PrecomputeAutoMap)
__global__ void k_automorph(u64* out, const u64* in, int n, int logn, u32 k) {
    pdl_sync();
    int p = blockIdx.x * blockDim.x + threadIdx.x; if (p >= n) return;
    size_t ro = (size_t)blockIdx.y * n;
    u32 j = __brev((u32)p) >> (32 - logn);
    u64 jt = 2ull * j + 1;
    u32 idx = (u32)(((jt * k) & (2ull * n - 1)) >> 1);
    u32 src = __brev(idx) >> (32 - logn);
    out[ro + p] = in[ro + src];
}
void launch_automorph(Context* c, u64* out, const u64* in, int rows, uint32_t k) {
    KL(c); ProfScope ps(c, KC_AUTOMORPH, 16.0 * rows * c->n); launch_pdl(c, k_automorph, grid_for(c, rows), 256, 0, out, in, c->n, c->log_n, k);
}

static u64 host_powmod(u64 a, u64 e, u64 q) {
    unsigned __int128 r = 1, b = a % q;
    while (e) { if (e & 1) r = r * b % q; b = b * b % q; e >>= 1; }
    return (u64)r;
}
// DropLastElementAndScale of `parts` polynomials with nl_in limbs each -> out [parts][nl_in-1][n]:
// INTT of the last limbs, then ONE forward NTT launch over all remaining limbs whose prologue switches the last
// limb's coefficients to the row modulus and whose epilogue forms (c_i - that) * q_last^-1.
// Tables of an integer scalar s (two's complement in s_hi:s_lo) for the scaled rescale: [0][M] pairs (s mod q_i, Shoup), [1][M] pairs
// ((n >> lrep)^-1 s mod q_i, Shoup).  The scalars of FLEXIBLEAUTOEXT's level adjustments are functions of the scaling factors
// alone, so a sorter meets the same few dozen values in every sort: the tables are cached per context (bounded).
static const u64* scaled_tables(Context* c, u64 s_lo, u64 s_hi, int lrep) {
    const auto key = std::make_tuple(s_lo, s_hi, lrep);
    auto it = c->scal_tabs.find(key);
    if (it != c->scal_tabs.end()) return it->second;
    if (c->scal_tabs.size() >= 4096) return nullptr;
    const int M = c->M();
    std::vector<u64> v(4 * (size_t)M);
    const __int128 sv = (__int128)(((unsigned __int128)s_hi << 64) | s_lo);
    for (int i = 0; i < M; i++) {
        const u64 q = c->moduli[i];
        __int128 r = sv % (__int128)q; if (r < 0) r += q;
        const u64 si = (u64)r;
        const u64 ninv = (u64)((unsigned __int128)c->h_mod[i].ninv * ((1ull << lrep) % q) % q);
        const u64 w = (u64)((unsigned __int128)ninv * si % q);
        v[2 * i] = si; v[2 * i + 1] = host_shoup(si, q);
        v[2 * M + 2 * i] = w; v[2 * M + 2 * i + 1] = host_shoup(w, q);
    }
    u64* d = nullptr;
    if (!c->dry) {
        if (cudaMalloc(&d, v.size() * sizeof(u64)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        cudaMemcpy(d, v.data(), v.size() * sizeof(u64), cudaMemcpyHostToDevice);   // synchronous: visible to every stream of the context
    } else d = reinterpret_cast<u64*>(8);   // dry contexts launch nothing; any non-null marker
    c->scal_tabs.emplace(key, d);
    return d;
}

static const u64* rescale_pairs_for(Context* c, int nl) {
    auto it = c->rescale_pairs.find(nl);
    if (it == c->rescale_pairs.end()) {
        const int nlo = nl - 1;
        std::vector<u64> v(2 * (size_t)c->M(), 0); u64 ql = c->moduli[nl - 1];
        for (int i = 0; i < nlo; i++) { u64 q = c->moduli[i]; v[2 * i] = host_powmod(ql % q, q - 2, q); v[2 * i + 1] = host_shoup(v[2 * i], q); }
        u64* d = nullptr;
        if (!c->dry) { cudaMalloc(&d, v.size() * sizeof(u64)); cudaMemcpy(d, v.data(), v.size() * sizeof(u64), cudaMemcpyHostToDevice); }
        it = c->rescale_pairs.emplace(nl, d).first;
    }
    return it->second;
}

bool launch_rescale_scaled(Context* c, u64* out, const u64* in, int parts, int nl_in, int nl_keep, u64 s_lo, u64 s_hi, int lrep) {
    const u64* tabs = scaled_tables(c, s_lo, s_hi, lrep);
    if (!tabs) return false;
    const int nl = nl_in;
    const size_t n = c->n >> lrep;
    u64* last = dev_alloc(c, (size_t)parts * n);
    // coefficients of s * (last limb): the scalar rides on the inverse transform's final factor
    launch_ntt_ex(c, last, in, parts, mods_range(nl - 1, nl), true, c->dry ? nullptr : tabs + 2 * (size_t)c->M(), RowMap{1, nl, nl - 1, 1}, RowMap{}, nullptr, lrep);
    NttFuse f;
    f.pro = PRO_SWITCHMOD; f.pro_mod = nl - 1;
    f.epi = EPI_RESCALE; f.epi_in = in; f.emap = RowMap{nl_keep, nl, 0, 1}; f.epi_pairs = rescale_pairs_for(c, nl); f.epi_pre = tabs;
    {
        ProfScope ps(c, KC_RESCALE, 0.0, 0);   // the fused launch is accounted to the NTT class
    }
    launch_ntt_ex(c, out, last, parts * nl_keep, mods_q(nl_keep), false, nullptr, RowMap{nl_keep, 1, 0, 0}, RowMap{}, &f, lrep);
    dev_free(c, last);
    return true;
}

void launch_rescale(Context* c, u64* out, const u64* in, int parts, int nl_in, int lrep) {
    const int nl = nl_in, nlo = nl_in - 1;
    const size_t n = c->n >> lrep;
    u64* last = dev_alloc(c, (size_t)parts * n);
    launch_ntt_ex(c, last, in, parts, mods_range(nl - 1, nl), true, nullptr, RowMap{1, nl, nl - 1, 1}, RowMap{}, nullptr, lrep);
    NttFuse f;
    f.pro = PRO_SWITCHMOD; f.pro_mod = nl - 1;
    f.epi = EPI_RESCALE; f.epi_in = in; f.emap = RowMap{nlo, nl, 0, 1}; f.epi_pairs = rescale_pairs_for(c, nl);
    c->stats.kernel_launches += 0;
    {
        ProfScope ps(c, KC_RESCALE, 0.0, 0);   // the fused launch is accounted to the NTT class
    }
    launch_ntt_ex(c, out, last, parts * nlo, mods_q(nlo), false, nullptr, RowMap{nlo, 1, 0, 0}, RowMap{}, &f, lrep);
    dev_free(c, last);
}

void launch_copy_rows(Context* c, u64* dst, const u64* src, size_t words) {
    KL(c); ProfScope ps(c, KC_COPY, 16.0 * words); cudaMemcpyAsync(dst, src, words * sizeof(u64), cudaMemcpyDeviceToDevice, c->stream);
}

__global__ void k_mont(u64* d, LimbMods mods, const ModInfo* mi, int n, int to) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y; const ModInfo m = mi[mods.m[row % mods.n]];
    size_t o = (size_t)row * n + x;
    if (to) { u64 v = d[o]; u64 lo = v * m.r2, hi = __umul64hi(v, m.r2); d[o] = mont_reduce(lo, hi, m.q, m.qinv); }
    else d[o] = mont_reduce(d[o], 0, m.q, m.qinv);
}
void launch_to_mont(Context* c, u64* data, int rows, const LimbMods& m) { KL(c); launch_pdl(c, k_mont, grid_for(c, rows), 256, 0, data, m, c->d_mod, c->n, 1); }
void launch_from_mont(Context* c, u64* data, int rows, const LimbMods& m) { KL(c); launch_pdl(c, k_mont, grid_for(c, rows), 256, 0, data, m, c->d_mod, c->n, 0); }

// ---- counter-based sampling (spec shared with oracle/ckks_prims.c orc_hash / orc_sample) ----
__device__ __forceinline__ u64 mix64(u64 x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31; return x;
}
__device__ __forceinline__ u64 hash3(u64 seed, u64 stream, u64 idx) {
    return mix64(mix64(seed + 0x9E3779B97F4A7C15ULL * (stream + 1)) ^ mix64(idx + 0xD1B54A32D192ED03ULL));
}
// production mode (Context::rng_secure): one ChaCha20 block per sample, keyed with 256 bits of OS entropy
__device__ __forceinline__ void rng_words(const RngKey& k, u64 stream, u64 idx, u64& w0, u64& w1) {
    if (!k.secure) { w0 = hash3(k.seed, stream, idx); w1 = 0; return; }
    uint32_t b[16];
    chacha20_block(k.k, stream, idx, b);
    w0 = (u64)b[0] | ((u64)b[1] << 32); w1 = (u64)b[2] | ((u64)b[3] << 32);
}
static RngKey rng_key_of(const Context* c) {
    RngKey k; for (int i = 0; i < 8; i++) k.k[i] = c->rng_key[i];
    k.seed = c->seed; k.secure = c->rng_secure ? 1 : 0;
    return k;
}
// kind 0: uniform ternary secret / encryption mask; kind 1: error, centred binomial of 2 x 20 bits (variance 10, sigma 3.16 --
// the stand-in for OpenFHE's discrete Gaussian of sigma 3.19)
__global__ void k_sample_small(u64* out, LimbMods mods, const ModInfo* mi, int n, const __grid_constant__ RngKey key, u64 stream, int kind) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y; u64 q = mi[mods.m[row % mods.n]].q;
    u64 h, h1, r;
    rng_words(key, stream, (u64)x, h, h1);
    if (kind == 0) { u64 t = key.secure ? (u64)((((unsigned __int128)h1 << 64) | h) % 3) : h % 3; r = t == 0 ? 0 : (t == 1 ? 1 : q - 1); }
    else { int e = __popcll(h & 0xFFFFF) - __popcll((h >> 20) & 0xFFFFF); r = e >= 0 ? (u64)e : q - (u64)(-e); }
    out[(size_t)row * n + x] = r;
}
void launch_sample_small(Context* c, u64* out, int rows, const LimbMods& m, uint64_t stream, int kind) {
    KL(c); launch_pdl(c, k_sample_small, grid_for(c, rows), 256, 0, out, m, c->d_mod, c->n, rng_key_of(c), stream, kind);
}
__global__ void k_sample_uniform(u64* out, LimbMods mods, const ModInfo* mi, int n, const __grid_constant__ RngKey key, u64 stream) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y, mid = mods.m[row % mods.n]; u64 q = mi[mid].q;
    u64 h, h1;
    rng_words(key, stream, (((u64)mid + 1) << 32) + (u64)x, h, h1);
    // test mode keeps the 64-bit residue the oracle shares; production reduces 128 bits (bias 2^-64 instead of up to 2^-4 for 60-bit q)
    out[(size_t)row * n + x] = key.secure ? (u64)((((unsigned __int128)h1 << 64) | h) % q) : h % q;
}
void launch_sample_uniform(Context* c, u64* out, int rows, const LimbMods& m, uint64_t stream) {
    KL(c); launch_pdl(c, k_sample_uniform, grid_for(c, rows), 256, 0, out, m, c->d_mod, c->n, rng_key_of(c), stream);
}

// evk part: b = -a*s_new + e (+ P*s_old on the digit's own limbs [lo,hi)); all [M][n] EVALUATION
__global__ void k_evk_combine(u64* b, const u64* a, const u64* e, const u64* sn, const u64* so, const ModInfo* mi, int n,
                              int lo, int hi, LimbScalars pm) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int row = blockIdx.y; const ModInfo m = mi[row];
    size_t o = (size_t)row * n + x;
    u64 r = sub_mod(e[o], barrett_mul(a[o], sn[o], m), m.q);
    if (row >= lo && row < hi) r = add_mod(r, shoup_mul(so[o], pm.v[row], pm.sh[row], m.q), m.q);
    b[o] = r;
}
void launch_evk_combine(Context* c, u64* b, const u64* a, const u64* e, const u64* s_new, const u64* s_old, int lo, int hi) {
    KL(c);
    LimbScalars pm;
    for (int i = 0; i < c->M(); i++) {
        unsigned __int128 p = 1; u64 q = c->moduli[i];
        for (int k = 0; k < c->size_p; k++) p = p * (c->moduli[c->size_q + k] % q) % q;
        pm.v[i] = (u64)p; pm.sh[i] = host_shoup((u64)p, q);
    }
    launch_pdl(c, k_evk_combine, grid_for(c, c->M()), 256, 0, b, a, e, s_new, s_old, c->d_mod, c->n, lo, hi, pm);
}

// Encrypt: c0 = b*v + e0 + m, c1 = a*v + e1 ; pk = [2][sizeQ][n]
__global__ void k_encrypt(u64* out, const u64* pt, const u64* pk, const u64* v, const u64* e0, const u64* e1,
                          const ModInfo* mi, int n, int nl, int size_q) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int l = blockIdx.y; const ModInfo m = mi[l];
    size_t o = (size_t)l * n + x;
    u64 vv = v[o];
    out[o] = add_mod(add_mod(barrett_mul(pk[o], vv, m), e0[o], m.q), pt[o], m.q);
    out[(size_t)nl * n + o] = add_mod(barrett_mul(pk[(size_t)size_q * n + o], vv, m), e1[o], m.q);
}
void launch_encrypt_combine(Context* c, u64* out, const u64* pt, const u64* v, const u64* e0, const u64* e1, int nl) {
    KL(c); launch_pdl(c, k_encrypt, grid_for(c, nl), 256, 0, out, pt, c->d_pk, v, e0, e1, c->d_mod, c->n, nl, c->size_q);
}

// Decrypt core on the first `use` limbs: out = c0 + c1*s (+ c2*s^2)
__global__ void k_decrypt(u64* out, const u64* d, const u64* s, const ModInfo* mi, int n, int nl, int parts) {
    pdl_sync();
    int x = blockIdx.x * blockDim.x + threadIdx.x; if (x >= n) return;
    int l = blockIdx.y; const ModInfo m = mi[l];
    size_t o = (size_t)l * n + x, P = (size_t)nl * n;
    u64 sv = s[o];
    u64 r = add_mod(d[o], barrett_mul(d[o + P], sv, m), m.q);
    if (parts == 3) r = add_mod(r, barrett_mul(barrett_mul(d[o + 2 * P], sv, m), sv, m), m.q);
    out[o] = r;
}
void launch_decrypt_combine(Context* c, u64* out, const Ct* ct, int use) {
    KL(c); launch_pdl(c, k_decrypt, grid_for(c, use), 256, 0, out, ct->d, c->d_secret, c->d_mod, c->n, ct->nl, ct->parts);
}
