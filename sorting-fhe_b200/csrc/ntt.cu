// ntt.cu -- batched negacyclic NTT / inverse NTT over all RNS limbs (kernel K1).
//
// Replaces OpenFHE ChineseRemainderTransformFTT::ForwardTransformToBitReverse /
// InverseTransformFromBitReverse, which every cc->EvalMult / EvalRotate / rescale of the
// reference reaches (src/sign.cpp:23-56, src/rotation.h:224, src/sort_algo.h:358).
// Semantics (pinned by oracle/ckks_prims.c orc_ntt_fwd/orc_ntt_inv): Cooley-Tukey, natural in ->
// bit-reversed out with table W[m+i] = psi^bitrev(m+i); inverse Gentleman-Sande with n^-1.
//
// Design (B200): the transform is bound by instruction issue on the 32-bit integer pipes (a 64-bit Shoup
// butterfly is ~25 SASS instructions), not by HBM, so the kernel is organised to spend instructions on
// butterflies only:
//   * a limb of n = 2^logn words is processed in <= 2 passes; a CTA owns a tile of 2^T words (T = min(11,
//     logn)); pass "A" (first logn-T stages) works on tiles of 2^a rows x 2^(T-a) contiguous columns, pass "B"
//     (last T stages) on contiguous chunks;
//   * every thread keeps 8 words in registers and runs up to 3 stages per round; the first round of a pass
//     loads straight from global memory (coalesced 256 B runs), the last one stores straight back, only the
//     rounds in between exchange through (XOR-swizzled, conflict-free) shared memory;
//   * butterflies are lazy (Harvey): no per-butterfly correction for moduli < 2^58 forward / < 2^46 inverse
//     (values may grow to (2s+1)q resp. 2^s q < 2^64); larger moduli use one conditional subtraction;
//   * twiddles are (w, floor(w 2^64 / q)) pairs fetched with one 16-byte load; rounds touch 1, 2 and 4
//     consecutive pairs per thread.
// Results are canonical residues, bit-identical to the oracle.
#include "engine.h"
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

#ifndef NTT_TB_MAX
#define NTT_TB_MAX 10
#endif
static constexpr int TB_MAX = NTT_TB_MAX;   // log2 of the largest tile (words) a CTA owns
typedef ulonglong2 tw_t;   // .x = w, .y = Shoup companion

// Shared-memory slot of tile-local element e: an XOR swizzle, linear over GF(2), that sends the 16 elements a half-warp
// touches in ANY round window (p = 0..7: element bits {6..3}, {0,6,5,4}, {1,0,6,5}, {2,1,0,6} or {3..0} vary) to 16
// different 8-byte bank pairs.  (The former padding e + (e >> 3) was 2-way conflicted for every window at p >= 4:
// ncu counted 786 432 excessive shared wavefronts per 32-limb pass, a third of the shared traffic.)
__device__ __forceinline__ int swz(int e) { return e ^ ((e >> 4) & 7) ^ (((0x96 >> (e & 7)) & 1) << 3); }
// slot of register k of thread u in the round at window p: swz is linear and the two index parts share no bits
__device__ __forceinline__ int slot_of(int u, int k, int p) {
    return swz(((u >> p) << (p + 3)) | (u & ((1 << p) - 1))) ^ swz(k << p);
}

// 256-bit global accesses (sm_100: LDG/STG.E.256).  In the round whose window sits at bit 0 a thread owns 8 consecutive
// words (64 bytes); eight 8-byte accesses at a 64-byte lane stride touch 32 sectors per request, a quarter filled each
// (ncu: 32 sectors and 16 wavefronts per store request) -- two 32-byte accesses write / read whole sectors.
__device__ __forceinline__ void ld256(const u64* p, u64& a, u64& b, u64& c, u64& d) {
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}
__device__ __forceinline__ void st256(u64* p, u64 a, u64 b, u64 c, u64 d) {
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}
__device__ __forceinline__ void load8(const u64* p, u64 (&v)[8]) { ld256(p, v[0], v[1], v[2], v[3]); ld256(p + 4, v[4], v[5], v[6], v[7]); }
__device__ __forceinline__ void store8(u64* p, const u64 (&v)[8]) { st256(p, v[0], v[1], v[2], v[3]); st256(p + 4, v[4], v[5], v[6], v[7]); }

// element index (tile-local) of register k of thread u in a round whose 3-bit window sits at bit p
__device__ __forceinline__ int local_index(int u, int k, int p) {
    return ((u >> p) << (p + 3)) | (k << p) | (u & ((1 << p) - 1));
}

// x * w mod q in [0, 2q) for any x < 2^64, computed as x * w + floor(x w' / 2^64) * (2^64 - q): with the negated modulus the low
// 64 bits are two chained IMAD.WIDE and four IMAD into the high word -- 6 instructions where `x * w - hi * q` compiles to 10 (two
// separate 64-bit products, a negation and a 64-bit add; profiles/r02_sass_ntt.txt).  Same value, bit for bit.
__device__ __forceinline__ u64 shoup_lazy_n(u64 x, u64 w, u64 wsh, u64 nq) {
    const u64 hi = __umul64hi(x, wsh);
    u64 T;
    asm("{\n\t.reg .u32 x0,x1,w0,w1,h0,h1,n0,n1,t0,t1;\n\t.reg .u64 a;\n\t"
        "mov.b64 {x0,x1}, %1;\n\tmov.b64 {w0,w1}, %2;\n\tmov.b64 {h0,h1}, %3;\n\tmov.b64 {n0,n1}, %4;\n\t"
        "mul.wide.u32 a, x0, w0;\n\t"
        "mad.wide.u32 a, h0, n0, a;\n\t"
        "mov.b64 {t0,t1}, a;\n\t"
        "mad.lo.u32 t1, x0, w1, t1;\n\t"
        "mad.lo.u32 t1, x1, w0, t1;\n\t"
        "mad.lo.u32 t1, h0, n1, t1;\n\t"
        "mad.lo.u32 t1, h1, n0, t1;\n\t"
        "mov.b64 %0, {t0,t1};\n\t}"
        : "=l"(T) : "l"(x), "l"(w), "l"(hi), "l"(nq));
    return T;
}

__device__ __forceinline__ void bf_fwd(u64& X, u64& Y, const tw_t w, const u64 nq, const u64 q2) {
    const u64 T = shoup_lazy_n(Y, w.x, w.y, nq);   // [0, 2q)
    const u64 x = X;
    X = x + T;
    Y = x - T + q2;
}
template <bool FIX>
__device__ __forceinline__ void bf_inv(u64& X, u64& Y, const tw_t w, const u64 nq, const u64 q2, const u64 cq) {
    const u64 x = X, y = Y;
    if (FIX) {                       // operands < 2q
        u64 s = x + y;
        if (s >= q2) s -= q2;
        X = s;
        Y = shoup_lazy_n(x - y + q2, w.x, w.y, nq);
    } else {                         // operands < cq = 2^t q (t = stages done so far), no correction
        X = x + y;
        Y = shoup_lazy_n(x - y + cq, w.x, w.y, nq);
    }
}

// One round on the 8 registers of a thread: butterfly stages j in [J0, J1) of the 3-bit window.
// s0 = global stage number of j = 0; ehi = all element-index bits above the window.
// Forward FIX rows (moduli of 59 / 60 bits, q < 2^60): a stage adds at most 2q to a value, so a round of <= 3 stages needs its inputs
// below 2^64 - 6q.  The fold at the head of a round is `r >= 2^63 ? r - 8q : r` -- a sign test on the high word instead of a 64-bit
// compare (r >= 2^63 implies r >= 8q; afterwards r < max(2^63, 2^64 - 8q), and both plus 6q stay below 2^64) -- and `nofold` skips it
// altogether in the first pass of a two-pass transform: canonical inputs grow to at most (2 * 7 + 1) q < 2^64 over its <= 7 stages.
template <bool INV, bool FIX, int J0, int J1>
__device__ __forceinline__ void round8(u64 (&r)[8], const tw_t* __restrict__ W, const u32 ehi, const int s0, const u64 q,
                                       int t_done, const bool nofold) {
    const u64 q2 = q << 1, nq = 0 - q;
    if (!INV) {
        if (FIX && !nofold) {
            const u64 q8 = q << 3;
#pragma unroll
            for (int k = 0; k < 8; k++) if ((long long)r[k] < 0) r[k] -= q8;
        }
#pragma unroll
        for (int j = J0; j < J1; j++) {
            const int bit = 4 >> j;
            const tw_t* Ws = W + ((1u << (s0 + j)) + (ehi << j));
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                bf_fwd(r[k], r[k | bit], __ldg(Ws + (k >> (3 - j))), nq, q2);
            }
        }
    } else {
#pragma unroll
        for (int j = J1 - 1; j >= J0; j--) {
            const int bit = 4 >> j;
            const tw_t* Ws = W + ((1u << (s0 + j)) + (ehi << j));
            const u64 cq = q << t_done;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                bf_inv<FIX>(r[k], r[k | bit], __ldg(Ws + (k >> (3 - j))), nq, q2, cq);
            }
            t_done++;
        }
    }
}

struct NttArgs {
    u64* dst;                // limb rows; logical row r lives at row_of(dmap, r)
    const u64* src;          // (== dst for in-place)
    RowMap smap, dmap;
    const ModInfo* mi;
    const tw_t* tw;          // [M][n] forward or inverse twiddle pairs
    const double* twd;       // [M][n] the same twiddles as doubles (rows of moduli <= FP_BITS bits), or nullptr
    const double* twb;       // [M][n/8] W[n/2 + 4e]: base of the last three stages' twiddles (full-size transforms), or nullptr
    const tw_t* scale;       // INV only: final factor (n^-1 [* extra]) as Shoup pair, indexed by modulus id
    int logn, T, a;          // T = tile bits, a = logn - T stages in pass A
    int twlog;               // log2 of the twiddle table row length (ring size; >= logn, see compact plaintexts)
    int pass;                // 0 = pass A (strided), 1 = pass B (contiguous)
    int final_pass;          // this pass produces the final output (canonicalise / scale)
    int first_pass;          // this pass reads the launch's input (prologue applies)
    NttFuse f;               // optional work fused into the first load / last store (engine.h)
};

// Per-row launch table, passed by value (constant bank): physical limb rows of the source / destination / epilogue input,
// modulus id, modulus, its reciprocal and which butterfly path the row takes.  Resolving the row maps on the host removes
// three integer divisions (~190 of the ~625 instructions a thread of a pass executed) and the dependent ModInfo load from
// the head of every CTA (ncu source page of round 2: prologue 225 instructions, 13 % of the stall samples on its first use).
#define NTT_MAX_ROWS 96
enum { ROW_FP = 0, ROW_INT = 1, ROW_INT_FIX = 2 };
struct NttRows {
    int src[NTT_MAX_ROWS], dst[NTT_MAX_ROWS], epi[NTT_MAX_ROWS];
    u64 q[NTT_MAX_ROWS];
    double qinv[NTT_MAX_ROWS];          // 1 / q, correctly rounded (== __drcp_rn)
    short mid[NTT_MAX_ROWS];
    short addrow[NTT_MAX_ROWS];         // EPI_MODDOWN: limb row inside the addend of this row's polynomial
    unsigned char kind[NTT_MAX_ROWS];
    unsigned char addsel[NTT_MAX_ROWS]; // EPI_MODDOWN: which addend (polynomial index)
};

// EVALUATION-format (bit-reversed) index that element e moves to under the automorphism whose INVERSE index is kinv
// (OpenFHE PrecomputeAutoMap gathers out[p] = in[F_k(p)]; F_kinv is the inverse permutation, so in[e] lands at F_kinv(e))
__device__ __forceinline__ size_t auto_dest(size_t e, u32 kinv, int logn) {
    const u32 j = __brev((u32)e) >> (32 - logn);
    const u32 idx = (u32)((((2ull * j + 1) * kinv) & ((2ull << logn) - 1)) >> 1);
    return (size_t)(__brev(idx) >> (32 - logn));
}

// Final store of a forward transform's canonical results v[8] (register k of thread u = tile-local element
// local_index(u, k, p)): the fused epilogue, then the store -- vectorised, scalar, or scattered through an automorphism.
template <bool FUSE, class GA>
__device__ __forceinline__ void epilogue_store(const NttArgs& A, const NttRows& R, const int row, const int logn, const bool vec8,
                                               const int u, const int p, const u64 q, const int mid, u64 (&v)[8], u64* dst, GA gaddr,
                                               const bool apply = true) {
    if (FUSE && apply && A.f.epi != EPI_NONE) {
        // EPI_RESCALE: out = (in - NTT(switched last limb)) * q_last^-1 (DropLastElementAndScale)
        // EPI_MODDOWN: out = (acc - NTT(converted P part)) * P^-1 [+ addend]    (ApproxModDown of a key switch)
        const u64* ein = A.f.epi_in + ((size_t)R.epi[row] << logn);
        const tw_t sc = reinterpret_cast<const tw_t*>(A.f.epi_pairs)[mid];
        u64 xin[8];
        if (vec8) load8(ein + gaddr(local_index(u, 0, 0)), xin);
        else {
#pragma unroll
            for (int k = 0; k < 8; k++) xin[k] = ein[gaddr(local_index(u, k, p))];
        }
        if (A.f.epi_pre) {   // scaled rescale: the minuend is s * in (the scalar product that preceded the rescale, never materialised)
            const tw_t ps = reinterpret_cast<const tw_t*>(A.f.epi_pre)[mid];
#pragma unroll
            for (int k = 0; k < 8; k++) xin[k] = shoup_mul(xin[k], ps.x, ps.y, q);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = shoup_mul(sub_mod(xin[k], v[k], q), sc.x, sc.y, q);
        if (A.f.epi == EPI_MODDOWN) {
            const u64* add = A.f.epi_add[R.addsel[row]];
            if (add) {
                add += (size_t)R.addrow[row] << logn;
                if (vec8) load8(add + gaddr(local_index(u, 0, 0)), xin);
                else {
#pragma unroll
                    for (int k = 0; k < 8; k++) xin[k] = add[gaddr(local_index(u, k, p))];
                }
#pragma unroll
                for (int k = 0; k < 8; k++) v[k] = add_mod(v[k], xin[k], q);
            }
        }
    }
    if (FUSE && A.f.auto_kinv) {
#pragma unroll
        for (int k = 0; k < 8; k++) dst[auto_dest(gaddr(local_index(u, k, p)), A.f.auto_kinv, logn)] = v[k];
    } else if (vec8) store8(dst + gaddr(local_index(u, 0, 0)), v);
    else {
#pragma unroll
        for (int k = 0; k < 8; k++) dst[gaddr(local_index(u, k, p))] = v[k];
    }
}

// rounds of one pass: window position p and stage range [J0,J1) per round (host fills, <= 4 rounds)
struct Sched { int nr; int p[4]; int j0[4]; int j1[4]; };

// rounds of a pass that butterflies `ns` consecutive tile-local bits starting at bit T-1 (the row bits of pass A are
// the top a bits of the tile index)
__host__ __device__ constexpr Sched make_sched(int T, int ns, bool inverse) {
    Sched s{};
    int ct = T - 1, rem = ns;
    bool first = true;
    while (rem > 0) {
        const int S = (first && rem % 3) ? rem % 3 : 3;           // a short round, if any, comes first: the last round always spans 3 stages
        int p = 0, j0 = 0, j1 = 0;
        if (S == 3) { p = ct - 2; j0 = 0; j1 = 3; }
        else if (first) { p = ct - 2; j0 = 0; j1 = S; }          // short first round: top S bits of the window
        else { p = ct - S + 1; j0 = 3 - S; j1 = 3; }             // short later round: low S bits of the window
        s.p[s.nr] = p; s.j0[s.nr] = j0; s.j1[s.nr] = j1; s.nr++;
        ct -= S; rem -= S; first = false;
    }
    if (inverse) {
        for (int i = 0; i < s.nr / 2; i++) {
            const int k = s.nr - 1 - i;
            int t = s.p[i]; s.p[i] = s.p[k]; s.p[k] = t;
            t = s.j0[i]; s.j0[i] = s.j0[k]; s.j0[k] = t;
            t = s.j1[i]; s.j1[i] = s.j1[k]; s.j1[k] = t;
        }
    }
    return s;
}

template <bool INV, bool FIX>
__device__ __forceinline__ void do_round(u64 (&r)[8], const tw_t* W, u32 ehi, int s0, u64 q, int t_done, int j0, int j1, bool nofold) {
    if (j0 == 0 && j1 == 3) round8<INV, FIX, 0, 3>(r, W, ehi, s0, q, t_done, nofold);
    else if (j0 == 1 && j1 == 3) round8<INV, FIX, 1, 3>(r, W, ehi, s0, q, t_done, nofold);
    else if (j0 == 2 && j1 == 3) round8<INV, FIX, 2, 3>(r, W, ehi, s0, q, t_done, nofold);
    else if (j0 == 0 && j1 == 2) round8<INV, FIX, 0, 2>(r, W, ehi, s0, q, t_done, nofold);
    else round8<INV, FIX, 0, 1>(r, W, ehi, s0, q, t_done, nofold);
}

// LOGN_C != 0: ring size, pass and round schedule are compile-time constants (all index arithmetic folds);
// LOGN_C == 0: generic runtime form.
template <bool INV, bool FIX, int LOGN_C, int PASS_C, bool FUSE>
__device__ __forceinline__ void ntt_tile(const NttArgs& A, const Sched& SR, const NttRows& R, const int row, const int tile,
                                         u64* sm, const u64 (&vin)[8]) {
    const int mid = R.mid[row];
    const ModInfo mi = A.mi[mid];
    constexpr int TC = LOGN_C < TB_MAX ? LOGN_C : TB_MAX;
    constexpr Sched SC = make_sched(TC ? TC : 3, LOGN_C ? (PASS_C == 0 ? LOGN_C - TC : TC) : 3, INV);
    const Sched& S = LOGN_C ? SC : SR;
    const u64 q = mi.q;
    const int logn = LOGN_C ? LOGN_C : A.logn, T = LOGN_C ? TC : A.T, a_ = LOGN_C ? LOGN_C - TC : A.a;
    const int cb = T - a_;                                // cb: column bits of a pass-A tile
    const tw_t* W = A.tw + ((size_t)mid << A.twlog);
    const u64* src = A.src + ((size_t)R.src[row] << logn);
    u64* dst = A.dst + ((size_t)R.dst[row] << logn);
    const int u = threadIdx.x;
    const bool strided = LOGN_C ? (PASS_C == 0) : (A.pass == 0);
    // global word offset of tile-local index L
    auto gaddr = [&](int L) -> size_t {
        return strided ? (((size_t)(L >> cb) << T) | (size_t)((tile << cb) | (L & ((1 << cb) - 1)))) : (((size_t)tile << T) | (size_t)L);
    };
    // position of tile-local bit 0 in the element index, and the index bits above the tile
    // pass B: E = tile << T | L.   pass A: E = (row bits = L >> cb) << T | (tile << cb | col): stages only touch
    // the row bits, whose position in E is T + (bit - cb).
    u64 r[8];
    int t_done = 0;   // inverse: number of GS stages already applied to the data (for the lazy offset)
    if (INV) t_done = strided ? T : 0;
#pragma unroll
    for (int ri = 0; ri < (LOGN_C ? SC.nr : SR.nr); ri++) {
        const int p = S.p[ri];
        if (ri == 0) {
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = vin[k];
            if (FUSE && !INV && A.first_pass && A.f.pro == PRO_SWITCHMOD) {
                // rescale: the source limb holds coefficients mod q_last; centred switch to this row's modulus
                const u64 ql = A.mi[A.f.pro_mod].q, half = ql >> 1;
                if (ql <= q) {      // a 40-bit limb dropped under the 60-bit q0: |centred value| < q, nothing to reduce
#pragma unroll
                    for (int k = 0; k < 8; k++) r[k] = r[k] > half ? q - (ql - r[k]) : r[k];
                } else {
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        const bool neg = r[k] > half;
                        u64 a = neg ? ql - r[k] : r[k];
                        a -= __umul64hi(a, mi.qrecip) * q;
                        if (a >= q) a -= q;
                        r[k] = neg ? (a ? q - a : 0) : a;
                    }
                }
            }
        } else {
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = sm[slot_of(u, k, p)];
        }
        // element-index bits above the window, and the global stage of window bit 2 (j = 0)
        const int L0 = local_index(u, 0, p);
        u32 ehi; int s0;
        if (strided) { ehi = (u32)(L0 >> (p + 3)); s0 = a_ - 1 - (p + 2 - cb); }
        else { ehi = (u32)(((size_t)tile << T | (size_t)L0) >> (p + 3)); s0 = logn - 1 - (p + 2); }
        do_round<INV, FIX>(r, W, ehi, s0, q, t_done, S.j0[ri], S.j1[ri], strided && A.first_pass && a_ <= 7);
        if (INV) t_done += S.j1[ri] - S.j0[ri];
        if (ri + 1 < (LOGN_C ? SC.nr : SR.nr)) {
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 8; k++) sm[slot_of(u, k, p)] = r[k];
        } else {
            if (A.final_pass) {
                if (!INV) {
                    if (FIX) {
                        // q >= 2^58: floor(r / q) < 64 is estimated from the high word alone, k = (hi32(r) * floor(2^90 / q)) >> 58 in
                        // {floor(r / q) - 1, floor(r / q)}, so r - k q lies in [0, 2q): one IMAD.HI and two multiply-adds with the negated
                        // modulus instead of a 64 x 64 high product and a 64-bit product (the host keeps floor(2^90 / q) in the row's
                        // otherwise unused FP64 reciprocal slot)
                        const u32 m32 = (u32)__double_as_longlong(R.qinv[row]);
                        const u64 nq = 0 - q;
#pragma unroll
                        for (int k = 0; k < 8; k++) {
                            const u32 kq = __umulhi((u32)(r[k] >> 32), m32) >> 26;
                            u64 v = r[k] + (u64)kq * (u32)nq;
                            v += (u64)(kq * (u32)(nq >> 32)) << 32;
                            r[k] = v >= q ? v - q : v;
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; k++) {   // any value < 2^64 -> [0, q)
                            u64 v = r[k] - __umul64hi(r[k], mi.qrecip) * q;
                            r[k] = v >= q ? v - q : v;
                        }
                    }
                    epilogue_store<FUSE>(A, R, row, logn, !strided && p == 0, u, p, q, mid, r, dst, gaddr);
                    continue;
                } else {
                    const tw_t sc = A.scale[mid];
#pragma unroll
                    for (int k = 0; k < 8; k++) { u64 v = shoup_lazy(r[k], sc.x, sc.y, q); r[k] = v >= q ? v - q : v; }
                }
            } else if (INV && FIX) {
                // between inverse passes the FIX path keeps values < 2q already
            }
            if (!strided && p == 0) store8(dst + gaddr(local_index(u, 0, 0)), r);
            else {
#pragma unroll
                for (int k = 0; k < 8; k++) dst[gaddr(local_index(u, k, p))] = r[k];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// FP64 butterflies for moduli of at most FP_BITS bits (the 40-bit scaling primes: 30 of the 32 Q limbs of the
// DirectSort card).  B200 issues DFMA at the same rate as IMAD (64 lanes/clk/SM), and a double carries a 53-bit
// product piece where an IMAD carries 32 bits, so x*w mod q costs 6 FP64 instructions instead of ~20 integer ones:
//     h = x*w (rounded), l = fma(x, w, -h) (exact low part), qh = rint(h / q), r = fma(-qh, q, h) + l
// Every intermediate is an integer below 2^53, hence exact; residues are centred, |r| <= q/2 + 1.  Values stay
// lazy (forward: +q/2 per stage; inverse: doubling, folded back every second round) and are canonicalised once on
// the final store, so the results are the same residues as the integer path, bit for bit.
#define FP_BITS 44
#ifndef NTT_FP_EPILOGUE
#define NTT_FP_EPILOGUE 1   /* 0: the integer epilogue on FP64 rows too (A/B builds) */
#endif
#define FP_MAGIC 6755399441055744.0   // 1.5 * 2^52: adding and subtracting it rounds to the nearest integer
#define FP_2P52 4503599627370496.0
__device__ __forceinline__ double fp_from_u64(u64 v) {   // v < 2^52
    return __dadd_rn(__longlong_as_double((long long)(v | 0x4330000000000000ULL)), -FP_2P52);
}
__device__ __forceinline__ u64 fp_to_u64(double y) {     // integer y in [0, 2^52)
    return (u64)__double_as_longlong(__dadd_rn(y, FP_2P52)) & 0x000FFFFFFFFFFFFFULL;
}
__device__ __forceinline__ double fp_mulmod(double a, double w, double q, double qinv) {
    const double h = __dmul_rn(a, w);
    const double l = __fma_rn(a, w, -h);
    const double qh = __dadd_rn(__fma_rn(h, qinv, FP_MAGIC), -FP_MAGIC);
    return __dadd_rn(__fma_rn(-qh, q, h), l);
}
__device__ __forceinline__ double fp_reduce(double x, double q, double qinv) {   // -> [-q/2 - 1, q/2 + 1]
    const double t = __dadd_rn(__fma_rn(x, qinv, FP_MAGIC), -FP_MAGIC);
    return __fma_rn(-t, q, x);
}
// canonical residue of an already reduced value, |y| < q (fp_mulmod returns |y| <= q/2 (1 + 2^-8) + 2^31 for lazy operands)
__device__ __forceinline__ u64 fp_canon_reduced(double y, double q) {
    if (y < 0.0) y = __dadd_rn(y, q);
    return fp_to_u64(y);
}
__device__ __forceinline__ u64 fp_canon(double x, double q, double qinv) {
    double y = fp_reduce(x, q, qinv);
    if (y < 0.0) y = __dadd_rn(y, q);
    return fp_to_u64(y);
}

template <bool INV, int J0, int J1>
__device__ __forceinline__ void round8_fp(double (&r)[8], const double* __restrict__ W, const u32 ehi, const int s0, const double q,
                                          const double qinv) {
    if (!INV) {
#pragma unroll
        for (int j = J0; j < J1; j++) {
            const int bit = 4 >> j;
            const double* Ws = W + ((1u << (s0 + j)) + (ehi << j));
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                const double T = fp_mulmod(r[k | bit], __ldg(Ws + (k >> (3 - j))), q, qinv);
                const double x = r[k];
                r[k] = __dadd_rn(x, T);
                r[k | bit] = __dadd_rn(x, -T);
            }
        }
    } else {
#pragma unroll
        for (int j = J1 - 1; j >= J0; j--) {
            const int bit = 4 >> j;
            const double* Ws = W + ((1u << (s0 + j)) + (ehi << j));
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                const double x = r[k], y = r[k | bit];
                r[k] = __dadd_rn(x, y);
                r[k | bit] = fp_mulmod(__dadd_rn(x, -y), __ldg(Ws + (k >> (3 - j))), q, qinv);
            }
        }
    }
}
// The last three stages of a transform (window at bit 0) use, per thread, the 1 + 2 + 4 twiddles
//   W[m/4 + e], W[m/2 + 2e + {0,1}], W[m + 4e + {0,1,2,3}]   (m = n/2, e = element index >> 3),
// 7/8 of all the twiddles a limb has.  With W[k] = psi^bitrev(k) they are all powers of b = W[m + 4e]:
//   W[m + 4e + c] = b * {1, W[1], W[2], W[3]}[c],   W[m/2 + 2e + c] = b^2 * {1, W[1]}[c],   W[m/4 + e] = b^4
// (W[1] = psi^(n/2) = sqrt(-1), W[2], W[3] the primitive 8th roots; the inverse table satisfies the same identities).
// So the round loads one double from the packed base table and spends six more multiplications, instead of
// streaming seven table entries per thread from L2/HBM.  The derived values are the same residues (centred instead of
// canonical), and the final canonicalisation makes the outputs identical.
template <bool INV>
__device__ __forceinline__ void round8_fp_ot(double (&r)[8], const double* __restrict__ W, const double b, const double q, const double qinv) {
    const double wi = __ldg(W + 1), we = __ldg(W + 2), we3 = __ldg(W + 3);
    const double b2 = fp_mulmod(b, b, q, qinv), b4 = fp_mulmod(b2, b2, q, qinv);
    const double t1[2] = {b2, fp_mulmod(b2, wi, q, qinv)};
    const double t2[4] = {b, fp_mulmod(b, wi, q, qinv), fp_mulmod(b, we, q, qinv), fp_mulmod(b, we3, q, qinv)};
    if (!INV) {
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int bit = 4 >> j;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                const double w = j == 0 ? b4 : j == 1 ? t1[k >> 2] : t2[k >> 1];
                const double T = fp_mulmod(r[k | bit], w, q, qinv);
                const double x = r[k];
                r[k] = __dadd_rn(x, T);
                r[k | bit] = __dadd_rn(x, -T);
            }
        }
    } else {
#pragma unroll
        for (int j = 2; j >= 0; j--) {
            const int bit = 4 >> j;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (k & bit) continue;
                const double w = j == 0 ? b4 : j == 1 ? t1[k >> 2] : t2[k >> 1];
                const double x = r[k], y = r[k | bit];
                r[k] = __dadd_rn(x, y);
                r[k | bit] = fp_mulmod(__dadd_rn(x, -y), w, q, qinv);
            }
        }
    }
}
template <bool INV>
__device__ __forceinline__ void do_round_fp(double (&r)[8], const double* W, u32 ehi, int s0, double q, double qinv, int j0, int j1) {
    if (j0 == 0 && j1 == 3) round8_fp<INV, 0, 3>(r, W, ehi, s0, q, qinv);
    else if (j0 == 1 && j1 == 3) round8_fp<INV, 1, 3>(r, W, ehi, s0, q, qinv);
    else if (j0 == 2 && j1 == 3) round8_fp<INV, 2, 3>(r, W, ehi, s0, q, qinv);
    else if (j0 == 0 && j1 == 2) round8_fp<INV, 0, 2>(r, W, ehi, s0, q, qinv);
    else round8_fp<INV, 0, 1>(r, W, ehi, s0, q, qinv);
}

// same tiling, passes and round schedule as ntt_tile; between the two passes of a transform the buffer holds the
// raw bits of the (lazy) doubles
template <bool INV, int LOGN_C, int PASS_C, bool FUSE>
__device__ __forceinline__ void ntt_tile_fp(const NttArgs& A, const Sched& SR, const NttRows& R, const int row, const int tile,
                                            u64* sm, const u64 (&vin)[8]) {
    const int mid = R.mid[row];
    constexpr int TC = LOGN_C < TB_MAX ? LOGN_C : TB_MAX;
    constexpr Sched SC = make_sched(TC ? TC : 3, LOGN_C ? (PASS_C == 0 ? LOGN_C - TC : TC) : 3, INV);
    const Sched& S = LOGN_C ? SC : SR;
    const u64 qi = R.q[row];
    const double q = (double)qi, qinv = R.qinv[row];
    const int logn = LOGN_C ? LOGN_C : A.logn, T = LOGN_C ? TC : A.T, a_ = LOGN_C ? LOGN_C - TC : A.a;
    const int cb = T - a_;
    const double* W = A.twd + ((size_t)mid << A.twlog);
    const u64* src = A.src + ((size_t)R.src[row] << logn);
    u64* dst = A.dst + ((size_t)R.dst[row] << logn);
    double* smd = reinterpret_cast<double*>(sm);
    const int u = threadIdx.x;
    const bool strided = LOGN_C ? (PASS_C == 0) : (A.pass == 0);
    auto gaddr = [&](int L) -> size_t {
        return strided ? (((size_t)(L >> cb) << T) | (size_t)((tile << cb) | (L & ((1 << cb) - 1)))) : (((size_t)tile << T) | (size_t)L);
    };
    double r[8];
    const int nr = LOGN_C ? SC.nr : SR.nr;
#pragma unroll
    for (int ri = 0; ri < (LOGN_C ? SC.nr : SR.nr); ri++) {
        const int p = S.p[ri];
        if (ri == 0) {
            u64 v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = vin[k];
            if (A.first_pass) {
                bool switched = false;
                if (FUSE && !INV && A.f.pro == PRO_SWITCHMOD) {   // rescale: centred switch of the q_last residues to this row's modulus
                    const u64 ql = A.mi[A.f.pro_mod].q, half = ql >> 1;
                    if (ql < 2 * qi && ql < (1ull << FP_BITS)) {
                        // the centred representative |c| <= q_last / 2 < q is a valid lazy operand as it is: the transform is linear
                        // modulo q and the final canonicalisation gives the residues that reducing c first would give -- one
                        // compare and one subtraction on the FP64 pipe instead of a 64-bit Barrett step and two negations per word
                        const double qld = (double)ql;
#pragma unroll
                        for (int k = 0; k < 8; k++) { const double c = fp_from_u64(v[k]); r[k] = v[k] > half ? __dadd_rn(c, -qld) : c; }
                        switched = true;
                    } else {
                        const u64 qrecip = A.mi[mid].qrecip;
#pragma unroll
                        for (int k = 0; k < 8; k++) {
                            const bool neg = v[k] > half;
                            u64 a = neg ? ql - v[k] : v[k];
                            a -= __umul64hi(a, qrecip) * qi;
                            if (a >= qi) a -= qi;
                            v[k] = neg ? (a ? qi - a : 0) : a;
                        }
                    }
                }
                if (!switched) {
#pragma unroll
                    for (int k = 0; k < 8; k++) r[k] = fp_from_u64(v[k]);
                }
            } else {
#pragma unroll
                for (int k = 0; k < 8; k++) r[k] = __longlong_as_double((long long)v[k]);
            }
        } else {
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = smd[slot_of(u, k, p)];
        }
        if (INV && (ri & 1) == 0 && !(ri == 0 && A.first_pass)) {   // fold the doubling of the GS sums back below q
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = fp_reduce(r[k], q, qinv);
        }
        const int L0 = local_index(u, 0, p);
        u32 ehi; int s0;
        if (strided) { ehi = (u32)(L0 >> (p + 3)); s0 = a_ - 1 - (p + 2 - cb); }
        else { ehi = (u32)(((size_t)tile << T | (size_t)L0) >> (p + 3)); s0 = logn - 1 - (p + 2); }
        if (!strided && p == 0 && S.j0[ri] == 0 && S.j1[ri] == 3 && A.twb && A.twlog == logn)
            round8_fp_ot<INV>(r, W, __ldg(A.twb + ((size_t)mid << (A.twlog - 3)) + ehi), q, qinv);
        else
            do_round_fp<INV>(r, W, ehi, s0, q, qinv, S.j0[ri], S.j1[ri]);
        if (ri + 1 < nr) {
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 8; k++) smd[slot_of(u, k, p)] = r[k];
        } else if (A.final_pass) {
            u64 v[8];
            if (!INV) {
                if (FUSE && A.f.epi != EPI_NONE && NTT_FP_EPILOGUE) {
                    // the fused epilogue on the FP64 pipe, straight from the lazy results: out = (in [* s] - r) * pair [+ addend] as exact
                    // floating-point modular products, canonicalised once -- instead of canonicalising r and then two (three) 64-bit
                    // Shoup products and conditional subtractions on the integer pipes; same residues
                    const bool vec8 = !strided && p == 0;
                    const u64* ein = A.f.epi_in + ((size_t)R.epi[row] << logn);
                    const double scd = (double)reinterpret_cast<const tw_t*>(A.f.epi_pairs)[mid].x;
                    u64 xin[8];
                    if (vec8) load8(ein + gaddr(local_index(u, 0, 0)), xin);
                    else {
#pragma unroll
                        for (int k = 0; k < 8; k++) xin[k] = ein[gaddr(local_index(u, k, p))];
                    }
                    double y[8];
#pragma unroll
                    for (int k = 0; k < 8; k++) y[k] = fp_from_u64(xin[k]);
                    if (A.f.epi_pre) {
                        const double psd = (double)reinterpret_cast<const tw_t*>(A.f.epi_pre)[mid].x;
#pragma unroll
                        for (int k = 0; k < 8; k++) y[k] = fp_mulmod(y[k], psd, q, qinv);
                    }
#pragma unroll
                    for (int k = 0; k < 8; k++) y[k] = fp_mulmod(__dadd_rn(y[k], -r[k]), scd, q, qinv);
                    const u64* add = A.f.epi == EPI_MODDOWN ? A.f.epi_add[R.addsel[row]] : nullptr;
                    if (add) {
                        add += (size_t)R.addrow[row] << logn;
                        if (vec8) load8(add + gaddr(local_index(u, 0, 0)), xin);
                        else {
#pragma unroll
                            for (int k = 0; k < 8; k++) xin[k] = add[gaddr(local_index(u, k, p))];
                        }
#pragma unroll
                        for (int k = 0; k < 8; k++) v[k] = fp_canon(__dadd_rn(y[k], fp_from_u64(xin[k])), q, qinv);
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; k++) v[k] = fp_canon_reduced(y[k], q);
                    }
                    epilogue_store<FUSE>(A, R, row, logn, vec8, u, p, qi, mid, v, dst, gaddr, false);
                    continue;
                }
#pragma unroll
                for (int k = 0; k < 8; k++) v[k] = fp_canon(r[k], q, qinv);
                epilogue_store<FUSE>(A, R, row, logn, !strided && p == 0, u, p, qi, mid, v, dst, gaddr, !NTT_FP_EPILOGUE);   // no epilogue left here
                continue;
            } else {
                const double sc = (double)A.scale[mid].x;
#pragma unroll
                for (int k = 0; k < 8; k++) v[k] = fp_canon_reduced(fp_mulmod(r[k], sc, q, qinv), q);   // the product is reduced already
            }
            if (!strided && p == 0) store8(dst + gaddr(local_index(u, 0, 0)), v);
            else {
#pragma unroll
                for (int k = 0; k < 8; k++) dst[gaddr(local_index(u, k, p))] = v[k];
            }
        } else {
            u64 v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = (u64)__double_as_longlong(r[k]);
            if (!strided && p == 0) store8(dst + gaddr(local_index(u, 0, 0)), v);
            else {
#pragma unroll
                for (int k = 0; k < 8; k++) dst[gaddr(local_index(u, k, p))] = v[k];
            }
        }
    }
}

// Pass B touches, per local stage sl, the contiguous run W[2^(a+sl) + tile 2^sl + [0, 2^sl)) of its row's twiddles exactly once:
// pull all of it (2^T - 1 entries, one 128-byte line per prefetch) into L1 ahead of the rounds.  The tables are written at context
// set-up, so the kernel issues these BEFORE griddepcontrol.wait, while its predecessor is still draining.
template <int LOGN_C, int PASS_C>
__device__ __forceinline__ void prefetch_twiddles(const NttArgs& A, const NttRows& R, const int row, const int tile) {
    constexpr int TC = LOGN_C < TB_MAX ? LOGN_C : TB_MAX;
    const int logn = LOGN_C ? LOGN_C : A.logn, T = LOGN_C ? TC : A.T, a_ = LOGN_C ? LOGN_C - TC : A.a;
    const bool strided = LOGN_C ? (PASS_C == 0) : (A.pass == 0);
    if (strided) return;
    const int mid = R.mid[row], u = threadIdx.x;
    if (R.kind[row] == ROW_FP) {
        const double* W = A.twd + ((size_t)mid << A.twlog);
        const bool ot = A.twb && A.twlog == logn && T >= 3;
        if (ot && u * 16 < (1 << (T - 3))) asm volatile("prefetch.global.L1 [%0];" ::"l"(A.twb + ((size_t)mid << (A.twlog - 3)) + ((size_t)tile << (T - 3)) + u * 16));
        for (int sl = 3; sl < (ot ? T - 3 : T); sl++) {
            const double* base = W + ((size_t)1 << (a_ + sl)) + ((size_t)tile << sl);
            for (int g = u * 16; g < (1 << sl); g += 16 * (1 << (T - 3)))
                asm volatile("prefetch.global.L1 [%0];" ::"l"(base + g));
        }
    } else {
        const tw_t* W = A.tw + ((size_t)mid << A.twlog);
        for (int sl = 3; sl < T; sl++) {
            const tw_t* base = W + ((size_t)1 << (a_ + sl)) + ((size_t)tile << sl);
            for (int g = u * 8; g < (1 << sl); g += 8 * (1 << (T - 3)))
                asm volatile("prefetch.global.L1 [%0];" ::"l"(base + g));
        }
    }
}

// first-round words of tile `tile` of limb row `row` (the same index pattern for the integer and the FP64 path)
template <bool INV, int LOGN_C, int PASS_C>
__device__ __forceinline__ void load_tile(const NttArgs& A, const Sched& SR, const NttRows& R, const int row, const int tile, u64 (&v)[8]) {
    constexpr int TC = LOGN_C < TB_MAX ? LOGN_C : TB_MAX;
    constexpr Sched SC = make_sched(TC ? TC : 3, LOGN_C ? (PASS_C == 0 ? LOGN_C - TC : TC) : 3, INV);
    const int logn = LOGN_C ? LOGN_C : A.logn, T = LOGN_C ? TC : A.T, a_ = LOGN_C ? LOGN_C - TC : A.a;
    const int cb = T - a_, p = LOGN_C ? SC.p[0] : SR.p[0];
    const bool strided = LOGN_C ? (PASS_C == 0) : (A.pass == 0);
    const u64* src = A.src + ((size_t)R.src[row] << logn);
    if (!strided && p == 0) load8(src + (((size_t)tile << T) | (size_t)(threadIdx.x * 8)), v);
    else {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int L = local_index(threadIdx.x, k, p);
            const size_t g = strided ? (((size_t)(L >> cb) << T) | (size_t)((tile << cb) | (L & ((1 << cb) - 1)))) : (((size_t)tile << T) | (size_t)L);
            v[k] = src[g];
        }
    }
}

// One CTA per work item id -> (row = id / tiles_per_row, tile = id % tiles_per_row).  (A persistent variant that
// prefetched the next tile's words into registers was measured slower on B200: 80 registers cut the resident CTAs
// from 4 to 3 per SM and 2048 tiles over 444 CTAs quantise badly -- 65 us vs 56 us for a 32-limb 2^17 transform.)
// KM: which butterfly paths the instantiation contains (bit 0 FP64, bit 1 integer, bit 2 integer with folds).  A launch whose rows
// are all of one kind (the special-prime rows of a ModUp at <= alpha limbs, the ModDown INTT, the rescale INTT of a 40-bit limb)
// runs a kernel a third of the size: the round-3 ncu source page of the mixed kernel has instruction fetch (no_instruction) as the
// top stall of sub-wave forward integer launches, 35-44 % of the warp samples, concentrated where the 97 KB kernel jumps to the
// integer path (profiles/r03_ks8_ntt_ncu_summary.txt).
template <bool INV, int LOGN_C, int PASS_C, int KM>
__global__ void __launch_bounds__((1 << TB_MAX) / 8 > 256 ? (1 << TB_MAX) / 8 : 256, TB_MAX >= 12 ? 2 : 4)
ntt_kernel(const __grid_constant__ NttArgs A, const __grid_constant__ Sched S, const __grid_constant__ NttRows R, int total,
           int tpr_log) {
    pdl_trigger();
    extern __shared__ u64 sm[];
    const int id = blockIdx.x;
    if (id >= total) return;
    const int row = id >> tpr_log, tile = id & ((1 << tpr_log) - 1);
    prefetch_twiddles<LOGN_C, PASS_C>(A, R, row, tile);
    constexpr bool FUSE = (KM & 8) != 0;   // bit 3: the launch carries a fused prologue / epilogue / automorphism (NttFuse)
    if (FUSE && !INV && A.f.epi != EPI_NONE && A.final_pass && (LOGN_C ? PASS_C == 1 : A.pass == 1)) {
        // the fused epilogue reads its minuend (and addend) tile only after the last round: pull the lines into L2 now (they were written
        // before this transform's first pass, so this too happens ahead of griddepcontrol.wait), the
        // ncu capture of round 2 had long-scoreboard on those loads as the top stall of the ModDown NTT (profiles/r02_ks_kernels_ncu_summary.txt)
        const int logn = LOGN_C ? LOGN_C : A.logn, T = LOGN_C ? (LOGN_C < TB_MAX ? LOGN_C : TB_MAX) : A.T;
        const int lines = (1 << T) / 16;                       // 128-byte lines of one tile
        if ((int)threadIdx.x < lines) {
            const size_t off = ((size_t)tile << T) + (size_t)threadIdx.x * 16;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(A.f.epi_in + ((size_t)R.epi[row] << logn) + off));
            if (A.f.epi == EPI_MODDOWN) {
                const u64* add = A.f.epi_add[R.addsel[row]];
                if (add) asm volatile("prefetch.global.L2 [%0];" ::"l"(add + ((size_t)R.addrow[row] << logn) + off));
            }
        }
    }
    pdl_wait();
    u64 v[8];
    load_tile<INV, LOGN_C, PASS_C>(A, S, R, row, tile, v);
    const int kind = R.kind[row];
    if ((KM & 7) == 1 || ((KM & 1) && kind == ROW_FP)) ntt_tile_fp<INV, LOGN_C, PASS_C, FUSE>(A, S, R, row, tile, sm, v);
    else if ((KM & 7) == 4 || ((KM & 4) && kind == ROW_INT_FIX)) ntt_tile<INV, true, LOGN_C, PASS_C, FUSE>(A, S, R, row, tile, sm, v);
    else if (KM & 2) ntt_tile<INV, false, LOGN_C, PASS_C, FUSE>(A, S, R, row, tile, sm, v);
}

template <bool INV, int LOGN_C, int PASS_C>
static void launch_one(Context* c, int total, int tpr_log, int threads, size_t smem, const NttArgs& A, const Sched& S, const NttRows& m) {
    static const bool mixed_only = getenv("CKKS_B200_NTT_MIXED_ONLY") != nullptr;   // A/B switch: always the three-path kernel
    int km = 0;
    const int rows = total >> tpr_log;
    for (int i = 0; i < rows; i++) km |= 1 << m.kind[i];
    // fused work this PASS executes: the prologue rides on the first pass, the epilogue / automorphism on the final one
    const bool fuse = !INV && ((A.first_pass && A.f.pro != PRO_NONE) || (A.final_pass && (A.f.epi != EPI_NONE || A.f.auto_kinv != 0)));
#define NTT_LAUNCH(KMV) launch_pdl(c, ntt_kernel<INV, LOGN_C, PASS_C, KMV>, total, threads, smem, A, S, m, total, tpr_log)
    if (LOGN_C != 17 || mixed_only) { NTT_LAUNCH(15); return; }
    if (INV) {          // inverse transforms never carry fused work
        if (km == 1) NTT_LAUNCH(1); else if (km == 4) NTT_LAUNCH(4); else if (km == 5) NTT_LAUNCH(5); else NTT_LAUNCH(7);
    } else if (fuse) {
        if (km == 1) NTT_LAUNCH(9); else if (km == 4) NTT_LAUNCH(12); else if (km == 5) NTT_LAUNCH(13); else NTT_LAUNCH(15);
    } else {
        if (km == 1) NTT_LAUNCH(1); else if (km == 4) NTT_LAUNCH(4); else if (km == 5) NTT_LAUNCH(5); else NTT_LAUNCH(7);
    }
#undef NTT_LAUNCH
}
template <bool INV, int PASS_C>
static void launch_pass(Context* c, int total, int tpr_log, int threads, size_t smem, const NttArgs& A, const Sched& S, const NttRows& m) {
    static const bool generic = getenv("CKKS_B200_NTT_GENERIC") != nullptr;   // A/B switch: the runtime-schedule (rolled-loop) form
    switch (generic ? 0 : A.logn) {   // compile-time schedules for the ring sizes of the reference's fixtures (+ their compact plaintexts)
        case 17: launch_one<INV, 17, PASS_C>(c, total, tpr_log, threads, smem, A, S, m); break;
        case 16: launch_one<INV, 16, PASS_C>(c, total, tpr_log, threads, smem, A, S, m); break;
        case 15: launch_one<INV, 15, PASS_C>(c, total, tpr_log, threads, smem, A, S, m); break;
        case 13: launch_one<INV, 13, PASS_C>(c, total, tpr_log, threads, smem, A, S, m); break;
        case 12: launch_one<INV, 12, PASS_C>(c, total, tpr_log, threads, smem, A, S, m); break;
        default: launch_one<INV, 0, 0>(c, total, tpr_log, threads, smem, A, S, m); break;
    }
}

static const u64* ninv_pairs_for(Context* c, int lrep) {
    if (lrep == 0) return c->d_ninv_pairs;
    auto it = c->ninv_pairs_rep.find(lrep);
    if (it != c->ninv_pairs_rep.end()) return it->second;
    const int M = c->M();
    std::vector<u64> v(2 * (size_t)M);
    for (int i = 0; i < M; i++) {   // (n >> lrep)^-1 = n^-1 * 2^lrep mod q
        const u64 q = c->moduli[i];
        const u64 w = (u64)((unsigned __int128)c->h_mod[i].ninv * ((1ull << lrep) % q) % q);
        v[2 * i] = w; v[2 * i + 1] = (u64)((((unsigned __int128)w) << 64) / q);
    }
    u64* d = nullptr;
    cudaMalloc(&d, v.size() * sizeof(u64));
    cudaMemcpy(d, v.data(), v.size() * sizeof(u64), cudaMemcpyHostToDevice);
    c->ninv_pairs_rep[lrep] = d;
    return d;
}

void launch_ntt_ex(Context* c, u64* dst, const u64* src, int rows, const LimbMods& m, bool inverse, const u64* scale_pairs,
                   RowMap smap, RowMap dmap, const NttFuse* fuse, int lrep) {
    const int logn = c->log_n - lrep, T = logn < TB_MAX ? logn : TB_MAX, a = logn - T;
    const int nchunks = (rows + NTT_MAX_ROWS - 1) / NTT_MAX_ROWS;
    c->stats.kernel_launches += ((logn > TB_MAX) ? 2 : 1) * nchunks;
    if (c->dry || rows == 0) return;
    ProfScope ps(c, KC_NTT, 16.0 * rows * (c->n >> lrep), ((logn > TB_MAX) ? 2 : 1) * nchunks);
    const int threads = (1 << T) / 8;
    const size_t smem = (size_t)((1 << T) < 16 ? 16 : (1 << T)) * sizeof(u64);
    const int tpr_log = logn - T;   // tiles per row (log2)
    NttArgs A;
    A.dst = dst; A.mi = c->d_mod; A.logn = logn; A.T = T; A.a = a; A.twlog = c->log_n;
    A.smap = smap; A.dmap = dmap;
    if (fuse) A.f = *fuse; else A.f = NttFuse{};
    A.tw = (const tw_t*)(inverse ? c->d_itw : c->d_tw);
    A.twd = inverse ? c->d_itwd : c->d_twd;
    A.twb = inverse ? c->d_itwb : c->d_twb;   // nullptr under CKKS_B200_NO_OT_TWIDDLES=1 (context.cu): last round's twiddles from the table
    A.scale = (const tw_t*)(scale_pairs ? scale_pairs : ninv_pairs_for(c, lrep));
    for (int r0 = 0; r0 < rows; r0 += NTT_MAX_ROWS) {
        const int cnt = rows - r0 < NTT_MAX_ROWS ? rows - r0 : NTT_MAX_ROWS, total = cnt << tpr_log;
        NttRows R1, R2;                                     // first pass: src map -> dst map; second pass: in place in dst
        // Launch order of the rows.  The FP64-butterfly rows (40-bit moduli) and the integer-butterfly rows (60-bit q0 and P
        // primes, ~2.5x the time of an FP64 row) use different pipes, but consecutive CTA ids walk one row at a time, so an SM
        // works on a single kind at any moment.  Interleaving the two kinds proportionally (CKKS_B200_NTT_INTERLEAVE=1) was
        // measured in round 2 and does NOT pay: key switch at l = 32 512.3 vs 510.1 us, whole sort 63.5 vs 62.8 ms end to end --
        // both kinds are bound by issue / dependency latency at ~55 % issue utilisation, not by their own pipe.  Row order stays.
        int order[NTT_MAX_ROWS];
        {
            static const bool no_mix = getenv("CKKS_B200_NTT_INTERLEAVE") == nullptr;
            int fp[NTT_MAX_ROWS], in_[NTT_MAX_ROWS], nfp = 0, nin = 0;
            for (int i = 0; i < cnt; i++) {
                const ModInfo& mi = c->h_mod[m.m[(r0 + i) % m.n]];
                if (mi.bits <= FP_BITS && A.twd) fp[nfp++] = i; else in_[nin++] = i;
            }
            if (no_mix || nfp == 0 || nin == 0) for (int i = 0; i < cnt; i++) order[i] = i;
            else {
                int a_ = 0, b_ = 0;
                for (int i = 0; i < cnt; i++) {       // Bresenham merge: after i rows, about i nin / cnt of them are integer rows
                    const bool take_int = b_ < nin && (a_ >= nfp || (long)(b_ + 1) * cnt <= (long)(i + 1) * nin);
                    order[i] = take_int ? in_[b_++] : fp[a_++];
                }
            }
        }
        for (int i = 0; i < cnt; i++) {
            const int r = r0 + order[i], mid = m.m[r % m.n];
            const ModInfo& mi = c->h_mod[mid];
            R1.src[i] = row_of(smap, r); R1.dst[i] = row_of(dmap, r); R1.epi[i] = fuse ? row_of(fuse->emap, r) : 0;
            R1.mid[i] = (short)mid; R1.q[i] = mi.q; R1.qinv[i] = 1.0 / (double)mi.q;
            R1.kind[i] = (mi.bits <= FP_BITS && A.twd) ? ROW_FP : ((inverse ? mi.bits > 46 : mi.bits > 58) ? ROW_INT_FIX : ROW_INT);
            if (R1.kind[i] == ROW_INT_FIX && !inverse) {   // floor(2^90 / q) < 2^32 for the final canonicalisation (ntt_tile), as raw bits
                const long long m32 = (long long)((((unsigned __int128)1) << 90) / mi.q);
                static_assert(sizeof(double) == sizeof(long long), "bit copy");
                memcpy(&R1.qinv[i], &m32, sizeof(double));
            }
            R1.addsel[i] = 0; R1.addrow[i] = 0;
            if (fuse && fuse->epi == EPI_MODDOWN) { R1.addsel[i] = (unsigned char)((r / fuse->epi_add_grp) & 1); R1.addrow[i] = (short)(r % fuse->epi_add_grp); }
        }
        R2 = R1;
        for (int i = 0; i < cnt; i++) R2.src[i] = R1.dst[i];
        A.first_pass = 1;
        if (!inverse) {
            A.src = src;
            if (a > 0) {
                A.pass = 0; A.final_pass = 0;
                if (fuse && fuse->first_in_place) {           // pass A in place in the (scratch) source, pass B source -> destination
                    NttRows R0 = R1;
                    for (int i = 0; i < cnt; i++) R0.dst[i] = R1.src[i];
                    A.dst = const_cast<u64*>(src);
                    launch_pass<false, 0>(c, total, tpr_log, threads, smem, A, make_sched(T, a, false), R0);
                    A.dst = dst; A.first_pass = 0;
                    A.pass = 1; A.final_pass = 1;
                    launch_pass<false, 1>(c, total, tpr_log, threads, smem, A, make_sched(T, T, false), R1);
                    continue;
                }
                launch_pass<false, 0>(c, total, tpr_log, threads, smem, A, make_sched(T, a, false), R1);
                A.src = dst; A.first_pass = 0;
            }
            A.pass = 1; A.final_pass = 1;
            launch_pass<false, 1>(c, total, tpr_log, threads, smem, A, make_sched(T, T, false), a > 0 ? R2 : R1);
        } else {
            A.src = src; A.pass = 1; A.final_pass = a == 0;
            launch_pass<true, 1>(c, total, tpr_log, threads, smem, A, make_sched(T, T, true), R1);
            if (a > 0) {
                A.src = dst; A.pass = 0; A.final_pass = 1; A.first_pass = 0;
                launch_pass<true, 0>(c, total, tpr_log, threads, smem, A, make_sched(T, a, true), R2);
            }
        }
    }
}

void launch_ntt(Context* c, u64* data, int rows, const LimbMods& m, bool inverse) {
    launch_ntt_ex(c, data, data, rows, m, inverse, nullptr, RowMap{}, RowMap{}, nullptr);
}
