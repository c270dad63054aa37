// engine.h -- internal types of the B200 CKKS engine (behind include/ckks_b200.h).
//
// Data layout in HBM: every polynomial bundle is limb-major uint64 [part][limb][n]; ciphertexts
// and keys stay in EVALUATION (bit-reversed negacyclic NTT) format, like OpenFHE's DCRTPoly.
// Modulus index space: 0..sizeQ-1 = Q chain, sizeQ..sizeQ+sizeP-1 = special primes P.
#pragma once
#include <chrono>
#include <stdexcept>
#include <string>
#include <cstdint>
#include <cstddef>
#include <deque>
#include <map>
#include <tuple>
#include <unordered_map>
#include <mutex>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "modarith.cuh"

#define MAX_LIMBS 96
#define MAX_DIGITS 4   // hybrid key switch digits (ConvBatch groups, DigitMap): validated in context_generate

struct LimbMods {          // passed by value to kernels: row r of a bundle uses modulus m[r % n]
    int n;
    short m[MAX_LIMBS];
};

// logical row r of a batched launch -> physical limb row: (r / grp) * stride + off + r % grp
struct RowMap { int grp = 1 << 30, stride = 0, off = 0, mul = 1; };
__host__ __device__ inline int row_of(const RowMap& m, int r) { return (r / m.grp) * m.stride + m.off + (r % m.grp) * m.mul; }

// work fused into an NTT launch: a prologue applied to the words as they are first loaded, an epilogue applied to
// the canonical results as they are stored
enum { PRO_NONE = 0, PRO_SWITCHMOD = 1 };
// EPI_RESCALE: out = (epi_in - NTT(..)) * pair.  EPI_MODDOWN (the ApproxModDown epilogue of a key switch, pair = P^-1 mod q_i):
// the same, plus an optional addend per polynomial, optionally stored through an automorphism.
enum { EPI_NONE = 0, EPI_RESCALE = 1, EPI_MODDOWN = 2 };
struct NttFuse {
    int pro = PRO_NONE, pro_mod = 0;      // PRO_SWITCHMOD: modulus id the source limb is reduced by
    int epi = EPI_NONE;
    const u64* epi_in = nullptr;          // minuend limbs, addressed through emap
    RowMap emap;
    const u64* epi_pairs = nullptr;       // [M] (factor, Shoup) pairs indexed by modulus id
    const u64* epi_pre = nullptr;         // EPI_RESCALE: optional [M] (s mod q_i, Shoup) pairs -- the minuend is multiplied by s first
                                          // (a scalar product folded into the rescale that follows it, evaluator.cu scaled_rescale_new)
    const u64* epi_add[2] = {nullptr, nullptr};   // EPI_MODDOWN: addend limbs [epi_add_grp][n] of logical rows r / epi_add_grp == 0, 1 (or nullptr)
    int epi_add_grp = 1 << 30;
    uint32_t auto_kinv = 0;               // != 0: result element e is stored at the EVALUATION-format image of e under X -> X^k (k kinv = 1 mod 2n)
    int first_in_place = 0;               // two-pass forward transform: run the first pass in place in src (scratch), the second src -> dst
};

struct ConvTables {        // ApproxSwitchCRTBasis tables for one (source basis -> target basis)
    int nin, nout;
    short in_ids[MAX_LIMBS], out_ids[MAX_LIMBS];
    u64* d_hatinv;         // [nin]   (Qhat_i^-1 mod q_i)
    u64* d_hatinv_sh;      // [nin]   Shoup companions
    u64* d_hatmod;         // [nin][nout]  (Qhat_i mod t_j) * 2^64 mod t_j  (Montgomery form)
    u64* d_hatplain;       // [nin][nout]  Qhat_i mod t_j (plain form: FP64 epilogue of k_conv_fp2 on targets of <= 42 bits)
    unsigned wide = 0;     // bit i: source modulus i has more than 42 bits
    bool split_ok = false; // the 32-bit-halves kernel's 64-bit column sums cannot overflow for this basis pair (keyswitch.cu)
};

struct KsTables {          // per number-of-Q-limbs tables of the hybrid key switch
    int nl, nparts;
    std::vector<ConvTables> modup;   // one per digit
    std::vector<int> dig_lo, dig_hi; // digit limb ranges [lo, hi)
    ConvTables moddown;              // P -> Q_l
    u64* d_pinv;                     // [nl] P^-1 mod q_i
    u64* d_pinv_sh;
    u64* d_pinv_pairs;               // [nl] interleaved (P^-1 mod q_i, Shoup) pairs: the fused ModDown epilogue of the NTT (EPI_MODDOWN)
    u64* d_modup_scale;              // [M] pairs: n^-1 * Qhat_i^-1 (digit of limb i) -- final factor of the ModUp INTT
    u64* d_moddown_scale;            // [M] pairs: n^-1 * Phat_k^-1 -- final factor of the ModDown INTT
};

struct EvalKey {           // hybrid evk: [part][2 (b, a)][sizeQ+sizeP][n], Montgomery form
    u64* d = nullptr;
    int nparts = 0;
};

// kernel classes for the per-class device timing bench.py reports (roofline object)
enum KClass { KC_NTT = 0, KC_BASECONV, KC_KSINNER, KC_MODDOWN, KC_POINTWISE, KC_RESCALE, KC_AUTOMORPH, KC_ENCODE, KC_COPY,
              KC_COUNT };
struct ProfRec { int cls; cudaEvent_t e0, e1; };
struct Prof {
    unsigned mask = 0;                       // bit per KClass; 0 = profiling off
    std::vector<ProfRec> pending;
    std::vector<cudaEvent_t> pool;
    double ms[KC_COUNT] = {0}, bytes[KC_COUNT] = {0};
    long launches[KC_COUNT] = {0};
};

struct Stats { std::map<std::pair<std::string, int>, long> ops; long kernel_launches = 0;
               long host_ns[4] = {0, 0, 0, 0}, host_calls[4] = {0, 0, 0, 0}; /* host time in launch / alloc / free / other API calls (CKKS_B200_HOST_PROF=1) */ };

struct Context {
    // parameters
    int log_n = 0, n = 0, mult_depth = 0, scaling_bits = 0, first_bits = 0, ext = 1;
    int size_q = 0, size_p = 0, alpha = 0, num_part_q = 0;
    uint64_t seed = 1;
    // seed == 0 at creation: production mode -- a 256-bit ChaCha20 key from getrandom(2) drives every sample (kernels.cu rng_*);
    // a non-zero seed selects the reproducible 64-bit counter hash the oracle shares (tests, bench; NOT secure)
    uint32_t rng_key[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool rng_secure = false;
    uint64_t uid = 0;      // process-unique id (ckks_ctx_info selector 8): lets caches keyed by the handle detect address reuse
    bool dry = false;      // metadata-only mode: no device allocations, no kernel launches
    bool enc_offload = true;   // plaintext level / scale adjustments run on the encode stream (CKKS_B200_NO_ENC_OFFLOAD=1 turns it off)
    bool host_prof = false;
    bool pdl = true;       // programmatic dependent launch on the engine's kernels (CKKS_B200_NO_PDL=1 turns it off)
    std::vector<u64> moduli, psi;           // Q then P
    std::vector<double> sf_real, sf_big;
    // device tables
    ModInfo* d_mod = nullptr;               // [M]
    std::vector<ModInfo> h_mod;
    u64 *d_tw = nullptr, *d_itw = nullptr;   // [M][n] (w, Shoup companion) pairs: psi / psi^-1 powers, bit-reversed
    double *d_twb = nullptr, *d_itwb = nullptr;   // [M][n/8] base twiddles W[n/2 + 4e] of the last stage (ntt.cu, on-the-fly twiddles)
    double *d_twd = nullptr, *d_itwd = nullptr;   // [M][n] the same twiddles as doubles (moduli <= 44 bits; FP64 butterflies)
    u64* d_ninv_pairs = nullptr;             // [M] (n^-1, Shoup companion): default final factor of the inverse NTT
    std::map<int, u64*> ninv_pairs_rep;      // lrep -> [M] ((n >> lrep)^-1, Shoup) for transforms of compact plaintexts
    double* d_ksi = nullptr;                // [2n+1] complex roots for the special FFT
    uint32_t* d_rot = nullptr;              // [n/2] 5^j mod 2n
    std::map<int, KsTables> ks;             // keyed by nl
    std::map<int, u64*> rescale_pairs;      // keyed by nl_in: device [M] (q_last^-1 mod q_i, Shoup) pairs
    // (integer scalar s, lrep) -> device [2][M] pairs: (s mod q_i, Shoup) and ((n >> lrep)^-1 s mod q_i, Shoup); kernels.cu scaled_tables
    std::map<std::tuple<u64, u64, int>, u64*> scal_tabs;
    // keys (device)
    u64* d_secret = nullptr;                // [M][n] EVALUATION
    u64* d_pk = nullptr;                    // [2][sizeQ][n]  (b, a)
    EvalKey relin;
    std::map<uint32_t, EvalKey> rot;        // automorphism index -> key
    uint64_t enc_counter = 0;
    double* enc_pinned = nullptr;           // pinned staging ring for MakeCKKSPackedPlaintext inputs
    std::vector<cudaEvent_t> enc_events;
    uint64_t enc_ring_pos = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    // plaintext encodes (small FFT + compact NTT: a few CTAs) run on their own high-priority stream so that they overlap
    // the main stream's full-width kernels; a consumer waits on the plaintext's `ready` event (evaluator.cu ct_wait)
    cudaStream_t enc_stream = nullptr;
    std::vector<cudaEvent_t> ev_pool;
    Stats stats;
    Prof prof;
    std::string last_error;
    // handle lifetime: ciphertexts may outlive ckks_ctx_destroy (shared_ptr / Python GC order); the context is
    // torn down when it is closed AND the last ciphertext handle is gone
    long live_cts = 0;
    // Caching arena over the stream-ordered pool (context.cu dev_alloc / dev_free): freed temporaries are kept in exact-size
    // bins PER STREAM and handed out again to allocations issued on the same stream -- reuse is then ordered by the stream
    // itself, and a sort in steady state makes no cudaMallocAsync / cudaFreeAsync call at all (1 900 of each per sort before).
    std::map<std::pair<cudaStream_t, size_t>, std::vector<void*>> arena_bins;
    std::unordered_map<void*, size_t> arena_sizes;     // live and cached buffers -> bytes
    size_t arena_cached = 0, arena_cap = 0;            // bytes held in the bins; beyond arena_cap a free goes back to the pool
    long arena_hits = 0, arena_misses = 0;
    bool closed = false;
    bool release_pending = false;   // closed and the last handle is gone: the C entry point that notices tears the context down after unlocking
    // Thread safety (SURVEY 8b "Threading": the reference calls cc-> from nested OpenMP regions, src/sort_algo.h:334,349,
    // 425,713,757): every extern "C" entry point holds this mutex for its whole duration, so evaluator state (the
    // stream swap of the encode path, stats, event pools, key-switch tables, the twin / ModUp caches written on const
    // operands) is never touched by two host threads at once.  The GPU work is asynchronous; the lock covers enqueueing.
    std::recursive_mutex mu;
    std::map<int, struct BootPlan*> boot;     // slots -> bootstrapping plan (bootstrap.cu)
    std::deque<const struct Ct*> modup_lru;   // handles that carry a cached ModUp, oldest first (bounded: evaluator.cu)
    int modup_cap = 8;
    int add_const_mode = 0;         // EvalAdd(ct, double) on noiseScaleDeg > 1: 0 = round(|x| * ct scale), 1 = OpenFHE CRTMult form
    int M() const { return size_q + size_p; }
};

struct Ct;
struct AdjTwin { double x; int drop; double sf; Ct* t; };   // evaluator.cu adjusted_twin()
struct Ct {                // ciphertext (parts 2/3) or plaintext (parts 1)
    Context* ctx;
    u64* d;
    int parts, nl;
    int level, deg;
    double sf;
    int slots;
    mutable Ct* resc = nullptr;   // cached rescaled twin (owned); see rescaled_twin() in evaluator.cu
    mutable Ct* lift = nullptr;   // plaintexts only: cached x 1.0 twin (noiseScaleDeg + 1, owned); see lifted_twin()
    mutable int refs = 1;         // ckks_ct_t handles that share this object (capi.cu: ckks_ct_clone is copy-on-write)
    mutable std::vector<AdjTwin> adj;   // cached level-adjusted twins LevelReduce(Rescale(this * x), drop) (owned, at most 4)
    mutable cudaEvent_t ready = nullptr;   // recorded on the encode stream; the first consumer makes the main stream wait on it
    mutable u64* modup = nullptr; // cached ModUp of c1 (owned): shared by all rotations of this ciphertext (hoisting)
    // Plaintexts are encoded on the encode stream, and the level / scale adjustments a multiplication applies to them
    // (lift, rescale, level drop -- OpenFHE's plaintext morphing) stay there too: enc_home marks data whose producer is
    // ordered on that stream; enc_reader is recorded after the last encode-stream kernel that READS this buffer, so
    // that it is not released (on the main stream) under it.
    bool enc_home = false;
    mutable cudaEvent_t enc_reader = nullptr;
    // Plaintexts only: log2 of the replication factor.  A plaintext with s slots is G(X^gap), gap = n / (2 s), so its
    // EVALUATION form is the (n / gap)-point transform of G with every value repeated gap times in a row
    // (ntt.cu).  Such a plaintext is stored compact, [limb][n >> lrep], and element x of the full form is d[x >> lrep].
    int lrep = 0;
};

void context_release(Context* c);
// RAII scope around the launches of one kernel class: CUDA events on the engine stream when that class is profiled
struct ProfScope {
    Context* c; int idx = -1;
    ProfScope(Context* c, int cls, double alg_bytes, int nlaunch = 1);
    ~ProfScope();
};
void prof_collect(Context* c);   // synchronises and folds pending event pairs into Prof::ms   // capi.cu: free device tables + delete (called once closed && live_cts == 0)

// ---- sampler key passed by value to the sampling kernels; ChaCha20 block function (RFC 8439) shared by host and device
struct RngKey { uint32_t k[8]; uint64_t seed; int secure; };
#ifdef __CUDACC__
#define CKKS_HD __host__ __device__ __forceinline__
#else
#define CKKS_HD inline
#endif
CKKS_HD uint32_t chacha_rotl(uint32_t v, int c) { return (v << c) | (v >> (32 - c)); }
// state: constants | key | counter = idx (64 bit) | nonce = stream (64 bit); out = the 16 words of one block
CKKS_HD void chacha20_block(const uint32_t key[8], uint64_t stream, uint64_t idx, uint32_t out[16]) {
    uint32_t x[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key[0], key[1], key[2], key[3], key[4], key[5], key[6], key[7],
                      (uint32_t)idx, (uint32_t)(idx >> 32), (uint32_t)stream, (uint32_t)(stream >> 32)};
    uint32_t w[16];
    for (int i = 0; i < 16; i++) w[i] = x[i];
#define CKKS_QR(a, b, c, d)                                                                         \
    w[a] += w[b]; w[d] = chacha_rotl(w[d] ^ w[a], 16); w[c] += w[d]; w[b] = chacha_rotl(w[b] ^ w[c], 12); \
    w[a] += w[b]; w[d] = chacha_rotl(w[d] ^ w[a], 8);  w[c] += w[d]; w[b] = chacha_rotl(w[b] ^ w[c], 7);
    for (int r = 0; r < 10; r++) {
        CKKS_QR(0, 4, 8, 12) CKKS_QR(1, 5, 9, 13) CKKS_QR(2, 6, 10, 14) CKKS_QR(3, 7, 11, 15)
        CKKS_QR(0, 5, 10, 15) CKKS_QR(1, 6, 11, 12) CKKS_QR(2, 7, 8, 13) CKKS_QR(3, 4, 9, 14)
    }
#undef CKKS_QR
    for (int i = 0; i < 16; i++) out[i] = w[i] + x[i];
}

// ---- device memory helpers (stream-ordered pool) ----
u64* dev_alloc(Context* c, size_t words);
void dev_free(Context* c, void* p);
void arena_trim(Context* c);

// ---- kernel launchers (ntt.cu / kernels.cu / keyswitch.cu / encode.cu) ----
LimbMods mods_q(int nl);                          // Q limbs 0..nl-1
LimbMods mods_qp(const Context* c, int nl);       // Q limbs 0..nl-1 then all P limbs
LimbMods mods_range(int lo, int hi);              // moduli lo..hi-1
void launch_ntt(Context* c, u64* data, int rows, const LimbMods& m, bool inverse);
// out-of-place capable form; scale_pairs (inverse only): [M] (factor, Shoup) pairs indexed by modulus id that
// replace the n^-1 factor (nullptr = n^-1)
// lrep > 0: the rows are compact plaintext limbs of n >> lrep words (same twiddle tables: the first stages of the
// n-point transform restricted to the multiples of 2^lrep ARE the (n >> lrep)-point transform)
void launch_ntt_ex(Context* c, u64* dst, const u64* src, int rows, const LimbMods& m, bool inverse, const u64* scale_pairs,
                   RowMap smap, RowMap dmap, const NttFuse* fuse = nullptr, int lrep = 0);

void launch_add(Context* c, u64* out, const u64* a, const u64* b, int rows, const LimbMods& m, int op /*0 add 1 sub*/);
void launch_neg(Context* c, u64* out, const u64* a, int rows, const LimbMods& m);
void launch_mul(Context* c, u64* out, const u64* a, const u64* b, int rows_out, int rows_b, const LimbMods& m, int lrep_b = 0);
void launch_mul_scalar(Context* c, u64* out, const u64* in, int rows, const LimbMods& m, const u64* h_scalars /*[m.n]*/, int lrep = 0);
// out[0] = (negate ? -in[0] : in[0]) + s ; out[p>0] = (negate ? -in[p] : in[p])
void launch_add_scalar(Context* c, u64* out, const u64* in, int parts, int nl, const u64* h_scalars, int negate /* 0 none, 1 negate first, 2 double first */);
// out[p] = (p < xparts ? x[p] : 0) +- (p < yparts ? y[p] : 0), p < max(xparts, yparts)
// lrep_x / lrep_y: the 1-part operand may be a compact plaintext
void launch_addsub_parts(Context* c, u64* out, const u64* x, int xparts, const u64* y, int yparts, int nl, bool sub,
                         int lrep_x = 0, int lrep_y = 0);
void launch_expand_rep(Context* c, u64* out, const u64* in, int rows, int lrep);   // compact -> full rows
// out = sum_i in_i * s_i with per-limb integer scalars h_scalars[i * nl + l] (EvalLinearWSum in one pass)
void launch_wsum(Context* c, u64* out, const u64* const* ins, const u64* h_scalars, int cnt, int parts, int nl);
#define WSUM_GROUPS 4
#define WSUM_MULTI_MAX_IN 64
struct WsumMultiArgs { int cnt; const u64* in[WSUM_MULTI_MAX_IN]; u64* out[WSUM_GROUPS]; };
void launch_wsum_multi(Context* c, u64* const* outs, int groups, const u64* const* ins, const u64* h_scalars, int cnt, int parts, int nl);
#define IP_MAX 15
struct IpArgs { int cnt; const u64* init; const u64* ct[IP_MAX]; const u64* pt[IP_MAX]; int lrep[IP_MAX]; };
void launch_ip_plain(Context* c, u64* out, const IpArgs& A, int nl);
#define IP_GROUPS 4
struct IpMultiArgs { int cnt, groups; const u64* init; const u64* ct[IP_MAX]; const u64* pt[IP_GROUPS][IP_MAX]; int lrep[IP_GROUPS][IP_MAX]; u64* out[IP_GROUPS]; };
void launch_ip_plain_multi(Context* c, const IpMultiArgs& A, int nl);
void launch_mod_reduce(Context* c, u64* d, int rows, int nl);
void launch_add_many(Context* c, u64* out, const u64* const* ins, int cnt, int rows, int nl);
void launch_rescale(Context* c, u64* out, const u64* in, int parts, int nl_in, int lrep = 0);
// DropLastElementAndScale of (s * in) restricted to the first nl_keep < nl_in limbs: what EvalMult(ct, const) -> rescale ->
// LevelReduce(nl_in - 1 - nl_keep) leaves, in one INTT + one fused NTT over parts * nl_keep rows.  s128 = the integer scalar
// (low, high words of its two's complement).  Returns false (nothing launched) when the scalar-table cache is full.
bool launch_rescale_scaled(Context* c, u64* out, const u64* in, int parts, int nl_in, int nl_keep, u64 s_lo, u64 s_hi, int lrep = 0);
void launch_tensor(Context* c, u64* d3, const u64* a, const u64* b, int nl);
void launch_automorph(Context* c, u64* out, const u64* in, int rows, uint32_t k);
#define MAX_CONV_IN 24
#define MAX_CONV_OUT 88
struct ConvGroup { int in_row0, nin, out_row0, nout; const u64* hatmod; /* [nin][nout] Montgomery form */ const u64* hatplain; /* plain form */
                   unsigned wide; /* bit i: source modulus i is wider than 42 bits (its residues have a third 21-bit piece, k_conv_fp) */ };
struct ConvBatch { int ng; ConvGroup g[4]; short out_ids[4][MAX_CONV_OUT]; bool split_ok = false; /* every group passes ConvTables::split_ok */ };

void launch_conv_batch(Context* c, u64* out, const u64* in, const ConvBatch& B);
void launch_ks_inner2(Context* c, u64* acc /*[2][nl+K][n]*/, const u64* src, const u64* conv, const EvalKey& k,
                      const KsTables& t);
void launch_moddown_final(Context* c, u64* out, const u64* acc, const u64* conv, const u64* add0, const u64* add1, int nl,
                          uint32_t auto_k, const KsTables& t);
void launch_copy_rows(Context* c, u64* dst, const u64* src, size_t words);
void launch_to_mont(Context* c, u64* data, int rows, const LimbMods& m);
void launch_from_mont(Context* c, u64* data, int rows, const LimbMods& m);
void launch_sample_small(Context* c, u64* out, int rows, const LimbMods& m, uint64_t stream, int kind);
void launch_sample_uniform(Context* c, u64* out, int rows, const LimbMods& m, uint64_t stream);
void launch_evk_combine(Context* c, u64* b, const u64* a, const u64* e, const u64* s_new, const u64* s_old,
                        int lo, int hi);
void launch_encode(Context* c, u64* out, const double* h_values, int nvals, int slots, double sf, int nl, int lrep = 0, int is_complex = 0);
void launch_decode(Context* c, double* h_out, const u64* m /*[<=2][n] coeff*/, int nlimbs, double sf, int slots);
void launch_decrypt_combine(Context* c, u64* out, const Ct* ct, int use);
void launch_encrypt_combine(Context* c, u64* out, const u64* pt, const u64* v, const u64* e0, const u64* e1, int nl);

const KsTables& get_ks_tables(Context* c, int nl);

// ---- programmatic dependent launch.  Every kernel of the engine starts with pdl_sync(): it lets the NEXT kernel of the
// stream be scheduled onto free SM slots right away and then blocks until the PREVIOUS grid has completed and its
// writes are visible, so consecutive dependent kernels lose the drain + launch gap (about 1-2 us on each of the ~6700
// launches of a sort) without any change to what they read or write.  A kernel launched without the attribute sees
// both instructions as no-ops.
struct HostTimer {   // accumulates host wall time of a CUDA API call class when CKKS_B200_HOST_PROF is set
    Context* c; int cls; std::chrono::steady_clock::time_point t0;
    HostTimer(Context* c_, int cls_) : c(c_), cls(cls_) { if (c->host_prof) t0 = std::chrono::steady_clock::now(); }
    ~HostTimer() {
        if (!c->host_prof) return;
        c->stats.host_ns[cls] += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
        c->stats.host_calls[cls]++;
    }
};
#ifdef __CUDACC__
// The two halves on their own: a kernel whose prologue only reads context tables (written at set-up, never by a kernel of the
// stream) can run it between pdl_trigger() and pdl_wait(), i.e. while its predecessor is still draining.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <class... KArgs, class... Args>
inline void launch_pdl(Context* c, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = c->pdl ? 1 : 0;
    HostTimer ht(c, 0);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
    if (e != cudaSuccess) throw std::runtime_error(std::string("kernel launch failed: ") + cudaGetErrorString(e));
}
#endif
