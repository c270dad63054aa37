// evaluator.cu -- host-side schedule of the CKKS evaluator over the CUDA kernels.
//
// This is the layer that sits where OpenFHE's LeveledSHECKKSRNS / KeySwitchHYBRID / PKERNS sit
// under the reference's `cc->` calls: FLEXIBLEAUTO(EXT)
// level/scale bookkeeping (AdjustLevelsAndDepthInPlace), lazy rescale, hybrid key switching for
// relinearisation and rotations, key generation, encode/encrypt/decrypt.  All polynomial
// arithmetic is launched on the GPU; the host only moves metadata.  With Context::dry set the
// same code runs metadata-only (no allocations or launches) so that level budgets and op traces
// can be checked on a machine without a GPU.  Bit-level behaviour is pinned by oracle/ckks_ops.py.
#include "engine.h"
#include "evaluator.h"
#include <cstdlib>
#include <memory>
#include <cmath>
#include <stdexcept>
#include <algorithm>

typedef __int128 i128;

static inline size_t words(const Context* c, int rows, int lrep = 0) { return (size_t)rows * (c->n >> lrep); }
static void count(Context* c, const char* name, int nl) { c->stats.ops[{name, nl}]++; }

Ct* ct_new(Context* c, int parts, int nl, int level, int deg, double sf, int slots, int lrep) {
    Ct* t = new Ct{c, nullptr, parts, nl, level, deg, sf, slots};
    t->lrep = lrep;
    t->d = dev_alloc(c, words(c, parts * nl, lrep));
    c->live_cts++;
    return t;
}
void ct_wait(const Ct* a) {
    if (!a->ready) return;
    Context* c = a->ctx;
    cudaStreamWaitEvent(c->stream, a->ready, 0);
    c->ev_pool.push_back(a->ready);          // everything queued on the main stream from here on is ordered after it
    a->ready = nullptr;
}
Ct* ct_clone(const Ct* a) {
    ct_wait(a);
    Ct* t = ct_new(a->ctx, a->parts, a->nl, a->level, a->deg, a->sf, a->slots, a->lrep);
    launch_copy_rows(a->ctx, t->d, a->d, words(a->ctx, a->parts * a->nl, a->lrep));
    return t;
}
void ct_drop_caches(const Ct* a);
static void drop_twin(const Ct* a) { ct_drop_caches(a); }
void ct_drop_caches(const Ct* a) {
    if (a->resc) { Ct* t = a->resc; a->resc = nullptr; ct_free(t); }
    if (a->lift) { Ct* t = a->lift; a->lift = nullptr; ct_free(t); }
    while (!a->adj.empty()) { Ct* t = a->adj.back().t; a->adj.pop_back(); ct_free(t); }
    if (a->modup) {
        dev_free(a->ctx, a->modup); a->modup = nullptr;
        auto& q = a->ctx->modup_lru;
        for (auto it = q.begin(); it != q.end(); ++it) if (*it == a) { q.erase(it); break; }
    }
}
static void wait_enc_reader(const Ct* a) {
    if (!a->enc_reader) return;
    Context* c = a->ctx;
    cudaStreamWaitEvent(c->stream, a->enc_reader, 0);
    c->ev_pool.push_back(a->enc_reader);
    a->enc_reader = nullptr;
}
static cudaEvent_t enc_event(Context* c) {
    cudaEvent_t ev;
    if (!c->ev_pool.empty()) { ev = c->ev_pool.back(); c->ev_pool.pop_back(); }
    else cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    cudaEventRecord(ev, c->enc_stream);
    return ev;
}
// true when an out-of-place op on plaintext `a` can be queued on the encode stream (its input is ordered there)
static bool enc_offload(const Ct* a) {
    Context* c = a->ctx;
    return a->parts == 1 && a->enc_home && c->enc_stream && !c->dry && c->stream != c->enc_stream && c->enc_offload;
}
// run fn (allocation + launches on c->stream) on the encode stream; the result carries its completion event, the source
// remembers the reader
template <class F> static Ct* on_enc_stream(const Ct* src, F&& fn) {
    Context* c = src->ctx;
    cudaStream_t main = c->stream;
    c->stream = c->enc_stream;
    Ct* r = nullptr;
    try { r = fn(); } catch (...) { c->stream = main; throw; }
    r->enc_home = true;
    r->ready = enc_event(c);
    if (src->enc_reader) c->ev_pool.push_back(src->enc_reader);   // an older reader on the same stream is subsumed
    src->enc_reader = enc_event(c);
    c->stream = main;
    return r;
}
void ct_free(Ct* a) {
    if (!a) return;
    Context* c = a->ctx;
    ct_wait(a);                               // the buffer is released on the main stream
    wait_enc_reader(a);
    drop_twin(a);
    dev_free(c, a->d);
    delete a;
    if (--c->live_cts == 0 && c->closed) c->release_pending = true;   // torn down by the C entry point once it has unlocked
}
// a takes over t's buffer and metadata (t is consumed)
static void ct_steal(Ct* a, Ct* t) {
    drop_twin(a);
    ct_wait(a); wait_enc_reader(a);
    dev_free(a->ctx, a->d);
    a->d = t->d; a->parts = t->parts; a->nl = t->nl; a->level = t->level; a->deg = t->deg; a->sf = t->sf; a->lrep = t->lrep;
    a->resc = t->resc; t->resc = nullptr;
    a->lift = t->lift; t->lift = nullptr;
    a->ready = t->ready; t->ready = nullptr; a->enc_reader = t->enc_reader; t->enc_reader = nullptr; a->enc_home = t->enc_home;
    t->d = nullptr;
    ct_free(t);
}

static void int_residues(const Context* c, i128 v, int nl, u64* out) {
    for (int i = 0; i < nl; i++) { i128 q = (i128)c->moduli[i]; i128 r = v % q; if (r < 0) r += q; out[i] = (u64)r; }
}

// ModReduceInternal -> DropLastElementAndScale, out of place
static Ct* rescale_new(const Ct* a) {
    if (enc_offload(a)) return on_enc_stream(a, [&] { return rescale_new(a); });
    Context* c = a->ctx; const int nl = a->nl, parts = a->parts;
    if (nl < 2) throw std::runtime_error("rescale: no limb left to drop (multiplicative depth exhausted)");
    count(c, "rescale", nl);
    Ct* out = ct_new(c, parts, nl - 1, a->level + 1, a->deg - 1, a->sf / (double)c->moduli[nl - 1], a->slots, a->lrep);
    launch_rescale(c, out->d, a->d, parts, nl, a->lrep);
    return out;
}
// the rescaled twin of a pristine ciphertext is computed once and cached on it: under FLEXIBLEAUTO the same
// operand is (lazily) rescaled again by every multiplication that consumes it
static const Ct* rescaled_twin(const Ct* a) {
    if (!a->resc) { a->resc = rescale_new(a); a->ctx->stats.ops[{"rescale_unique", a->nl}]++; }
    else count(a->ctx, "rescale_cached", a->nl);
    return a->resc;
}

// Operand of an evaluator op: a read-only view of a caller's ciphertext (or of its cached twin) until an
// adjustment has to change the data, at which point it owns a private temporary.
struct Opnd {
    const Ct* v;        // current value
    Ct* own = nullptr;  // == v when the value is a private temporary
    explicit Opnd(const Ct* src) : v(src) { ct_wait(src); }
    ~Opnd() { if (own) ct_free(own); }
    Opnd(const Opnd&) = delete;
    void set_own(Ct* t) { if (own) ct_free(own); own = t; v = t; }
    Ct* release() { Ct* t = own; own = nullptr; return t; }   // caller takes the temporary (or nullptr)
};

static Ct* mult_const_new(const Ct* a, double x) {   // EvalMultCore(ct, double), out of place
    if (enc_offload(a)) return on_enc_stream(a, [&] { return mult_const_new(a, x); });
    Context* c = a->ctx;
    const double sf = c->sf_real[a->level];
    i128 v = (i128)(x * sf + 0.5);
    u64 s[MAX_LIMBS]; int_residues(c, v, a->nl, s);
    Ct* r = ct_new(c, a->parts, a->nl, a->level, a->deg + 1, a->sf * sf, a->slots, a->lrep);
    launch_mul_scalar(c, r->d, a->d, a->parts * a->nl, mods_q(a->nl), s, a->lrep);
    return r;
}
static Ct* level_reduce_new(const Ct* a, int levels) {
    if (enc_offload(a)) return on_enc_stream(a, [&] { return level_reduce_new(a, levels); });
    Context* c = a->ctx; const int nl2 = a->nl - levels;
    if (nl2 < 1) throw std::runtime_error("level_reduce: not enough limbs");
    Ct* r = ct_new(c, a->parts, nl2, a->level + levels, a->deg, a->sf, a->slots, a->lrep);
    for (int p = 0; p < a->parts; p++)
        launch_copy_rows(c, c->dry ? nullptr : r->d + words(c, p * nl2, a->lrep), c->dry ? nullptr : a->d + words(c, p * a->nl, a->lrep),
                         words(c, nl2, a->lrep));
    return r;
}
// A pristine plaintext that has to gain a scale degree (EvalMult(ct of degree 2, pt): AdjustLevelsAndDepth multiplies
// the plaintext by 1.0) gets that twin cached on the handle like the rescaled one: the sorters' masks are the same
// handles in every sort (host/openfhe_shim.h plaintext cache), so the adjustment is computed once per context.
static const Ct* lifted_twin(const Ct* a) {
    if (!a->lift) a->lift = mult_const_new(a, 1.0);
    else count(a->ctx, "lift_cached", a->nl);
    return a->lift;
}
static void op_mult_const(Opnd& o, double x) {
    if (!o.own && o.v->parts == 1 && x == 1.0) { o.v = lifted_twin(o.v); ct_wait(o.v); return; }
    o.set_own(mult_const_new(o.v, x));
}
static void op_rescale(Opnd& o) { if (o.own) o.set_own(rescale_new(o.v)); else o.v = rescaled_twin(o.v); }
// EvalMultCore(ct, x) -> ModReduceInternal -> LevelReduceInternal(drop) in one step (kernels.cu launch_rescale_scaled): the scalar
// rides on the inverse transform's final factor and on the rescale epilogue's minuend, and only the limbs that survive the level
// reduction are computed.  Limb for limb the same residues as the three separate steps (every step is exact modular arithmetic).
// DirectSort N=128 meets this adjustment 128 times per sort (EvalAdd / EvalMult of two depth-2 operands at different levels),
// 18 of them from 32 limbs down to 2: 150 scalar-product launches and the transforms of ~1 400 discarded limbs per sort.
static Ct* scaled_rescale_new(const Ct* a, double x, int drop) {
    if (enc_offload(a)) return on_enc_stream(a, [&] { return scaled_rescale_new(a, x, drop); });
    Context* c = a->ctx; const int nl = a->nl, keep = nl - 1 - drop;
    const double sf = c->sf_real[a->level];
    const i128 v = (i128)(x * sf + 0.5);                  // same rounding as mult_const_new
    count(c, "rescale", nl); count(c, "rescale_scaled", nl);
    Ct* out = ct_new(c, a->parts, keep, a->level + 1 + drop, a->deg, (a->sf * sf) / (double)c->moduli[nl - 1], a->slots, a->lrep);
    if (!launch_rescale_scaled(c, out->d, a->d, a->parts, nl, keep, (u64)(unsigned __int128)v, (u64)((unsigned __int128)v >> 64), a->lrep)) {
        ct_free(out);
        throw std::runtime_error("internal: scalar-table cache exhausted");
    }
    return out;
}
static bool can_fuse_adjust(const Opnd& o, int drop) {
    static const bool off = getenv("CKKS_B200_NO_FUSED_ADJUST") != nullptr;   // A/B switch: the three separate steps
    return !off && o.v->parts >= 2 && o.v->nl - 1 - drop >= 1 && o.v->ctx->scal_tabs.size() < 4000;   // plaintexts keep their cached twins
}
static void op_level_reduce(Opnd& o, int levels);
// The adjusted twin of a pristine operand is cached on it like the rescaled one: a sorter adds the same accumulator seed (the
// context's zero ciphertext, a rank vector) to partial results at one and the same deeper level over and over.  Keyed by the
// scalar, the level drop and the scaling factor the caller assigns afterwards (AdjustLevelsAndDepth sets it to the other operand's).
// Clones made through ckks_ct_clone share the handle's data and therefore its twins (capi.cu), so the reference's
// `getZero()->Clone()` seeds hit the cache as well.
static const Ct* adjusted_twin(const Ct* a, double x, int drop, double sf) {
    for (const AdjTwin& e : a->adj)
        if (e.x == x && e.drop == drop && e.sf == sf) { count(a->ctx, "adjust_cached", a->nl); return e.t; }
    Ct* t = scaled_rescale_new(a, x, drop);
    t->sf = sf;
    if (a->adj.size() >= 4) { Ct* old = a->adj.front().t; a->adj.erase(a->adj.begin()); ct_free(old); }
    a->adj.push_back(AdjTwin{x, drop, sf, t});
    return t;
}
static void op_set_sf(Opnd& o, double sf);
// a <- LevelReduce(Rescale(a * x), drop) with the scaling factor set to sf
static void op_mult_rescale_reduce(Opnd& o, double x, int drop, double sf) {
    if (can_fuse_adjust(o, drop)) {
        if (!o.own) { o.v = adjusted_twin(o.v, x, drop, sf); ct_wait(o.v); return; }
        o.set_own(scaled_rescale_new(o.v, x, drop));
    } else { op_mult_const(o, x); op_rescale(o); op_level_reduce(o, drop); }
    op_set_sf(o, sf);
}
static void op_level_reduce(Opnd& o, int levels) { if (levels > 0) o.set_own(level_reduce_new(o.v, levels)); }
static void op_set_sf(Opnd& o, double sf) {
    if (o.v->sf == sf) return;
    if (!o.own) o.set_own(ct_clone(o.v));
    o.own->sf = sf;
}

// full-length copy of a compact plaintext (ops without a compact form use this)
static Ct* expand_new(const Ct* a) {
    Context* c = a->ctx;
    ct_wait(a);
    Ct* r = ct_new(c, a->parts, a->nl, a->level, a->deg, a->sf, a->slots);
    launch_expand_rep(c, r->d, a->d, a->parts * a->nl, a->lrep);
    return r;
}
Ct* ct_full(const Ct* a) { return a->lrep ? expand_new(a) : ct_clone(a); }
static void op_full(Opnd& o) { if (o.v->lrep) o.set_own(expand_new(o.v)); }

// LeveledSHECKKSRNS::AdjustLevelsAndDepthInPlace on operand views
static void adjust_levels_and_depth(Opnd& A, Opnd& B) {
    Context* c = A.v->ctx;
    if (A.v->level == B.v->level) {
        if (A.v->deg < B.v->deg) op_mult_const(A, 1.0);
        else if (B.v->deg < A.v->deg) op_mult_const(B, 1.0);
        return;
    }
    Opnd& a = A.v->level < B.v->level ? A : B;     // a: lower level (more limbs), brought to b
    Opnd& b = A.v->level < B.v->level ? B : A;
    const int l1 = a.v->level, l2 = b.v->level;
    const double scf = c->sf_real[l1], bsf = b.v->sf;
    if (a.v->deg == 2) {
        if (b.v->deg == 2) {
            double q1 = (double)c->moduli[a.v->nl - 1];
            op_mult_rescale_reduce(a, bsf / a.v->sf * q1 / scf, l2 - l1 - 1, bsf);
        } else if (l1 + 1 == l2) {
            op_rescale(a);
        } else {
            double q1 = (double)c->moduli[a.v->nl - 1];
            {   // the intermediate's scaling factor is its natural one: an owned temporary, no twin
                const double x = c->sf_big[l2 - 1] / a.v->sf * q1 / scf; const int drop = l2 - l1 - 2;
                if (can_fuse_adjust(a, drop)) a.set_own(scaled_rescale_new(a.v, x, drop));
                else { op_mult_const(a, x); op_rescale(a); op_level_reduce(a, drop); }
            }
            op_rescale(a);
            op_set_sf(a, bsf);
        }
    } else {
        if (b.v->deg == 2) {
            op_mult_const(a, bsf / a.v->sf / scf);
            op_level_reduce(a, l2 - l1);
            op_set_sf(a, bsf);
        } else {
            op_mult_const(a, c->sf_big[l2 - 1] / a.v->sf / scf);
            op_level_reduce(a, l2 - l1 - 1);
            op_rescale(a);
            op_set_sf(a, bsf);
        }
    }
}
static void adjust_to_one(Opnd& a, Opnd& b) {
    adjust_levels_and_depth(a, b);
    if (a.v->deg == 2) { op_rescale(a); op_rescale(b); }
}

// ---- in-place forms (EvalLinearWSumMutable and the Chebyshev evaluator mutate their operands)
static void commit(Ct* a, Opnd& o) {
    if (o.v == a) return;
    if (o.own) { ct_steal(a, o.release()); return; }
    // o.v is one of a's cached twins: a becomes that twin
    Ct* t;
    if (o.v == a->resc) { t = a->resc; a->resc = nullptr; }
    else if (o.v == a->lift) { t = a->lift; a->lift = nullptr; }
    else if (a->lift && o.v == a->lift->resc) { t = a->lift->resc; a->lift->resc = nullptr; }
    else if ([&] {
                 for (size_t i = 0; i < a->adj.size(); i++) {
                     Ct* w = a->adj[i].t;
                     if (w == o.v) { t = w; a->adj.erase(a->adj.begin() + i); return true; }
                     if (w->resc && w->resc == o.v) { t = w->resc; w->resc = nullptr; return true; }
                 }
                 return false;
             }()) {}
    else throw std::runtime_error("internal: operand view is not a twin of its ciphertext");
    ct_steal(a, t);
}
void mult_core_const(Ct* a, double x) { Opnd o(a); op_mult_const(o, x); commit(a, o); }
void rescale(Ct* a) { Opnd o(a); op_rescale(o); commit(a, o); }
void level_reduce(Ct* a, int levels) { Opnd o(a); op_level_reduce(o, levels); commit(a, o); }
void adjust_levels_and_depth(Ct* a, Ct* b) {
    if (a == b) return;
    Opnd x(a), y(b);
    adjust_levels_and_depth(x, y);
    commit(a, x); commit(b, y);
}

// ModUp of one polynomial (EvalKeySwitchPrecomputeCore): INTT (x n^-1 Qhat^-1 fused) -> basis conversion of every
// digit -> NTT of all converted limbs.  Returns the converted complements [sum_j nout_j][n] (caller frees).
u64* key_switch_modup(Context* c, const u64* src, int nl) {
    const KsTables& t = get_ks_tables(c, nl);
    const int K = c->size_p, next = nl + K;
    if (K > MAX_CONV_IN || c->alpha > MAX_CONV_IN || next > MAX_CONV_OUT) throw std::runtime_error("key switch: parameter set too large");
    count(c, "modup", nl);
    u64* coef = dev_alloc(c, words(c, nl));
    launch_ntt_ex(c, coef, src, nl, mods_q(nl), true, t.d_modup_scale, RowMap{}, RowMap{});
    int conv_rows = 0;
    for (int j = 0; j < t.nparts; j++) conv_rows += t.modup[j].nout;
    u64* conv = dev_alloc(c, words(c, conv_rows));
    {
        ConvBatch B; B.ng = t.nparts; B.split_ok = true;
        for (int j = 0; j < t.nparts; j++) B.split_ok = B.split_ok && t.modup[j].split_ok;
        LimbMods m; m.n = 0;
        int row0 = 0;
        for (int j = 0; j < t.nparts; j++) {
            const ConvTables& ct = t.modup[j];
            B.g[j] = ConvGroup{t.dig_lo[j], ct.nin, row0, ct.nout, ct.d_hatmod, ct.d_hatplain, ct.wide};
            for (int i = 0; i < ct.nout; i++) { B.out_ids[j][i] = ct.out_ids[i]; if (m.n < MAX_LIMBS) m.m[m.n++] = ct.out_ids[i]; }
            row0 += ct.nout;
        }
        launch_conv_batch(c, conv, coef, B);
        if (conv_rows <= MAX_LIMBS) launch_ntt(c, conv, conv_rows, m, false);
        else {
            for (int j = 0, r0 = 0; j < t.nparts; r0 += t.modup[j].nout, j++) {
                LimbMods mj; mj.n = t.modup[j].nout;
                for (int i = 0; i < mj.n; i++) mj.m[i] = t.modup[j].out_ids[i];
                launch_ntt(c, c->dry ? nullptr : conv + words(c, r0), mj.n, mj, false);
            }
        }
    }
    dev_free(c, coef);
    return conv;
}

// KeySwitchHYBRID::KeySwitchCore (+ optional addends): out[h] = add_h + ModDown(sum_j d_j * evk_j[h]);
// auto_k != 0 additionally applies the automorphism X -> X^k to both outputs while writing them.
// modup != nullptr: precomputed ModUp of src (hoisting: EvalFastRotationPrecompute / the per-ciphertext cache)
void key_switch(Context* c, const u64* src, int nl, const EvalKey& key, u64* out, const u64* add0, const u64* add1,
                uint32_t auto_k, const u64* modup) {
    count(c, "keyswitch", nl);
    if (!c->dry && !key.d) throw std::runtime_error("key switch: evaluation key not generated");
    const KsTables& t = get_ks_tables(c, nl);
    const int K = c->size_p, next = nl + K;
    u64* own_conv = modup ? nullptr : key_switch_modup(c, src, nl);
    const u64* conv = modup ? modup : own_conv;
    // ---- inner product with the evaluation key over Q_l u P
    u64* acc = dev_alloc(c, words(c, 2 * next));
    launch_ks_inner2(c, acc, src, conv, key, t);
    if (own_conv) dev_free(c, own_conv);
    // ---- ModDown of both polynomials: INTT of the P limbs (x n^-1 Phat^-1) -> conversion to Q_l -> NTT -> epilogue
    u64* pp = dev_alloc(c, words(c, 2 * K));
    {
        LimbMods mp; mp.n = K;
        for (int k = 0; k < K; k++) mp.m[k] = (short)(c->size_q + k);
        launch_ntt_ex(c, pp, acc, 2 * K, mp, true, t.d_moddown_scale, RowMap{K, next, nl}, RowMap{});
    }
    u64* conv2 = dev_alloc(c, words(c, 2 * nl));
    {
        ConvBatch B; B.ng = 2; B.split_ok = t.moddown.split_ok;
        for (int h = 0; h < 2; h++) {
            B.g[h] = ConvGroup{h * K, K, h * nl, nl, t.moddown.d_hatmod, t.moddown.d_hatplain, t.moddown.wide};
            for (int i = 0; i < nl; i++) B.out_ids[h][i] = (short)i;
        }
        launch_conv_batch(c, conv2, pp, B);
    }
    static const bool old_moddown = getenv("CKKS_B200_OLD_MODDOWN") != nullptr;   // A/B: separate epilogue kernel (round 1)
    if (old_moddown) {
        launch_ntt(c, conv2, 2 * nl, mods_q(nl), false);
        launch_moddown_final(c, out, acc, conv2, add0, add1, nl, auto_k, t);
    } else {
        // the ApproxModDown epilogue rides on the last pass of the NTT of the converted limbs: (acc - NTT(conv)) P^-1 + addend,
        // stored through the automorphism -- one launch and one round trip of 2 nl limbs less than NTT + k_moddown_final
        NttFuse f;
        f.epi = EPI_MODDOWN; f.epi_in = acc; f.emap = RowMap{nl, next, 0, 1}; f.epi_pairs = t.d_pinv_pairs;
        f.epi_add[0] = add0; f.epi_add[1] = add1; f.epi_add_grp = nl;
        f.first_in_place = 1;
        if (auto_k) {   // inverse of auto_k modulo 2n (Newton iteration on the 2-adic inverse)
            const uint32_t m = 2u * (uint32_t)c->n;
            uint32_t x = auto_k;
            for (int it = 0; it < 5; it++) x *= 2u - auto_k * x;
            f.auto_kinv = x & (m - 1);
            if (((uint64_t)f.auto_kinv * auto_k) % m != 1) throw std::runtime_error("internal: automorphism inverse");
        }
        launch_ntt_ex(c, out, conv2, 2 * nl, mods_q(nl), false, nullptr, RowMap{}, RowMap{}, &f);
    }
    dev_free(c, pp); dev_free(c, conv2); dev_free(c, acc);
}

// ------------------------------------------------------------------ public evaluator ops
// The integer EvalAdd(ct, double) adds to c0.  Context::add_const_mode 0: round(|x| * ct scale).  Mode 1: OpenFHE 1.1.4
// LeveledSHECKKSRNS::GetElementForEvalAddOrSub as recalled (unverifiable offline, SURVEY 8c'-7): round(|x| * sf_level),
// then multiplied by round(sf_level) mod q_i once per extra noiseScaleDeg (CRTMult); at level 0 under FLEXIBLEAUTOEXT the
// big scaling factor is used and the degree loop is skipped (same integer as mode 0).  Oracle twin: ckks_ops._add_const.
static void add_const_residues(const Ct* a, double x, u64* s) {
    Context* c = a->ctx;
    if (c->add_const_mode == 0 || a->deg <= 1 || (c->ext && a->level == 0)) {
        i128 v = (i128)(std::fabs(x) * a->sf + 0.5);
        if (x < 0) v = -v;
        int_residues(c, v, a->nl, s);
        return;
    }
    const double sf = c->sf_real[a->level];
    const i128 v0 = (i128)(std::fabs(x) * sf + 0.5), f = (i128)(sf + 0.5);
    for (int i = 0; i < a->nl; i++) {
        const i128 q = (i128)c->moduli[i];
        i128 r = v0 % q;
        for (int d = 1; d < a->deg; d++) r = (i128)((unsigned __int128)r * (unsigned __int128)(f % q) % (unsigned __int128)q);
        if (x < 0 && r) r = q - r;
        s[i] = (u64)r;
    }
}
// EvalAdd(ct, double): add the scaled constant (signed) to c0
Ct* eval_add_const(const Ct* a_in, double x) {
    Opnd fa(a_in); op_full(fa); const Ct* a = fa.v;
    Context* c = a->ctx;
    u64 s[MAX_LIMBS]; add_const_residues(a, x, s);
    Ct* r = ct_new(c, a->parts, a->nl, a->level, a->deg, a->sf, a->slots);
    launch_add_scalar(c, r->d, a->d, a->parts, a->nl, s, 0);
    return r;
}
// EvalAdd(EvalAdd(ct, ct), double) in one pass: the T_2i = 2 T_i^2 - 1 steps of the Chebyshev recurrences (cheby.cu)
Ct* eval_double_add_const(const Ct* a_in, double x) {
    Opnd fa(a_in); op_full(fa); const Ct* a = fa.v;
    Context* c = a->ctx;
    count(c, "add", a->nl);
    u64 s[MAX_LIMBS]; add_const_residues(a, x, s);
    Ct* r = ct_new(c, a->parts, a->nl, a->level, a->deg, a->sf, a->slots);
    launch_add_scalar(c, r->d, a->d, a->parts, a->nl, s, 2);
    return r;
}
// EvalSub(double, ct) = -ct + x in one pass
Ct* eval_const_sub(double x, const Ct* a_in) {
    Opnd fa(a_in); op_full(fa); const Ct* a = fa.v;
    Context* c = a->ctx;
    u64 s[MAX_LIMBS]; add_const_residues(a, x, s);
    Ct* r = ct_new(c, a->parts, a->nl, a->level, a->deg, a->sf, a->slots);
    launch_add_scalar(c, r->d, a->d, a->parts, a->nl, s, 1);
    return r;
}

Ct* eval_negate(const Ct* a_in) {
    Opnd fa(a_in); op_full(fa); const Ct* a = fa.v;
    Ct* r = ct_new(a->ctx, a->parts, a->nl, a->level, a->deg, a->sf, a->slots);
    launch_neg(a->ctx, r->d, a->d, a->parts * a->nl, mods_q(a->nl));
    return r;
}

// EvalAdd / EvalSub over (ct,ct) (ct,pt) (pt,ct): sign=+1 add, -1 sub (a - b)
Ct* eval_add(const Ct* a, const Ct* b, int sign) {
    Context* c = a->ctx;
    Opnd x(a), y(b);
    adjust_levels_and_depth(x, y);
    if (x.v->parts == 1 && y.v->parts == 1) { op_full(x); op_full(y); }   // plaintext +- plaintext: no compact form
    const Ct *X = x.v, *Y = y.v;
    ct_wait(X); ct_wait(Y);                               // adjusted plaintexts come from the encode stream
    const int nl = X->nl, parts = std::max(X->parts, Y->parts);
    count(c, "add", nl);
    Ct* r = ct_new(c, parts, nl, X->level, X->deg, X->sf, std::max(a->slots, b->slots));
    // out[p] = (p < X.parts ? X[p] : 0) +- (p < Y.parts ? Y[p] : 0)
    launch_addsub_parts(c, r->d, X->d, X->parts, Y->d, Y->parts, nl, sign < 0, X->lrep, Y->lrep);
    return r;
}

Ct* eval_mult_const(const Ct* a, double x) {
    Opnd o(a);
    if (o.v->deg == 2) op_rescale(o);
    return mult_const_new(o.v, x);
}

// sum_i cts[i] * w[i] for operands that are already aligned (same level, limbs, noiseScaleDeg 1): the per-term
// EvalMult(ct, double) + EvalAdd chain of EvalLinearWSumMutable in one pass.  Returns nullptr when not applicable.
Ct* eval_wsum_aligned(const std::vector<Ct*>& cts, const std::vector<double>& w) {
    const Ct* f = cts[0]; Context* c = f->ctx;
    if (cts.size() < 2 || f->deg != 1 || f->lrep) return nullptr;
    for (const Ct* g : cts)
        if (g->parts != f->parts || g->nl != f->nl || g->level != f->level || g->deg != 1 || g->sf != f->sf || g->lrep) return nullptr;
    const double sf = c->sf_real[f->level];
    std::vector<u64> sc(cts.size() * (size_t)f->nl);
    std::vector<const u64*> ins(cts.size());
    int slots = 0;
    for (size_t i = 0; i < cts.size(); i++) {
        i128 v = (i128)(w[i] * sf + 0.5);                 // same rounding as mult_const_new
        int_residues(c, v, f->nl, sc.data() + i * f->nl);
        ins[i] = cts[i]->d; slots = std::max(slots, cts[i]->slots);
        count(c, "mult_const", f->nl);
        if (i) count(c, "add", f->nl);
    }
    Ct* r = ct_new(c, f->parts, f->nl, f->level, 2, f->sf * sf, slots);
    launch_wsum(c, r->d, ins.data(), sc.data(), (int)cts.size(), f->parts, f->nl);
    return r;
}

// `groups` weighted sums over a common, already aligned input list: outs[g] = sum_{i < cnt[g]} cts[i] * w[g][i], each exactly what
// eval_wsum_aligned(cts[0..cnt[g]), w[g]) returns, in one pass over the inputs (kernels.cu k_wsum_multi).  Returns false (nothing
// done) when the inputs are not aligned.
bool eval_wsum_aligned_multi(const std::vector<Ct*>& cts, const std::vector<std::vector<double>>& w, Ct** outs) {
    const int groups = (int)w.size();
    if (groups < 1 || groups > WSUM_GROUPS) return false;
    int cmax = 0;
    for (const auto& g : w) cmax = std::max(cmax, (int)g.size());
    if (cmax < 2 || cmax > (int)cts.size() || cmax > WSUM_MULTI_MAX_IN) return false;
    for (const auto& g : w) if (g.size() < 2) return false;     // eval_wsum_aligned declines single terms
    const Ct* f = cts[0]; Context* c = f->ctx;
    if (f->deg != 1 || f->lrep) return false;
    for (int i = 0; i < cmax; i++) {
        const Ct* g = cts[i];
        if (g->parts != f->parts || g->nl != f->nl || g->level != f->level || g->deg != 1 || g->sf != f->sf || g->lrep) return false;
    }
    const double sf = c->sf_real[f->level];
    std::vector<u64> sc((size_t)groups * cmax * f->nl, 0);
    std::vector<const u64*> ins(cmax);
    for (int i = 0; i < cmax; i++) { ct_wait(cts[i]); ins[i] = cts[i]->d; }
    std::vector<u64*> od(groups);
    for (int g = 0; g < groups; g++) {
        int slots = 0;
        for (size_t i = 0; i < w[g].size(); i++) {
            const i128 v = (i128)(w[g][i] * sf + 0.5);                 // same rounding as mult_const_new
            int_residues(c, v, f->nl, sc.data() + ((size_t)g * cmax + i) * f->nl);
            slots = std::max(slots, cts[i]->slots);
            count(c, "mult_const", f->nl);
            if (i) count(c, "add", f->nl);
        }
        outs[g] = ct_new(c, f->parts, f->nl, f->level, 2, f->sf * sf, slots);
        od[g] = outs[g]->d;
    }
    launch_wsum_multi(c, od.data(), groups, ins.data(), sc.data(), cmax, f->parts, f->nl);
    return true;
}

// T = init; for i: T = EvalAdd(T, EvalMult(cts[i], pts[i]))   (vecRotsOpt / blindRotationOptN of the reference,
// src/sort_algo.h:336-345, 567-577).  Every pair is adjusted exactly as EvalMult would (lazy rescale of the ciphertext
// through its cached twin, lift + rescale of the plaintext), the accumulator exactly as the first EvalAdd would; when
// the adjusted products then share level, limbs and scale -- the sorters' case -- the cnt multiplications and additions
// are one fused pass.  Anything else runs the chain itself.  Both give the same limbs.
static Ct* inner_product_chain(const Ct* init, const Ct* const* cts, const Ct* const* pts, int cnt) {
    init->ctx->stats.ops[{"inner_product_chain", init->nl}]++;
    Ct* T = ct_clone(init);
    for (int i = 0; i < cnt; i++) {
        Ct* pr = eval_mult(cts[i], pts[i]);
        Ct* nx = eval_add(T, pr, 1);
        ct_free(pr); ct_free(T); T = nx;
    }
    return T;
}
Ct* eval_inner_product_plain(const Ct* init, const Ct* const* cts, const Ct* const* pts, int cnt) {
    Context* c = init->ctx;
    bool ok = cnt >= 1 && cnt <= IP_MAX && init->parts == 2 && init->lrep == 0;
    for (int i = 0; i < cnt && ok; i++) ok = cts[i]->parts == 2 && pts[i]->parts == 1 && cts[i]->lrep == 0;
    if (!ok) return inner_product_chain(init, cts, pts, cnt);
    std::vector<std::unique_ptr<Opnd>> X(cnt), Y(cnt);
    for (int i = 0; i < cnt; i++) {
        X[i].reset(new Opnd(cts[i])); Y[i].reset(new Opnd(pts[i]));
        adjust_to_one(*X[i], *Y[i]);
    }
    const Ct* f = X[0]->v;
    const int nl = f->nl, L = f->level;
    const double psf = f->sf * Y[0]->v->sf;
    for (int i = 0; i < cnt && ok; i++) {
        const Ct *a = X[i]->v, *p = Y[i]->v;
        ok = a->nl == nl && a->level == L && p->nl == nl && p->level == L && a->deg + p->deg == 2 && a->sf * p->sf == psf && nl <= 64;
    }
    // the accumulator must be the operand the first EvalAdd adjusts (it is the shallower one in the sorters)
    ok = ok && (init->level < L || (init->level == L && init->deg <= 2));
    if (!ok) return inner_product_chain(init, cts, pts, cnt);
    Ct prod{c, nullptr, 2, nl, L, 2, psf, 0};              // metadata of the first product; never touched by the adjustment
    Opnd T(init), P(&prod);
    adjust_levels_and_depth(T, P);
    if (P.v != &prod || P.own || T.v->nl != nl || T.v->level != L || T.v->deg != 2) throw std::runtime_error("EvalInnerProductPlain: adjustment invariant broken");
    IpArgs A; A.cnt = cnt; A.init = T.v->d;
    int slots = init->slots;
    for (int i = 0; i < cnt; i++) {
        ct_wait(X[i]->v); ct_wait(Y[i]->v);
        A.ct[i] = X[i]->v->d; A.pt[i] = Y[i]->v->d; A.lrep[i] = Y[i]->v->lrep;
        slots = std::max(slots, cts[i]->slots);
        count(c, "mult_pt", nl); count(c, "add", nl);
    }
    c->stats.ops[{"inner_product_fused", nl}]++;
    Ct* r = ct_new(c, 2, nl, L, 2, T.v->sf, slots);
    launch_ip_plain(c, r->d, A, nl);
    return r;
}

// `groups` inner products over the SAME ciphertexts and accumulator seed: outs[g] = init + sum_i cts[i] * pts[g * cnt + i].
// vecRotsOpt / blindRotationOptN mask the same np baby steps with num_partition / np different mask sets (src/sort_algo.h:331-352,
// 563-580); one pass that keeps several accumulators per coefficient reads the baby steps once per pass instead of once per
// output (8 x 62 limbs = 0.5 GB per read at ring 2^17).  Each output has the limbs the single call would give it; operands
// that do not align fall back to one call per group.
void eval_inner_product_plain_multi(const Ct* init, const Ct* const* cts, const Ct* const* pts, int cnt, int groups, Ct** outs) {
    Context* c = init->ctx;
    auto fallback = [&] { for (int g = 0; g < groups; g++) outs[g] = eval_inner_product_plain(init, cts, pts + (size_t)g * cnt, cnt); };
    bool ok = groups >= 2 && cnt >= 1 && cnt <= IP_MAX && init->parts == 2 && init->lrep == 0;
    for (int i = 0; i < cnt && ok; i++) ok = cts[i]->parts == 2 && cts[i]->lrep == 0;
    for (int i = 0; i < cnt * groups && ok; i++) ok = pts[i]->parts == 1;
    if (!ok) return fallback();
    std::vector<std::unique_ptr<Opnd>> X((size_t)cnt * groups), Y((size_t)cnt * groups);
    for (int k = 0; k < cnt * groups; k++) {
        X[k].reset(new Opnd(cts[k % cnt])); Y[k].reset(new Opnd(pts[k]));
        adjust_to_one(*X[k], *Y[k]);
    }
    const Ct* f = X[0]->v;
    const int nl = f->nl, L = f->level;
    const double psf = f->sf * Y[0]->v->sf;
    for (int k = 0; k < cnt * groups && ok; k++) {
        const Ct *a = X[k]->v, *p = Y[k]->v;
        ok = a == X[k % cnt]->v && a->nl == nl && a->level == L && p->nl == nl && p->level == L && a->deg + p->deg == 2 && a->sf * p->sf == psf && nl <= 64;
    }
    ok = ok && (init->level < L || (init->level == L && init->deg <= 2));
    if (!ok) { X.clear(); Y.clear(); return fallback(); }
    Ct prod{c, nullptr, 2, nl, L, 2, psf, 0};
    Opnd T(init), P(&prod);
    adjust_levels_and_depth(T, P);
    if (P.v != &prod || P.own || T.v->nl != nl || T.v->level != L || T.v->deg != 2) throw std::runtime_error("EvalInnerProductPlainMulti: adjustment invariant broken");
    int slots = init->slots;
    for (int i = 0; i < cnt; i++) slots = std::max(slots, cts[i]->slots);
    for (int k = 0; k < cnt * groups; k++) { ct_wait(X[k]->v); ct_wait(Y[k]->v); count(c, "mult_pt", nl); count(c, "add", nl); }
    // accumulators per pass (CKKS_B200_IP_GROUPS, A/B).  With the 128-bit integer multiply-accumulate on every limb two sets per pass
    // were the optimum (DirectSort N=128 per sort: 51.98 ms with one call per set, 51.76 / 51.49 / 51.65 / 52.20 ms with 1 / 2 / 3 / 4
    // sets per pass -- integer-issue bound beyond two); with the FP64 products on the 40-bit limbs (k_ip_plain_multi) four sets per
    // pass stay memory-bound: 49.13 / 48.99 / 49.00 ms with 2 / 3 / 4
    static const int chunk = getenv("CKKS_B200_IP_GROUPS") ? std::max(1, std::min(IP_GROUPS, atoi(getenv("CKKS_B200_IP_GROUPS")))) : IP_GROUPS;
    for (int g0 = 0; g0 < groups; g0 += chunk) {
        IpMultiArgs A; A.cnt = cnt; A.groups = std::min(chunk, groups - g0); A.init = T.v->d;
        for (int i = 0; i < cnt; i++) A.ct[i] = X[i]->v->d;
        for (int g = 0; g < A.groups; g++) {
            outs[g0 + g] = ct_new(c, 2, nl, L, 2, T.v->sf, slots);
            A.out[g] = outs[g0 + g]->d;
            for (int i = 0; i < cnt; i++) { const Ct* p = Y[(size_t)(g0 + g) * cnt + i]->v; A.pt[g][i] = p->d; A.lrep[g][i] = p->lrep; }
            c->stats.ops[{"inner_product_fused", nl}]++;
        }
        launch_ip_plain_multi(c, A, nl);
    }
}

// EvalMult(ct,ct) with relinearisation / EvalMult(ct,pt)
Ct* eval_mult(const Ct* a, const Ct* b) {
    Context* c = a->ctx;
    Opnd x(a), y(b);
    if (a == b) { if (x.v->deg == 2) op_rescale(x); y.v = x.v; }   // EvalSquare: one operand, rescaled once
    else adjust_to_one(x, y);
    const Ct *X = x.v, *Y = y.v;
    ct_wait(X); ct_wait(Y);
    const int nl = X->nl;
    Ct* r;
    if (X->parts == 2 && Y->parts == 2) {
        count(c, "mult", nl);
        u64* d3 = dev_alloc(c, words(c, 3 * nl));
        launch_tensor(c, d3, X->d, Y->d, nl);
        r = ct_new(c, 2, nl, X->level, X->deg + Y->deg, X->sf * Y->sf, std::max(a->slots, b->slots));
        key_switch(c, c->dry ? nullptr : d3 + words(c, 2 * nl), nl, c->relin, r->d, d3, c->dry ? nullptr : d3 + words(c, nl), 0);
        dev_free(c, d3);
    } else {
        const Ct *ct = X->parts == 2 ? X : Y, *pt = X->parts == 2 ? Y : X;
        if (ct->parts != 2 || pt->parts != 1) throw std::runtime_error("eval_mult: unsupported operand shapes");
        count(c, "mult_pt", nl);
        r = ct_new(c, 2, nl, X->level, X->deg + Y->deg, X->sf * Y->sf, ct->slots);
        launch_mul(c, r->d, ct->d, pt->d, 2 * nl, nl, mods_q(nl), pt->lrep);
    }
    return r;
}

uint32_t auto_index(const Context* c, int r) {
    const u64 m = 2ull * c->n;
    if (r == 0) return 1;
    u64 g = 5;
    if (r < 0) {  // 5^-1 mod m by exponentiation in (Z/m)^*: order divides m/2
        u64 e = m / 2 - 1, b = 5, acc = 1;   // 5^(ord-1); ord of 5 is m/4, and m/2-1 = 2*(m/4)-1 == -1 mod ord
        while (e) { if (e & 1) acc = acc * b % m; b = b * b % m; e >>= 1; }
        g = acc;
    }
    u64 res = 1, base = g; unsigned e = (unsigned)std::abs(r);
    while (e) { if (e & 1) res = res * base % m; base = base * base % m; e >>= 1; }
    return (uint32_t)res;
}

// EvalRotate -> EvalAutomorphism: key switch c1, then permute both polynomials (the permutation is applied by
// the mod-down epilogue while it writes the result)
Ct* eval_rotate(const Ct* a, int r) {
    Context* c = a->ctx;
    if (a->parts != 2) throw std::runtime_error("EvalRotate: ciphertext must have 2 components");
    if (r == 0) return ct_clone(a);   // EvalRotate / EvalAtIndex(ct, 0) returns a copy (no key involved)
    return eval_automorphism(a, auto_index(c, r), ("EvalRotate: no rotation key for index " + std::to_string(r)).c_str());
}
// complex conjugation of the slots: the automorphism X -> X^(2n-1) (bootstrapping splits real and imaginary parts with it)
Ct* eval_conjugate(const Ct* a) {
    if (a->parts != 2) throw std::runtime_error("EvalConjugate: ciphertext must have 2 components");
    return eval_automorphism(a, 2u * (uint32_t)a->ctx->n - 1u, "EvalConjugate: no conjugation key (EvalBootstrapKeyGen generates it)");
}
Ct* eval_automorphism(const Ct* a, uint32_t k, const char* what) {
    Context* c = a->ctx;
    auto it = c->rot.find(k);
    if (it == c->rot.end()) throw std::runtime_error(what);
    count(c, "rotate", a->nl);
    const int nl = a->nl;
    Ct* out = ct_new(c, 2, nl, a->level, a->deg, a->sf, a->slots);
    // hoisting: the ModUp of c1 is computed on the first rotation of a ciphertext and kept on the handle, so the
    // baby-step rotations of one input (sort_algo.h:425-431, 733-736) share it.  Same limbs as EvalRotate: ModUp does
    // not depend on the rotation index (EvalFastRotation == EvalRotate bit for bit).
    const u64* c1 = c->dry ? nullptr : a->d + words(c, nl);
    if (!a->modup) {
        u64* m = key_switch_modup(c, c1, nl);
        a->modup = c->dry ? (u64*)1 : m;   // dry: sentinel (dev_free ignores it)
        // the cache is (beta (l + K) - l) limbs per handle (92 MB at l = 32): keep it on the most recently rotated handles only
        c->modup_lru.push_back(a);
        while ((int)c->modup_lru.size() > c->modup_cap) {
            const Ct* old = c->modup_lru.front(); c->modup_lru.pop_front();
            if (old != a && old->modup) { dev_free(c, old->modup); old->modup = nullptr; }
        }
    } else count(c, "modup_reused", nl);
    key_switch(c, c1, nl, it->second, out->d, a->d, nullptr, k, a->modup);
    return out;
}

// EvalFastRotationPrecompute (rotation.h:286): the ModUp digits of c1 as a handle of their own (parts == 0)
Ct* fast_rotation_precompute(const Ct* a) {
    Context* c = a->ctx;
    if (a->parts != 2) throw std::runtime_error("EvalFastRotationPrecompute: ciphertext must have 2 components");
    Ct* p = new Ct{c, nullptr, 0, a->nl, a->level, a->deg, a->sf, a->slots};
    c->live_cts++;
    p->d = key_switch_modup(c, c->dry ? nullptr : a->d + words(c, a->nl), a->nl);
    return p;
}
// EvalFastRotation(ct, index, m, digits) (rotation.h:344-347)
Ct* eval_fast_rotate(const Ct* a, int r, const Ct* pre) {
    Context* c = a->ctx;
    if (a->parts != 2 || pre->parts != 0 || pre->nl != a->nl) throw std::runtime_error("EvalFastRotation: digits do not match the ciphertext");
    if (r == 0) return ct_clone(a);
    uint32_t k = auto_index(c, r);
    auto it = c->rot.find(k);
    if (it == c->rot.end()) throw std::runtime_error("EvalFastRotation: no rotation key for index " + std::to_string(r));
    count(c, "rotate", a->nl);
    const int nl = a->nl;
    Ct* out = ct_new(c, 2, nl, a->level, a->deg, a->sf, a->slots);
    key_switch(c, c->dry ? nullptr : a->d + words(c, nl), nl, it->second, out->d, a->d, nullptr, k, c->dry ? (const u64*)1 : pre->d);
    return out;
}

// ------------------------------------------------------------------ keys / encrypt / decrypt
enum { STREAM_SECRET = 1, STREAM_PK_A = 2, STREAM_PK_E = 3 };
static uint64_t stream_evk(uint64_t tag, int part, int which) { return (1ull << 32) + tag * 64 + part * 2 + which; }
static uint64_t stream_enc(uint64_t cnt, int which) { return (2ull << 32) + 3 * cnt + which; }

static void sample_small_ntt(Context* c, u64* out, const LimbMods& m, uint64_t stream, int kind) {
    launch_sample_small(c, out, m.n, m, stream, kind);
    launch_ntt(c, out, m.n, m, false);
}

void keygen(Context* c) {
    const int M = c->M(), sq = c->size_q;
    if (!c->d_secret) c->d_secret = dev_alloc(c, words(c, M));
    sample_small_ntt(c, c->d_secret, mods_range(0, M), STREAM_SECRET, 0);
    if (!c->d_pk) c->d_pk = dev_alloc(c, words(c, 2 * sq));
    u64* a = c->dry ? nullptr : c->d_pk + words(c, sq);
    launch_sample_uniform(c, a, sq, mods_q(sq), STREAM_PK_A);
    u64* e = dev_alloc(c, words(c, sq));
    sample_small_ntt(c, e, mods_q(sq), STREAM_PK_E, 1);
    launch_mul(c, c->d_pk, a, c->d_secret, sq, sq, mods_q(sq));      // a*s
    launch_add(c, c->d_pk, e, c->d_pk, sq, mods_q(sq), 1);           // e - a*s
    dev_free(c, e);
}

static void switch_key_gen(Context* c, EvalKey& key, const u64* s_old, const u64* s_new, uint64_t tag) {
    const int M = c->M();
    key.nparts = c->num_part_q;
    if (!key.d) key.d = dev_alloc(c, words(c, key.nparts * 2 * M));
    u64* e = dev_alloc(c, words(c, M));
    for (int part = 0; part < key.nparts; part++) {
        u64* b = c->dry ? nullptr : key.d + words(c, (part * 2 + 0) * M);
        u64* a = c->dry ? nullptr : key.d + words(c, (part * 2 + 1) * M);
        launch_sample_uniform(c, a, M, mods_range(0, M), stream_evk(tag, part, 0));
        sample_small_ntt(c, e, mods_range(0, M), stream_evk(tag, part, 1), 1);
        int lo = c->alpha * part, hi = std::min(c->alpha * (part + 1), c->size_q);
        launch_evk_combine(c, b, a, e, s_new, s_old, lo, hi);
        launch_to_mont(c, b, 2 * M, mods_range(0, M));   // b and a rows are adjacent
    }
    dev_free(c, e);
}

void relin_keygen(Context* c) {
    if (!c->d_secret && !c->dry) throw std::runtime_error("EvalMultKeyGen: no secret key");
    const int M = c->M();
    u64* s2 = dev_alloc(c, words(c, M));
    launch_mul(c, s2, c->d_secret, c->d_secret, M, M, mods_range(0, M));
    switch_key_gen(c, c->relin, s2, c->d_secret, 0);
    dev_free(c, s2);
}

void conj_keygen(Context* c) {
    if (!c->d_secret && !c->dry) throw std::runtime_error("EvalBootstrapKeyGen: no secret key");
    const uint32_t k = 2u * (uint32_t)c->n - 1u;              // its own inverse mod 2n
    if (c->rot.count(k)) return;
    const int M = c->M();
    u64* sp = dev_alloc(c, words(c, M));
    launch_automorph(c, sp, c->d_secret, M, k);
    EvalKey key;
    switch_key_gen(c, key, c->d_secret, sp, k);
    c->rot[k] = key;
    dev_free(c, sp);
}

void rot_keygen(Context* c, const int* indices, int count_) {
    if (!c->d_secret && !c->dry) throw std::runtime_error("EvalRotateKeyGen: no secret key");
    const int M = c->M(); const u64 m = 2ull * c->n;
    u64* sp = dev_alloc(c, words(c, M));
    for (int t = 0; t < count_; t++) {
        uint32_t k = auto_index(c, indices[t]);
        if (c->rot.count(k)) continue;
        uint32_t kinv = auto_index(c, -indices[t]);
        if ((u64)k * kinv % m != 1) throw std::runtime_error("internal: automorphism inverse");
        launch_automorph(c, sp, c->d_secret, M, kinv);
        EvalKey key;
        switch_key_gen(c, key, c->d_secret, sp, k);
        c->rot[k] = key;
    }
    dev_free(c, sp);
}

// bootstrapping's ModRaise: the single limb of `a` (mod q0) is read as centred integers and reduced into every limb of
// `level`; the result is declared noiseScaleDeg 1 at that level's scaling factor.  INTT of the q0 rows, then one forward
// NTT launch whose prologue switches the q0 coefficients to each row's modulus (the rescale path's PRO_SWITCHMOD).
Ct* mod_raise(const Ct* a, int level) {
    Context* c = a->ctx;
    if (a->nl != 1 || a->deg != 1 || a->lrep) throw std::runtime_error("ModRaise: expects one limb at noiseScaleDeg 1");
    if (level < 0 || level >= c->size_q) throw std::runtime_error("ModRaise: bad level");
    ct_wait(a);
    const int nl = c->size_q - level, parts = a->parts;
    count(c, "mod_raise", nl);
    u64* coef = dev_alloc(c, words(c, parts));
    LimbMods m0; m0.n = 1; m0.m[0] = 0;
    launch_ntt_ex(c, coef, a->d, parts, m0, true, nullptr, RowMap{}, RowMap{});
    Ct* out = ct_new(c, parts, nl, level, 1, c->sf_real[level], a->slots);
    NttFuse f; f.pro = PRO_SWITCHMOD; f.pro_mod = 0;
    launch_ntt_ex(c, out->d, coef, parts * nl, mods_q(nl), false, nullptr, RowMap{nl, 1, 0, 0}, RowMap{}, &f);
    dev_free(c, coef);
    return out;
}

// sum of ciphertexts; operands that already share level, limbs and scale (the common case) are added in one pass
Ct* eval_add_many(const std::vector<const Ct*>& cts) {
    if (cts.empty()) throw std::runtime_error("EvalAddMany: empty list");
    const Ct* f = cts[0];
    bool same = cts.size() <= 32 && f->parts == 2 && f->lrep == 0;
    for (size_t i = 1; i < cts.size() && same; i++) {
        const Ct* g = cts[i];
        same = g->parts == f->parts && g->nl == f->nl && g->level == f->level && g->deg == f->deg && g->sf == f->sf && g->lrep == 0;
    }
    if (same && cts.size() > 2) {
        int slots = 0; std::vector<const u64*> ins;
        for (const Ct* g : cts) { ct_wait(g); ins.push_back(g->d); slots = std::max(slots, g->slots); }
        Ct* r = ct_new(f->ctx, f->parts, f->nl, f->level, f->deg, f->sf, slots);
        f->ctx->stats.ops[{"add", f->nl}] += (long)cts.size() - 1;
        launch_add_many(f->ctx, r->d, ins.data(), (int)cts.size(), f->parts * f->nl, f->nl);
        return r;
    }
    std::vector<Ct*> cur; std::vector<bool> owned;
    for (const Ct* g : cts) { cur.push_back(const_cast<Ct*>(g)); owned.push_back(false); }
    while (cur.size() > 1) {   // binary tree, same association as oracle EvalAddMany
        std::vector<Ct*> nxt; std::vector<bool> no;
        for (size_t i = 0; i + 1 < cur.size(); i += 2) {
            nxt.push_back(eval_add(cur[i], cur[i + 1], 1)); no.push_back(true);
            if (owned[i]) ct_free(cur[i]);
            if (owned[i + 1]) ct_free(cur[i + 1]);
        }
        if (cur.size() % 2) { nxt.push_back(cur.back()); no.push_back(owned.back()); }
        cur.swap(nxt); owned.swap(no);
    }
    return owned[0] ? cur[0] : ct_clone(cur[0]);
}

Ct* encode(Context* c, const double* values, int nvals, int scale_deg, int level, int slots, int is_complex) {
    if (slots <= 0) { slots = 1; while (slots < nvals) slots <<= 1; }
    if (nvals > slots || slots > c->n / 2 || (slots & (slots - 1))) throw std::runtime_error("MakeCKKSPackedPlaintext: bad slot count");
    if (level < 0 || level >= c->size_q) throw std::runtime_error("MakeCKKSPackedPlaintext: bad level");
    double sf; int deg;
    if (c->ext && level == 0) { sf = c->sf_big[0]; deg = 2; }
    else { sf = std::pow(c->sf_real[level], (double)scale_deg); deg = scale_deg; }
    const int nl = c->size_q - level;
    // compact form: the plaintext is G(X^gap), gap = n / (2 slots); keep at least 2^8 words per limb
    int lrep = 0;
    while ((2 * slots << lrep) < c->n && c->log_n - lrep > 8) lrep++;
    count(c, "encode", nl);
    if (c->dry || !c->enc_stream) {
        Ct* pt = ct_new(c, 1, nl, level, deg, sf, slots, lrep);
        launch_encode(c, pt->d, values, nvals, slots, sf, nl, lrep, is_complex);
        return pt;
    }
    // the whole encode (allocation, FFT, expansion, NTT) is queued on the encode stream
    cudaStream_t main = c->stream;
    c->stream = c->enc_stream;
    Ct* pt = nullptr;
    try {
        pt = ct_new(c, 1, nl, level, deg, sf, slots, lrep);
        launch_encode(c, pt->d, values, nvals, slots, sf, nl, lrep, is_complex);
        pt->ready = enc_event(c);
        pt->enc_home = true;
    } catch (...) { c->stream = main; throw; }
    c->stream = main;
    return pt;
}

Ct* encrypt(Context* c, const Ct* pt_in) {
    if (pt_in->parts != 1) throw std::runtime_error("Encrypt: operand is not a plaintext");
    Opnd fp(pt_in); op_full(fp); const Ct* pt = fp.v;
    if (!c->d_pk && !c->dry) throw std::runtime_error("Encrypt: no public key");
    const int nl = pt->nl; uint64_t cnt = c->enc_counter++;
    u64* t = dev_alloc(c, words(c, 3 * nl));
    u64 *v = t, *e0 = c->dry ? nullptr : t + words(c, nl), *e1 = c->dry ? nullptr : t + words(c, 2 * nl);
    sample_small_ntt(c, v, mods_q(nl), stream_enc(cnt, 0), 0);
    sample_small_ntt(c, e0, mods_q(nl), stream_enc(cnt, 1), 1);
    sample_small_ntt(c, e1, mods_q(nl), stream_enc(cnt, 2), 1);
    Ct* ct = ct_new(c, 2, nl, pt->level, pt->deg, pt->sf, pt->slots);
    launch_encrypt_combine(c, ct->d, pt->d, v, e0, e1, nl);
    dev_free(c, t);
    return ct;
}

void decrypt(const Ct* ct_in, double* out, int out_len) {
    Opnd fc(ct_in); op_full(fc); const Ct* ct = fc.v;
    Context* c = ct->ctx;
    const int use = std::min(2, ct->nl);
    u64* m = dev_alloc(c, words(c, use));
    if (ct->parts == 1) launch_copy_rows(c, m, ct->d, words(c, use));
    else {
        if (!c->d_secret && !c->dry) throw std::runtime_error("Decrypt: no secret key");
        launch_decrypt_combine(c, m, ct, use);
    }
    launch_ntt(c, m, use, mods_q(use), true);
    std::vector<double> tmp(ct->slots);
    launch_decode(c, tmp.data(), m, use, ct->sf, ct->slots);
    dev_free(c, m);
    for (int i = 0; i < out_len; i++) out[i] = i < ct->slots ? tmp[i] : 0.0;
}
