// capi.cu -- extern "C" boundary (include/ckks_b200.h) over the evaluator.
#include "../../include/ckks_b200.h"
#include "evaluator.h"
#include <cstring>
#include <sstream>
#include <stdexcept>

struct ckks_ctx { Context c; };
struct ckks_ct { mutable Ct* p; };

static thread_local std::string g_err;
extern "C" const char* ckks_last_error(void) { return g_err.c_str(); }

#define API_BEGIN try {
#define API_END                                                             \
    return 0; }                                                             \
    catch (const std::exception& e) { g_err = e.what(); return 1; }         \
    catch (...) { g_err = "unknown error"; return 2; }

static void check_cuda(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}
static void require_device(bool dry) {
    if (dry) return;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        throw std::runtime_error("no CUDA device: the CKKS engine has no CPU fallback (use dry_run=1 for metadata only)");
    }
}
static ckks_ct* box(Ct* p) { return new ckks_ct{p}; }
// Scope of one C entry point: holds the context mutex (engine.h) and, when the entry point released the last handle of a
// closed context, tears the context down after unlocking.
struct Locked {
    Context* c; std::unique_lock<std::recursive_mutex> lk;
    explicit Locked(Context* c_) : c(c_), lk(c_->mu) {}
    explicit Locked(const Context* c_) : Locked(const_cast<Context*>(c_)) {}
    ~Locked() {
        const bool rel = c->release_pending;
        if (rel) c->release_pending = false;
        lk.unlock();
        if (rel) context_release(c);
    }
};
static Context* C(ckks_ctx_t* x) { if (!x) throw std::runtime_error("null context"); return &x->c; }
static const Ct* G(const ckks_ct_t* x) { if (!x || !x->p) throw std::runtime_error("null ciphertext handle"); return x->p; }

void context_release(Context* c) {
    if (!c->dry) {
        if (c->enc_stream) cudaStreamSynchronize(c->enc_stream);
        cudaStreamSynchronize(c->stream);
        arena_trim(c);
        for (auto e : c->ev_pool) cudaEventDestroy(e);
        cudaFree(c->d_mod); cudaFree(c->d_tw); cudaFree(c->d_itw); cudaFree(c->d_ninv_pairs); cudaFree(c->d_twd); cudaFree(c->d_itwd); cudaFree(c->d_twb); cudaFree(c->d_itwb);
        for (auto& kv : c->ninv_pairs_rep) cudaFree(kv.second);
        cudaFree(c->d_ksi); cudaFree(c->d_rot);
        if (c->enc_pinned) cudaFreeHost(c->enc_pinned);
        for (auto e : c->enc_events) if (e) cudaEventDestroy(e);
        for (auto e : c->prof.pool) cudaEventDestroy(e);
        for (auto& kv : c->ks) {
            for (auto& t : kv.second.modup) { cudaFree(t.d_hatinv); cudaFree(t.d_hatinv_sh); cudaFree(t.d_hatmod); cudaFree(t.d_hatplain); }
            cudaFree(kv.second.moddown.d_hatinv); cudaFree(kv.second.moddown.d_hatinv_sh); cudaFree(kv.second.moddown.d_hatmod); cudaFree(kv.second.moddown.d_hatplain);
            cudaFree(kv.second.d_pinv); cudaFree(kv.second.d_pinv_sh); cudaFree(kv.second.d_pinv_pairs);
            cudaFree(kv.second.d_modup_scale); cudaFree(kv.second.d_moddown_scale);
        }
        for (auto& kv : c->rescale_pairs) cudaFree(kv.second);
        for (auto& kv : c->scal_tabs) cudaFree(kv.second);
        dev_free(c, c->d_secret); dev_free(c, c->d_pk); dev_free(c, c->relin.d);
        for (auto& kv : c->rot) dev_free(c, kv.second.d);
        cudaStreamSynchronize(c->stream);
        if (c->own_stream) cudaStreamDestroy(c->stream);
        if (c->enc_stream) cudaStreamDestroy(c->enc_stream);
        cudaGetLastError();
    }
    delete reinterpret_cast<ckks_ctx*>(c);   // Context is the first (only) member of ckks_ctx
}

extern "C" {

int ckks_ctx_create(ckks_ctx_t** out, int log_n, int mult_depth, int scaling_bits, int first_bits, int num_large_digits,
                    int scaling_technique, uint64_t seed, int dry_run) {
    API_BEGIN
    require_device(dry_run != 0);
    if (!dry_run) cudaGetLastError();   // do not inherit a stale error from earlier work in this process
    ckks_ctx* x = new ckks_ctx();
    try { context_generate(&x->c, log_n, mult_depth, scaling_bits, first_bits, num_large_digits, scaling_technique, seed, dry_run); }
    catch (...) { delete x; throw; }
    if (!dry_run) check_cuda("context tables");
    *out = x;
    API_END
}

int ckks_ctx_create_from_moduli(ckks_ctx_t** out, int log_n, int mult_depth, int scaling_bits, int first_bits, int num_large_digits,
                                int scaling_technique, uint64_t seed, int dry_run, const uint64_t* moduli, const uint64_t* roots, int count) {
    API_BEGIN
    require_device(dry_run != 0);
    if (!moduli || count < 1) throw std::runtime_error("ctx_create_from_moduli: no moduli given");
    if (!dry_run) cudaGetLastError();
    ckks_ctx* x = new ckks_ctx();
    try { context_generate(&x->c, log_n, mult_depth, scaling_bits, first_bits, num_large_digits, scaling_technique, seed, dry_run, moduli, roots, count); }
    catch (...) { delete x; throw; }
    if (!dry_run) check_cuda("context tables");
    *out = x;
    API_END
}
static void h2d(Context* c, u64* dev, const uint64_t* host, size_t words);
static void import_rows(Context* c, u64** slot, const uint64_t* host, size_t words_, size_t expect, const char* what) {
    if (words_ != expect) throw std::runtime_error(std::string(what) + ": size mismatch");
    if (!*slot) *slot = dev_alloc(c, expect);
    h2d(c, *slot, host, expect);
}
int ckks_import_secret(ckks_ctx_t* x, const uint64_t* host, size_t words_) {
    API_BEGIN Context* c = C(x); Locked lk_(c); import_rows(c, &c->d_secret, host, words_, (size_t)c->M() * c->n, "import_secret"); API_END
}
int ckks_import_public_key(ckks_ctx_t* x, const uint64_t* host, size_t words_) {
    API_BEGIN Context* c = C(x); Locked lk_(c); import_rows(c, &c->d_pk, host, words_, (size_t)2 * c->size_q * c->n, "import_public_key"); API_END
}
int ckks_ctx_destroy(ckks_ctx_t* x) {
    API_BEGIN
    if (!x) return 0;
    Context* c = &x->c;
    Locked lk_(c);
    if (c->closed) return 0;
    boot_release(c);                                    // cached diagonal plaintexts are handles of this context
    c->closed = true;
    if (c->live_cts == 0) c->release_pending = true;   // otherwise the last ckks_ct_free finishes the teardown
    API_END
}

int ckks_ctx_set_stream(ckks_ctx_t* x, void* s) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (c->dry) return 0;
    cudaStreamSynchronize(c->stream);
    arena_trim(c);                       // the cached temporaries are ordered on the old stream
    if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
    c->stream = (cudaStream_t)s;
    API_END
}
int ckks_ctx_sync(ckks_ctx_t* x) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (!c->dry) {
        if (c->enc_stream) cudaStreamSynchronize(c->enc_stream);
        cudaError_t e = cudaStreamSynchronize(c->stream); if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); check_cuda("sync");
    }
    API_END
}
int ckks_ctx_info(ckks_ctx_t* x, int what, int64_t* out) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    switch (what) {
        case 0: *out = c->n; break;
        case 1: *out = c->size_q; break;
        case 2: *out = c->size_p; break;
        case 3: *out = c->alpha; break;
        case 4: *out = c->num_part_q; break;
        case 5: *out = 2 * (int64_t)c->n; break;
        case 6: *out = c->stats.kernel_launches; break;
        case 7: *out = (int64_t)((c->relin.d ? 1 : 0) + (int64_t)c->rot.size()) * c->num_part_q * 2 * c->M() * c->n * 8; break;
        case 8: *out = (int64_t)c->uid; break;
        case 9: *out = (int64_t)(intptr_t)c->stream; break;
        case 10: { size_t fr = 0, tot = 0; if (c->dry || cudaMemGetInfo(&fr, &tot) != cudaSuccess) fr = 0; *out = (int64_t)fr; break; }
        default: throw std::runtime_error("ckks_ctx_info: unknown selector");
    }
    API_END
}
int ckks_ctx_set_option(ckks_ctx_t* x, int what, int64_t value) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    switch (what) {
        case 0: if (value != 0 && value != 1) throw std::runtime_error("add_const_mode must be 0 or 1"); c->add_const_mode = (int)value; break;
        case 1: if (value < 0) throw std::runtime_error("modup cache size must be >= 0"); c->modup_cap = (int)value; break;
        case 2: if (value < 0) throw std::runtime_error("arena capacity must be >= 0 MiB"); c->arena_cap = (size_t)value << 20;
                if (c->arena_cached > c->arena_cap && !c->dry) arena_trim(c);
                break;
        default: throw std::runtime_error("ckks_ctx_set_option: unknown option");
    }
    API_END
}
int ckks_ctx_get_moduli(ckks_ctx_t* x, uint64_t* moduli, uint64_t* roots, int cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (cap < c->M()) throw std::runtime_error("ckks_ctx_get_moduli: buffer too small");
    for (int i = 0; i < c->M(); i++) { moduli[i] = c->moduli[i]; if (roots) roots[i] = c->psi[i]; }
    API_END
}
int ckks_ctx_get_scaling_factors(ckks_ctx_t* x, double* sf_real, double* sf_big, int cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (cap < c->size_q) throw std::runtime_error("ckks_ctx_get_scaling_factors: buffer too small");
    for (int i = 0; i < c->size_q; i++) sf_real[i] = c->sf_real[i];
    if (sf_big) for (size_t i = 0; i < c->sf_big.size(); i++) sf_big[i] = c->sf_big[i];
    API_END
}

int ckks_keygen(ckks_ctx_t* x) { API_BEGIN Locked lk_(C(x)); keygen(C(x)); API_END }
int ckks_eval_mult_keygen(ckks_ctx_t* x) { API_BEGIN Locked lk_(C(x)); relin_keygen(C(x)); API_END }
int ckks_eval_rotate_keygen(ckks_ctx_t* x, const int32_t* idx, int count) { API_BEGIN Locked lk_(C(x)); rot_keygen(C(x), idx, count); API_END }

static void d2h(Context* c, uint64_t* host, const u64* dev, size_t words) {
    if (c->dry) { memset(host, 0, words * 8); return; }
    cudaMemcpyAsync(host, dev, words * 8, cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) throw std::runtime_error(std::string("device->host copy: ") + cudaGetErrorString(e));
}
static void h2d(Context* c, u64* dev, const uint64_t* host, size_t words) {
    if (c->dry) return;
    cudaMemcpyAsync(dev, host, words * 8, cudaMemcpyHostToDevice, c->stream);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) throw std::runtime_error(std::string("host->device copy: ") + cudaGetErrorString(e));
}

int ckks_export_secret(ckks_ctx_t* x, uint64_t* host, size_t cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c); size_t w = (size_t)c->M() * c->n;
    if (cap < w) throw std::runtime_error("export_secret: buffer too small");
    if (!c->d_secret && !c->dry) throw std::runtime_error("export_secret: no secret key");
    d2h(c, host, c->d_secret, w);
    API_END
}
int ckks_export_public_key(ckks_ctx_t* x, uint64_t* host, size_t cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c); size_t w = (size_t)2 * c->size_q * c->n;
    if (cap < w) throw std::runtime_error("export_public_key: buffer too small");
    if (!c->d_pk && !c->dry) throw std::runtime_error("export_public_key: no public key");
    d2h(c, host, c->d_pk, w);
    API_END
}
static EvalKey* find_key(Context* c, int is_rot, int32_t index, bool create) {
    if (!is_rot) return &c->relin;
    uint32_t k = is_rot == 2 ? (uint32_t)index : auto_index(c, index);   // 2: `index` is the automorphism index itself
    auto it = c->rot.find(k);
    if (it == c->rot.end()) { if (!create) return nullptr; return &c->rot[k]; }
    return &it->second;
}
int ckks_export_eval_key(ckks_ctx_t* x, int is_rot, int32_t index, uint64_t* host, size_t cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    EvalKey* k = find_key(c, is_rot, index, false);
    if (!k || (!k->d && !c->dry)) throw std::runtime_error("export_eval_key: key not present");
    int rows = c->num_part_q * 2 * c->M(); size_t w = (size_t)rows * c->n;
    if (cap < w) throw std::runtime_error("export_eval_key: buffer too small");
    u64* tmp = dev_alloc(c, w);
    launch_copy_rows(c, tmp, k->d, w);
    launch_from_mont(c, tmp, rows, mods_range(0, c->M()));
    d2h(c, host, tmp, w);
    dev_free(c, tmp);
    API_END
}
int ckks_list_automorphism_keys(ckks_ctx_t* x, uint32_t* ks, int cap, int* count) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if ((int)c->rot.size() > cap) throw std::runtime_error("list_automorphism_keys: buffer too small");
    int i = 0;
    for (auto& kv : c->rot) ks[i++] = kv.first;
    *count = i;
    API_END
}
int ckks_import_eval_key(ckks_ctx_t* x, int is_rot, int32_t index, const uint64_t* host, size_t words_) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    int rows = c->num_part_q * 2 * c->M(); size_t w = (size_t)rows * c->n;
    if (words_ != w) throw std::runtime_error("import_eval_key: size mismatch");
    EvalKey* k = find_key(c, is_rot, index, true);
    k->nparts = c->num_part_q;
    if (!k->d) k->d = dev_alloc(c, w);
    h2d(c, k->d, host, w);
    launch_to_mont(c, k->d, rows, mods_range(0, c->M()));
    API_END
}

int ckks_make_packed_plaintext(ckks_ctx_t* x, const double* v, int count, int scale_deg, int level, int slots, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x)); *out = box(encode(C(x), v, count, scale_deg, level, slots)); API_END
}
int ckks_encrypt(ckks_ctx_t* x, const ckks_ct_t* pt, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(encrypt(C(x), G(pt))); API_END }
int ckks_decrypt(ckks_ctx_t* x, const ckks_ct_t* ct, double* v, int count) { API_BEGIN Locked lk_(C(x)); decrypt(G(ct), v, count); check_cuda("decrypt"); API_END }

int ckks_ct_free(ckks_ct_t* ct) { API_BEGIN if (ct) { if (ct->p) { Locked lk_(ct->p->ctx); if (--ct->p->refs == 0) ct_free(ct->p); } delete ct; } API_END }
// Clone is copy-on-write: every evaluator entry point is out of place on its operands, so a clone shares the source's limbs (and its
// cached rescaled / adjusted twins and ModUp digits) until one of the three entry points that change a handle -- SetSlots, the
// mod-reduce after an in-place all-reduce, or handing out the device pointer -- is applied to it.  The sorters seed every accumulator
// with `getZero()->Clone()` (reference src/sort_algo.h:563-566): 64 MB copies at ring 2^17, each followed by a level adjustment of
// the same limbs.  CKKS_B200_EAGER_CLONE=1 restores the copying clone (A/B).
static Ct* unshared(const ckks_ct_t* ct) {
    Ct* p = const_cast<Ct*>(G(ct));
    if (p->refs > 1) { Ct* q = ct_clone(p); p->refs--; ct->p = q; return q; }
    return p;
}
int ckks_ct_clone(const ckks_ct_t* ct, ckks_ct_t** out) {
    API_BEGIN
    static const bool eager = getenv("CKKS_B200_EAGER_CLONE") != nullptr;
    const Ct* p = G(ct); Locked lk_(p->ctx);
    if (eager) *out = box(ct_clone(p));
    else { p->refs++; *out = box(const_cast<Ct*>(p)); }
    API_END
}
int ckks_ct_set_slots(ckks_ct_t* ct, int slots) {
    API_BEGIN
    Locked lk_(G(ct)->ctx);
    if (slots <= 0 || (slots & (slots - 1)) || slots > G(ct)->ctx->n / 2) throw std::runtime_error("SetSlots: bad slot count");
    if (G(ct)->slots == slots) return 0;
    Ct* p = unshared(ct);
    struct Rec { static void set(const Ct* t, int s) {      // cached twins carry the slot count of their source
        const_cast<Ct*>(t)->slots = s;
        if (t->resc) set(t->resc, s);
        if (t->lift) set(t->lift, s);
        for (const AdjTwin& e : t->adj) set(e.t, s);
    } };
    Rec::set(p, slots);
    API_END
}
int ckks_ct_get(const ckks_ct_t* ct, int what, int64_t* out) {
    API_BEGIN
    const Ct* p = G(ct); Locked lk_(p->ctx);
    switch (what) {
        case 0: *out = p->slots; break; case 1: *out = p->level; break; case 2: *out = p->deg; break;
        case 3: *out = p->nl; break; case 4: *out = p->parts; break;
        default: throw std::runtime_error("ckks_ct_get: unknown selector");
    }
    API_END
}
int ckks_ct_get_scaling_factor(const ckks_ct_t* ct, double* out) { API_BEGIN Locked lk_(G(ct)->ctx); *out = G(ct)->sf; API_END }
int ckks_ct_export(const ckks_ct_t* ct, uint64_t* host, size_t cap) {
    API_BEGIN
    const Ct* p = G(ct); Locked lk_(p->ctx); size_t w = (size_t)p->parts * p->nl * p->ctx->n;
    if (cap < w) throw std::runtime_error("ct_export: buffer too small");
    ct_wait(p);
    if (p->lrep) {   // compact plaintext: export the full-length limbs
        Ct* f = ct_full(p);
        try { d2h(p->ctx, host, f->d, w); } catch (...) { ct_free(f); throw; }
        ct_free(f);
    } else d2h(p->ctx, host, p->d, w);
    API_END
}
int ckks_ctx_set_encryption_counter(ckks_ctx_t* x, uint64_t counter) { API_BEGIN Locked lk_(C(x)); C(x)->enc_counter = counter; API_END }
int ckks_ct_device_ptr(const ckks_ct_t* ct, void** dev_ptr, size_t* words) {
    API_BEGIN
    Locked lk_(G(ct)->ctx);
    const Ct* p = unshared(ct);                 // the caller may write through the pointer (shard.py all-reduces in place)
    if (p->lrep) throw std::runtime_error("ct_device_ptr: compact plaintexts have no full-length device form");
    ct_wait(p);
    *dev_ptr = p->d; *words = (size_t)p->parts * p->nl * p->ctx->n;
    API_END
}
int ckks_ct_mod_reduce(ckks_ctx_t* x, ckks_ct_t* ct) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c); Ct* p = unshared(ct);
    if (p->lrep) throw std::runtime_error("ct_mod_reduce: compact plaintext");
    if (p->resc) { ct_free(p->resc); p->resc = nullptr; }            // the limbs are about to change: drop derived caches
    if (p->modup) { dev_free(c, p->modup); p->modup = nullptr; }
    launch_mod_reduce(c, p->d, p->parts * p->nl, p->nl);
    API_END
}
int ckks_ct_import(ckks_ctx_t* x, const uint64_t* host, int polys, int limbs, int level, int deg, double sf, int slots, ckks_ct_t** out) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (polys < 1 || polys > 3 || limbs < 1 || limbs + level != c->size_q) throw std::runtime_error("ct_import: shape does not match context");
    Ct* p = ct_new(c, polys, limbs, level, deg, sf, slots);
    try { h2d(c, p->d, host, (size_t)polys * limbs * c->n); } catch (...) { ct_free(p); throw; }
    *out = box(p);
    API_END
}

int ckks_eval_add(ckks_ctx_t* x, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_add(G(a), G(b), 1)); API_END }
int ckks_eval_sub(ckks_ctx_t* x, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_add(G(a), G(b), -1)); API_END }
int ckks_eval_add_const(ckks_ctx_t* x, const ckks_ct_t* a, double v, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_add_const(G(a), v)); API_END }
int ckks_eval_sub_const(ckks_ctx_t* x, const ckks_ct_t* a, double v, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_add_const(G(a), -v)); API_END }
int ckks_eval_const_sub(ckks_ctx_t* x, double v, const ckks_ct_t* a, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x));
    *out = box(eval_const_sub(v, G(a)));
    API_END
}
int ckks_eval_negate(ckks_ctx_t* x, const ckks_ct_t* a, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_negate(G(a))); API_END }
int ckks_eval_inner_product_plain(ckks_ctx_t* x, const ckks_ct_t* init, const ckks_ct_t* const* cts, const ckks_ct_t* const* pts, int count,
                                  ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x));
    if (count < 0) throw std::runtime_error("EvalInnerProductPlain: negative count");
    std::vector<const Ct*> a(count), p(count);
    for (int i = 0; i < count; i++) { a[i] = G(cts[i]); p[i] = G(pts[i]); }
    *out = box(eval_inner_product_plain(G(init), a.data(), p.data(), count));
    API_END
}
int ckks_eval_inner_product_plain_multi(ckks_ctx_t* x, const ckks_ct_t* init, const ckks_ct_t* const* cts, const ckks_ct_t* const* pts,
                                        int count, int groups, ckks_ct_t** outs) {
    API_BEGIN Locked lk_(C(x));
    if (count < 1 || groups < 1) throw std::runtime_error("EvalInnerProductPlainMulti: count and groups must be positive");
    std::vector<const Ct*> a(count), p((size_t)count * groups);
    for (int i = 0; i < count; i++) a[i] = G(cts[i]);
    for (int i = 0; i < count * groups; i++) p[i] = G(pts[i]);
    std::vector<Ct*> r(groups, nullptr);
    try { eval_inner_product_plain_multi(G(init), a.data(), p.data(), count, groups, r.data()); }
    catch (...) { for (Ct* t : r) if (t) ct_free(t); throw; }
    for (int g = 0; g < groups; g++) outs[g] = box(r[g]);
    API_END
}
int ckks_eval_add_many(ckks_ctx_t* x, const ckks_ct_t* const* cts, int count, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x));
    if (count < 1) throw std::runtime_error("EvalAddMany: empty list");
    std::vector<const Ct*> v;
    for (int i = 0; i < count; i++) v.push_back(G(cts[i]));
    *out = box(eval_add_many(v));
    API_END
}
int ckks_eval_mult(ckks_ctx_t* x, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_mult(G(a), G(b))); API_END }
int ckks_eval_mult_const(ckks_ctx_t* x, const ckks_ct_t* a, double v, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_mult_const(G(a), v)); API_END }
int ckks_eval_square(ckks_ctx_t* x, const ckks_ct_t* a, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_mult(G(a), G(a))); API_END }
int ckks_eval_rotate(ckks_ctx_t* x, const ckks_ct_t* a, int32_t index, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_rotate(G(a), index)); API_END }
int ckks_fast_rotation_precompute(ckks_ctx_t* x, const ckks_ct_t* a, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(fast_rotation_precompute(G(a))); API_END }
int ckks_fast_rotation(ckks_ctx_t* x, const ckks_ct_t* a, int32_t index, const ckks_ct_t* precomp, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x)); *out = box(eval_fast_rotate(G(a), index, G(precomp))); API_END
}
int ckks_has_rotation_key(ckks_ctx_t* x, int32_t index) {
    try { Context* c = C(x); Locked lk_(c); return c->rot.count(auto_index(c, index)) ? 1 : 0; } catch (...) { return 0; }
}
int ckks_eval_chebyshev_series_ps(ckks_ctx_t* x, const ckks_ct_t* ct, const double* coeffs, int count, double a, double b, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x));
    if (count < 2) throw std::runtime_error("EvalChebyshevSeriesPS: need at least 2 coefficients");
    *out = box(eval_chebyshev_series_ps(G(ct), std::vector<double>(coeffs, coeffs + count), a, b));
    API_END
}
int ckks_eval_chebyshev_series(ckks_ctx_t* x, const ckks_ct_t* ct, const double* coeffs, int count, double a, double b, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x));
    if (count < 2) throw std::runtime_error("EvalChebyshevSeries: need at least 2 coefficients");
    *out = box(eval_chebyshev_series(G(ct), std::vector<double>(coeffs, coeffs + count), a, b));
    API_END
}
int ckks_eval_poly_linear(ckks_ctx_t* x, const ckks_ct_t* ct, const double* coeffs, int count, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x)); *out = box(eval_poly_linear(G(ct), std::vector<double>(coeffs, coeffs + count))); API_END
}

int ckks_make_packed_plaintext_complex(ckks_ctx_t* x, const double* re_im, int count, int scale_deg, int level, int slots, ckks_ct_t** out) {
    API_BEGIN Locked lk_(C(x)); *out = box(encode(C(x), re_im, count, scale_deg, level, slots, 1)); API_END
}
int ckks_eval_conjugate(ckks_ctx_t* x, const ckks_ct_t* a, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(eval_conjugate(G(a))); API_END }
int ckks_mod_raise(ckks_ctx_t* x, const ckks_ct_t* a, int level, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(mod_raise(G(a), level)); API_END }
int ckks_eval_bootstrap_setup(ckks_ctx_t* x, int budget_c2s, int budget_s2c, int dim1_c2s, int dim1_s2c, int slots) {
    API_BEGIN Locked lk_(C(x)); (void)dim1_c2s; (void)dim1_s2c; boot_setup(C(x), budget_c2s, budget_s2c, slots); API_END
}
int ckks_eval_bootstrap_keygen(ckks_ctx_t* x, int slots) { API_BEGIN Locked lk_(C(x)); boot_keygen(C(x), slots); API_END }
int ckks_eval_bootstrap(ckks_ctx_t* x, const ckks_ct_t* a, ckks_ct_t** out) { API_BEGIN Locked lk_(C(x)); *out = box(boot_eval(G(a))); API_END }

int ckks_stats_json(ckks_ctx_t* x, char* buf, size_t cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    std::ostringstream os;
    os << "{\"kernel_launches\": " << c->stats.kernel_launches << ", \"host_ns\": [" << c->stats.host_ns[0] << ", " << c->stats.host_ns[1] << ", " << c->stats.host_ns[2] << ", " << c->stats.host_ns[3]
       << "], \"host_calls\": [" << c->stats.host_calls[0] << ", " << c->stats.host_calls[1] << ", " << c->stats.host_calls[2] << ", " << c->stats.host_calls[3] << "], \"ops\": [";
    bool first = true;
    for (auto& kv : c->stats.ops) { os << (first ? "" : ", ") << "[\"" << kv.first.first << "\", " << kv.first.second << ", " << kv.second << "]"; first = false; }
    os << "]}";
    std::string s = os.str();
    if (s.size() + 1 > cap) throw std::runtime_error("stats_json: buffer too small");
    memcpy(buf, s.c_str(), s.size() + 1);
    API_END
}
int ckks_stats_reset(ckks_ctx_t* x) { API_BEGIN Locked lk_(C(x)); C(x)->stats = Stats(); API_END }

int ckks_profile_enable(ckks_ctx_t* x, unsigned mask) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    prof_collect(c);
    c->prof.mask = mask;
    API_END
}
int ckks_profile_json(ckks_ctx_t* x, char* buf, size_t cap) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    prof_collect(c);
    static const char* names[KC_COUNT] = {"ntt", "baseconv", "ks_inner", "moddown", "pointwise", "rescale", "automorph",
                                          "encode", "copy"};
    std::ostringstream os;
    os.precision(17);
    os << "{";
    for (int i = 0; i < KC_COUNT; i++) {
        os << (i ? ", " : "") << "\"" << names[i] << "\": [" << c->prof.ms[i] << ", " << c->prof.launches[i] << ", "
           << c->prof.bytes[i] << "]";
        c->prof.ms[i] = 0; c->prof.launches[i] = 0; c->prof.bytes[i] = 0;
    }
    os << "}";
    std::string s = os.str();
    if (s.size() + 1 > cap) throw std::runtime_error("profile_json: buffer too small");
    memcpy(buf, s.c_str(), s.size() + 1);
    API_END
}

int ckks_rng_block(const uint32_t* key8, uint64_t stream, uint64_t index, uint32_t* out16) {
    API_BEGIN
    if (!key8 || !out16) throw std::runtime_error("rng_block: null buffer");
    chacha20_block(key8, stream, index, out16);          // the host instance of the function the sampling kernels run
    API_END
}
int ckks_ntt_host(ckks_ctx_t* x, uint64_t* data, int rows, const int32_t* mod_ids, int inverse) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (rows < 1 || rows > MAX_LIMBS) throw std::runtime_error("ntt_host: bad row count");
    LimbMods m; m.n = rows;
    for (int i = 0; i < rows; i++) { if (mod_ids[i] < 0 || mod_ids[i] >= c->M()) throw std::runtime_error("ntt_host: bad modulus id"); m.m[i] = (short)mod_ids[i]; }
    size_t w = (size_t)rows * c->n;
    u64* d = dev_alloc(c, w);
    h2d(c, d, data, w);
    launch_ntt(c, d, rows, m, inverse != 0);
    d2h(c, data, d, w);
    dev_free(c, d);
    if (!c->dry) check_cuda("ntt");
    API_END
}
int ckks_ntt_dev(ckks_ctx_t* x, void* dev, int rows, int first_mod, int inverse) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (rows < 1 || first_mod < 0 || first_mod + rows > c->M()) throw std::runtime_error("ntt_dev: bad limb range");
    launch_ntt(c, (u64*)dev, rows, mods_range(first_mod, first_mod + rows), inverse != 0);
    API_END
}
int ckks_key_switch_dev(ckks_ctx_t* x, const void* src, int limbs, int is_rot, int32_t index, void* out) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    EvalKey* k = find_key(c, is_rot, index, false);
    if (!k) throw std::runtime_error("key_switch_dev: key not present");
    key_switch(c, (const u64*)src, limbs, *k, (u64*)out, nullptr, nullptr, 0);
    API_END
}

// ---------------------------------------------------------------- allocation-free key switch (no reference counterpart)
// ckks_key_switch_dev_noalloc: the schedule of key_switch() (evaluator.cu) with every temporary -- INTT output, ModUp
// complements, evk inner-product accumulators, ModDown P limbs and their conversion -- carved out of ONE file-scope device
// array instead of the stream-ordered pool, for callers that may not allocate (CUDA graph capture, serving loops).
// The array is shared by every call on a device and keeps the previous call's contents; every region is fully written
// before it is read, so nothing needs clearing, but two calls that overlap on the GPU (different streams) race.
#define NOALLOC_SCRATCH_WORDS (1u << 21)   /* 16 MiB per GPU that loads the library, held until exit */
static __device__ __align__(256) u64 g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];
static std::mutex g_noalloc_mu;
static u64* g_noalloc_ptr[64];             // device address of the array, per device ordinal (the host symbol is NOT an address)

static u64* noalloc_scratch_of_current_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) throw std::runtime_error(std::string("key_switch_dev_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) throw std::runtime_error("key_switch_dev_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> g(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {             // set-up, once per device, with that device current
        u64* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) throw std::runtime_error(std::string("key_switch_dev_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_noalloc_ptr[dev] = p;
    }
    return g_noalloc_ptr[dev];
}

int ckks_key_switch_dev_noalloc(ckks_ctx_t* x, const void* src, int limbs, int is_rot, int32_t index, void* out) {
    API_BEGIN
    Context* c = C(x); Locked lk_(c);
    if (c->dry) throw std::runtime_error("key_switch_dev_noalloc: needs a device (dry-run context)");
    if (!src || !out) throw std::runtime_error("key_switch_dev_noalloc: null buffer");
    if (limbs < 1 || limbs > c->size_q) throw std::runtime_error("key_switch_dev_noalloc: bad limb count");
    const EvalKey* k = find_key(c, is_rot, index, false);
    if (!k || !k->d) throw std::runtime_error("key_switch_dev_noalloc: key not present");
    if (c->prof.mask) throw std::runtime_error("key_switch_dev_noalloc: per-class profiling creates CUDA events; disable it (ckks_profile_enable(ctx, 0))");
    auto it = c->ks.find(limbs);
    if (it == c->ks.end())
        throw std::runtime_error("key_switch_dev_noalloc: the constant tables of this limb count are allocated on first use; run "
                                 "ckks_key_switch_dev (or any EvalMult / EvalRotate at this level) once during set-up");
    const KsTables& t = it->second;
    const int nl = limbs, K = c->size_p, next = nl + K;
    const size_t n = (size_t)c->n;
    if (K > MAX_CONV_IN || c->alpha > MAX_CONV_IN || next > MAX_CONV_OUT) throw std::runtime_error("key switch: parameter set too large");
    int conv_rows = 0;
    for (int j = 0; j < t.nparts; j++) conv_rows += t.modup[j].nout;
    const size_t rows = (size_t)nl + conv_rows + 2 * (size_t)next + 2 * (size_t)K + 2 * (size_t)nl;
    if (rows * n > NOALLOC_SCRATCH_WORDS)
        throw std::runtime_error("key_switch_dev_noalloc: needs " + std::to_string(rows) + " scratch rows of " + std::to_string(n) +
                                 " words = " + std::to_string(rows * n * 8) + " bytes, beyond the static scratch of " +
                                 std::to_string((size_t)NOALLOC_SCRATCH_WORDS * 8) + " bytes; use ckks_key_switch_dev");
    u64* coef = noalloc_scratch_of_current_device();   // [nl][n]        INTT of src (x n^-1 Qhat^-1)
    u64* conv = coef + (size_t)nl * n;                  // [conv_rows][n] ModUp complements of every digit
    u64* acc = conv + (size_t)conv_rows * n;            // [2][nl+K][n]   sum_j d_j * evk_j
    u64* pp = acc + 2 * (size_t)next * n;               // [2][K][n]      INTT of the P limbs (x n^-1 Phat^-1)
    u64* conv2 = pp + 2 * (size_t)K * n;                // [2][nl][n]     their conversion to Q_l
    c->stats.ops[{"keyswitch_noalloc", nl}]++;
    // ---- ModUp
    launch_ntt_ex(c, coef, (const u64*)src, nl, mods_q(nl), true, t.d_modup_scale, RowMap{}, RowMap{});
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
        else
            for (int j = 0, r0 = 0; j < t.nparts; r0 += t.modup[j].nout, j++) {
                LimbMods mj; mj.n = t.modup[j].nout;
                for (int i = 0; i < mj.n; i++) mj.m[i] = t.modup[j].out_ids[i];
                launch_ntt(c, conv + (size_t)r0 * n, mj.n, mj, false);
            }
    }
    // ---- inner product with the evaluation key, ModDown of both polynomials
    launch_ks_inner2(c, acc, (const u64*)src, conv, *k, t);
    {
        LimbMods mp; mp.n = K;
        for (int i = 0; i < K; i++) mp.m[i] = (short)(c->size_q + i);
        launch_ntt_ex(c, pp, acc, 2 * K, mp, true, t.d_moddown_scale, RowMap{K, next, nl}, RowMap{});
    }
    {
        ConvBatch B; B.ng = 2; B.split_ok = t.moddown.split_ok;
        for (int h = 0; h < 2; h++) {
            B.g[h] = ConvGroup{h * K, K, h * nl, nl, t.moddown.d_hatmod, t.moddown.d_hatplain, t.moddown.wide};
            for (int i = 0; i < nl; i++) B.out_ids[h][i] = (short)i;
        }
        launch_conv_batch(c, conv2, pp, B);
    }
    launch_ntt(c, conv2, 2 * nl, mods_q(nl), false);
    launch_moddown_final(c, (u64*)out, acc, conv2, nullptr, nullptr, nl, 0, t);
    check_cuda("key_switch_dev_noalloc");
    API_END
}
}  // extern "C"
