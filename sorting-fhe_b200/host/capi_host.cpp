// capi_host.cpp -- extern "C" boundary of the sorter layer (include/sortfhe_host.h).
#include "../../include/sortfhe_host.h"
#include <cstring>
#include <map>
#include <mutex>
#include "sort_algo.h"
#include "mehp24.h"
#include "kway.h"

namespace {
thread_local std::string g_err;
typedef Ciphertext<DCRTPoly> CT;

#define HOST_BEGIN try {
#define HOST_END return 0; } catch (const std::exception& e) { g_err = e.what(); return 1; } catch (...) { g_err = "unknown error"; return 2; }

// One persistent CryptoContext wrapper per (engine context, batch size): it carries the encoded-plaintext cache
// (openfhe_shim.h), which has to outlive the individual C calls.  The engine's process-unique context id guards against a
// new context reusing the address of a destroyed one; sortfhe_ctx_release drops the wrappers (and with them the cached
// plaintext handles that keep the engine context alive) -- callers invoke it before ckks_ctx_destroy.
std::mutex g_reg_mu;
struct RegEntry { int64_t uid = 0; std::map<int, CryptoContext<DCRTPoly>> by_batch; };
std::map<ckks_ctx_t*, RegEntry> g_reg;
CryptoContext<DCRTPoly> borrow(ckks_ctx_t* c, int batch) {
    if (!c) throw std::runtime_error("null context");
    int64_t uid = 0;
    ckks_check(ckks_ctx_info(c, 8, &uid), "ctx_info");
    std::lock_guard<std::mutex> g(g_reg_mu);
    RegEntry& e = g_reg[c];
    if (e.uid != uid) { e.by_batch.clear(); e.uid = uid; }
    auto& w = e.by_batch[batch];
    if (!w) w = std::make_shared<CryptoContextImpl<DCRTPoly>>(c, false, (uint32_t)batch);
    return w;
}
CT borrow(const ckks_ct_t* h) {
    if (!h) throw std::runtime_error("null ciphertext handle");
    return std::make_shared<CiphertextImpl<DCRTPoly>>(const_cast<ckks_ct_t*>(h), false);
}
ckks_ct_t* release(const CT& c) {   // hand a result to the caller (clone if the sorter returned a borrowed input)
    if (!c->owned) { ckks_ct_t* o = nullptr; ckks_check(ckks_ct_clone(c->h, &o), "clone"); return o; }
    c->owned = false;
    return c->h;
}
SignConfig make_cfg(int n, int dg, int df) { return SignConfig(CompositeSignConfig(n, dg, df)); }

// run f.template operator()<N>() for the runtime array size n
template <class F> auto dispatch_n(int n, F&& f) {
    switch (n) {
        case 4: return f(std::integral_constant<int, 4>());
        case 8: return f(std::integral_constant<int, 8>());
        case 16: return f(std::integral_constant<int, 16>());
        case 32: return f(std::integral_constant<int, 32>());
        case 64: return f(std::integral_constant<int, 64>());
        case 128: return f(std::integral_constant<int, 128>());
        case 256: return f(std::integral_constant<int, 256>());
        case 512: return f(std::integral_constant<int, 512>());
        case 1024: return f(std::integral_constant<int, 1024>());
        default: throw std::runtime_error("array size must be a power of two in [4, 1024]");
    }
}

template <int N> std::unique_ptr<DirectSort<N>> make_sorter(ckks_ctx_t* ctx, const int32_t* idx, int cnt) {
    auto cc = borrow(ctx, N);
    auto key = std::make_shared<KeyImpl>(KeyImpl{cc.get()});
    auto enc = std::make_shared<Encryption>(cc, key);
    return std::unique_ptr<DirectSort<N>>(new DirectSort<N>(cc, key, std::vector<int>(idx, idx + cnt), enc));
}
}  // namespace

extern "C" {
const char* sortfhe_last_error(void) { return g_err.c_str(); }

int sortfhe_ctx_release(ckks_ctx_t* ctx) {
    HOST_BEGIN
    std::map<int, CryptoContext<DCRTPoly>> drop;
    {
        std::lock_guard<std::mutex> g(g_reg_mu);
        auto it = g_reg.find(ctx);
        if (it != g_reg.end()) { drop.swap(it->second.by_batch); g_reg.erase(it); }
    }
    drop.clear();   // frees the cached plaintext handles outside the registry lock
    HOST_END
}
int sortfhe_plaintext_cache_stats(ckks_ctx_t* ctx, int64_t* entries, int64_t* bytes, int64_t* hits, int64_t* misses) {
    HOST_BEGIN
    *entries = *bytes = *hits = *misses = 0;
    std::lock_guard<std::mutex> g(g_reg_mu);
    auto it = g_reg.find(ctx);
    if (it != g_reg.end())
        for (auto& kv : it->second.by_batch) {
            std::lock_guard<std::mutex> g2(kv.second->pt_mu);
            *entries += (int64_t)kv.second->pt_cache.size(); *bytes += (int64_t)kv.second->pt_bytes;
            *hits += (int64_t)kv.second->pt_hits; *misses += (int64_t)kv.second->pt_misses;
        }
    HOST_END
}

int sortfhe_direct_sort_params(int n, int ring_dim, int* mult_depth, int32_t* rotations, int cap, int* count) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        CCParams<CryptoContextCKKSRNS> p; p.SetRingDim((uint32_t)ring_dim);
        std::vector<int> r;
        DirectSort<N>::getSizeParameters(p, r);
        if ((int)r.size() > cap) throw std::runtime_error("direct_sort_params: rotation buffer too small");
        *mult_depth = (int)p.GetMultiplicativeDepth(); *count = (int)r.size();
        for (size_t i = 0; i < r.size(); i++) rotations[i] = r[i];
        return 0;
    });
    HOST_END
}

int sortfhe_direct_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sn, int dg, int df, const int32_t* idx, int cnt,
                        ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        SignConfig cfg = make_cfg(sn, dg, df);
        *out = release(s->sort(borrow(input), SignFunc::CompositeSign, cfg));
        return 0;
    });
    HOST_END
}
int sortfhe_construct_rank(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sn, int dg, int df, const int32_t* idx, int cnt,
                           ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        SignConfig cfg = make_cfg(sn, dg, df);
        *out = release(s->constructRank(borrow(input), SignFunc::CompositeSign, cfg));
        return 0;
    });
    HOST_END
}
int sortfhe_rotation_index_check(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n, const int32_t* idx,
                                 int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        *out = release(s->rotationIndexCheckN(borrow(rank), borrow(input)));
        return 0;
    });
    HOST_END
}

int sortfhe_hybrid_sort_params(int n, int variant, int ring_dim, int* mult_depth, int32_t* rotations, int cap, int* count) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        CCParams<CryptoContextCKKSRNS> p; p.SetRingDim((uint32_t)ring_dim);
        std::vector<int> r;
        DirectSort<N>::getHybridParameters(variant, p, r);
        if ((int)r.size() > cap) throw std::runtime_error("hybrid_sort_params: rotation buffer too small");
        *mult_depth = (int)p.GetMultiplicativeDepth(); *count = (int)r.size();
        for (size_t i = 0; i < r.size(); i++) rotations[i] = r[i];
        return 0;
    });
    HOST_END
}
int sortfhe_hybrid_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int variant, int sn, int dg, int df, const int32_t* idx,
                        int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    if (variant < 0 || variant > 2) throw std::runtime_error("hybrid sort: variant must be 0, 1 or 2");
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        SignConfig cfg = make_cfg(sn, dg, df);
        typedef typename DirectSort<N>::HybridIndicator K;
        *out = release(s->sort_hybrid(borrow(input), SignFunc::CompositeSign, cfg,
                                      variant == 0 ? K::SincOrCompare : variant == 1 ? K::Mehp24 : K::Sinc));
        return 0;
    });
    HOST_END
}
int sortfhe_hybrid_index_check(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n, int variant,
                               const int32_t* idx, int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    if (variant < 0 || variant > 2) throw std::runtime_error("hybrid sort: variant must be 0, 1 or 2");
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        typedef typename DirectSort<N>::HybridIndicator K;
        *out = release(s->rotationIndexCheckHybrid(borrow(rank), borrow(input),
                                                   variant == 0 ? K::SincOrCompare : variant == 1 ? K::Mehp24 : K::Sinc));
        return 0;
    });
    HOST_END
}

int sortfhe_compare(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, int sn, int dg, int df, ckks_ct_t** out) {
    HOST_BEGIN
    SignConfig cfg = make_cfg(sn, dg, df);
    Comparison comp;
    *out = release(comp.compare(borrow(ctx, 0), borrow(a), borrow(b), SignFunc::CompositeSign, cfg));
    HOST_END
}
int sortfhe_sign(ckks_ctx_t* ctx, const ckks_ct_t* x, int sn, int dg, int df, ckks_ct_t** out) {
    HOST_BEGIN
    *out = release(sign(borrow(x), borrow(ctx, 0), SignFunc::CompositeSign, make_cfg(sn, dg, df)));
    HOST_END
}
int sortfhe_sign_func(ckks_ctx_t* ctx, const ckks_ct_t* x, int func, int sn, int dg, int df, ckks_ct_t** out) {
    HOST_BEGIN
    if (func < 0 || func > 3) throw std::runtime_error("sign_func: func must be 0..3 (SignFunc order of src/sign.h:6-11)");
    *out = release(sign(borrow(x), borrow(ctx, 0), (SignFunc)func, make_cfg(sn, dg, df)));
    HOST_END
}
int sortfhe_indicator(ckks_ctx_t* ctx, const ckks_ct_t* x, double c, int sn, int dg, int df, ckks_ct_t** out) {
    HOST_BEGIN
    SignConfig cfg = make_cfg(sn, dg, df);
    Comparison comp;
    *out = release(comp.indicator(borrow(ctx, 0), borrow(x), c, SignFunc::CompositeSign, cfg));
    HOST_END
}

int sortfhe_rotate(ckks_ctx_t* ctx, const ckks_ct_t* x, int n, int rotation, const int32_t* idx, int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        RotationComposer<N> rc(borrow(ctx, 0), nullptr, std::vector<int>(idx, idx + cnt));
        *out = release(rc.rotate(borrow(x), rotation));
        return 0;
    });
    HOST_END
}
int sortfhe_tree_rotate(ckks_ctx_t* ctx, const ckks_ct_t* x, int n, int start, int end, int algo, const int32_t* idx, int cnt,
                        ckks_ct_t** outs /*[end - start + 1]*/) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        RotationTree<N> tree(borrow(ctx, 0), std::vector<int>(idx, idx + cnt),
                             algo == 0 ? DecomposeAlgo::NAF : algo == 1 ? DecomposeAlgo::BNAF : DecomposeAlgo::BINARY);
        tree.buildTree(start, end);
        CT in = borrow(x);
        for (int r = start; r <= end; r++) outs[r - start] = nullptr;
        for (int r = start; r <= end; r++) {
            ckks_ct_t* o = nullptr;                     // results may be shared tree nodes (or the input): hand out clones
            ckks_check(ckks_ct_clone(tree.treeRotate(in, r)->h, &o), "clone");
            outs[r - start] = o;
        }
        return 0;
    });
    HOST_END
}
int sortfhe_decompose(int n, int rotation, int wrap, int algo, const int32_t* idx, int cnt, int32_t* steps, int cap, int* count) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        Decomposer<N> d(std::vector<int>(idx, idx + cnt));
        auto st = d.decompose(rotation, wrap, algo == 0 ? DecomposeAlgo::NAF : algo == 1 ? DecomposeAlgo::BNAF : DecomposeAlgo::BINARY);
        if ((int)st.size() > cap) throw std::runtime_error("decompose: step buffer too small");
        *count = (int)st.size();
        for (size_t i = 0; i < st.size(); i++) steps[i] = st[i].stepSize;
        return 0;
    });
    HOST_END
}

int sortfhe_sinc_coefficients(int n, int doubled, double* coeffs, int cap, int* count) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        const std::vector<double>& c = doubled ? selectDoubledSincCoefficients<N>() : selectCoefficients<N>();
        *count = (int)c.size();
        if (coeffs) {
            if ((int)c.size() > cap) throw std::runtime_error("sinc_coefficients: buffer too small");
            memcpy(coeffs, c.data(), c.size() * sizeof(double));
        }
        return 0;
    });
    HOST_END
}

// ---- shardable halves of DirectSort (multi-GPU: SURVEY 8e row 2)
int sortfhe_num_batches(int n, int ring_dim, int* num_batch) {
    HOST_BEGIN
    const int P = std::min(n, ring_dim / 2 / n);
    if (P < 1) throw std::runtime_error("num_batches: array does not fit the ring");
    *num_batch = n / P;
    HOST_END
}
int sortfhe_construct_rank_partial(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sn, int dg, int df, const int32_t* idx, int cnt,
                                   int b_lo, int b_hi, int seed, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        SignConfig cfg = make_cfg(sn, dg, df);
        *out = release(s->constructRankPartial(borrow(input), SignFunc::CompositeSign, cfg, b_lo, b_hi, seed != 0));
        return 0;
    });
    HOST_END
}
int sortfhe_finish_rank(ckks_ctx_t* ctx, const ckks_ct_t* partial, int n, const int32_t* idx, int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        *out = release(s->finishRank(borrow(partial)->Clone()));
        return 0;
    });
    HOST_END
}
int sortfhe_index_check_partial(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n, const int32_t* idx, int cnt,
                                int b_lo, int b_hi, int seed, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        *out = release(s->rotationIndexCheckPartial(borrow(rank), borrow(input), b_lo, b_hi, seed != 0));
        return 0;
    });
    HOST_END
}
int sortfhe_finish_index_check(ckks_ctx_t* ctx, const ckks_ct_t* partial, int n, const int32_t* idx, int cnt, ckks_ct_t** out) {
    HOST_BEGIN
    dispatch_n(n, [&](auto tag) {
        constexpr int N = decltype(tag)::value;
        auto s = make_sorter<N>(ctx, idx, cnt);
        *out = release(s->finishIndexCheck(borrow(partial)->Clone()));
        return 0;
    });
    HOST_END
}

// ---- MEHP24 (reference src/mehp24/)
int sortfhe_mehp24_params(int n, int* mult_depth, int32_t* rotations, int cap, int* count) {
    HOST_BEGIN
    std::vector<int32_t> r = mehp24::utils::getRotationIndices((size_t)n);
    if ((int)r.size() > cap) throw std::runtime_error("mehp24_params: rotation buffer too small");
    *mult_depth = mehp24::utils::fixtureDepth(n); *count = (int)r.size();
    for (size_t i = 0; i < r.size(); i++) rotations[i] = r[i];
    HOST_END
}
int sortfhe_mehp24_sort_fg(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sn, int dg, int df, int dg_i, int df_i, ckks_ct_t** out) {
    HOST_BEGIN
    if (n < 2 || (n & (n - 1))) throw std::runtime_error("mehp24 sortFG: vector length must be a power of two");
    auto cc = borrow(ctx, n * n);                       // batch size N^2 (tests/mehp24/Mehp24SortTest.cpp:36-37)
    SignConfig cfg = make_cfg(sn, dg, df);
    std::unique_ptr<Comparison> comp(new Comparison());
    *out = release(mehp24::sortFG(borrow(input), (size_t)n, SignFunc::CompositeSign, cfg, comp, (uint32_t)dg_i, (uint32_t)df_i, cc));
    HOST_END
}
int sortfhe_mehp24_sort_large(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, int sn, int dg, int df, int dg_i, int df_i,
                              ckks_ct_t** out) {
    HOST_BEGIN
    if (sub_len < 2 || (sub_len & (sub_len - 1)) || total_len < sub_len || total_len % sub_len)
        throw std::runtime_error("mehp24 sortLargeArrayFG: sub length must be a power of two dividing the total length");
    auto cc = borrow(ctx, sub_len * sub_len);            // one sub_len x sub_len matrix per ciphertext
    SignConfig cfg = make_cfg(sn, dg, df);
    std::unique_ptr<Comparison> comp(new Comparison());
    *out = release(mehp24::sortLargeArrayFG(borrow(input), (size_t)total_len, (size_t)sub_len, SignFunc::CompositeSign, cfg, comp,
                                            (uint32_t)dg_i, (uint32_t)df_i, cc));
    HOST_END
}
// multi-ciphertext sortFG in stages (mehp24.h namespace blocks): the shardable halves behind shard.py's mehp24_sort_large_sharded
namespace {
std::vector<CT> borrow_all(const ckks_ct_t* const* h, int cnt, int first = 0) {
    std::vector<CT> v((size_t)cnt);
    for (int i = first; i < cnt; i++) v[(size_t)i] = borrow(h[i]);
    return v;
}
void check_blocks(int nc, int sub_len) {
    if (nc < 1 || sub_len < 2 || (sub_len & (sub_len - 1))) throw std::runtime_error("mehp24 blocks: sub length must be a power of two, at least one block");
}
}  // namespace
// the overloads that do not go through Comparison (mehp24_sort.cpp:23-47, 49-217, 219-249, 285-443, 607-623)
int sortfhe_mehp24_sort_fg_adv(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int dg_c, int df_c, int dg_i, int df_i, ckks_ct_t** out) {
    HOST_BEGIN
    if (n < 2 || (n & (n - 1))) throw std::runtime_error("mehp24 sortFG: vector length must be a power of two");
    auto cc = borrow(ctx, n * n);
    *out = release(mehp24::sortFG(borrow(input), (size_t)n, (uint32_t)dg_c, (uint32_t)df_c, (uint32_t)dg_i, (uint32_t)df_i, cc));
    HOST_END
}
int sortfhe_mehp24_sort_large_adv(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, int dg_c, int df_c, int dg_i, int df_i,
                                  ckks_ct_t** out) {
    HOST_BEGIN
    if (sub_len < 2 || (sub_len & (sub_len - 1)) || total_len < sub_len || total_len % sub_len)
        throw std::runtime_error("mehp24 sortLargeArrayFG: sub length must be a power of two dividing the total length");
    auto cc = borrow(ctx, sub_len * sub_len);
    *out = release(mehp24::sortLargeArrayFG(borrow(input), (size_t)total_len, (size_t)sub_len, (uint32_t)dg_c, (uint32_t)df_c, (uint32_t)dg_i,
                                            (uint32_t)df_i, cc));
    HOST_END
}
int sortfhe_mehp24_sort_cheb(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, double left_bound_c, double right_bound_c, int degree_c, int degree_i,
                             ckks_ct_t** out) {
    HOST_BEGIN
    if (n < 2 || (n & (n - 1))) throw std::runtime_error("mehp24 sort: vector length must be a power of two");
    auto cc = borrow(ctx, n * n);
    *out = release(mehp24::sort(borrow(input), (size_t)n, left_bound_c, right_bound_c, (uint32_t)degree_c, (uint32_t)degree_i, cc));
    HOST_END
}
int sortfhe_mehp24_sort_cheb_multi(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, double left_bound_c, double right_bound_c,
                                   int degree_c, int degree_i, ckks_ct_t** out) {
    HOST_BEGIN
    check_blocks(nc, sub_len);
    auto cc = borrow(ctx, sub_len * sub_len);
    std::vector<CT> r = mehp24::sort(borrow_all(parts, nc), (size_t)sub_len, left_bound_c, right_bound_c, (uint32_t)degree_c, (uint32_t)degree_i, cc);
    for (size_t i = 0; i < r.size(); i++) out[i] = release(r[i]);
    HOST_END
}
int sortfhe_mehp24_split(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, ckks_ct_t** out) {
    HOST_BEGIN
    if (sub_len < 2 || (sub_len & (sub_len - 1)) || total_len < sub_len || total_len % sub_len)
        throw std::runtime_error("mehp24 splitCiphertext: sub length must be a power of two dividing the total length");
    auto cc = borrow(ctx, sub_len * sub_len);
    std::vector<CT> parts = mehp24::splitCiphertext(borrow(input), (size_t)total_len, (size_t)sub_len, cc);
    for (size_t i = 0; i < parts.size(); i++) out[i] = release(parts[i]);
    HOST_END
}
int sortfhe_mehp24_combine(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, ckks_ct_t** out) {
    HOST_BEGIN
    check_blocks(nc, sub_len);
    auto cc = borrow(ctx, sub_len * sub_len);
    *out = release(mehp24::combineCiphertext(borrow_all(parts, nc), (size_t)sub_len, cc));
    HOST_END
}
int sortfhe_mehp24_block_compare_partial(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, int sn, int dg, int df,
                                         int u_lo, int u_hi, ckks_ct_t** cv_out, ckks_ct_t** ch_out) {
    HOST_BEGIN
    check_blocks(nc, sub_len);
    if (u_lo < 0 || u_hi < u_lo || u_hi > (int)mehp24::blocks::numCompareUnits((size_t)nc))
        throw std::runtime_error("mehp24 block compare: unit range outside [0, nc (nc + 1) / 2]");
    auto cc = borrow(ctx, sub_len * sub_len);
    SignConfig cfg = make_cfg(sn, dg, df);
    std::unique_ptr<Comparison> comp(new Comparison());
    std::vector<CT> replR, replC, Cv, Ch;
    mehp24::blocks::replicas(cc, borrow_all(parts, nc), (size_t)sub_len, replR, replC);
    mehp24::blocks::comparePartial(cc, replR, replC, (size_t)u_lo, (size_t)u_hi, SignFunc::CompositeSign, cfg, comp, Cv, Ch, true);
    ch_out[0] = nullptr;
    for (int j = 0; j < nc; j++) { cv_out[j] = release(Cv[(size_t)j]); if (j) ch_out[j] = release(Ch[(size_t)j]); }
    HOST_END
}
int sortfhe_mehp24_block_ranks(ckks_ctx_t* ctx, const ckks_ct_t* const* cv, const ckks_ct_t* const* ch, int nc, int sub_len, ckks_ct_t** s_out) {
    HOST_BEGIN
    check_blocks(nc, sub_len);
    auto cc = borrow(ctx, sub_len * sub_len);
    std::vector<CT> s = mehp24::blocks::ranks(cc, borrow_all(cv, nc), borrow_all(ch, nc, 1), (size_t)sub_len);
    for (int j = 0; j < nc; j++) s_out[j] = release(s[(size_t)j]);
    HOST_END
}
int sortfhe_mehp24_block_sorted(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, const ckks_ct_t* const* s, int nc, int sub_len, int j,
                                int dg_i, int df_i, ckks_ct_t** out) {
    HOST_BEGIN
    check_blocks(nc, sub_len);
    if (j < 0 || j >= nc) throw std::runtime_error("mehp24 block sorted: target block outside [0, nc)");
    auto cc = borrow(ctx, sub_len * sub_len);
    std::vector<CT> replR, replC;
    mehp24::blocks::replicas(cc, borrow_all(parts, nc), (size_t)sub_len, replR, replC);
    *out = release(mehp24::blocks::sortedBlock(cc, replR, borrow_all(s, nc), (size_t)sub_len, (size_t)j, (uint32_t)dg_i, (uint32_t)df_i));
    HOST_END
}
int sortfhe_mehp24_matrix_op(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int op, int mask_output, ckks_ct_t** out) {
    HOST_BEGIN
    using namespace mehp24::utils;
    auto cc = borrow(ctx, n * n); CT c = borrow(input), r;
    switch (op) {
        case 0: r = replicateRow(cc, c, n); break;
        case 1: r = replicateColumn(cc, c, n); break;
        case 2: r = sumRows(cc, c, n, mask_output != 0); break;
        case 3: r = sumColumns(cc, c, n, mask_output != 0); break;
        case 4: r = transposeRow(cc, c, n, mask_output != 0); break;
        case 5: r = transposeColumn(cc, c, n, mask_output != 0); break;
        default: throw std::runtime_error("mehp24_matrix_op: unknown op");
    }
    *out = release(r);
    HOST_END
}
int sortfhe_mehp24_indicator_adv(ckks_ctx_t* ctx, const ckks_ct_t* x, double b, int dg, int df, ckks_ct_t** out) {
    HOST_BEGIN
    *out = release(mehp24::utils::indicatorAdv(borrow(ctx, 0), borrow(x), b, (size_t)dg, (size_t)df));
    HOST_END
}

// ---- k-way network (reference src/k-way/, src/kway_adapter.h)
int sortfhe_kway_params(int n, int* mult_depth, int* scaling_mod_size, int* first_mod_size, int32_t* rotations, int cap, int* count) {
    HOST_BEGIN
    CCParams<CryptoContextCKKSRNS> p; std::vector<int> r;
    kwaySort::adapterParameters(n, p, r);
    if ((int)r.size() > cap) throw std::runtime_error("kway_params: rotation buffer too small");
    *mult_depth = (int)p.multDepth; *scaling_mod_size = (int)p.scalingModSize; *first_mod_size = (int)p.firstModSize;
    *count = (int)r.size();
    for (size_t i = 0; i < r.size(); i++) rotations[i] = r[i];
    HOST_END
}
int sortfhe_kway_sort_type(int k, int M, int stage, int* m, int* log_dist, int* slope, int* stage_count, int64_t* rotate_distance) {
    HOST_BEGIN
    auto t = kwaySort::sortType(k, M, stage);
    *m = std::get<0>(t); *log_dist = std::get<1>(t); *slope = std::get<2>(t);
    *stage_count = M + M * (M - 1) / 2 * ((k + 1) / 2);
    *rotate_distance = kwaySort::getRotateDistance(k, *log_dist, *slope);
    HOST_END
}
int sortfhe_kway_gen_indices(int num_slots, int k, int M, int m, int log_dist, int slope, int32_t* out /*[2][num_slots]*/) {
    HOST_BEGIN
    auto r = kwaySort::genIndices(num_slots, k, M, m, log_dist, slope);
    for (int h = 0; h < 2; h++) for (int i = 0; i < num_slots; i++) out[(size_t)h * num_slots + i] = r[h][i];
    HOST_END
}
int sortfhe_kway_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int num_slots, int k, int M, int sn, int dg, int df, int mult_depth,
                      int first_stage, int last_stage, ckks_ct_t** out) {
    HOST_BEGIN
    int batch = 1; while (batch < num_slots) batch <<= 1;   // SetBatchSize(next_power_of_two(N)), kway_adapter.h:44
    auto cc = borrow(ctx, batch);
    kwaySort::Sorter s(cc, nullptr, num_slots, k, M);
    SignConfig cfg(CompositeSignConfig(sn, dg, df), mult_depth);
    CT work = borrow(input)->Clone(), res;           // KWayAdapter::sort works on a copy (kway_adapter.h:66-72)
    s.sorter(work, res, cfg, first_stage, last_stage);
    *out = release(res);
    HOST_END
}
}  // extern "C"
