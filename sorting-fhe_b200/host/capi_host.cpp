// capi_host.cpp -- extern "C" boundary of the sorter layer (include/sortfhe_host.h).
#include "../../include/sortfhe_host.h"
#include <cstring>
#include "sort_algo.h"

namespace {
thread_local std::string g_err;
typedef Ciphertext<DCRTPoly> CT;

#define HOST_BEGIN try {
#define HOST_END return 0; } catch (const std::exception& e) { g_err = e.what(); return 1; } catch (...) { g_err = "unknown error"; return 2; }

CryptoContext<DCRTPoly> borrow(ckks_ctx_t* c, int batch) { return std::make_shared<CryptoContextImpl>(c, false, (uint32_t)batch); }
CT borrow(const ckks_ct_t* h) {
    if (!h) throw std::runtime_error("null ciphertext handle");
    return std::make_shared<CiphertextImpl>(const_cast<ckks_ct_t*>(h), false);
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
}  // extern "C"
