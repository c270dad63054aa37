// cheby.cu -- Paterson-Stockmeyer Chebyshev evaluator and linear-combination helpers (K10).
//
// Host-side schedule over the primitive evaluator ops; restates OpenFHE 1.1.4
// AdvancedSHECKKSRNS::EvalChebyshevSeriesPS / InnerEvalChebyshevPS / EvalLinearWSumMutable /
// EvalPolyLinear and ckksrns-utils (Degree, ComputeDegreesPS, LongDivisionChebyshev), which the
// reference calls at src/sort_algo.h:727-728 (doubled-sinc indicator), src/sign.cpp:76 (g_4) and
// src/mehp24/mehp24_utils.cpp:256-259.  Same operation order as oracle/cheby.py so that both
// produce identical doubles and identical ciphertext limbs.
#include "evaluator.h"
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <memory>
#include <stdexcept>

typedef std::shared_ptr<Ct> CtP;
static CtP wrap(Ct* p) { return CtP(p, [](Ct* q) { ct_free(q); }); }
static CtP add(const CtP& a, const CtP& b) { return wrap(eval_add(a.get(), b.get(), 1)); }
static CtP sub(const CtP& a, const CtP& b) { return wrap(eval_add(a.get(), b.get(), -1)); }
static CtP addc(const CtP& a, double x) { return wrap(eval_add_const(a.get(), x)); }
static CtP mul(const CtP& a, const CtP& b) { return wrap(eval_mult(a.get(), b.get())); }
static CtP mulc(const CtP& a, double x) { return wrap(eval_mult_const(a.get(), x)); }
static CtP dbl_addc(const CtP& a, double x) { return wrap(eval_double_add_const(a.get(), x)); }   // addc(add(a, a), x) in one pass

static int degree(const std::vector<double>& c, size_t len) {
    for (size_t i = len; i > 0; i--) if (c[i - 1] != 0) return (int)i - 1;
    return 0;
}
static int degree(const std::vector<double>& c) { return degree(c, c.size()); }

void compute_degrees_ps(int n, int& kb, int& mb) {
    long best = -1;
    for (int k = 1; k <= n; k++) {
        int mmax = (int)(std::ceil(std::log2((double)n / (double)k) + 1) + 1);
        for (int m = 1; m <= mmax; m++) {
            if ((long)n - (long)k * ((1L << m) - 1) < 0) {
                if (std::fabs(std::floor(std::log2((double)k)) - std::floor(std::log2(std::sqrt((double)n / 2.0)))) <= 1) {
                    long cost = k + 2 * m + (1L << (m - 1)) - 4;
                    if (best < 0 || cost < best) { best = cost; kb = k; mb = m; }
                }
            }
        }
    }
    if (best < 0) throw std::runtime_error("ComputeDegreesPS: no decomposition");
}

void long_division_chebyshev(const std::vector<double>& f, const std::vector<double>& g, std::vector<double>& q,
                             std::vector<double>& r) {
    r = f;
    std::vector<double> ga = g;
    int n = degree(r), k = degree(ga);
    if (n < k) { q.assign(1, 0.0); return; }
    r.resize(n + 1); ga.resize(k + 1);
    r[0] *= 0.5; ga[0] *= 0.5;
    q.assign(n - k + 1, 0.0);
    while (n > k) {
        double c = 2.0 * r[n] / ga[k];
        q[n - k] = c;
        for (int i = 0; i <= k; i++) {
            double t = (c * ga[i]) * 0.5;
            r[n - k + i] -= t;
            r[std::abs(n - k - i)] -= t;
        }
        r[n] = 0.0;
        n = degree(r, n);
        r.resize(n + 1);
    }
    if (n == k) {
        double c = r[n] / ga[k];
        q[0] = c;
        for (int i = 0; i <= k; i++) r[i] -= c * ga[i];
        r[n] = 0.0;
        n = n > 0 ? degree(r, n) : 0;
        r.resize(n + 1);
    }
    r[0] *= 2.0; q[0] *= 2.0;
}

Ct* eval_linear_wsum_mutable(std::vector<Ct*>& cts, const std::vector<double>& w) {
    size_t mi = 0;
    for (size_t i = 1; i < cts.size(); i++) if (cts[i]->level > cts[mi]->level) mi = i;
    for (size_t i = 0; i < cts.size(); i++) if (i != mi) adjust_levels_and_depth(cts[i], cts[mi]);
    if (cts[mi]->deg == 2) for (size_t i = 0; i < cts.size(); i++) rescale(cts[i]);
    if (Ct* fused = eval_wsum_aligned(cts, w)) return fused;
    CtP acc = wrap(eval_mult_const(cts[0], w[0]));
    for (size_t i = 1; i < cts.size(); i++) { CtP t = wrap(eval_mult_const(cts[i], w[i])); acc = add(acc, t); }
    return ct_clone(acc.get());
}

// ---- batched leaves.  Every weighted sum of one Paterson-Stockmeyer evaluation combines a prefix of the same baby-step basis
// T_1..T_{k-1}, and which sums are formed, in which order and with which weights depends on the coefficients alone.  ps_plan walks
// the recursion of ps_core on the coefficients only and lists them; wsum_precompute then forms them in that order -- so that the
// in-place adjustments EvalLinearWSumMutable applies to the T_i happen in OpenFHE's order -- but up to four at a time in one pass
// over the inputs (evaluator.cu eval_wsum_aligned_multi); ps_core's wsum() calls pick the results up.  DirectSort N=128's degree-848
// doubled-sinc indicator has 61 such sums over up to 31 inputs of 16 limbs: 61 passes over 15 GB become 16 over 5.
struct WsumItem { int cnt; std::vector<double> w; };
struct WsumQueue { std::vector<WsumItem> items; std::vector<CtP> res; size_t next = 0; };
static thread_local WsumQueue* g_wq = nullptr;

static CtP wsum(std::vector<CtP>& T, int cnt, const std::vector<double>& coeffs) {
    std::vector<Ct*> cts(cnt); std::vector<double> w(cnt);
    for (int i = 0; i < cnt; i++) { cts[i] = T[i].get(); w[i] = coeffs[i + 1]; }
    if (g_wq) {
        // bitwise comparison: degenerate series (a zero divisor in the Chebyshev long division) carry NaN weights in both walks
        if (g_wq->next >= g_wq->items.size() || g_wq->items[g_wq->next].cnt != cnt ||
            std::memcmp(g_wq->items[g_wq->next].w.data(), w.data(), sizeof(double) * (size_t)cnt) != 0)
            throw std::runtime_error("internal: Paterson-Stockmeyer plan and evaluation disagree");
        return g_wq->res[g_wq->next++];
    }
    return wrap(eval_linear_wsum_mutable(cts, w));
}

static CtP ps_leaf(const std::vector<double>& coeffs, int k, std::vector<CtP>& T, int doublings) {
    std::vector<double> cc(k, 0.0);
    for (int i = 0; i < k && i < (int)coeffs.size(); i++) cc[i] = coeffs[i];
    int d = degree(cc);
    CtP acc;
    if (d > 0) acc = wsum(T, d, coeffs);
    CtP s = T[k - 1];
    for (int i = 0; i < doublings; i++) s = add(s, s);
    acc = acc ? add(acc, s) : s;
    return addc(acc, coeffs[0] / 2.0);
}

static void plan_wsum(std::vector<WsumItem>& items, int cnt, const std::vector<double>& coeffs) {
    WsumItem it; it.cnt = cnt; it.w.resize(cnt);
    for (int i = 0; i < cnt; i++) it.w[i] = coeffs[i + 1];
    items.push_back(std::move(it));
}
static void plan_leaf(const std::vector<double>& coeffs, int k, std::vector<WsumItem>& items) {
    std::vector<double> cc(k, 0.0);
    for (int i = 0; i < k && i < (int)coeffs.size(); i++) cc[i] = coeffs[i];
    const int d = degree(cc);
    if (d > 0) plan_wsum(items, d, coeffs);
}
// the coefficient arithmetic of ps_core, nothing else; `uses_t0` reports a T_1 * constant product formed before the first sum
// (it would see T_1 before the first sum's in-place rescale: the plan is then not used)
static void ps_plan(const std::vector<double>& f, int k, int m, std::vector<WsumItem>& items, bool& t0_before_first_sum) {
    const int k2m2k = k * (1 << (m - 1)) - k;
    std::vector<double> tkm(k2m2k + k + 1, 0.0); tkm.back() = 1.0;
    std::vector<double> q, r;
    long_division_chebyshev(f, tkm, q, r);
    std::vector<double> r2 = r;
    if (k2m2k - degree(r) <= 0) {
        if ((int)r2.size() < k2m2k + 1) r2.resize(k2m2k + 1, 0.0);
        r2[k2m2k] -= 1.0;
        r2.resize(degree(r2) + 1);
    } else {
        r2.resize(k2m2k + 1, 0.0);
        r2.back() = -1.0;
    }
    std::vector<double> cq, s;
    long_division_chebyshev(r2, q, cq, s);
    std::vector<double> s2 = s;
    s2.resize(k2m2k + 1, 0.0);
    s2.back() = 1.0;
    const int dc = degree(cq);
    if (dc == 1 && cq[1] != 1 && items.empty()) t0_before_first_sum = true;
    if (dc >= 2) plan_wsum(items, dc, cq);
    if (degree(q) > k) ps_plan(q, k, m - 1, items, t0_before_first_sum); else plan_leaf(q, k, items);
    if (degree(s2) > k) ps_plan(s2, k, m - 1, items, t0_before_first_sum); else plan_leaf(s2, k, items);
}
// true when EvalLinearWSumMutable on T[0..cnt) would leave its operands as they are (aligned, noiseScaleDeg 1)
static bool wsum_inputs_settled(const std::vector<CtP>& T, int cnt) {
    const Ct* f = T[0].get();
    for (int i = 0; i < cnt; i++) {
        const Ct* g = T[i].get();
        if (g->deg != 1 || g->level != f->level || g->nl != f->nl || g->sf != f->sf || g->parts != f->parts || g->lrep) return false;
    }
    return true;
}
static void wsum_precompute(std::vector<CtP>& T, WsumQueue& Q) {
    const size_t n = Q.items.size();
    Q.res.resize(n);
    size_t i = 0;
    while (i < n) {
        size_t j = i;
        if (wsum_inputs_settled(T, Q.items[i].cnt)) {
            while (j < n && j - i < WSUM_GROUPS && Q.items[j].cnt >= 2 && wsum_inputs_settled(T, Q.items[j].cnt)) j++;
        }
        if (j - i >= 2) {
            int cmax = 0;
            std::vector<std::vector<double>> w;
            for (size_t t = i; t < j; t++) { cmax = std::max(cmax, Q.items[t].cnt); w.push_back(Q.items[t].w); }
            std::vector<Ct*> cts(cmax);
            for (int t = 0; t < cmax; t++) cts[t] = T[t].get();
            Ct* outs[WSUM_GROUPS] = {nullptr, nullptr, nullptr, nullptr};
            if (eval_wsum_aligned_multi(cts, w, outs)) {
                for (size_t t = i; t < j; t++) Q.res[t] = wrap(outs[t - i]);
                i = j;
                continue;
            }
        }
        std::vector<Ct*> cts(Q.items[i].cnt);
        for (int t = 0; t < Q.items[i].cnt; t++) cts[t] = T[t].get();
        Q.res[i] = wrap(eval_linear_wsum_mutable(cts, Q.items[i].w));     // also the step that settles the inputs
        i++;
    }
}

static CtP ps_core(const std::vector<double>& f, int k, int m, std::vector<CtP>& T, std::vector<CtP>& T2) {
    const int k2m2k = k * (1 << (m - 1)) - k;
    std::vector<double> tkm(k2m2k + k + 1, 0.0); tkm.back() = 1.0;
    std::vector<double> q, r;
    long_division_chebyshev(f, tkm, q, r);
    std::vector<double> r2 = r;
    if (k2m2k - degree(r) <= 0) {
        if ((int)r2.size() < k2m2k + 1) r2.resize(k2m2k + 1, 0.0);
        r2[k2m2k] -= 1.0;
        r2.resize(degree(r2) + 1);
    } else {
        r2.resize(k2m2k + 1, 0.0);
        r2.back() = -1.0;
    }
    std::vector<double> cq, s;
    long_division_chebyshev(r2, q, cq, s);
    std::vector<double> s2 = s;
    s2.resize(k2m2k + 1, 0.0);
    s2.back() = 1.0;
    const int dc = degree(cq);
    CtP cu;
    if (dc >= 1) {
        if (dc == 1) cu = (cq[1] != 1) ? mulc(T[0], cq[1]) : T[0];
        else cu = wsum(T, dc, cq);
        cu = addc(cu, cq[0] / 2.0);
    }
    CtP qu, su;
    if (degree(q) > k) qu = ps_core(q, k, m - 1, T, T2);
    else qu = ps_leaf(q, k, T, (int)std::lround(std::log2(q.back())));
    if (degree(s2) > k) su = ps_core(s2, k, m - 1, T, T2);
    else su = ps_leaf(s2, k, T, 0);
    CtP res = cu ? add(T2[m - 1], cu) : addc(T2[m - 1], cq[0] / 2.0);
    res = mul(res, qu);
    return add(res, su);
}

Ct* eval_chebyshev_series_ps(const Ct* x, const std::vector<double>& coefficients, double a, double b) {
    const int n = degree(coefficients);
    std::vector<double> f2(coefficients.begin(), coefficients.begin() + n + 1);
    int k = 0, m = 0;
    compute_degrees_ps(n, k, m);
    std::vector<CtP> T(k);
    if (std::fabs(a + 1.0) < 1e-10 && std::fabs(b - 1.0) < 1e-10) T[0] = wrap(ct_clone(x));
    else {
        double alpha = 2.0 / (b - a), beta = 2.0 * a / (b - a);
        CtP t = wrap(eval_mult_const(x, alpha));
        T[0] = addc(t, -1.0 - beta);
    }
    CtP y = wrap(ct_clone(T[0].get()));
    for (int i = 2; i <= k; i++) {
        if ((i & (i - 1)) == 0 || i % 2 == 0) {
            CtP sq = mul(T[i / 2 - 1], T[i / 2 - 1]);
            T[i - 1] = dbl_addc(sq, -1.0);
        } else {
            CtP pr = mul(T[i / 2 - 1], T[i / 2]);
            T[i - 1] = sub(add(pr, pr), y);
        }
    }
    for (int i = 1; i < k; i++) adjust_levels_and_depth(T[i - 1].get(), T[k - 1].get());
    std::vector<CtP> T2(m);
    T2[0] = T[k - 1];
    for (int i = 1; i < m; i++) {
        CtP sq = mul(T2[i - 1], T2[i - 1]);
        T2[i] = dbl_addc(sq, -1.0);
    }
    CtP T2km1 = T2[0];
    for (int i = 1; i < m; i++) {
        CtP pr = mul(T2km1, T2[i]);
        T2km1 = sub(add(pr, pr), T2[0]);
    }
    const int k2m2k = k * (1 << (m - 1)) - k;
    f2.resize(2 * k2m2k + k + 1, 0.0);
    f2.back() = 1.0;
    static const bool no_batch = getenv("CKKS_B200_NO_WSUM_BATCH") != nullptr;   // A/B switch: one pass per weighted sum
    WsumQueue Q;
    bool t0_first = false;
    if (!no_batch) ps_plan(f2, k, m, Q.items, t0_first);
    const bool use_plan = !no_batch && !t0_first && Q.items.size() >= 2;
    CtP res;
    if (use_plan) {
        wsum_precompute(T, Q);
        struct Guard { WsumQueue* prev; Guard(WsumQueue* q) : prev(g_wq) { g_wq = q; } ~Guard() { g_wq = prev; } } guard(&Q);
        res = ps_core(f2, k, m, T, T2);
        if (Q.next != Q.items.size()) throw std::runtime_error("internal: Paterson-Stockmeyer plan longer than the evaluation");
    } else res = ps_core(f2, k, m, T, T2);
    res = sub(res, T2km1);
    return ct_clone(res.get());
}

// EvalChebyshevSeriesLinear: all T_1..T_n by the doubling recurrences, one weighted sum, + c_0 / 2.  OpenFHE takes
// this route for degree < 5 (EvalChebyshevSeries), which none of the reference's call sites reach.
Ct* eval_chebyshev_series_linear(const Ct* x, const std::vector<double>& coefficients, double a, double b) {
    const int k = (int)coefficients.size() - 1;
    if (k < 1) throw std::runtime_error("EvalChebyshevSeriesLinear: degree must be >= 1");
    std::vector<CtP> T(k);
    if (std::fabs(a + 1.0) < 1e-10 && std::fabs(b - 1.0) < 1e-10) T[0] = wrap(ct_clone(x));
    else {
        double alpha = 2.0 / (b - a), beta = 2.0 * a / (b - a);
        T[0] = addc(wrap(eval_mult_const(x, alpha)), -1.0 - beta);
    }
    CtP y = wrap(ct_clone(T[0].get()));
    for (int i = 2; i <= k; i++) {
        if (i % 2 == 0) {
            CtP sq = mul(T[i / 2 - 1], T[i / 2 - 1]);
            T[i - 1] = dbl_addc(sq, -1.0);
        } else {
            CtP pr = mul(T[i / 2 - 1], T[i / 2]);
            T[i - 1] = sub(add(pr, pr), y);
        }
    }
    for (int i = 1; i < k; i++) adjust_levels_and_depth(T[i - 1].get(), T[k - 1].get());
    CtP acc = wsum(T, k, coefficients);
    return ct_clone(addc(acc, coefficients[0] / 2.0).get());
}

Ct* eval_chebyshev_series(const Ct* x, const std::vector<double>& coefficients, double a, double b) {
    return degree(coefficients) < 5 ? eval_chebyshev_series_linear(x, coefficients, a, b)
                                    : eval_chebyshev_series_ps(x, coefficients, a, b);
}

Ct* eval_poly_linear(const Ct* x, const std::vector<double>& coeffs) {
    const int n = (int)coeffs.size() - 1;
    if (n < 1) throw std::runtime_error("EvalPolyLinear: degree must be >= 1");
    std::vector<int> need(n + 1, 0);
    auto hb = [](int v) { int b = 0; while ((1 << (b + 1)) <= v) b++; return 1 << b; };
    for (int i = n; i > 0; i--) {
        if ((i & (i - 1)) == 0) need[i] = 1;
        else if (coeffs[i] != 0) {
            need[i] = 1;
            int rem = i % hb(i);
            need[rem] = 1;
            while (rem & (rem - 1)) { rem = rem % hb(rem); need[rem] = 1; }
        }
    }
    std::vector<CtP> pw(n + 1);
    pw[1] = wrap(ct_clone(x));
    for (int i = 2; i <= n; i++) {
        if ((i & (i - 1)) == 0) pw[i] = mul(pw[i / 2], pw[i / 2]);
        else if (need[i]) { int p2 = hb(i); pw[i] = mul(pw[p2], pw[i - p2]); }
    }
    CtP res = mulc(pw[n], coeffs[n]);
    for (int i = 1; i < n; i++) if (coeffs[i] != 0) res = add(res, mulc(pw[i], coeffs[i]));
    res = addc(res, coeffs[0]);
    return ct_clone(res.get());
}
