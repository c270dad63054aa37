// bootstrap.cu -- CKKS bootstrapping: EvalBootstrapSetup / EvalBootstrapKeyGen / EvalBootstrap.
//
// Reached from the reference at src/k-way/EvalUtils.cpp:57-86 (checkLevelAndBoot -> EvalBootstrap :76), src/sign.cpp:164-171
// and src/kway_adapter.h:55-61 (level budgets {4,4} / {5,5}); fixtures tests/k-way/KWaySort2Test.cpp:47-49,
// tests/k-way/EvalUtilsTest.cpp:34-35,132-158.  OpenFHE 1.1.4's FHECKKSRNS is not available offline; this is the published
// algorithm [CHKKS18, HK20, BMTH21] with OpenFHE's parameterisation for UNIFORM_TERNARY secrets (K = 512, R = 6 double-angle
// steps) and sparse packing (slots <= ring / 4).  Host-side schedule only -- every step is one of the engine's kernels:
//
//   adjust    constant mult + rescale down to the last limb q0: plaintext integer m q0 2^-corr
//   ModRaise  INTT(q0 rows) -> NTT with the switch-modulus prologue into every limb of the raise level (evaluator.cu)
//   SubSum    log2(ring / 2 slots) rotations by slots 2^j
//   C2S       U^-1 as levelBudget[0] groups of inverse special-FFT layers: sum_d diag_d (.) Rot(ct, d), the rotations of one
//             input sharing its ModUp (hoisting); the last group is evaluated with G and -i G, u + conj(u) yields Re and Im
//   EvalMod   Paterson-Stockmeyer Chebyshev series (degree 119) of cos((2 pi K u - pi/2) / 2^R), then R x (c <- 2 c^2 - 1)
//   S2C       U as levelBudget[1] groups; the first one applied to both parts with G and i G
//
// The oracle twin is oracle/bootstrap.py: same op sequence and the same constants -- the FFT-layer products and the
// Chebyshev coefficients are formed from the same elementary double operations in the same order, so the encoded
// diagonals (hence every limb of the result) agree bit for bit.
#include "engine.h"
#include "evaluator.h"
#include <cmath>
#include <map>
#include <memory>
#include <set>
#include <tuple>

static const int K_UNIFORM = 512, R_UNIFORM = 6, CHEB_DEGREE = 119, CORRECTION = 10;

struct BootDiag { int d; std::vector<double> v; /* interleaved (re, im), slots pairs */ };
struct BootGroup { std::vector<BootDiag> main, alt; };
struct BootPlan {
    int slots = 0, log_s = 0, raise_level = 0, budget[2] = {0, 0};
    std::vector<BootGroup> c2s, s2c;
    std::vector<double> cheb;
    std::map<std::tuple<int, int, int, int, int>, Ct*> pts;   // (direction, group, which, d, level) -> encoded diagonal
};

static std::vector<int> layer_groups(int log_s, int budget) {
    const int b = std::max(1, std::min(budget, log_s));
    const int base = log_s / b, extra = log_s % b;
    std::vector<int> g(b);
    for (int i = 0; i < b; i++) g[i] = base + (i < extra ? 1 : 0);
    return g;
}

// dense (re, im) of the product of the special-FFT butterfly layers `lens` (application order), row-major s x s.
// Forward layer: [u; v] -> [u + w v; u - w v]; inverse layer: [a; b] -> [(a + b) / 2; (a - b) w' / 2].
static void group_matrix(int s, const std::vector<int>& lens, const std::vector<double>& ksi, const std::vector<uint32_t>& rot, int M,
                         bool inverse, std::vector<double>& re, std::vector<double>& im) {
    re.assign((size_t)s * s, 0.0); im.assign((size_t)s * s, 0.0);
    for (int i = 0; i < s; i++) re[(size_t)i * s + i] = 1.0;
    for (int len : lens) {
        const int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (int i = 0; i < s; i += len)
            for (int j = 0; j < lenh; j++) {
                const int idx = (inverse ? (lenq - (int)(rot[j] % (uint32_t)lenq)) : (int)(rot[j] % (uint32_t)lenq)) * gap;
                const double wr = ksi[2 * (size_t)idx], wi = ksi[2 * (size_t)idx + 1];
                double *pr = &re[(size_t)(i + j) * s], *pi = &im[(size_t)(i + j) * s];
                double *qr = &re[(size_t)(i + j + lenh) * s], *qi = &im[(size_t)(i + j + lenh) * s];
                for (int col = 0; col < s; col++) {
                    const double ar = pr[col], ai = pi[col], br = qr[col], bi = qi[col];
                    if (!inverse) {
                        const double vr = br * wr - bi * wi, vi = br * wi + bi * wr;
                        pr[col] = ar + vr; pi[col] = ai + vi; qr[col] = ar - vr; qi[col] = ai - vi;
                    } else {
                        const double ur = ar + br, ui = ai + bi, dr = ar - br, di = ai - bi;
                        pr[col] = ur * 0.5; pi[col] = ui * 0.5;
                        qr[col] = (dr * wr - di * wi) * 0.5; qi[col] = (dr * wi + di * wr) * 0.5;
                    }
                }
            }
    }
}

// generalised diagonals diag_d[i] = G[i, (i + d) mod s] * (sr + i si); all-zero diagonals dropped
static std::vector<BootDiag> diagonals(int s, const std::vector<double>& re, const std::vector<double>& im, double sr, double si) {
    std::vector<BootDiag> out;
    for (int d = 0; d < s; d++) {
        bool any = false;
        for (int i = 0; i < s && !any; i++) {
            const size_t o = (size_t)i * s + (size_t)((i + d) % s);
            any = re[o] != 0.0 || im[o] != 0.0;
        }
        if (!any) continue;
        BootDiag g; g.d = d; g.v.resize(2 * (size_t)s);
        for (int i = 0; i < s; i++) {
            const size_t o = (size_t)i * s + (size_t)((i + d) % s);
            const double cr = re[o], ci = im[o];
            g.v[2 * (size_t)i] = cr * sr - ci * si;
            g.v[2 * (size_t)i + 1] = cr * si + ci * sr;
        }
        out.push_back(std::move(g));
    }
    return out;
}

// c_i = (2 / n) sum_j f(u_j) cos(step i (j + 1/2)), u_j = cos(step (j + 1/2)): plain left-to-right sums (oracle twin:
// bootstrap.boot_cheb_coefficients)
static std::vector<double> boot_cheb_coefficients() {
    const int n = CHEB_DEGREE + 1;
    const double step = M_PI / (double)n;
    std::vector<double> f(n), c(n);
    for (int j = 0; j < n; j++) {
        const double u = std::cos(step * ((double)j + 0.5));
        f[j] = std::cos((2.0 * M_PI * (double)K_UNIFORM * u - 0.5 * M_PI) / (double)(1 << R_UNIFORM));
    }
    for (int i = 0; i < n; i++) {
        double acc = 0.0;
        for (int j = 0; j < n; j++) acc += f[j] * std::cos(step * (double)i * ((double)j + 0.5));
        c[i] = acc * 2.0 / (double)n;
    }
    return c;
}

void boot_release(Context* c) {
    for (auto& kv : c->boot) {
        for (auto& p : kv.second->pts) ct_free(p.second);
        delete kv.second;
    }
    c->boot.clear();
}

void boot_setup(Context* c, int budget_c2s, int budget_s2c, int slots) {
    if (slots <= 0) slots = c->n / 4;
    if (slots < 2 || (slots & (slots - 1)) || slots > c->n / 4)
        throw std::runtime_error("EvalBootstrapSetup: sparse packing only (2 <= slots <= ring/4, power of two)");
    if (budget_c2s < 1 || budget_s2c < 1) throw std::runtime_error("EvalBootstrapSetup: level budgets must be >= 1");
    auto old = c->boot.find(slots);
    if (old != c->boot.end()) { for (auto& p : old->second->pts) ct_free(p.second); delete old->second; c->boot.erase(old); }
    std::unique_ptr<BootPlan> P(new BootPlan());
    P->slots = slots; P->budget[0] = budget_c2s; P->budget[1] = budget_s2c;
    while ((1 << P->log_s) < slots) P->log_s++;
    P->raise_level = c->ext ? 1 : 0;
    const int M = 2 * c->n;
    std::vector<double> ksi(2 * (size_t)(M + 1));
    for (int k = 0; k <= M; k++) { const double ang = 2.0 * M_PI * (double)k / (double)M; ksi[2 * (size_t)k] = std::cos(ang); ksi[2 * (size_t)k + 1] = std::sin(ang); }
    std::vector<uint32_t> rot(c->n / 2); { uint64_t f = 1; for (int j = 0; j < c->n / 2; j++) { rot[j] = (uint32_t)f; f = f * 5 % (uint64_t)M; } }
    std::vector<int> lens_fwd(P->log_s);
    for (int i = 0; i < P->log_s; i++) lens_fwd[i] = 2 << i;
    const std::vector<int> gC = layer_groups(P->log_s, budget_c2s), gS = layer_groups(P->log_s, budget_s2c);
    const double g = (double)(c->n / (2 * slots)), q0 = (double)c->moduli[0], sf_r = c->sf_real[P->raise_level];
    const double tot_c = sf_r / (g * q0 * (double)K_UNIFORM) * 0.5;
    const double tot_s = (double)(1 << CORRECTION) / (2.0 * M_PI);
    const double sc = std::pow(tot_c, 1.0 / (double)gC.size()), ss = std::pow(tot_s, 1.0 / (double)gS.size());
    std::vector<double> re, im;
    int pos = P->log_s;
    for (size_t k = 0; k < gC.size(); k++) {
        std::vector<int> lens;
        for (int t = 0; t < gC[k]; t++) lens.push_back(lens_fwd[pos - 1 - t]);
        pos -= gC[k];
        group_matrix(slots, lens, ksi, rot, M, true, re, im);
        BootGroup G; G.main = diagonals(slots, re, im, sc, 0.0);
        if (k + 1 == gC.size()) G.alt = diagonals(slots, re, im, 0.0, -sc);
        P->c2s.push_back(std::move(G));
    }
    pos = 0;
    for (size_t k = 0; k < gS.size(); k++) {
        std::vector<int> lens;
        for (int t = 0; t < gS[k]; t++) lens.push_back(lens_fwd[pos + t]);
        pos += gS[k];
        group_matrix(slots, lens, ksi, rot, M, false, re, im);
        BootGroup G; G.main = diagonals(slots, re, im, ss, 0.0);
        if (k == 0) G.alt = diagonals(slots, re, im, 0.0, ss);
        P->s2c.push_back(std::move(G));
    }
    P->cheb = boot_cheb_coefficients();
    c->boot[slots] = P.release();
}

static BootPlan* plan_of(Context* c, int slots) {
    auto it = c->boot.find(slots);
    if (it == c->boot.end()) throw std::runtime_error("EvalBootstrap: EvalBootstrapSetup was not called for " + std::to_string(slots) + " slots");
    return it->second;
}

void boot_keygen(Context* c, int slots) {
    BootPlan* P = plan_of(c, slots <= 0 ? c->n / 4 : slots);
    std::set<int> idx;
    for (int j = 1, g = c->n / (2 * P->slots); j < g; j <<= 1) idx.insert(P->slots * j);
    for (auto* grp : {&P->c2s, &P->s2c})
        for (auto& G : *grp)
            for (auto& d : G.main) if (d.d) idx.insert(d.d);
    std::vector<int> v(idx.begin(), idx.end());
    rot_keygen(c, v.data(), (int)v.size());
    conj_keygen(c);
}

// encoded diagonal, cached per (direction, group, which, d, level): the plaintexts depend on the context only
static const Ct* diag_plain(Context* c, BootPlan* P, int dir, int grp, int which, const BootDiag& d, int level) {
    auto key = std::make_tuple(dir, grp, which, d.d, level);
    auto it = P->pts.find(key);
    if (it != P->pts.end()) return it->second;
    Ct* pt = encode(c, d.v.data(), P->slots, 1, level, P->slots, 1);
    P->pts[key] = pt;
    return pt;
}

// sum_d diag_d (.) Rot(ct, d): one level
static Ct* linear(Context* c, BootPlan* P, const Ct* in, int dir, int grp, int which, const std::vector<BootDiag>& diags) {
    Ct* own = nullptr;
    const Ct* ct = in;
    if (ct->deg == 2) { own = ct_clone(ct); rescale(own); ct = own; }
    std::vector<Ct*> terms;
    try {
        for (const BootDiag& d : diags) {
            Ct* r = d.d == 0 ? nullptr : eval_rotate(ct, d.d);
            try { terms.push_back(eval_mult(r ? r : ct, diag_plain(c, P, dir, grp, which, d, ct->level))); }
            catch (...) { if (r) ct_free(r); throw; }
            if (r) ct_free(r);
        }
        std::vector<const Ct*> v(terms.begin(), terms.end());
        Ct* out = eval_add_many(v);
        for (Ct* t : terms) ct_free(t);
        if (own) ct_free(own);
        return out;
    } catch (...) {
        for (Ct* t : terms) ct_free(t);
        if (own) ct_free(own);
        throw;
    }
}

static Ct* eval_mod(BootPlan* P, const Ct* x) {
    Ct* c = eval_chebyshev_series_ps(x, P->cheb, -1.0, 1.0);
    for (int i = 0; i < R_UNIFORM; i++) {
        Ct* sq = eval_mult(c, c);
        Ct* two = eval_add(sq, sq, 1);
        Ct* nx = eval_add_const(two, -1.0);
        ct_free(sq); ct_free(two); ct_free(c);
        c = nx;
    }
    return c;
}

struct CtHold {   // frees a temporary on scope exit
    Ct* p; explicit CtHold(Ct* q = nullptr) : p(q) {} ~CtHold() { if (p) ct_free(p); }
    CtHold(const CtHold&) = delete; void reset(Ct* q) { if (p) ct_free(p); p = q; } Ct* release() { Ct* q = p; p = nullptr; return q; }
};

Ct* boot_eval(const Ct* in) {
    Context* c = in->ctx;
    if (in->parts != 2) throw std::runtime_error("EvalBootstrap: ciphertext must have 2 components");
    BootPlan* P = plan_of(c, in->slots);
    c->stats.ops[{"bootstrap", in->nl}]++;
    CtHold ct(ct_clone(in));
    if (ct.p->deg == 2) rescale(ct.p);
    if (ct.p->nl < 2) throw std::runtime_error("EvalBootstrap: no limb left to drop (level == multiplicative depth)");
    level_reduce(ct.p, ct.p->nl - 2);
    const int last = c->size_q - 1;
    mult_core_const(ct.p, (double)c->moduli[0] / (double)(1 << CORRECTION) / c->sf_real[last]);
    rescale(ct.p);                                                     // one limb: plaintext integer m q0 2^-corr
    ct.reset(mod_raise(ct.p, P->raise_level));
    ct.p->slots = P->slots;
    for (int j = 1, g = c->n / (2 * P->slots); j < g; j <<= 1) {       // SubSum
        CtHold r(eval_rotate(ct.p, P->slots * j));
        ct.reset(eval_add(ct.p, r.p, 1));
    }
    const int nc = (int)P->c2s.size(), ns = (int)P->s2c.size();
    for (int k = 0; k + 1 < nc; k++) ct.reset(linear(c, P, ct.p, 0, k, 0, P->c2s[k].main));
    CtHold u(linear(c, P, ct.p, 0, nc - 1, 0, P->c2s[nc - 1].main));
    CtHold v(linear(c, P, ct.p, 0, nc - 1, 1, P->c2s[nc - 1].alt));
    CtHold re, im;
    { CtHold cu(eval_conjugate(u.p)); re.reset(eval_add(u.p, cu.p, 1)); }
    { CtHold cv(eval_conjugate(v.p)); im.reset(eval_add(v.p, cv.p, 1)); }
    re.reset(eval_mod(P, re.p));
    im.reset(eval_mod(P, im.p));
    CtHold a(linear(c, P, re.p, 1, 0, 0, P->s2c[0].main));
    CtHold b(linear(c, P, im.p, 1, 0, 1, P->s2c[0].alt));
    ct.reset(eval_add(a.p, b.p, 1));
    for (int k = 1; k < ns; k++) ct.reset(linear(c, P, ct.p, 1, k, 0, P->s2c[k].main));
    ct.p->slots = in->slots;
    return ct.release();
}
