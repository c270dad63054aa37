#include <cmath>
#include <cstdlib>
// sign.cpp -- compositeSign<3>, compositeSign<4> and the dispatcher (reference src/sign.cpp:9-185,635-651).
// Polynomials: g_3/f_3 degree-7 odd (sign.cpp:18-21,41-44), g_4 degree-27 Chebyshev series (sign.cpp:68-75),
// f_4 degree-15 odd (sign.cpp:81-88).  The evaluation ORDER is the reference's, because under
// FLEXIBLEAUTOEXT the order decides where rescales land and therefore every limb of the result.
#include "sign.h"

namespace {
typedef Ciphertext<DCRTPoly> CT;
typedef CryptoContext<DCRTPoly> CC;

const double kG3[4] = {4589.0 / 1024, -16577.0 / 1024, 25614.0 / 1024, -12860.0 / 1024};
const double kF3[4] = {35.0 / 16, -35.0 / 16, 21.0 / 16, -5.0 / 16};
const double kF4[8] = {3.14208984375, -7.33154296875, 13.19677734375, -15.71044921875,
                       12.21923828125, -5.99853515625, 1.69189453125, -0.20947265625};
const std::vector<double> kG4 = {
    0.0, 1.077117252745569, 0.0, -0.36166113998402755, 0.0, 0.2137420717859748, 0.0, -0.15635204788780485,
    0.0, 0.11749645501187332, 0.0, -0.10074154666447852, 0.0, 0.08002086947825496, 0.0, -0.07533558758484624,
    0.0, 0.059514472116534836, 0.0, -0.06146663712787884, 0.0, 0.04570084927999001, 0.0, -0.05403683682999072,
    0.0, 0.03364293851188723, 0.0, -0.054459493266273494};

// y = c1 x + (c3 x) x^2 + ((c5 x) + (c7 x) x^2) x^4   -- depth 3, 2 squares + 3 relinearised products
CT odd7(const CC& cc, const CT& x, const double* c) {
    CT x2 = cc->EvalSquare(x);
    CT x4 = cc->EvalSquare(x2);
    CT y = cc->EvalMult(x, c[0]);
    y = cc->EvalAdd(y, cc->EvalMultAndRelinearize(cc->EvalMult(x, c[1]), x2));
    CT c5x = cc->EvalMult(x, c[2]);
    CT c7x3 = cc->EvalMultAndRelinearize(cc->EvalMult(x, c[3]), x2);
    return cc->EvalAdd(y, cc->EvalMultAndRelinearize(cc->EvalAdd(c5x, c7x3), x4));
}

// degree-15 odd polynomial, depth 4: 3 squares + 7 relinearised products (sign.cpp:90-156)
CT f4(const CC& cc, const CT& x) {
    CT x2 = cc->EvalSquare(x);
    CT cx[8];
    for (int i = 0; i < 8; i++) cx[i] = cc->EvalMult(x, kF4[i]);
    CT x4 = cc->EvalSquare(x2);
    CT c3x3 = cc->EvalMultAndRelinearize(cx[1], x2);
    CT c7x3 = cc->EvalMultAndRelinearize(cx[3], x2);
    CT c11x3 = cc->EvalMultAndRelinearize(cx[5], x2);
    CT c15x3 = cc->EvalMultAndRelinearize(cx[7], x2);
    CT x8 = cc->EvalSquare(x4);
    CT y = cc->EvalAdd(cx[0], c3x3);
    y = cc->EvalAdd(y, cc->EvalMultAndRelinearize(cc->EvalAdd(cx[2], c7x3), x4));
    CT t1 = cc->EvalAdd(cx[4], c11x3);
    CT t2 = cc->EvalAdd(cx[6], c15x3);
    t1 = cc->EvalAdd(t1, cc->EvalMultAndRelinearize(t2, x4));
    return cc->EvalAdd(y, cc->EvalMultAndRelinearize(t1, x8));
}

template <int n> struct Poly;
template <> struct Poly<3> {
    static CT g(const CC& cc, const CT& x) { return odd7(cc, x, kG3); }
    static CT f(const CC& cc, const CT& x) { return odd7(cc, x, kF3); }
    static constexpr int depth_g = 3, depth_f = 3;
};
template <> struct Poly<4> {
    static CT g(const CC& cc, const CT& x) { return cc->EvalChebyshevSeriesPS(x, kG4, -1.0, 1.0); }
    static CT f(const CC& cc, const CT& x) { return f4(cc, x); }
    static constexpr int depth_g = 4, depth_f = 4;
};

// sign.cpp:164-171: bootstrap when the remaining depth cannot hold the next polynomial
void lazyBootstrap(const CC& cc, CT& x, int needed, const SignConfig& Cfg) {
    if (Cfg.multDepth - (int)x->GetLevel() < needed + 2) x = cc->EvalBootstrap(x);
}
}  // namespace

template <int n>
Ciphertext<DCRTPoly> compositeSign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, const SignConfig& Cfg) {
    lazyBootstrap(cc, x, Poly<n>::depth_g, Cfg);
    CT y = Poly<n>::g(cc, x);                      // first g is unconditional (sign.cpp:174)
    for (int i = 1; i < Cfg.compos.dg; i++) { lazyBootstrap(cc, y, Poly<n>::depth_g, Cfg); y = Poly<n>::g(cc, y); }
    for (int i = 0; i < Cfg.compos.df; i++) { lazyBootstrap(cc, y, Poly<n>::depth_f, Cfg); y = Poly<n>::f(cc, y); }
    return y;
}
template Ciphertext<DCRTPoly> compositeSign<3>(Ciphertext<DCRTPoly>, CryptoContext<DCRTPoly>, const SignConfig&);
template Ciphertext<DCRTPoly> compositeSign<4>(Ciphertext<DCRTPoly>, CryptoContext<DCRTPoly>, const SignConfig&);

// Chebyshev-interpolated alternatives (sign.cpp:610-633): a degree-119 step function and a degree-1006 steep tanh
Ciphertext<DCRTPoly> naive_discrete_sign(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc) {
    return cc->EvalChebyshevFunction([](double v) -> double { return v >= 0 ? 1.0 : 0.0; }, a, -1.0, 1.0, 119);
}
Ciphertext<DCRTPoly> tanh(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc) {
    return cc->EvalChebyshevFunction([](double v) -> double { return std::tanh(v * (double)RAND_MAX); }, a, -1.0, 1.0, 1006);
}

Ciphertext<DCRTPoly> sign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, SignFunc func, const SignConfig& Cfg) {
    switch (func) {
        case SignFunc::CompositeSign: break;
        case SignFunc::Tanh: return tanh(x, cc);
        case SignFunc::NaiveDiscrete: return naive_discrete_sign(x, cc);
        default:
            throw std::runtime_error("sign: SignFunc::SignumPolycircuit (a hard-coded degree-1023 table of a third-party circuit) is not "
                                     "provided; use CompositeSign, Tanh or NaiveDiscrete");
    }
    switch (Cfg.compos.n) {
        case 3: return compositeSign<3>(x, cc, Cfg);
        case 4: return compositeSign<4>(x, cc, Cfg);
        default: throw std::runtime_error("sign: composite sign supports n = 3 or 4");
    }
}
