// comparison.h -- Comparison::compare / indicator and the Sinc<N> plaintext functions
// (reference src/comparison.h:11-103, src/comparison.cpp:4-40).
#pragma once
#include <cmath>
#include <memory>
#include "sign.h"

template <int N> struct Sinc {
    // sin(pi N x) / (pi N x), periodic in the array size (comparison.h:29-36)
    static double scaled_sinc(double x) {
        if (std::abs(x) < 1e-10) return 1.0;
        return std::sin(M_PI * N * x) / (M_PI * N * x);
    }
    // scaled_sinc(x) + scaled_sinc(x + 1/2): the second lobe catches negative index differences (:57-78)
    static double doubled_sinc(double x) { return scaled_sinc(x) + scaled_sinc(x + 0.5); }
};

class Comparison {
    std::shared_ptr<Encryption> m_enc;
  public:
    Comparison(std::shared_ptr<Encryption> enc) : m_enc(enc) {}
    Comparison() : m_enc(nullptr) {}
    // (sign(a - b) + 1) / 2   (comparison.cpp:4-22)
    Ciphertext<DCRTPoly> compare(const CryptoContext<DCRTPoly>& cc, const Ciphertext<DCRTPoly>& a,
                                 const Ciphertext<DCRTPoly>& b, SignFunc SignFunc, SignConfig& Cfg) {
        auto s = sign(cc->EvalSub(a, b), cc, SignFunc, Cfg);
        return cc->EvalMult(cc->EvalAdd(s, 1.0), 0.5);
    }
    // 1 where |x| < c, else 0 (MEHP24; comparison.cpp:24-40): ((sign(x+c)+1)/2) * (1 - (sign(x-c)+1)/2)
    Ciphertext<DCRTPoly> indicator(const CryptoContext<DCRTPoly>& cc, const Ciphertext<DCRTPoly>& x, const double c,
                                   SignFunc SignFunc, SignConfig& Cfg) {
        auto s1 = sign(cc->EvalAdd(x, c), cc, SignFunc, Cfg);
        auto s2 = sign(cc->EvalSub(x, c), cc, SignFunc, Cfg);
        auto c1 = cc->EvalMult(cc->EvalAdd(s1, 1.0), 0.5);
        auto c2 = cc->EvalMult(cc->EvalAdd(s2, 1.0), 0.5);
        return cc->EvalMult(c1, cc->EvalSub(1.0, c2));
    }
};
