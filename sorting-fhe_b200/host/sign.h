// sign.h -- composite-polynomial sign with the reference's interface (reference src/sign.h:6-41).
#pragma once
#include "encryption.h"

enum class SignFunc { CompositeSign, SignumPolycircuit, Tanh, NaiveDiscrete };

struct CompositeSignConfig {
    int n, dg, df;
    CompositeSignConfig(int n, int dg, int df) : n(n), dg(dg), df(df) {}
};

struct SignConfig {
    CompositeSignConfig compos;
    int multDepth;
    SignConfig() : compos(0, 0, 0), multDepth(0) {}
    SignConfig(CompositeSignConfig c) : compos(c), multDepth(100) {}            // no bootstrap (sentinel depth)
    SignConfig(CompositeSignConfig c, int depth) : compos(c), multDepth(depth) {}  // bootstrap enabled
};

template <int n>
Ciphertext<DCRTPoly> compositeSign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, const SignConfig& Cfg);

Ciphertext<DCRTPoly> naive_discrete_sign(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc);
Ciphertext<DCRTPoly> tanh(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc);
Ciphertext<DCRTPoly> sign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, SignFunc func, const SignConfig& Cfg);
