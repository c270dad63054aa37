// encryption.h -- client-side helpers with the reference's names (reference src/encryption.h:57-100,
// src/encryption.cpp:5-33): Encryption::encryptInput = encode + encrypt, DebugEncryption adds decrypt.
#pragma once
#include <cmath>
#include <memory>
#include <stdexcept>
#include <vector>
#include "openfhe_shim.h"

using namespace lbcrypto;

#define PRINT_PT(enc, ct)   // reference src/encryption.h:34-46 (decrypt-and-print hook; compiled out)

class Encryption {
  public:
    Encryption(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> pk) : m_cc(cc), m_PublicKey(pk) {}
    virtual ~Encryption() = default;
    // src/encryption.cpp:5-12 (the size assertion becomes an exception: no abort() inside a library)
    Ciphertext<DCRTPoly> encryptInput(std::vector<double> input) {
        if (input.size() > m_cc->GetRingDimension() / 2)
            throw std::runtime_error("Input size is larger than the maximum available batch size");
        return m_cc->Encrypt(m_PublicKey, m_cc->MakeCKKSPackedPlaintext(input));
    }
    virtual std::vector<double> getPlaintext(const Ciphertext<DCRTPoly>&, double = 1e-10) const {
        throw std::runtime_error("Decryption not available in base Encryption class");
    }
    virtual Plaintext getDecrypt(const Ciphertext<DCRTPoly>&) const {
        throw std::runtime_error("Decryption not available in base Encryption class");
    }
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
};

class DebugEncryption : public Encryption {
  public:
    DebugEncryption(CryptoContext<DCRTPoly> cc, KeyPair<DCRTPoly> kp)
        : Encryption(cc, kp.publicKey), m_PrivateKey(kp.secretKey) {}
    std::vector<double> getPlaintext(const Ciphertext<DCRTPoly>& ct, double threshold = 1e-10) const override {
        std::vector<double> r = getDecrypt(ct)->GetRealPackedValue();
        for (auto& v : r) if (std::abs(v) < threshold) v = 0.0;
        return r;
    }
    Plaintext getDecrypt(const Ciphertext<DCRTPoly>& ct) const override {
        Plaintext p; m_cc->Decrypt(m_PrivateKey, ct, &p); return p;
    }
  private:
    PrivateKey<DCRTPoly> m_PrivateKey;
};
