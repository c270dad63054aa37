// serial.h -- lbcrypto::Serial / SerType and the context's key (de)serialisation members, as the reference uses them
// (src/sort.h:31-102: DeserializeFromFile for the context, public key and input ciphertext; DeserializeEvalMultKey /
// DeserializeEvalAutomorphismKey from streams; SerializeToFile for the output; CLI src/main.cpp:14-41).
//
// WHAT IS IMPLEMENTED, AND WHAT IS NOT.  OpenFHE 1.1.4 writes cereal PortableBinary archives of its own
// class hierarchy.  No OpenFHE build and no sample file exist offline, so the member-by-member schema of OpenFHE's classes
// cannot be reproduced verifiably.  This header therefore provides
//   (1) the cereal PortableBinary FRAMING layer (PbOut / PbIn below): leading endianness byte, little-endian PODs, 64-bit
//       size prefixes for strings and vectors, the shared_ptr id word (MSB = first occurrence) and class-version words;
//   (2) on top of it a flat object schema of this engine ("SFHE flat v1", documented in INTEGRATION.md section 4) holding
//       exactly the payload the engine needs -- parameters, moduli AND roots of unity (never re-derived on import), key and
//       ciphertext limbs in EVALUATION format with level / noiseScaleDeg / scaling factor / slots;
//   (3) tools/openfhe_dump.cpp: a one-page program for a machine that HAS OpenFHE 1.1.4, which loads OpenFHE's own BINARY
//       files (or live objects) and writes them in that flat schema -- the bridge through which OpenFHE-generated contexts,
//       keys and ciphertexts reach this engine and limb-level parity can be pinned.
// A file that does not start with the flat magic is rejected with a message that says so; nothing is guessed.
#pragma once
#include <cstring>
#include <fstream>
#include <istream>
#include <ostream>
#include <sstream>
#include "openfhe_shim.h"

namespace lbcrypto {

enum class SerType { JSON, BINARY };

// ---- cereal PortableBinary framing
class PbOut {
    std::ostream& os;
  public:
    explicit PbOut(std::ostream& s) : os(s) { const uint8_t little = 1; os.write((const char*)&little, 1); }   // archive header: endianness flag
    template <class T> void pod(const T& v) { os.write((const char*)&v, sizeof v); }                                 // little-endian host assumed (x86-64 / aarch64-le)
    void size(uint64_t n) { pod(n); }                                                                                // cereal::size_type = uint64
    void str(const std::string& s) { size(s.size()); os.write(s.data(), (std::streamsize)s.size()); }
    template <class T> void vec(const std::vector<T>& v) { size(v.size()); os.write((const char*)v.data(), (std::streamsize)(v.size() * sizeof(T))); }
    void words(const uint64_t* p, size_t n) { size(n); os.write((const char*)p, (std::streamsize)(n * 8)); }
    void ptr_first(uint32_t id) { pod((uint32_t)(id | 0x80000000u)); }   // shared_ptr: id with MSB set on first occurrence, data follows
    void ptr_again(uint32_t id) { pod(id); }
    void class_version(uint32_t v) { pod(v); }
    bool good() const { return os.good(); }
};
class PbIn {
    std::istream& is;
  public:
    explicit PbIn(std::istream& s) : is(s) {
        uint8_t little = 0; is.read((char*)&little, 1);
        if (!is || little != 1) throw std::runtime_error("portable binary archive: missing header or big-endian archive");
    }
    template <class T> T pod() { T v; is.read((char*)&v, sizeof v); if (!is) throw std::runtime_error("portable binary archive: truncated"); return v; }
    uint64_t size() { return pod<uint64_t>(); }
    std::string str() { uint64_t n = size(); if (n > (1u << 20)) throw std::runtime_error("portable binary archive: bad string length"); std::string s(n, 0); is.read(&s[0], (std::streamsize)n); if (!is) throw std::runtime_error("portable binary archive: truncated"); return s; }
    template <class T> std::vector<T> vec(uint64_t max_elems) {
        uint64_t n = size();
        if (n > max_elems) throw std::runtime_error("portable binary archive: vector longer than the context allows");
        std::vector<T> v(n); is.read((char*)v.data(), (std::streamsize)(n * sizeof(T)));
        if (!is) throw std::runtime_error("portable binary archive: truncated");
        return v;
    }
    uint32_t ptr(bool* first) { uint32_t id = pod<uint32_t>(); *first = (id & 0x80000000u) != 0; return id & 0x7FFFFFFFu; }
    uint32_t class_version() { return pod<uint32_t>(); }
};

// ---- SFHE flat v1 object kinds
namespace flat {
static const char kMagic[9] = "SFHEFLT1";
enum Kind : uint32_t { CONTEXT = 1, PUBLIC_KEY = 2, EVAL_MULT_KEYS = 3, EVAL_AUTO_KEYS = 4, CIPHERTEXT = 5 };
inline void begin(PbOut& a, Kind k) { a.str(std::string(kMagic, 8)); a.class_version(1); a.pod((uint32_t)k); }
inline void expect(PbIn& a, Kind k, const char* what) {
    std::string m;
    try { m = a.str(); } catch (...) { m.clear(); }
    if (m != std::string(kMagic, 8))
        throw std::runtime_error(std::string(what) + ": not an SFHE flat v1 file.  OpenFHE's native cereal archives are converted with "
                                 "tools/openfhe_dump.cpp on a machine that has OpenFHE 1.1.4 (host/serial.h)");
    if (a.class_version() != 1) throw std::runtime_error(std::string(what) + ": unsupported flat schema version");
    if (a.pod<uint32_t>() != (uint32_t)k) throw std::runtime_error(std::string(what) + ": the file holds a different kind of object");
}
}  // namespace flat

// the context the next key / ciphertext file is attached to (OpenFHE resolves this through CryptoContextFactory)
inline CryptoContext<DCRTPoly>& serial_current_context() { static CryptoContext<DCRTPoly> cc; return cc; }

namespace Serial {

// ---------------------------------------------------------------- context
inline void write_context(std::ostream& os, const CryptoContext<DCRTPoly>& cc) {
    if (!cc->hasParams) throw std::runtime_error("SerializeToFile: this context wrapper does not know its parameters (borrowed handle)");
    PbOut a(os); flat::begin(a, flat::CONTEXT);
    const auto& p = cc->params;
    int logn = 0; while ((1u << logn) < cc->GetRingDimension()) logn++;
    a.pod((uint32_t)logn); a.pod(p.multDepth); a.pod(p.scalingModSize); a.pod(p.firstModSize); a.pod(p.numLargeDigits);
    a.pod((uint32_t)(p.scalTech == FLEXIBLEAUTOEXT ? 1 : 0)); a.pod(p.batchSize);
    const int M = (int)(cc->info(1) + cc->info(2));
    std::vector<uint64_t> mod(M), root(M);
    ckks_check(ckks_ctx_get_moduli(cc->ctx, mod.data(), root.data(), M), "get_moduli");
    a.pod((uint32_t)cc->info(1)); a.pod((uint32_t)cc->info(2));
    a.vec(mod); a.vec(root);
}
inline CryptoContext<DCRTPoly> read_context(std::istream& is) {
    PbIn a(is); flat::expect(a, flat::CONTEXT, "DeserializeFromFile(CryptoContext)");
    CCParams<CryptoContextCKKSRNS> p;
    const uint32_t logn = a.pod<uint32_t>();
    p.multDepth = a.pod<uint32_t>(); p.scalingModSize = a.pod<uint32_t>(); p.firstModSize = a.pod<uint32_t>(); p.numLargeDigits = a.pod<uint32_t>();
    p.scalTech = a.pod<uint32_t>() ? FLEXIBLEAUTOEXT : FLEXIBLEAUTO; p.batchSize = a.pod<uint32_t>();
    if (logn < 8 || logn > 17) throw std::runtime_error("DeserializeFromFile(CryptoContext): bad ring dimension");
    p.ringDim = 1u << logn;
    const uint32_t sq = a.pod<uint32_t>(), sp = a.pod<uint32_t>();
    auto mod = a.vec<uint64_t>(256), root = a.vec<uint64_t>(256);
    if (mod.size() != sq + sp || root.size() != mod.size()) throw std::runtime_error("DeserializeFromFile(CryptoContext): moduli / roots count mismatch");
    ckks_ctx_t* c = nullptr;
    ckks_check(ckks_ctx_create_from_moduli(&c, (int)logn, (int)p.multDepth, (int)p.scalingModSize, (int)p.firstModSize, (int)p.numLargeDigits,
                                           p.scalTech == FLEXIBLEAUTOEXT ? 1 : 0, /*seed: OS entropy*/ 0, 0, mod.data(), root.data(), (int)mod.size()),
               "DeserializeFromFile(CryptoContext)");
    auto cc = std::make_shared<CryptoContextImpl<DCRTPoly>>(c, true, p.batchSize);
    cc->params = p; cc->hasParams = true;
    return cc;
}

// ---------------------------------------------------------------- public key: [2 (b, a)][sizeQ][n]
inline void write_public_key(std::ostream& os, const CryptoContext<DCRTPoly>& cc) {
    PbOut a(os); flat::begin(a, flat::PUBLIC_KEY);
    const size_t sq = (size_t)cc->info(1), n = cc->GetRingDimension();
    std::vector<uint64_t> buf(2 * sq * n);
    ckks_check(ckks_export_public_key(cc->ctx, buf.data(), buf.size()), "export_public_key");
    a.pod((uint32_t)sq); a.pod((uint32_t)n); a.vec(buf);
}
inline void read_public_key(std::istream& is, const CryptoContext<DCRTPoly>& cc) {
    PbIn a(is); flat::expect(a, flat::PUBLIC_KEY, "DeserializeFromFile(PublicKey)");
    const size_t sq = a.pod<uint32_t>(), n = a.pod<uint32_t>();
    if (sq != (size_t)cc->info(1) || n != cc->GetRingDimension()) throw std::runtime_error("DeserializeFromFile(PublicKey): shape does not match the context");
    auto buf = a.vec<uint64_t>(2 * sq * n);
    ckks_check(ckks_import_public_key(cc->ctx, buf.data(), buf.size()), "import_public_key");
}

// ---------------------------------------------------------------- evaluation keys: [digits][2 (b, a)][M][n] each
inline void write_one_key(PbOut& a, const CryptoContext<DCRTPoly>& cc, int is_rot, int32_t index) {
    const size_t dg = (size_t)cc->info(4), M = (size_t)(cc->info(1) + cc->info(2)), n = cc->GetRingDimension();
    std::vector<uint64_t> buf(dg * 2 * M * n);
    ckks_check(ckks_export_eval_key(cc->ctx, is_rot, index, buf.data(), buf.size()), "export_eval_key");
    a.pod((uint32_t)dg); a.pod((uint32_t)M); a.pod((uint32_t)n); a.vec(buf);
}
inline void read_one_key(PbIn& a, const CryptoContext<DCRTPoly>& cc, int is_rot, int32_t index) {
    const size_t dg = a.pod<uint32_t>(), M = a.pod<uint32_t>(), n = a.pod<uint32_t>();
    if (dg != (size_t)cc->info(4) || M != (size_t)(cc->info(1) + cc->info(2)) || n != cc->GetRingDimension())
        throw std::runtime_error("evaluation key: shape does not match the context");
    auto buf = a.vec<uint64_t>(dg * 2 * M * n);
    ckks_check(ckks_import_eval_key(cc->ctx, is_rot, index, buf.data(), buf.size()), "import_eval_key");
}
inline bool write_eval_keys(const CryptoContext<DCRTPoly>& cc, std::ostream& os, bool automorphism) {
    PbOut a(os); flat::begin(a, automorphism ? flat::EVAL_AUTO_KEYS : flat::EVAL_MULT_KEYS);
    if (!automorphism) { a.pod((uint32_t)1); write_one_key(a, cc, 0, 0); return a.good(); }
    std::vector<uint32_t> ks(4096); int cnt = 0;
    ckks_check(ckks_list_automorphism_keys(cc->ctx, ks.data(), (int)ks.size(), &cnt), "list_automorphism_keys");
    a.pod((uint32_t)cnt);
    for (int i = 0; i < cnt; i++) { a.pod(ks[i]); write_one_key(a, cc, 2, (int32_t)ks[i]); }   // keyed by automorphism index, as OpenFHE's map
    return a.good();
}
inline bool read_eval_keys(const CryptoContext<DCRTPoly>& cc, std::istream& is, bool automorphism) {
    PbIn a(is); flat::expect(a, automorphism ? flat::EVAL_AUTO_KEYS : flat::EVAL_MULT_KEYS, automorphism ? "DeserializeEvalAutomorphismKey" : "DeserializeEvalMultKey");
    const uint32_t cnt = a.pod<uint32_t>();
    if (cnt > 4096) throw std::runtime_error("evaluation key map: implausible entry count");
    for (uint32_t i = 0; i < cnt; i++) {
        if (!automorphism) read_one_key(a, cc, 0, 0);
        else { const uint32_t k = a.pod<uint32_t>(); read_one_key(a, cc, 2, (int32_t)k); }
    }
    return true;
}

// ---------------------------------------------------------------- ciphertext
inline void write_ciphertext(std::ostream& os, const Ciphertext<DCRTPoly>& ct) {
    PbOut a(os); flat::begin(a, flat::CIPHERTEXT);
    const uint32_t polys = (uint32_t)ct->get(4), limbs = (uint32_t)ct->get(3);
    auto cc = ct->GetCryptoContext();
    const size_t n = cc->GetRingDimension();
    std::vector<uint64_t> buf((size_t)polys * limbs * n);
    ckks_check(ckks_ct_export(ct->h, buf.data(), buf.size()), "ct_export");
    a.pod(polys); a.pod(limbs); a.pod((uint32_t)n); a.pod((uint32_t)ct->GetLevel()); a.pod((uint32_t)ct->GetNoiseScaleDeg());
    a.pod((uint32_t)ct->GetSlots()); a.pod(ct->GetScalingFactor()); a.vec(buf);
}
inline Ciphertext<DCRTPoly> read_ciphertext(std::istream& is, const CryptoContext<DCRTPoly>& cc) {
    PbIn a(is); flat::expect(a, flat::CIPHERTEXT, "DeserializeFromFile(Ciphertext)");
    const uint32_t polys = a.pod<uint32_t>(), limbs = a.pod<uint32_t>(), n = a.pod<uint32_t>(), level = a.pod<uint32_t>(), deg = a.pod<uint32_t>(),
                   slots = a.pod<uint32_t>();
    const double sf = a.pod<double>();
    if (n != cc->GetRingDimension() || polys < 1 || polys > 3 || limbs > 256) throw std::runtime_error("DeserializeFromFile(Ciphertext): shape does not match the context");
    auto buf = a.vec<uint64_t>((uint64_t)polys * limbs * n);
    return cc->ImportCiphertext(buf.data(), (int)polys, (int)limbs, (int)level, (int)deg, sf, (int)slots);
}

// ---------------------------------------------------------------- the file API of src/sort.h:34-99 (false on failure, like OpenFHE)
template <class F> inline bool guarded(F&& f) { try { f(); return true; } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; } }
inline bool DeserializeFromFile(const std::string& path, CryptoContext<DCRTPoly>& cc, const SerType&) {
    return guarded([&] { std::ifstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path); cc = read_context(f); serial_current_context() = cc; });
}
inline bool DeserializeFromFile(const std::string& path, PublicKey<DCRTPoly>& pk, const SerType&) {
    return guarded([&] {
        auto cc = serial_current_context(); if (!cc) throw std::runtime_error("DeserializeFromFile(PublicKey): no context has been deserialised");
        std::ifstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path);
        read_public_key(f, cc); pk = std::make_shared<KeyImpl>(KeyImpl{cc.get()}); });
}
inline bool DeserializeFromFile(const std::string& path, Ciphertext<DCRTPoly>& ct, const SerType&) {
    return guarded([&] {
        auto cc = serial_current_context(); if (!cc) throw std::runtime_error("DeserializeFromFile(Ciphertext): no context has been deserialised");
        std::ifstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path);
        ct = read_ciphertext(f, cc); });
}
inline bool SerializeToFile(const std::string& path, const CryptoContext<DCRTPoly>& cc, const SerType&) {
    return guarded([&] { std::ofstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path); write_context(f, cc); });
}
inline bool SerializeToFile(const std::string& path, const PublicKey<DCRTPoly>& pk, const SerType&) {
    return guarded([&] {
        if (!pk || !pk->cc) throw std::runtime_error("SerializeToFile(PublicKey): empty key");
        std::ofstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path);
        write_public_key(f, pk->cc->weak_from_this().lock()); });
}
inline bool SerializeToFile(const std::string& path, const Ciphertext<DCRTPoly>& ct, const SerType&) {
    return guarded([&] { std::ofstream f(path, std::ios::binary); if (!f) throw std::runtime_error("cannot open " + path); write_ciphertext(f, ct); });
}
}  // namespace Serial

// used by the CryptoContextImpl members (openfhe_shim.h) once this header is included
inline bool serial_keys_io(CryptoContextImpl<DCRTPoly>& cc, std::istream* in, std::ostream* out, bool automorphism) {
    auto self = cc.weak_from_this().lock();
    if (!self) throw std::runtime_error("key serialisation needs a shared CryptoContext");
    return Serial::guarded([&] { if (in) Serial::read_eval_keys(self, *in, automorphism); else Serial::write_eval_keys(self, *out, automorphism); });
}

}  // namespace lbcrypto
