// openfhe_shim.h -- the lbcrypto value types the reference's sorter headers use, re-created on top of
// the C-ABI of the B200 engine (include/ckks_b200.h).
//
// The reference has no FFI seam: its headers call lbcrypto::CryptoContextImpl<DCRTPoly> members after
// `#include "openfhe.h"` (reference src/sort_algo.h:3-12, src/comparison.h:3-9, src/sign.h:3-4,
// src/rotation.h:1-10, src/sort.h:1-13).  This header provides exactly those names -- CryptoContext<DCRTPoly>,
// Ciphertext<DCRTPoly>, Plaintext, KeyPair/PublicKey/PrivateKey, CCParams<CryptoContextCKKSRNS>,
// GenCryptoContext -- with the same call signatures and shared_ptr value semantics, each member forwarding
// to one extern "C" entry point.  OpenFHE exceptions become std::runtime_error carrying ckks_last_error().
// No arithmetic happens on the host: every member ends in a CUDA kernel launch inside libckks_b200.so.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <deque>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/ckks_b200.h"

namespace lbcrypto {

struct DCRTPoly {};               // element tag (all polynomials live on the device)
struct CryptoContextCKKSRNS {};   // scheme tag for CCParams
enum PKESchemeFeature { PKE = 1, KEYSWITCH = 2, LEVELEDSHE = 4, ADVANCEDSHE = 8, FHE = 16 };
enum SecurityLevel { HEStd_128_classic, HEStd_192_classic, HEStd_256_classic, HEStd_NotSet };
enum ScalingTechnique { FIXEDMANUAL, FIXEDAUTO, FLEXIBLEAUTO, FLEXIBLEAUTOEXT };
enum KeySwitchTechnique { BV, HYBRID };
enum SecretKeyDist { GAUSSIAN, UNIFORM_TERNARY, SPARSE_TERNARY };

inline void ckks_check(int rc, const char* what) {
    if (rc != 0) throw std::runtime_error(std::string(what) + ": " + ckks_last_error());
}

template <class Scheme> class CCParams;
template <> class CCParams<CryptoContextCKKSRNS> {
  public:
    uint32_t multDepth = 1, scalingModSize = 50, firstModSize = 60, batchSize = 0, ringDim = 0, numLargeDigits = 0;
    SecurityLevel security = HEStd_128_classic;
    ScalingTechnique scalTech = FLEXIBLEAUTOEXT;
    uint64_t seed = 0;   // engine extension: seed of the counter-based sampler; 0 = keyed from OS entropy (the default, as
                         // OpenFHE's PRNG); a fixed non-zero seed gives reproducible keys for parity tests and is not secure
    void SetSeed(uint64_t s) { seed = s; }
    void SetMultiplicativeDepth(uint32_t d) { multDepth = d; }
    void SetScalingModSize(uint32_t b) { scalingModSize = b; }
    void SetFirstModSize(uint32_t b) { firstModSize = b; }
    void SetBatchSize(uint32_t b) { batchSize = b; }
    void SetRingDim(uint32_t r) { ringDim = r; }
    void SetSecurityLevel(SecurityLevel s) { security = s; }
    void SetScalingTechnique(ScalingTechnique t) { scalTech = t; }
    void SetNumLargeDigits(uint32_t n) { numLargeDigits = n; }
    void SetKeySwitchTechnique(KeySwitchTechnique) {}
    void SetSecretKeyDist(SecretKeyDist) {}
    uint32_t GetMultiplicativeDepth() const { return multDepth; }
    uint32_t GetRingDim() const { return ringDim; }
    uint32_t GetBatchSize() const { return batchSize; }
};

template <class Element> class CryptoContextImpl;

// Ciphertext<DCRTPoly> / encoded Plaintext: owns (or borrows) one ckks_ct_t handle.  A class template on the element tag,
// like OpenFHE's, so that function templates over Ciphertext<Element> (reference src/mehp24/mehp24_utils.h:32-88) deduce.
template <class Element> class CiphertextImpl {
  public:
    ckks_ct_t* h = nullptr;
    bool owned = true;
    std::weak_ptr<CryptoContextImpl<Element>> cc;   // weak: the context's plaintext cache holds ciphertext handles
    std::shared_ptr<CryptoContextImpl<Element>> GetCryptoContext() const {
        auto p = cc.lock();
        if (!p) throw std::runtime_error("GetCryptoContext: the ciphertext's context is gone");
        return p;
    }
    CiphertextImpl(ckks_ct_t* handle, bool own = true) : h(handle), owned(own) {}
    CiphertextImpl(const CiphertextImpl&) = delete;
    ~CiphertextImpl() { if (h && owned) ckks_ct_free(h); }
    int64_t get(int what) const { int64_t v = 0; ckks_check(ckks_ct_get(h, what, &v), "ct_get"); return v; }
    size_t GetSlots() const { return (size_t)get(0); }
    size_t GetLevel() const { return (size_t)get(1); }
    size_t GetNoiseScaleDeg() const { return (size_t)get(2); }
    double GetScalingFactor() const { double s = 0; ckks_check(ckks_ct_get_scaling_factor(h, &s), "ct_sf"); return s; }
    // SetSlots mutates metadata through const references exactly like the reference relies on
    // (reference src/sort_algo.h:429,434,447,501,614,711)
    void SetSlots(size_t s) const { ckks_check(ckks_ct_set_slots(h, (int)s), "SetSlots"); }
    std::shared_ptr<CiphertextImpl> Clone() const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_ct_clone(h, &o), "Clone");
        auto p = std::make_shared<CiphertextImpl>(o);
        p->cc = cc;
        return p;
    }
};
template <class E> using Ciphertext = std::shared_ptr<CiphertextImpl<E>>;
template <class E> using ConstCiphertext = std::shared_ptr<const CiphertextImpl<E>>;

// Plaintext: either an encoded device polynomial (MakeCKKSPackedPlaintext) or a decrypted value vector.
class PlaintextImpl {
  public:
    std::shared_ptr<CiphertextImpl<DCRTPoly>> enc;   // encoded form (1 polynomial)
    std::vector<double> values;            // decoded form (Decrypt)
    double logPrecision = 0;
    void SetLength(size_t n) { values.resize(n); }
    const std::vector<double>& GetRealPackedValue() const { return values; }
    std::vector<double> GetCKKSPackedValueReal() const { return values; }
    double GetLogPrecision() const { return logPrecision; }
    size_t GetLevel() const { return enc ? enc->GetLevel() : 0; }
    size_t GetSlots() const { return enc ? enc->GetSlots() : values.size(); }
};
typedef std::shared_ptr<PlaintextImpl> Plaintext;
typedef std::shared_ptr<const PlaintextImpl> ConstPlaintext;

struct KeyImpl { CryptoContextImpl<DCRTPoly>* cc; };
template <class E> using PublicKey = std::shared_ptr<KeyImpl>;
template <class E> using PrivateKey = std::shared_ptr<KeyImpl>;
template <class E> struct KeyPair {
    PublicKey<E> publicKey; PrivateKey<E> secretKey;
    bool good() const { return publicKey && secretKey; }
};

struct EncodingParamsImpl { uint32_t batch; uint32_t GetBatchSize() const { return batch; } };

// CryptoContext<DCRTPoly>: one engine context (owned when created by GenCryptoContext, borrowed when
// wrapped around a handle that arrived through the C entry points of sortfhe_host.h).
template <class Element> class CryptoContextImpl : public std::enable_shared_from_this<CryptoContextImpl<Element>> {
    typedef std::shared_ptr<CiphertextImpl<Element>> CT;
    CT wrap(ckks_ct_t* o) const {
        auto p = std::make_shared<CiphertextImpl<Element>>(o);
        p->cc = std::const_pointer_cast<CryptoContextImpl>(this->weak_from_this().lock());
        return p;
    }
    typedef int (*BinFn)(ckks_ctx_t*, const ckks_ct_t*, const ckks_ct_t*, ckks_ct_t**);
    typedef int (*CstFn)(ckks_ctx_t*, const ckks_ct_t*, double, ckks_ct_t**);
    CT bin(BinFn f, const CT& a, const CT& b, const char* name) const {
        if (!a || !b) throw std::runtime_error(std::string(name) + ": null ciphertext");
        ckks_ct_t* o = nullptr; ckks_check(f(ctx, a->h, b->h, &o), name); return wrap(o);
    }
    CT cst(CstFn f, const CT& a, double c, const char* name) const {
        if (!a) throw std::runtime_error(std::string(name) + ": null ciphertext");
        ckks_ct_t* o = nullptr; ckks_check(f(ctx, a->h, c, &o), name); return wrap(o);
    }
    static const CT& enc_of(const Plaintext& p) {
        if (!p || !p->enc) throw std::runtime_error("plaintext is not encoded");
        return p->enc;
    }

  public:
    ckks_ctx_t* ctx = nullptr;
    bool owned = false;
    uint32_t batchSize = 0;
    CCParams<CryptoContextCKKSRNS> params;   // as given to GenCryptoContext / read from a context file (host/serial.h)
    bool hasParams = false;
    // (de)serialisation of the evaluation-key maps (reference src/sort.h:52,63); defined by host/serial.h
    template <class ST> bool DeserializeEvalMultKey(std::istream& s, const ST&) { return serial_keys_io(*this, &s, nullptr, false); }
    template <class ST> bool DeserializeEvalAutomorphismKey(std::istream& s, const ST&) { return serial_keys_io(*this, &s, nullptr, true); }
    template <class ST> bool SerializeEvalMultKey(std::ostream& s, const ST&, const std::string& = "") { return serial_keys_io(*this, nullptr, &s, false); }
    template <class ST> bool SerializeEvalAutomorphismKey(std::ostream& s, const ST&, const std::string& = "") { return serial_keys_io(*this, nullptr, &s, true); }
    // raw-limb landing hook of the importer
    std::shared_ptr<CiphertextImpl<Element>> ImportCiphertext(const uint64_t* limbs, int polys, int nl, int level, int deg, double sf, int slots) const {
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_ct_import(ctx, limbs, polys, nl, level, deg, sf, slots, &o), "ct_import");
        return wrap(o);
    }
    CryptoContextImpl(ckks_ctx_t* c, bool own, uint32_t batch) : ctx(c), owned(own), batchSize(batch) {}
    CryptoContextImpl(const CryptoContextImpl&) = delete;
    ~CryptoContextImpl() { if (ctx && owned) ckks_ctx_destroy(ctx); }

    int64_t info(int what) const { int64_t v = 0; ckks_check(ckks_ctx_info(ctx, what, &v), "ctx_info"); return v; }
    void Enable(PKESchemeFeature) {}
    uint32_t GetRingDimension() const { return (uint32_t)info(0); }
    uint32_t GetCyclotomicOrder() const { return (uint32_t)info(5); }
    std::shared_ptr<EncodingParamsImpl> GetEncodingParams() const {
        return std::make_shared<EncodingParamsImpl>(EncodingParamsImpl{batchSize ? batchSize : GetRingDimension() / 2});
    }

    // ---- keys (tests/DirectSortTest.cpp:40-57)
    KeyPair<DCRTPoly> KeyGen() {
        ckks_check(ckks_keygen(ctx), "KeyGen");
        auto k = std::make_shared<KeyImpl>(KeyImpl{this});
        return KeyPair<DCRTPoly>{k, k};
    }
    void EvalMultKeyGen(const PrivateKey<DCRTPoly>&) { ckks_check(ckks_eval_mult_keygen(ctx), "EvalMultKeyGen"); }
    void EvalRotateKeyGen(const PrivateKey<DCRTPoly>&, const std::vector<int>& idx) {
        std::vector<int32_t> v(idx.begin(), idx.end());
        ckks_check(ckks_eval_rotate_keygen(ctx, v.data(), (int)v.size()), "EvalRotateKeyGen");
    }
    void ClearEvalMultKeys() {}
    void ClearEvalAutomorphismKeys() {}

    // ---- encode / encrypt / decrypt
    // Encoded-plaintext cache.  The sorters encode the same 0/1 masks and index patterns in every sort (258 per
    // DirectSort<128>::sort: reference src/sort_algo.h:341,356,573,706,718); they depend on (N, ring, level, slots) only.
    // An encoded Plaintext is immutable, so MakeCKKSPackedPlaintext hands out the SAME handle for the same
    // (values, scaleDeg, level, slots) -- bit-identical by construction, and the level / scale adjustments the engine
    // caches on a plaintext handle (rescaled and lifted twins) are reused with it.  Entries are found by a 128-bit hash
    // of the values and confirmed by comparing the values themselves; the keyed form below skips generating and hashing
    // the vector.  FIFO eviction above SORTFHE_PT_CACHE_MB of device bytes (default: half of the free device memory at first use); SORTFHE_PT_CACHE_MB=0 turns
    // the cache off.  Guarded by a mutex: the reference encodes from OpenMP threads (sort_algo.h:349-362).
    struct PtKey {
        uint64_t h[2]; uint32_t scaleDeg, level, slots, len;
        bool operator<(const PtKey& o) const {
            return std::tie(h[0], h[1], scaleDeg, level, slots, len) < std::tie(o.h[0], o.h[1], o.scaleDeg, o.level, o.slots, o.len);
        }
    };
    struct PtEntry { std::vector<double> values; Plaintext pt; size_t bytes; };
    mutable std::mutex pt_mu;
    mutable std::map<PtKey, PtEntry> pt_cache;
    mutable std::deque<PtKey> pt_order;
    mutable size_t pt_bytes = 0, pt_hits = 0, pt_misses = 0;
    mutable size_t pt_cap = (size_t)-1;
    size_t pt_cap_bytes() const {   // SORTFHE_PT_CACHE_MB, else half of the device memory that is free when the cache is first used
        if (pt_cap == (size_t)-1) {
            const char* e = getenv("SORTFHE_PT_CACHE_MB");
            if (e) pt_cap = (size_t)atoll(e) << 20;
            else { const int64_t fr = info(10); pt_cap = fr > 0 ? (size_t)fr / 2 : (size_t)16384 << 20; }
        }
        return pt_cap;
    }
    static void hash128(const void* data, size_t bytes, uint64_t out[2]) {   // two multiply-xorshift lanes over 8-byte words
        const uint64_t* w = (const uint64_t*)data; const size_t nw = bytes / 8;
        uint64_t a = 0x9E3779B97F4A7C15ull ^ bytes, b = 0xC2B2AE3D27D4EB4Full + bytes;
        for (size_t i = 0; i < nw; i++) {
            a = (a ^ w[i]) * 0xFF51AFD7ED558CCDull; a ^= a >> 32;
            b = (b + w[i]) * 0xC4CEB9FE1A85EC53ull; b ^= b >> 29;
        }
        out[0] = a; out[1] = b;
    }
    Plaintext encodeNow(const std::vector<double>& v, size_t scaleDeg, uint32_t level, uint32_t s) const {
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_make_packed_plaintext(ctx, v.data(), (int)v.size(), (int)scaleDeg, (int)level, (int)s, &o),
                   "MakeCKKSPackedPlaintext");
        auto p = std::make_shared<PlaintextImpl>();
        p->enc = wrap(o);
        return p;
    }
    void ptInsert(const PtKey& k, PtEntry&& e) const {   // pt_mu held
        const size_t cap = pt_cap_bytes();
        while (!pt_order.empty() && pt_bytes + e.bytes > cap) {
            auto it = pt_cache.find(pt_order.front());
            if (it != pt_cache.end()) { pt_bytes -= it->second.bytes; pt_cache.erase(it); }
            pt_order.pop_front();
        }
        if (e.bytes > cap) return;
        pt_bytes += e.bytes;
        pt_order.push_back(k);
        pt_cache.emplace(k, std::move(e));
    }
    size_t ptDeviceBytes(uint32_t level, uint32_t slots) const {
        // encoded form plus its two cached twins; compact storage keeps min(ring, 2 * slots) (>= 256) words per limb
        const size_t ring = (size_t)info(0), words = std::min(ring, std::max<size_t>(256, 2 * (size_t)slots));
        return (size_t)3 * (size_t)std::max<int64_t>(1, info(1) - (int64_t)level) * words * 8;
    }
    Plaintext MakeCKKSPackedPlaintext(const std::vector<double>& v, size_t scaleDeg = 1, uint32_t level = 0,
                                      const void* params = nullptr, uint32_t slots = 0) const {
        (void)params;
        uint32_t s = slots ? slots : (batchSize ? batchSize : GetRingDimension() / 2);
        if (pt_cap_bytes() == 0) return encodeNow(v, scaleDeg, level, s);
        PtKey k; hash128(v.data(), v.size() * sizeof(double), k.h);
        k.scaleDeg = (uint32_t)scaleDeg; k.level = level; k.slots = s; k.len = (uint32_t)v.size();
        std::lock_guard<std::mutex> g(pt_mu);
        auto it = pt_cache.find(k);
        if (it != pt_cache.end() && it->second.values == v) { pt_hits++; return it->second.pt; }
        pt_misses++;
        Plaintext p = encodeNow(v, scaleDeg, level, s);
        if (it == pt_cache.end()) ptInsert(k, PtEntry{v, p, ptDeviceBytes(level, s)});
        return p;
    }
    // Keyed form for generated patterns: `key` names the pattern (generator tag + its integer arguments); the vector is
    // only generated on a miss.  Extension used by host/sort_algo.h; the result is the handle the plain call would return.
    template <class Gen>
    Plaintext MakeCKKSPackedPlaintextKeyed(const std::vector<int64_t>& key, Gen&& gen, size_t scaleDeg, uint32_t level,
                                           uint32_t slots) const {
        if (pt_cap_bytes() == 0) return encodeNow(gen(), scaleDeg, level, slots);
        PtKey k; hash128(key.data(), key.size() * sizeof(int64_t), k.h);
        k.h[0] ^= 0x5bd1e9955bd1e995ull;                  // keyed entries live in their own hash space
        k.scaleDeg = (uint32_t)scaleDeg; k.level = level; k.slots = slots; k.len = (uint32_t)key.size() | 0x80000000u;
        std::lock_guard<std::mutex> g(pt_mu);
        auto it = pt_cache.find(k);
        if (it != pt_cache.end() && it->second.values.size() == key.size() &&
            std::equal(key.begin(), key.end(), it->second.values.begin(), [](int64_t a, double b) { return (double)a == b; })) {
            pt_hits++;
            return it->second.pt;
        }
        pt_misses++;
        Plaintext p = encodeNow(gen(), scaleDeg, level, slots);
        if (it == pt_cache.end()) ptInsert(k, PtEntry{std::vector<double>(key.begin(), key.end()), p, ptDeviceBytes(level, slots)});
        return p;
    }
    void ClearPlaintextCache() const { std::lock_guard<std::mutex> g(pt_mu); pt_cache.clear(); pt_order.clear(); pt_bytes = 0; }
    CT Encrypt(const PublicKey<DCRTPoly>&, const Plaintext& pt) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_encrypt(ctx, enc_of(pt)->h, &o), "Encrypt"); return wrap(o);
    }
    void Decrypt(const PrivateKey<DCRTPoly>&, const CT& ct, Plaintext* out) const {
        auto p = std::make_shared<PlaintextImpl>();
        p->values.assign(ct->GetSlots(), 0.0);
        ckks_check(ckks_decrypt(ctx, ct->h, p->values.data(), (int)p->values.size()), "Decrypt");
        *out = p;
    }

    // ---- additions (reference: ~160 call sites, e.g. sort_algo.h:358,454,708,720; comparison.cpp:13,19)
    CT EvalAdd(const CT& a, const CT& b) const { return bin(ckks_eval_add, a, b, "EvalAdd"); }
    CT EvalAdd(const CT& a, const Plaintext& p) const { return bin(ckks_eval_add, a, enc_of(p), "EvalAdd"); }
    CT EvalAdd(const Plaintext& p, const CT& a) const { return bin(ckks_eval_add, a, enc_of(p), "EvalAdd"); }
    CT EvalAdd(const CT& a, double c) const { return cst(ckks_eval_add_const, a, c, "EvalAdd"); }
    CT EvalAdd(double c, const CT& a) const { return cst(ckks_eval_add_const, a, c, "EvalAdd"); }
    CT EvalSub(const CT& a, const CT& b) const { return bin(ckks_eval_sub, a, b, "EvalSub"); }
    CT EvalSub(const CT& a, const Plaintext& p) const { return bin(ckks_eval_sub, a, enc_of(p), "EvalSub"); }
    CT EvalSub(const Plaintext& p, const CT& a) const { return bin(ckks_eval_sub, enc_of(p), a, "EvalSub"); }
    CT EvalSub(const CT& a, double c) const { return cst(ckks_eval_sub_const, a, c, "EvalSub"); }
    CT EvalSub(double c, const CT& a) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_const_sub(ctx, c, a->h, &o), "EvalSub"); return wrap(o);
    }
    template <class B> void EvalAddInPlace(CT& a, const B& b) const { a = EvalAdd(a, b); }
    template <class B> void EvalSubInPlace(CT& a, const B& b) const { a = EvalSub(a, b); }
    CT EvalNegate(const CT& a) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_negate(ctx, a->h, &o), "EvalNegate"); return wrap(o);
    }
    void EvalNegateInPlace(CT& a) const { a = EvalNegate(a); }
    CT EvalAddMany(const std::vector<CT>& v) const {
        std::vector<const ckks_ct_t*> hs; for (auto& c : v) hs.push_back(c->h);
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_add_many(ctx, hs.data(), (int)hs.size(), &o), "EvalAddMany");
        return wrap(o);
    }
    // Extension (not an OpenFHE call): T = init; T = EvalAdd(T, EvalMult(cts[i], pts[i])) for every i, as one engine call
    // (the masked sums of src/sort_algo.h:336-345, 567-577).  Same result as the chain.
    CT EvalInnerProductPlain(const CT& init, const std::vector<CT>& cts, const std::vector<Plaintext>& pts) const {
        if (cts.size() != pts.size()) throw std::runtime_error("EvalInnerProductPlain: size mismatch");
        std::vector<const ckks_ct_t*> a, p;
        for (auto& c : cts) a.push_back(c->h);
        for (auto& q : pts) p.push_back(enc_of(q)->h);
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_eval_inner_product_plain(ctx, init->h, a.data(), p.data(), (int)a.size(), &o), "EvalInnerProductPlain");
        CT r = wrap(o);
        return r;
    }
    // outs[g] = EvalInnerProductPlain(init, cts, pts[g]) for every mask set g, as one engine call that reads the ciphertexts once per
    // pass of four outputs (ckks_eval_inner_product_plain_multi)
    std::vector<CT> EvalInnerProductPlainMulti(const CT& init, const std::vector<CT>& cts, const std::vector<std::vector<Plaintext>>& pts) const {
        std::vector<const ckks_ct_t*> a, p;
        for (auto& c : cts) a.push_back(c->h);
        for (auto& g : pts) {
            if (g.size() != cts.size()) throw std::runtime_error("EvalInnerProductPlainMulti: size mismatch");
            for (auto& q : g) p.push_back(enc_of(q)->h);
        }
        // SORTFHE_NO_IP_MULTI=1: one engine call per mask set (A/B; measured with the integer accumulate: 51.98 vs 51.49 ms per
        // DirectSort N=128 sort on B200 at two sets per pass; the FP64 accumulate of the 40-bit limbs made four sets per pass the default)
        static const bool single = getenv("SORTFHE_NO_IP_MULTI") != nullptr;
        if (single) {
            std::vector<CT> r;
            for (auto& g : pts) r.push_back(EvalInnerProductPlain(init, cts, g));
            return r;
        }
        std::vector<ckks_ct_t*> o(pts.size(), nullptr);
        ckks_check(ckks_eval_inner_product_plain_multi(ctx, init->h, a.data(), p.data(), (int)a.size(), (int)pts.size(), o.data()), "EvalInnerProductPlainMulti");
        std::vector<CT> r;
        for (ckks_ct_t* h : o) r.push_back(wrap(h));
        return r;
    }
    // ---- multiplications (sign.cpp:23-56,100-156; comparison.cpp:19,35-38; sort_algo.h:358,576,723,730)
    CT EvalMult(const CT& a, const CT& b) const { return bin(ckks_eval_mult, a, b, "EvalMult"); }
    CT EvalMult(const CT& a, const Plaintext& p) const { return bin(ckks_eval_mult, a, enc_of(p), "EvalMult"); }
    CT EvalMult(const Plaintext& p, const CT& a) const { return bin(ckks_eval_mult, a, enc_of(p), "EvalMult"); }
    CT EvalMult(const CT& a, double c) const { return cst(ckks_eval_mult_const, a, c, "EvalMult"); }
    CT EvalMult(double c, const CT& a) const { return cst(ckks_eval_mult_const, a, c, "EvalMult"); }
    void EvalMultInPlace(CT& a, double c) const { a = EvalMult(a, c); }
    CT EvalMultAndRelinearize(const CT& a, const CT& b) const { return bin(ckks_eval_mult, a, b, "EvalMultAndRelinearize"); }
    CT EvalSquare(const CT& a) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_square(ctx, a->h, &o), "EvalSquare"); return wrap(o);
    }
    // ---- rotations (rotation.h:224,231; mehp24_utils.h:32-40)
    CT EvalRotate(const CT& a, int32_t index) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_rotate(ctx, a->h, index, &o), "EvalRotate"); return wrap(o);
    }
    // EvalFastRotationPrecompute / EvalFastRotation (rotation.h:286,344-347): the digits are an opaque handle
    CT EvalFastRotationPrecompute(const CT& a) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_fast_rotation_precompute(ctx, a->h, &o), "EvalFastRotationPrecompute"); return wrap(o);
    }
    CT EvalFastRotation(const CT& a, int32_t index, uint32_t /*cyclotomicOrder*/, const CT& digits) const {
        ckks_ct_t* o = nullptr; ckks_check(ckks_fast_rotation(ctx, a->h, index, digits->h, &o), "EvalFastRotation"); return wrap(o);
    }
    bool HasRotationKey(int32_t index) const { return ckks_has_rotation_key(ctx, index) != 0; }
    // ---- polynomial evaluators (sort_algo.h:727-728, sign.cpp:76, mehp24_utils.cpp:256-259)
    CT EvalChebyshevSeriesPS(const CT& x, const std::vector<double>& coeffs, double a, double b) const {
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_eval_chebyshev_series_ps(ctx, x->h, coeffs.data(), (int)coeffs.size(), a, b, &o), "EvalChebyshevSeriesPS");
        return wrap(o);
    }
    CT EvalChebyshevSeries(const CT& x, const std::vector<double>& coeffs, double a, double b) const {
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_eval_chebyshev_series(ctx, x->h, coeffs.data(), (int)coeffs.size(), a, b, &o), "EvalChebyshevSeries");
        return wrap(o);
    }
    // Chebyshev interpolation at the degree+1 Chebyshev nodes of [a, b] (the reference's EvalChebyshevFunction call
    // sites: mehp24_utils.cpp:109-163, sign.cpp:616-633); c_0 is stored doubled, the series evaluators halve it
    static std::vector<double> EvalChebyshevCoefficients(const std::function<double(double)>& func, double a, double b,
                                                         uint32_t degree) {
        const uint32_t cnt = degree + 1;
        const double half = 0.5 * (b - a), mid = 0.5 * (b + a), step = M_PI / cnt;
        std::vector<double> f(cnt), c(cnt);
        for (uint32_t j = 0; j < cnt; j++) f[j] = func(std::cos(step * (j + 0.5)) * half + mid);
        for (uint32_t i = 0; i < cnt; i++) {
            double acc = 0;
            for (uint32_t j = 0; j < cnt; j++) acc += f[j] * std::cos(step * i * (j + 0.5));
            c[i] = acc * 2.0 / cnt;
        }
        return c;
    }
    CT EvalChebyshevFunction(const std::function<double(double)>& func, const CT& x, double a, double b, uint32_t degree) const {
        return EvalChebyshevSeries(x, EvalChebyshevCoefficients(func, a, b, degree), a, b);
    }
    CT EvalPolyLinear(const CT& x, const std::vector<double>& coeffs) const {
        ckks_ct_t* o = nullptr;
        ckks_check(ckks_eval_poly_linear(ctx, x->h, coeffs.data(), (int)coeffs.size(), &o), "EvalPolyLinear");
        return wrap(o);
    }
    // CKKS bootstrapping (EvalUtils.cpp:76, sign.cpp:168; set-up: tests/k-way/KWaySort2Test.cpp:47-49).  Sparse packing
    // (slots <= ring / 4); dim1 is accepted and ignored; numIterations / precision of the Meta-BTS variant are ignored.
    void EvalBootstrapSetup(std::vector<uint32_t> levelBudget = {5, 4}, std::vector<uint32_t> dim1 = {0, 0}, uint32_t slots = 0,
                            uint32_t /*correctionFactor*/ = 0, bool /*precompute*/ = true) {
        if (levelBudget.size() != 2) throw std::runtime_error("EvalBootstrapSetup: levelBudget needs two entries");
        dim1.resize(2, 0);
        ckks_check(ckks_eval_bootstrap_setup(ctx, (int)levelBudget[0], (int)levelBudget[1], (int)dim1[0], (int)dim1[1], (int)slots),
                   "EvalBootstrapSetup");
    }
    void EvalBootstrapKeyGen(const PrivateKey<DCRTPoly>&, uint32_t slots) {
        ckks_check(ckks_eval_bootstrap_keygen(ctx, (int)slots), "EvalBootstrapKeyGen");
    }
    CT EvalBootstrap(const CT& a, uint32_t /*numIterations*/ = 1, uint32_t /*precision*/ = 0) const {
        if (!a) throw std::runtime_error("EvalBootstrap: null ciphertext");
        ckks_ct_t* o = nullptr; ckks_check(ckks_eval_bootstrap(ctx, a->h, &o), "EvalBootstrap"); return wrap(o);
    }
};
template <class E> using CryptoContext = std::shared_ptr<CryptoContextImpl<E>>;

// Ciphertext operators of OpenFHE's ciphertext.h (used by reference src/mehp24/mehp24_sort.cpp:43,106,119,145)
template <class E> Ciphertext<E> operator+(const Ciphertext<E>& a, const Ciphertext<E>& b) { return a->GetCryptoContext()->EvalAdd(a, b); }
template <class E> Ciphertext<E> operator-(const Ciphertext<E>& a, const Ciphertext<E>& b) { return a->GetCryptoContext()->EvalSub(a, b); }
template <class E> Ciphertext<E> operator*(const Ciphertext<E>& a, const Ciphertext<E>& b) { return a->GetCryptoContext()->EvalMult(a, b); }
template <class E> Ciphertext<E> operator-(const Ciphertext<E>& a) { return a->GetCryptoContext()->EvalNegate(a); }
template <class E> const Ciphertext<E>& operator+=(Ciphertext<E>& a, const Ciphertext<E>& b) { return a = a + b; }
template <class E> const Ciphertext<E>& operator-=(Ciphertext<E>& a, const Ciphertext<E>& b) { return a = a - b; }
template <class E> const Ciphertext<E>& operator*=(Ciphertext<E>& a, const Ciphertext<E>& b) { return a = a * b; }

// GenCryptoContext(CCParams<CryptoContextCKKSRNS>) -- tests/DirectSortTest.cpp:29-44
inline CryptoContext<DCRTPoly> GenCryptoContext(const CCParams<CryptoContextCKKSRNS>& p) {
    if (p.ringDim == 0 || (p.ringDim & (p.ringDim - 1))) throw std::runtime_error("GenCryptoContext: SetRingDim(2^k) is required");
    int logn = 0; while ((1u << logn) < p.ringDim) logn++;
    int tech = p.scalTech == FLEXIBLEAUTOEXT ? 1 : 0;
    if (p.scalTech != FLEXIBLEAUTOEXT && p.scalTech != FLEXIBLEAUTO)
        throw std::runtime_error("GenCryptoContext: only FLEXIBLEAUTO / FLEXIBLEAUTOEXT are implemented");
    ckks_ctx_t* c = nullptr;
    ckks_check(ckks_ctx_create(&c, logn, (int)p.multDepth, (int)p.scalingModSize, (int)p.firstModSize, (int)p.numLargeDigits,
                               tech, p.seed, 0), "GenCryptoContext");
    auto cc = std::make_shared<CryptoContextImpl<DCRTPoly>>(c, true, p.batchSize);
    cc->params = p; cc->hasParams = true;
    return cc;
}

}  // namespace lbcrypto
