// dropin_ref.cpp -- drives the reference's UNMODIFIED sources on the B200 engine (tests/test_dropin.py).
//
// Compiled by tests/dropin/build.sh with  -I sorting-fhe_b200/host/compat -I /root/reference/src  so that
// `#include "openfhe.h"` inside the reference resolves to the forwarding header, and linked with the reference's own
// sign.cpp, comparison.cpp, encryption.cpp, mehp24/*.cpp, k-way/*.cpp and libckks_b200.so.  Everything below main()'s
// set-up is the reference's orchestration (DirectSort<N>::sort with its OpenMP regions enabled, mehp24::sortFG,
// Comparison::compare); the limbs it produces are written to <outdir>/<case>.bin and compared by the Python test with
// what the re-expressed host layer (sorting-fhe_b200/host/) produces on the same seed.
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "openfhe.h"

#include "comparison.h"
#include "encryption.h"
#include "kway_adapter.h"
#include "mehp24/mehp24_sort.h"
#include "mehp24/mehp24_utils.h"
#include "sign.h"
#include "sort.h"        // the reference's SortContext<N> file API (src/sort.h), over host/serial.h
#include "sort_algo.h"

using namespace lbcrypto;

static void dump(const std::string& path, const Ciphertext<DCRTPoly>& ct, CryptoContext<DCRTPoly> cc) {
    int64_t parts = 0, nl = 0, level = 0, deg = 0, slots = 0;
    ckks_ct_get(ct->h, 4, &parts); ckks_ct_get(ct->h, 3, &nl); ckks_ct_get(ct->h, 1, &level); ckks_ct_get(ct->h, 2, &deg);
    ckks_ct_get(ct->h, 0, &slots);
    const size_t words = (size_t)parts * nl * cc->GetRingDimension();
    std::vector<uint64_t> buf(words);
    ckks_check(ckks_ct_export(ct->h, buf.data(), words), "export");
    std::ofstream f(path, std::ios::binary);
    int64_t hdr[6] = {parts, nl, level, deg, slots, (int64_t)cc->GetRingDimension()};
    f.write((const char*)hdr, sizeof hdr);
    f.write((const char*)buf.data(), words * 8);
}

static std::vector<double> input(int n, int seed) {   // a permutation of {0, 1/n, ...} (tests/utils.h:28-50 shape), fixed
    std::vector<double> v(n);
    for (int i = 0; i < n; i++) v[i] = (double)((i * (2 * seed + 3) + seed) % n) / n;
    return v;
}

template <int N> static void run_direct_sort(const std::string& out, int logRing) {
    CCParams<CryptoContextCKKSRNS> parameters;
    std::vector<int> rotations;
    DirectSort<N>::getSizeParameters(parameters, rotations);     // reference src/sort_algo.h:87-201
    parameters.SetSecurityLevel(HEStd_NotSet);
    parameters.SetRingDim(1u << logRing);
    parameters.SetSeed(1);                                        // engine extension: reproducible keys for the parity check
    auto cc = GenCryptoContext(parameters);
    cc->Enable(PKE); cc->Enable(KEYSWITCH); cc->Enable(LEVELEDSHE); cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    cc->EvalRotateKeyGen(keys.secretKey, rotations);
    cc->EvalMultKeyGen(keys.secretKey);
    auto enc = std::make_shared<DebugEncryption>(cc, keys);
    auto ct = enc->encryptInput(input(N, 1));
    auto sorter = std::make_unique<DirectSort<N>>(cc, keys.publicKey, rotations, enc);
    SignConfig Cfg = SignConfig(CompositeSignConfig(3, 2, 2));
    auto res = sorter->sort(ct, SignFunc::CompositeSign, Cfg);    // reference src/sort_algo.h:752-774, OpenMP regions live
    dump(out + "/direct_sort_" + std::to_string(N) + ".bin", res, cc);
    Plaintext pt; cc->Decrypt(keys.secretKey, res, &pt);
    std::cout << "direct_sort " << N << " level " << res->GetLevel() << " :";
    for (int i = 0; i < N; i++) std::cout << " " << pt->GetRealPackedValue()[i];
    std::cout << std::endl;
}

static void run_mehp24(const std::string& out, int logRing) {
    constexpr int N = 4;
    CCParams<CryptoContextCKKSRNS> parameters;
    parameters.SetSecurityLevel(HEStd_NotSet);
    parameters.SetRingDim(1u << logRing);
    parameters.SetBatchSize(N * N);
    parameters.SetMultiplicativeDepth(31);                        // tests/mehp24/Mehp24SortTest.cpp:39-41
    parameters.SetScalingModSize(40);
    parameters.SetSeed(1);
    auto cc = GenCryptoContext(parameters);
    cc->Enable(PKE); cc->Enable(KEYSWITCH); cc->Enable(LEVELEDSHE); cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    cc->EvalMultKeyGen(keys.secretKey);
    cc->EvalRotateKeyGen(keys.secretKey, mehp24::utils::getRotationIndices(N));
    auto enc = std::make_shared<DebugEncryption>(cc, keys);
    std::unique_ptr<Comparison> comp = std::make_unique<Comparison>(enc);
    auto ct = enc->encryptInput(input(N, 2));
    SignConfig Cfg = SignConfig(CompositeSignConfig(3, 2, 2));
    uint32_t dg_i = (log2(N) + 1) / 2, df_i = 2;                 // Mehp24SortTest.cpp:137-138
    auto res = mehp24::sortFG(ct, N, SignFunc::CompositeSign, Cfg, comp, dg_i, df_i, cc);
    dump(out + "/mehp24_4.bin", res, cc);
    std::cout << "mehp24 level " << res->GetLevel() << std::endl;
}

static void run_compare(const std::string& out) {                // tests/CompareTest.cpp:13-62 verbatim
    CCParams<CryptoContextCKKSRNS> parameters;
    parameters.SetMultiplicativeDepth(50);
    parameters.SetScalingModSize(59);
    parameters.SetBatchSize(4);
    parameters.SetRingDim(1 << 12);
    parameters.SetSecurityLevel(HEStd_NotSet);
    parameters.SetSeed(1);
    auto cc = GenCryptoContext(parameters);
    cc->Enable(PKE); cc->Enable(KEYSWITCH); cc->Enable(LEVELEDSHE); cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    cc->EvalMultKeyGen(keys.secretKey);
    auto enc = std::make_shared<Encryption>(cc, keys.publicKey);
    auto comp = std::make_unique<Comparison>(enc);
    auto cA = enc->encryptInput({1.0, 5.0, 3.0, 4.0});
    auto cB = enc->encryptInput({2.0, 4.0, 3.0, 3.0});
    auto Cfg = SignConfig(CompositeSignConfig(4, 3, 3));
    auto res = comp->compare(cc, cA, cB, SignFunc::CompositeSign, Cfg);
    dump(out + "/compare.bin", res, cc);
    Plaintext pt; cc->Decrypt(keys.secretKey, res, &pt);
    std::cout << "compare :";
    for (int i = 0; i < 4; i++) std::cout << " " << pt->GetRealPackedValue()[i];
    std::cout << std::endl;
}

// src/sort.h end to end: context, public key, relinearisation key, rotation keys and the input ciphertext go through files
// (Serial::SerializeToFile / SerializeEval*Key -> the reference's SortContext<4> constructor -> eval -> deserializeOutput).
// The imported context keeps the file's moduli and roots; its encryption randomness is fresh, so the check is on values.
static void run_serial(const std::string& out) {
    constexpr int N = 4;
    CCParams<CryptoContextCKKSRNS> parameters;
    std::vector<int> rotations;
    DirectSort<N>::getSizeParameters(parameters, rotations);
    parameters.SetMultiplicativeDepth(45);                         // SortContext::eval uses Sign(4,3,3) (src/sort.h:93)
    parameters.SetSecurityLevel(HEStd_NotSet);
    parameters.SetRingDim(1u << 10);
    parameters.SetSeed(1);
    auto cc = GenCryptoContext(parameters);
    cc->Enable(PKE); cc->Enable(KEYSWITCH); cc->Enable(LEVELEDSHE); cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    cc->EvalRotateKeyGen(keys.secretKey, rotations);
    cc->EvalMultKeyGen(keys.secretKey);
    auto enc = std::make_shared<DebugEncryption>(cc, keys);
    const std::vector<double> x = input(N, 3);
    auto ct = enc->encryptInput(x);
    const std::string d = out + "/";
    bool ok = Serial::SerializeToFile(d + "cc.bin", cc, SerType::BINARY) && Serial::SerializeToFile(d + "pk.bin", keys.publicKey, SerType::BINARY) &&
              Serial::SerializeToFile(d + "in.bin", ct, SerType::BINARY);
    { std::ofstream f(d + "mult.bin", std::ios::binary); ok = ok && cc->SerializeEvalMultKey(f, SerType::BINARY); }
    { std::ofstream f(d + "rot.bin", std::ios::binary); ok = ok && cc->SerializeEvalAutomorphismKey(f, SerType::BINARY); }
    if (!ok) throw std::runtime_error("serialisation failed");
    std::vector<uint64_t> sk((size_t)(cc->info(1) + cc->info(2)) * cc->GetRingDimension());
    ckks_check(ckks_export_secret(cc->ctx, sk.data(), sk.size()), "export_secret");
    // in-memory reference result
    auto sorter = std::make_unique<DirectSort<N>>(cc, keys.publicKey, rotations, enc);
    auto Cfg = SignConfig(CompositeSignConfig(4, 3, 3));
    auto direct = sorter->sort(ct, SignFunc::CompositeSign, Cfg);
    Plaintext p0; cc->Decrypt(keys.secretKey, direct, &p0);
    // the reference's file-driven path, unmodified
    SortContext<N> sc(d + "cc.bin", d + "pk.bin", d + "mult.bin", d + "rot.bin", d + "in.bin", d + "out.bin");
    sc.eval(SortAlgo::DirectSort, rotations);
    sc.deserializeOutput();
    Ciphertext<DCRTPoly> back;
    if (!Serial::DeserializeFromFile(d + "out.bin", back, SerType::BINARY)) throw std::runtime_error("cannot read the output file back");
    ckks_check(ckks_import_secret(sc.m_cc->ctx, sk.data(), sk.size()), "import_secret");   // test only: decrypt in the imported context
    Plaintext p1; sc.m_cc->Decrypt(nullptr, back, &p1);
    std::ofstream r(out + "/serial.txt");
    r.precision(17);
    for (int i = 0; i < N; i++) r << x[i] << " " << p0->GetRealPackedValue()[i] << " " << p1->GetRealPackedValue()[i] << "\n";
    std::cout << "serial level " << back->GetLevel() << " imported ring " << sc.m_cc->GetRingDimension() << std::endl;
}

// mehp24_sort.h:37-40 declares the single-ciphertext compareAdv overload without its df_i parameter, so the function that
// mehp24_sort.cpp:219-221 defines is not reachable through the header; declare the defined signature here.
namespace mehp24 {
Ciphertext<DCRTPoly> sortFG(Ciphertext<DCRTPoly> c, const size_t vectorLength, uint32_t dg_c, uint32_t df_c, uint32_t dg_i, uint32_t df_i,
                            CryptoContext<DCRTPoly> m_cc);
}
// The MEHP24 overloads that do not go through Comparison, from the reference's unmodified mehp24_sort.cpp:
// sortFG with compareAdv / indicatorAdv (:219-249), sortLargeArrayFG over two blocks (:285-443, 607-623), and `sort` with
// the Chebyshev predicates, single ciphertext (:23-47) and two blocks (:49-217).
static void run_mehp24_overloads(const std::string& out, int logRing) {
    constexpr int N = 4;
    CCParams<CryptoContextCKKSRNS> parameters;
    parameters.SetSecurityLevel(HEStd_NotSet);
    parameters.SetRingDim(1u << logRing);
    parameters.SetBatchSize(N * N);
    parameters.SetMultiplicativeDepth(45);
    parameters.SetScalingModSize(40);
    parameters.SetSeed(1);
    auto cc = GenCryptoContext(parameters);
    cc->Enable(PKE); cc->Enable(KEYSWITCH); cc->Enable(LEVELEDSHE); cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    cc->EvalMultKeyGen(keys.secretKey);
    std::vector<int32_t> rots = mehp24::utils::getRotationIndices(N);
    rots.push_back(N); rots.push_back(-N);                        // splitCiphertext / combineCiphertext of two blocks
    cc->EvalRotateKeyGen(keys.secretKey, rots);
    auto enc = std::make_shared<DebugEncryption>(cc, keys);
    auto ct = enc->encryptInput(input(N, 2));
    dump(out + "/mehp24_adv_4.bin", mehp24::sortFG(ct, (size_t)N, (uint32_t)3, (uint32_t)2, (uint32_t)2, (uint32_t)2, cc), cc);
    dump(out + "/mehp24_cheb_4.bin", mehp24::sort(ct, N, -1.0, 1.0, 59, 119), cc);
    auto ct8 = enc->encryptInput(input(2 * N, 3));
    dump(out + "/mehp24_large_adv_8.bin", mehp24::sortLargeArrayFG(ct8, 2 * N, N, 3, 2, 2, 2, cc), cc);
    auto parts = mehp24::utils::splitCiphertext(ct8, 2 * N, N, cc);
    auto sorted = mehp24::sort(parts, N, -1.0, 1.0, 119, 247);   // 8 values 1/8 apart need the higher degrees (0.03 vs 0.14 max error)
    dump(out + "/mehp24_cheb_multi_8_0.bin", sorted[0], cc);
    dump(out + "/mehp24_cheb_multi_8_1.bin", sorted[1], cc);
}

int main(int argc, char** argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: dropin_ref <outdir>\n"); return 2; }
    const std::string out = argv[1];
    try {
        run_direct_sort<4>(out, 10);
        run_direct_sort<8>(out, 10);
        run_mehp24(out, 10);
        run_mehp24_overloads(out, 10);
        run_compare(out);
        run_serial(out);
    } catch (const std::exception& e) { std::fprintf(stderr, "dropin_ref: %s\n", e.what()); return 1; }
    std::cout << "dropin ok" << std::endl;
    return 0;
}
