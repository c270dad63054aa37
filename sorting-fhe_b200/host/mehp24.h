// mehp24.h -- MEHP24 ranking sort ("Efficient Ranking, Order Statistics, and Sorting under CKKS"), the
// single-ciphertext sortFG path and its matrix-in-slots utilities, with the reference's names
// (reference src/mehp24/mehp24_utils.{h,cpp}:21-261, src/mehp24/mehp24_sort.cpp:251-282).
//
// A length-N vector lives in an N x N slot matrix (row-major, N^2 slots).  Orchestration only: every cc-> call is
// a CUDA kernel schedule inside libckks_b200.so.  The reference reaches the context through
// c->GetCryptoContext(); here the context is an explicit first argument.
#pragma once
#include <cmath>
#include <vector>
#include "comparison.h"

namespace mehp24 {
namespace utils {

typedef Ciphertext<DCRTPoly> CT;
typedef CryptoContext<DCRTPoly> CC;

inline size_t log2ceil(size_t x) { size_t l = 0; while (((size_t)1 << l) < x) l++; return l; }

// operator<< / operator>> of mehp24_utils.h:32-40
inline CT rotl(const CC& cc, const CT& a, int32_t index) { return cc->EvalRotate(a, index); }
inline CT rotr(const CC& cc, const CT& a, int32_t index) { return cc->EvalRotate(a, -index); }
// ciphertext * vector<double>: one-argument MakeCKKSPackedPlaintext (level 0, batch-size slots), mehp24_utils.h:66-72
inline CT mulVec(const CC& cc, const CT& a, const std::vector<double>& v) { return cc->EvalMult(a, cc->MakeCKKSPackedPlaintext(v)); }
inline CT addVec(const CC& cc, const CT& a, const std::vector<double>& v) { return cc->EvalAdd(a, cc->MakeCKKSPackedPlaintext(v)); }

inline CT maskRow(const CC& cc, const CT& c, size_t n, size_t row) {        // mehp24_utils.cpp:21-30
    std::vector<double> m(n * n, 0.0);
    for (size_t i = 0; i < n; i++) m[n * row + i] = 1.0;
    return mulVec(cc, c, m);
}
inline CT maskColumn(const CC& cc, const CT& c, size_t n, size_t col) {     // :32-42
    std::vector<double> m(n * n, 0.0);
    for (size_t i = 0; i < n; i++) m[n * i + col] = 1.0;
    return mulVec(cc, c, m);
}
inline CT replicateRow(const CC& cc, CT c, size_t n) {                      // :44-50  row 0 -> every row
    for (size_t i = 0; i < log2ceil(n); i++) c = cc->EvalAdd(c, rotr(cc, c, 1 << (log2ceil(n) + i)));
    return c;
}
inline CT replicateColumn(const CC& cc, CT c, size_t n) {                   // :52-58  column 0 -> every column
    for (size_t i = 0; i < log2ceil(n); i++) c = cc->EvalAdd(c, rotr(cc, c, 1 << i));
    return c;
}
inline CT sumRows(const CC& cc, CT c, size_t n, bool maskOutput = false, size_t outputRow = 0) {   // :60-69
    for (size_t i = 0; i < log2ceil(n); i++) c = cc->EvalAdd(c, rotr(cc, c, 1 << (log2ceil(n) + i)));
    return maskOutput ? maskRow(cc, c, n, outputRow) : c;
}
inline CT sumColumns(const CC& cc, CT c, size_t n, bool maskOutput = false) {                       // :71-80
    for (size_t i = 0; i < log2ceil(n); i++) c = cc->EvalAdd(c, rotl(cc, c, 1 << i));
    return maskOutput ? maskColumn(cc, c, n, 0) : c;
}
inline CT transposeRow(const CC& cc, CT c, size_t n, bool maskOutput = false) {                     // :82-91
    for (size_t i = 1; i <= log2ceil(n); i++) c = cc->EvalAdd(c, rotr(cc, c, (int32_t)(n * (n - 1) / ((size_t)1 << i))));
    return maskOutput ? maskColumn(cc, c, n, 0) : c;
}
inline CT transposeColumn(const CC& cc, CT c, size_t n, bool maskOutput = false) {                  // :93-103
    for (size_t i = 1; i <= log2ceil(n); i++) c = cc->EvalAdd(c, rotl(cc, c, (int32_t)(n * (n - 1) / ((size_t)1 << i))));
    return maskOutput ? maskRow(cc, c, n, 0) : c;
}

// signAdv (:246-261): dg x g_3, (df - 1) x f_3, then f_3 rescaled to (1 + sign) / 2, all through EvalPolyLinear
inline CT signAdv(const CC& cc, CT c, size_t dg, size_t df) {
    const std::vector<double> f3 = {0, 35.0 / 16.0, 0, -35.0 / 16.0, 0, 21.0 / 16.0, 0, -5.0 / 16.0};
    const std::vector<double> f3_final = {0.5, 35.0 / 32.0, 0, -35.0 / 32.0, 0, 21.0 / 32.0, 0, -5.0 / 32.0};
    const std::vector<double> g3 = {0, 4589.0 / 1024.0, 0, -16577.0 / 1024.0, 0, 25614.0 / 1024.0, 0, -12860.0 / 1024.0};
    for (size_t d = 0; d < dg; d++) c = cc->EvalPolyLinear(c, g3);
    for (size_t d = 0; d + 1 < df; d++) c = cc->EvalPolyLinear(c, f3);
    return cc->EvalPolyLinear(c, f3_final);
}
// Chebyshev-interpolated predicates on c1 - c2 in [a, b] (:105-163); `error` is the dead zone around zero
inline CT equal(const CC& cc, const CT& c1, const CT& c2, double a, double b, uint32_t degree, double error = 0.00001) {
    return cc->EvalChebyshevFunction([error](double v) -> double { return (v > error || v < -error) ? 0.0 : 1.0; },
                                     cc->EvalSub(c1, c2), a, b, degree);
}
inline CT compare(const CC& cc, const CT& c1, const CT& c2, double a, double b, uint32_t degree, double error = 0.00001) {
    return cc->EvalChebyshevFunction([error](double v) -> double { return v > error ? 1.0 : v >= -error ? 0.5 : 0.0; },
                                     cc->EvalSub(c1, c2), a, b, degree);
}
inline CT compareGt(const CC& cc, const CT& c1, const CT& c2, double a, double b, uint32_t degree, double error = 0.00001) {
    return cc->EvalChebyshevFunction([error](double v) -> double { return v > error ? 1.0 : 0.0; }, cc->EvalSub(c1, c2), a, b,
                                     degree);
}
inline CT compareAdv(const CC& cc, const CT& c1, const CT& c2, size_t dg, size_t df) { return signAdv(cc, cc->EvalSub(c1, c2), dg, df); }
inline CT indicator(const CC& cc, const CT& c, double a1, double b1, double a, double b, uint32_t degree) {
    return cc->EvalChebyshevFunction([a1, b1](double v) -> double { return (v < a1 || v > b1) ? 0.0 : 1.0; }, c, a, b, degree);
}

// indicatorAdv (:166-174): 1 where |x| < 1/2 on x in (-b, b)
inline CT indicatorAdv(const CC& cc, const CT& c, double b, size_t dg, size_t df) {
    CT tmp = cc->EvalMult(c, 1.0 / b);
    CT c1 = signAdv(cc, cc->EvalAdd(tmp, 0.5 / b), dg, df);
    CT c2 = signAdv(cc, cc->EvalSub(tmp, 0.5 / b), dg, df);
    return cc->EvalMult(c1, cc->EvalSub(1.0, c2));
}
// getRotationIndices (:186-213)
inline std::vector<int32_t> getRotationIndices(size_t matrixSize) {
    int sz = (int)matrixSize;
    std::vector<int32_t> idx;
    if (matrixSize > 256) {
        for (size_t i = 0; i < matrixSize / 256; i++) { idx.push_back((int32_t)(i * 256)); idx.push_back(-(int32_t)(i * 256)); }
        sz = 256;
    }
    const int lg = (int)log2ceil((size_t)sz);
    for (int i = 0; i < lg; i++) {
        idx.push_back(1 << i); idx.push_back(-(1 << i));
        idx.push_back(-(1 << (lg + i)));
        const int32_t t = sz * (sz - 1) / (1 << (i + 1));
        idx.push_back(t); idx.push_back(-t);
    }
    return idx;
}
// multiplicative depth of the fixture per N (tests/mehp24/Mehp24SortTest.cpp:39-72)
inline int fixtureDepth(int n) {
    switch (n) { case 4: return 31; case 8: case 16: return 35; case 32: case 64: return 42; case 128: return 46;
                 case 256: return 49; case 512: return 57; case 1024: return 60; case 2048: return 64; default: return 0; }
}

}  // namespace utils

// sortFG, single ciphertext, Comparison-based overload (mehp24_sort.cpp:251-282)
inline Ciphertext<DCRTPoly> sortFG(Ciphertext<DCRTPoly> c, const size_t vectorLength, SignFunc SignFunc, SignConfig& Cfg,
                                   std::unique_ptr<Comparison>& comp, uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> m_cc) {
    using namespace utils;
    const size_t n = vectorLength;
    CT VR = replicateRow(m_cc, c, n);                                               // every row = the vector
    CT VC = replicateColumn(m_cc, transposeRow(m_cc, c, n, true), n);               // every column = the vector
    CT C = comp->compare(m_cc, VR, VC, SignFunc, Cfg);                              // C[i][j] = (v_j > v_i)
    CT R = sumRows(m_cc, C, n);                                                     // rank of v_j (+1/2) in every row
    std::vector<double> subMask(n * n);
    for (size_t i = 0; i < n; i++) for (size_t j = 0; j < n; j++) subMask[i * n + j] = -1.0 * (double)i - 0.5;
    CT M = indicatorAdv(m_cc, addVec(m_cc, R, subMask), (double)n, dg_i, df_i);     // M[i][j] = [rank_j == i]
    CT tmp = m_cc->EvalMult(M, VR);                                                 // (computed twice in the reference, :275-277)
    CT S = sumColumns(m_cc, m_cc->EvalMult(M, VR), n, true);                        // column 0: sorted values
    (void)tmp;
    return transposeColumn(m_cc, S, n, true);                                       // back to row 0
}

// splitCiphertext / combineCiphertext (mehp24_utils.cpp:266-302): block i of subLength values moved to the front / back
inline std::vector<Ciphertext<DCRTPoly>> splitCiphertext(const Ciphertext<DCRTPoly>& c, size_t totalLength, size_t subLength,
                                                         CryptoContext<DCRTPoly>& cc) {
    const size_t numParts = totalLength / subLength;
    std::vector<Ciphertext<DCRTPoly>> result(numParts);
    for (size_t i = 0; i < numParts; i++) {
        std::vector<double> mask(totalLength, 0.0);
        for (size_t j = 0; j < subLength; j++) mask[i * subLength + j] = 1.0;
        auto part = cc->EvalMult(c, cc->MakeCKKSPackedPlaintext(mask));
        if (i > 0) part = cc->EvalRotate(part, (int32_t)(i * subLength));
        result[i] = part;
    }
    return result;
}
inline Ciphertext<DCRTPoly> combineCiphertext(const std::vector<Ciphertext<DCRTPoly>>& parts, size_t subLength, CryptoContext<DCRTPoly>& cc) {
    auto result = parts[0];
    for (size_t i = 1; i < parts.size(); i++) result = cc->EvalAdd(result, cc->EvalRotate(parts[i], -(int32_t)(i * subLength)));
    return result;
}

// sortFG over several ciphertexts of subVectorLength values each, Comparison-based overload (mehp24_sort.cpp:445-605):
// block (j, k) of the comparison matrix from one compare (the lower triangle is 1 - upper), ranks from row sums of the
// column blocks plus transposed column sums of the row blocks, then one indicator per (target block, source block).
// The reference runs the blocks under OpenMP with critical-section accumulation; the sums are order-independent
// (all addends of one accumulator sit at the same level and scale, so EvalAdd is a limb-wise modular addition).
// The four stages are separate functions so that the block comparisons and the target blocks can be split over GPUs
// (sorting-fhe_b200/shard.py mehp24_sort_large_sharded); sortFG below is their composition on full ranges.
namespace blocks {
using utils::CC; using utils::CT;
inline void accumulate(const CC& cc, CT& acc, const CT& x) { acc = acc ? cc->EvalAdd(acc, x) : x; }
// replR[j]: every row = block j; replC[j]: every column = block j  (:459-482)
inline void replicas(const CC& cc, const std::vector<CT>& c, size_t n, std::vector<CT>& replR, std::vector<CT>& replC) {
    replR.resize(c.size()); replC.resize(c.size());
    for (size_t j = 0; j < c.size(); j++) replR[j] = utils::replicateRow(cc, c[j], n);
    for (size_t j = 0; j < c.size(); j++) replC[j] = utils::replicateColumn(cc, utils::transposeRow(cc, c[j], n, true), n);
}
inline size_t numCompareUnits(size_t nc) { return nc * (nc + 1) / 2; }
// comparison units u in [u_lo, u_hi) of the upper triangle (j <= k, row-major): Cv[j] += C_jk, Ch[k] += 1 - C_jk  (:484-522).
// With `fill`, accumulators this range never touched become exact zeros at the comparisons' level and scale (x - x), so
// that every shard hands the same set of same-shaped partial sums to the limb all-reduce.
template <class CompareFn>
inline void comparePartialWith(const CC& cc, const std::vector<CT>& replR, const std::vector<CT>& replC, size_t u_lo, size_t u_hi,
                               CompareFn compareBlocks, std::vector<CT>& Cv, std::vector<CT>& Ch, bool fill) {
    const size_t nc = replR.size();
    Cv.assign(nc, CT()); Ch.assign(nc, CT());
    CT any;
    size_t u = 0;
    for (size_t j = 0; j < nc; j++)
        for (size_t k = j; k < nc; k++, u++) {
            if (u < u_lo || u >= u_hi) continue;
            CT Cjk = compareBlocks(replR[j], replC[k]);
            accumulate(cc, Cv[j], Cjk);
            if (j != k) accumulate(cc, Ch[k], cc->EvalSub(1.0, Cjk));
            any = Cjk;
        }
    if (!fill) return;
    if (!any) throw std::runtime_error("mehp24 block compare: empty unit range (more shards than block comparisons)");
    CT zero = cc->EvalSub(any, any);
    for (size_t j = 0; j < nc; j++) {
        if (!Cv[j]) Cv[j] = zero->Clone();
        if (j > 0 && !Ch[j]) Ch[j] = zero->Clone();
    }
}
inline void comparePartial(const CC& cc, const std::vector<CT>& replR, const std::vector<CT>& replC, size_t u_lo, size_t u_hi,
                           SignFunc SignFunc, SignConfig& Cfg, std::unique_ptr<Comparison>& comp, std::vector<CT>& Cv,
                           std::vector<CT>& Ch, bool fill) {
    comparePartialWith(cc, replR, replC, u_lo, u_hi,
                       [&](const CT& a, const CT& b) { return comp->compare(cc, a, b, SignFunc, Cfg); }, Cv, Ch, fill);
}
// s[j] = sumRows(Cv[j]) (+ the transposed column sums of Ch[j] replicated over the rows, j >= 1)  (:524-552)
inline std::vector<CT> ranks(const CC& cc, const std::vector<CT>& Cv, const std::vector<CT>& Ch, size_t n) {
    const size_t nc = Cv.size();
    std::vector<CT> s(nc);
    for (size_t j = 0; j < nc; j++) accumulate(cc, s[j], utils::sumRows(cc, Cv[j], n));
    for (size_t j = 1; j < nc; j++) {
        CT shj = utils::sumColumns(cc, Ch[j], n, true);
        shj = utils::transposeColumn(cc, shj, n, true);
        accumulate(cc, s[j], utils::replicateRow(cc, shj, n));
    }
    return s;
}
// target block j: sum over the source blocks k of indicator(rank_k - (j n + r) - 1/2) * values_k, column sums, back to row 0  (:554-603)
// indicatorOf(x) = [|x| < 1/2]; backToRow: mask the column sums and move them to row 0 (the FG overloads do, `sort` does not)
template <class IndicatorFn>
inline CT sortedBlockWith(const CC& cc, const std::vector<CT>& replR, const std::vector<CT>& s, size_t n, size_t j, IndicatorFn indicatorOf,
                          bool backToRow) {
    const size_t nc = replR.size();
    std::vector<double> subMask(n * n);
    for (size_t r = 0; r < n; r++) for (size_t q = 0; q < n; q++) subMask[r * n + q] = -1.0 * (double)(j * n + r) - 0.5;
    CT subSorted;
    for (size_t k = 0; k < nc; k++) accumulate(cc, subSorted, cc->EvalMult(indicatorOf(utils::addVec(cc, s[k], subMask)), replR[k]));
    if (!backToRow) return utils::sumColumns(cc, subSorted, n);
    return utils::transposeColumn(cc, utils::sumColumns(cc, subSorted, n, true), n, true);
}
inline CT sortedBlock(const CC& cc, const std::vector<CT>& replR, const std::vector<CT>& s, size_t n, size_t j, uint32_t dg_i,
                      uint32_t df_i) {
    const double vectorLength = (double)(n * replR.size());
    return sortedBlockWith(cc, replR, s, n, j, [&](const CT& x) { return utils::indicatorAdv(cc, x, vectorLength, dg_i, df_i); }, true);
}
}  // namespace blocks

inline std::vector<Ciphertext<DCRTPoly>> sortFG(const std::vector<Ciphertext<DCRTPoly>>& c, const size_t subVectorLength,
                                                SignFunc SignFunc, SignConfig& Cfg, std::unique_ptr<Comparison>& comp,
                                                uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> m_cc) {
    using utils::CT;
    const size_t nc = c.size(), n = subVectorLength;
    std::vector<CT> replR, replC, Cv, Ch;
    blocks::replicas(m_cc, c, n, replR, replC);
    blocks::comparePartial(m_cc, replR, replC, 0, blocks::numCompareUnits(nc), SignFunc, Cfg, comp, Cv, Ch, false);
    std::vector<CT> s = blocks::ranks(m_cc, Cv, Ch, n);
    std::vector<CT> result(nc);
    for (size_t j = 0; j < nc; j++) result[j] = blocks::sortedBlock(m_cc, replR, s, n, j, dg_i, df_i);
    return result;
}

// sortLargeArrayFG (mehp24_sort.cpp:625-645): split -> multi-ciphertext sortFG -> combine
inline Ciphertext<DCRTPoly> sortLargeArrayFG(Ciphertext<DCRTPoly> c, const size_t totalLength, const size_t subLength, SignFunc SignFunc,
                                             SignConfig& Cfg, std::unique_ptr<Comparison>& comp, uint32_t dg_i, uint32_t df_i,
                                             CryptoContext<DCRTPoly> cc) {
    std::vector<Ciphertext<DCRTPoly>> parts = splitCiphertext(c, totalLength, subLength, cc);
    return combineCiphertext(sortFG(parts, subLength, SignFunc, Cfg, comp, dg_i, df_i, cc), subLength, cc);
}


// ---- the overloads that do not go through Comparison (SURVEY 8 A19 "other overloads"): Chebyshev-interpolated predicates
// (`sort`) and the f/g-composition predicates of mehp24_utils (`sortFG` with dg_c / df_c).

// sort, single ciphertext (mehp24_sort.cpp:23-47): result in column 0 of every row block (no masking, no transposition)
inline Ciphertext<DCRTPoly> sort(Ciphertext<DCRTPoly> c, const size_t vectorLength, double leftBoundC, double rightBoundC, uint32_t degreeC,
                                 uint32_t degreeI, CryptoContext<DCRTPoly> m_cc) {
    using namespace utils;
    const size_t n = vectorLength;
    CT VR = replicateRow(m_cc, c, n);
    CT VC = replicateColumn(m_cc, transposeRow(m_cc, c, n, true), n);
    CT C = compare(m_cc, VR, VC, leftBoundC, rightBoundC, degreeC);
    CT R = sumRows(m_cc, C, n);
    std::vector<double> subMask(n * n);
    for (size_t i = 0; i < n; i++) for (size_t j = 0; j < n; j++) subMask[i * n + j] = -1.0 * (double)i - 0.5;
    CT M = indicator(m_cc, addVec(m_cc, R, subMask), -0.5, 0.5, -1.0 * (double)n, 1.0 * (double)n, degreeI);
    return sumColumns(m_cc, m_cc->EvalMult(M, VR), n);
}
// sort over several ciphertexts (mehp24_sort.cpp:49-217): the block structure of the multi-ciphertext sortFG with the
// Chebyshev predicates; the indicator's interval is +-1.01 vectorLength (:178-180)
inline std::vector<Ciphertext<DCRTPoly>> sort(const std::vector<Ciphertext<DCRTPoly>>& c, const size_t subVectorLength, double leftBoundC,
                                              double rightBoundC, uint32_t degreeC, uint32_t degreeI, CryptoContext<DCRTPoly> m_cc) {
    using utils::CT;
    const size_t nc = c.size(), n = subVectorLength;
    const double vectorLength = (double)(n * nc);
    std::vector<CT> replR, replC, Cv, Ch;
    blocks::replicas(m_cc, c, n, replR, replC);
    blocks::comparePartialWith(m_cc, replR, replC, 0, blocks::numCompareUnits(nc),
                               [&](const CT& a, const CT& b) { return utils::compare(m_cc, a, b, leftBoundC, rightBoundC, degreeC); }, Cv, Ch, false);
    std::vector<CT> s = blocks::ranks(m_cc, Cv, Ch, n);
    std::vector<CT> result(nc);
    for (size_t j = 0; j < nc; j++)
        result[j] = blocks::sortedBlockWith(m_cc, replR, s, n, j, [&](const CT& x) {
            return utils::indicator(m_cc, x, -0.5, 0.5, -1.01 * vectorLength, 1.01 * vectorLength, degreeI); }, false);
    return result;
}
// sortFG, single ciphertext, compareAdv overload (mehp24_sort.cpp:219-249)
inline Ciphertext<DCRTPoly> sortFG(Ciphertext<DCRTPoly> c, const size_t vectorLength, uint32_t dg_c, uint32_t df_c, uint32_t dg_i, uint32_t df_i,
                                   CryptoContext<DCRTPoly> m_cc) {
    using namespace utils;
    const size_t n = vectorLength;
    CT VR = replicateRow(m_cc, c, n);
    CT VC = replicateColumn(m_cc, transposeRow(m_cc, c, n, true), n);
    CT C = compareAdv(m_cc, VR, VC, dg_c, df_c);
    CT R = sumRows(m_cc, C, n);
    std::vector<double> subMask(n * n);
    for (size_t i = 0; i < n; i++) for (size_t j = 0; j < n; j++) subMask[i * n + j] = -1.0 * (double)i - 0.5;
    CT M = indicatorAdv(m_cc, addVec(m_cc, R, subMask), (double)n, dg_i, df_i);
    CT tmp = m_cc->EvalMult(M, VR);                                                 // computed twice in the reference (:243-244)
    CT S = sumColumns(m_cc, m_cc->EvalMult(M, VR), n, true);
    (void)tmp;
    return transposeColumn(m_cc, S, n, true);
}
// sortFG over several ciphertexts, compareAdv overload (mehp24_sort.cpp:285-443)
inline std::vector<Ciphertext<DCRTPoly>> sortFG(const std::vector<Ciphertext<DCRTPoly>>& c, const size_t subVectorLength, uint32_t dg_c,
                                                uint32_t df_c, uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> m_cc) {
    using utils::CT;
    const size_t nc = c.size(), n = subVectorLength;
    std::vector<CT> replR, replC, Cv, Ch;
    blocks::replicas(m_cc, c, n, replR, replC);
    blocks::comparePartialWith(m_cc, replR, replC, 0, blocks::numCompareUnits(nc),
                               [&](const CT& a, const CT& b) { return utils::compareAdv(m_cc, a, b, dg_c, df_c); }, Cv, Ch, false);
    std::vector<CT> s = blocks::ranks(m_cc, Cv, Ch, n);
    std::vector<CT> result(nc);
    for (size_t j = 0; j < nc; j++) result[j] = blocks::sortedBlock(m_cc, replR, s, n, j, dg_i, df_i);
    return result;
}
// sortLargeArrayFG, compareAdv overload (mehp24_sort.cpp:607-623)
inline Ciphertext<DCRTPoly> sortLargeArrayFG(Ciphertext<DCRTPoly> c, const size_t totalLength, const size_t subLength, uint32_t dg_c,
                                             uint32_t df_c, uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> cc) {
    std::vector<Ciphertext<DCRTPoly>> parts = splitCiphertext(c, totalLength, subLength, cc);
    return combineCiphertext(sortFG(parts, subLength, dg_c, df_c, dg_i, df_i, cc), subLength, cc);
}

// the reference's signatures without a context argument (the context is the ciphertext's own)
inline Ciphertext<DCRTPoly> sort(Ciphertext<DCRTPoly> c, const size_t vectorLength, double leftBoundC, double rightBoundC, uint32_t degreeC,
                                 uint32_t degreeI) {
    return sort(c, vectorLength, leftBoundC, rightBoundC, degreeC, degreeI, c->GetCryptoContext());
}
inline std::vector<Ciphertext<DCRTPoly>> sort(const std::vector<Ciphertext<DCRTPoly>>& c, const size_t subVectorLength, double leftBoundC,
                                              double rightBoundC, uint32_t degreeC, uint32_t degreeI) {
    return sort(c, subVectorLength, leftBoundC, rightBoundC, degreeC, degreeI, c.at(0)->GetCryptoContext());
}
inline std::vector<Ciphertext<DCRTPoly>> sortFG(const std::vector<Ciphertext<DCRTPoly>>& c, const size_t subVectorLength, uint32_t dg_c,
                                                uint32_t df_c, uint32_t dg_i, uint32_t df_i) {
    return sortFG(c, subVectorLength, dg_c, df_c, dg_i, df_i, c.at(0)->GetCryptoContext());
}

}  // namespace mehp24
