// kway.h -- the k-way sorting network of HKC+21 with the reference's names (namespace kwaySort): Masking
// (src/k-way/Masking.cpp:26-167), EvalUtils (src/k-way/EvalUtils.cpp:18-147), SortUtils (src/k-way/SortUtils.cpp:5-82,
// 202-236), Sorter (src/k-way/Sorter.cpp:8-70,180-262,284-404) and KWayAdapter's parameter card
// (src/kway_adapter.h:41-64).  k = 2, 3 and 5 (BASELINE config 4 is k = 2; k = 5 adds the 4-/5-sorters, SortUtils.cpp:80-200,
// 239-433, Sorter.cpp:71-178, 264-282).  checkLevelAndBoot calls cc->EvalBootstrap exactly where the reference does
// (EvalUtils.cpp:57-86); the engine's bootstrapping (csrc/bootstrap.cu) refreshes the ciphertext there, so the whole
// 28-stage k = 2, N = 128 network runs at KWayAdapter's depth-40 card.  Orchestration only: every cc-> call is a CUDA
// kernel schedule inside libckks_b200.so.
#pragma once
#include <cmath>
#include <tuple>
#include <vector>
#include "comparison.h"

namespace kwaySort {

typedef Ciphertext<DCRTPoly> CT;

// ---------------------------------------------------------------- Masking (pure index logic)
inline long ipow(long b, long e) { long r = 1; while (e-- > 0) r *= b; return r; }

// stage -> (m, logDist, slope)   (Masking.cpp:26-47)
inline std::tuple<int, int, int> sortType(int k, int M, int stage) {
    (void)M;
    const int up = (k + 1) / 2;
    int r = 0;
    while (stage >= (r + 1 + r * (r + 1) / 2 * up)) r++;
    const int n = stage - (r + r * (r - 1) / 2 * up);
    const int m = (n + up - 1) / up;
    const int slope = n == 0 ? 0 : ((n - 1) % up) + 1;
    return std::make_tuple(m, r - m, slope);
}

// role table [2][numSlots]: row 0 = size of the group a slot is sorted in, row 1 = its position   (Masking.cpp:49-143)
inline std::vector<std::vector<int>> genIndices(long numSlots, long k, long M, long m, long logDist, long slope) {
    std::vector<std::vector<int>> res(2, std::vector<int>((size_t)numSlots, 0));
    const long km = ipow(k, m), dist = ipow(k, logDist), next = ipow(k, m + 1), total = ipow(k, M);
    auto diagonal = [&](long start, long row, long col) {
        long loc = 1;
        while (row < km && col >= 0) {
            for (long d = 0; d < dist; d++) {
                const long here = start + dist * (col + k * row) + d;
                res[0][here] = (int)loc;
                if (row == km - 1 || col - slope < 0) {
                    for (long i = 0; i < loc; i++) {
                        const long h2 = start + dist * ((col + i * slope) + k * (row - i)) + d;
                        res[1][h2] = (int)(loc - i);
                        res[0][h2] += (int)i;
                    }
                }
            }
            loc++; row++; col -= slope;
        }
    };
    for (long start = 0; start < total; start += dist * next) {
        if (slope == 0) {
            for (long s = 0; s < km; s++)
                for (long col = 0; col < k; col++)
                    for (long d = 0; d < dist; d++) {
                        const long here = start + dist * (s + km * col) + d;
                        res[0][here] = (int)k; res[1][here] = (int)(col + 1);
                    }
        } else if (slope > k / 2) {
            for (long t = 0; t < km - 1; t++) {
                const long col = k - k / 2;
                for (long loc = 1; loc < k; loc++)
                    for (long d = 0; d < dist; d++) {
                        const long here = start + dist * (col + k * t + loc - 1) + d;
                        res[0][here] = (int)(k - 1); res[1][here] = (int)loc;
                    }
            }
        } else {
            for (long t = slope; t < k; t++) diagonal(start, 0, t);
            for (long s = 1; s < km - 1; s++)
                for (long t = k - slope; t < k; t++) diagonal(start, s, t);
        }
    }
    return res;
}

inline void genMask(const std::vector<std::vector<int>>& indices, long index0, long index1, std::vector<double>& mask) {   // :145-155
    mask.assign(indices[0].size(), 0.0);
    for (size_t i = 0; i < indices[0].size(); i++)
        if (indices[0][i] == index0 && indices[1][i] == index1) mask[i] = 1.0;
}

inline long getRotateDistance(long k, long logDist, long slope) {   // :157-166
    const long dist = ipow(k, logDist);
    return (slope == 0 || slope == k / 2 + 1) ? dist : dist * (k - slope);
}

// ---------------------------------------------------------------- EvalUtils (EvalUtils.cpp:18-147)
class EvalUtils {
  protected:
    CryptoContext<DCRTPoly> m_cc;
    std::shared_ptr<Encryption> m_enc;

  public:
    EvalUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc) : m_cc(cc), m_enc(enc) {}
    virtual ~EvalUtils() = default;

    void multByInt(CT& ctxt, long coeff, CT& ctxt_out) {   // double-and-add (:18-43)
        CT origin = coeff < 0 ? m_cc->EvalNegate(ctxt) : ctxt;
        if (coeff < 0) coeff = -coeff;
        ctxt_out = origin;
        int top = 0;
        while ((coeff >> (top + 1)) > 0) top++;
        for (int i = top; i > 0; i--) {
            ctxt_out = m_cc->EvalAdd(ctxt_out, ctxt_out);
            if ((coeff >> (i - 1)) & 1) ctxt_out = m_cc->EvalAdd(ctxt_out, origin);
        }
    }
    // bootstrap when fewer than level + 1 levels are left (:57-86)
    void checkLevelAndBoot(CT& ctxt, int level, int multDepth, bool = false) {
        if (multDepth - (int)ctxt->GetLevel() < level + 1) ctxt = m_cc->EvalBootstrap(ctxt);
    }
    void checkLevelAndBoot2(CT& a, CT& b, long level, long multDepth, bool v = false) {
        checkLevelAndBoot(a, (int)level, (int)multDepth, v); checkLevelAndBoot(b, (int)level, (int)multDepth, v);
    }
    void flipCtxt(CT& ctxt) { ctxt = m_cc->EvalAdd(m_cc->EvalNegate(ctxt), 1.0); }                     // 1 - x   (:96-99)
    void flipCtxt(CT& ctxt, Plaintext& mask) { ctxt = m_cc->EvalAdd(m_cc->EvalNegate(ctxt), mask); }   // mask - x (:101-104)
    // rotations by r as a chain over the set bits of r, lowest first (:106-147)
    void leftRotate(CT& ctxt, long r, CT& ctxt_out) { rotateBits(ctxt, r, +1, ctxt_out); }
    void rightRotate(CT& ctxt, long r, CT& ctxt_out) { rotateBits(ctxt, r, -1, ctxt_out); }

  private:
    void rotateBits(CT& ctxt, long r, int dir, CT& ctxt_out) {
        CT cur = ctxt;
        for (long p = 1; r > 0; r >>= 1, p <<= 1)
            if (r & 1) cur = m_cc->EvalRotate(cur, (int32_t)(dir * p));
        ctxt_out = cur;
    }
};

// ---------------------------------------------------------------- SortUtils (SortUtils.cpp:5-82, 202-236)
class SortUtils : public EvalUtils {
  protected:
    long m_numSlots, m_k, m_M;

  public:
    SortUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M)
        : EvalUtils(cc, enc), m_numSlots(numSlots), m_k(k), m_M(M) {}

    void fcnL(CT& a, CT& b, CT& comp, CT& out) { out = m_cc->EvalAdd(m_cc->EvalMult(m_cc->EvalSub(a, b), comp), b); }   // (a>b)(a-b)+b
    void compareMax(CT& a, CT& b, CT& comp, CT& out) { fcnL(a, b, comp, out); }
    void compareMin(CT& a, CT& b, CT& comp, CT& out) { fcnL(b, a, comp, out); }
    void twoSorter(CT& a, CT& b, CT& comp, CT* out) {        // [min, max]
        fcnL(a, b, comp, out[1]);
        out[0] = m_cc->EvalSub(m_cc->EvalAdd(a, b), out[1]);
    }
    void twoSorter(CT* ctxt, CT& comp, CT* out) { twoSorter(ctxt[0], ctxt[1], comp, out); }
    // ctxt = [a, b, c], comp = [(a>b), (a>c), (b>c)] -> [min, mid, max]   (:59-82)
    void threeSorter(CT* ctxt, CT* comp, CT* sorted) {
        CT ab[2], abVsC[2];
        twoSorter(ctxt[0], ctxt[1], comp[0], ab);
        twoSorter(comp[1], comp[2], comp[0], abVsC);
        compareMax(ab[1], ctxt[2], abVsC[1], sorted[2]);
        compareMin(ab[0], ctxt[2], abVsC[0], sorted[0]);
        CT sum = m_cc->EvalAdd(m_cc->EvalAdd(ctxt[0], ctxt[1]), ctxt[2]);
        sorted[1] = m_cc->EvalSub(m_cc->EvalSub(sum, sorted[0]), sorted[2]);
    }
    // ctxt = [a, b, c, d], comp = [(a>b), (a>c), (a>d), (b>c), (b>d), (c>d)] -> ascending   (:80-129)
    void fourSorter(CT* ctxt, CT* comp, CT* sorted) {
        CT ab[2], cd[2], abVsC[2], abVsD[2], hiVs[2], loVs[2], left, right;
        twoSorter(ctxt[0], ctxt[1], comp[0], ab);
        twoSorter(ctxt[2], ctxt[3], comp[5], cd);
        twoSorter(comp[1], comp[3], comp[0], abVsC);        // min / max(a, b) against c
        twoSorter(comp[2], comp[4], comp[0], abVsD);        // ... against d
        twoSorter(abVsC[1], abVsD[1], comp[5], hiVs);       // max(a, b) against min / max(c, d)
        twoSorter(abVsC[0], abVsD[0], comp[5], loVs);       // min(a, b) against min / max(c, d)
        compareMax(ab[1], cd[1], hiVs[1], sorted[3]);
        compareMax(ab[0], cd[1], loVs[1], left);
        compareMax(ab[1], cd[0], hiVs[0], right);
        compareMax(left, right, hiVs[1], sorted[2]);
        compareMin(ab[0], cd[0], loVs[0], sorted[0]);
        sorted[1] = ctxt[0];
        for (int i = 1; i < 4; i++) sorted[1] = m_cc->EvalAdd(sorted[1], ctxt[i]);
        for (int i = 0; i < 4; i++) if (i != 1) sorted[1] = m_cc->EvalSub(sorted[1], sorted[i]);
    }
    // ctxt = [a..e], comp = [a>b a>c a>d a>e b>c b>d b>e c>d c>e d>e] -> ascending   (:131-200)
    void fiveSorter(CT* ctxt, CT* comp, CT* sorted) {
        CT abc[3] = {ctxt[0], ctxt[1], ctxt[2]}, abcComp[3] = {comp[0], comp[1], comp[4]}, abcSorted[3], de[2];
        threeSorter(abc, abcComp, abcSorted);
        twoSorter(ctxt[3], ctxt[4], comp[9], de);
        CT vsD[3] = {comp[2], comp[5], comp[7]}, vsE[3] = {comp[3], comp[6], comp[8]}, vsDs[3], vsEs[3];
        threeSorter(vsD, abcComp, vsDs);
        threeSorter(vsE, abcComp, vsEs);
        CT hi[2], mid[2], lo[2], left, right;
        twoSorter(vsDs[2], vsEs[2], comp[9], hi);           // max(a,b,c) against min / max(d, e)
        twoSorter(vsDs[1], vsEs[1], comp[9], mid);
        twoSorter(vsDs[0], vsEs[0], comp[9], lo);
        compareMax(abcSorted[2], de[1], hi[1], sorted[4]);
        compareMin(abcSorted[0], de[0], lo[0], sorted[0]);
        compareMax(abcSorted[1], de[1], mid[1], left);
        compareMax(abcSorted[2], de[0], hi[0], right);
        compareMax(left, right, hi[1], sorted[3]);
        compareMin(abcSorted[1], de[0], mid[0], left);
        compareMin(abcSorted[0], de[1], lo[1], right);
        compareMin(left, right, lo[0], sorted[1]);
        sorted[2] = ctxt[0];
        for (int i = 1; i < 5; i++) sorted[2] = m_cc->EvalAdd(sorted[2], ctxt[i]);
        for (int i = 0; i < 5; i++) if (i != 2) sorted[2] = m_cc->EvalSub(sorted[2], sorted[i]);
    }
    void slotMatching2(CT& ctxt, CT& comp, std::vector<std::vector<int>>&, long shift, CT* out, CT& comp_out) {   // :202-210
        out[0] = ctxt;
        leftRotate(ctxt, shift, out[1]);
        comp_out = comp;
    }
    void slotMatching3(CT& ctxt, CT& comp, std::vector<std::vector<int>>& indices, long shift, CT* out, CT* comp_out) {   // :212-236
        std::vector<double> mask3;
        genMask(indices, 3, 1, mask3);
        Plaintext mask3Plain = m_cc->MakeCKKSPackedPlaintext(mask3);
        for (int i = 0; i < 3; i++) leftRotate(ctxt, i * shift, out[i]);
        comp_out[1] = comp;
        leftRotate(comp, shift, comp_out[0]);
        leftRotate(comp, 2 * shift, comp_out[2]);
        flipCtxt(comp_out[0], mask3Plain);
        flipCtxt(comp_out[2], mask3Plain);
    }
    // 4 elements spaced `shift` apart and their six comparison bits out of the two partner comparisons (:239-284; the
    // masked products the reference computes first at :259-264 are overwritten at :267-273 and left out here)
    void slotMatching4(CT& ctxt, CT& comp1, CT& comp2, std::vector<std::vector<int>>& indices, long shift, CT* arr, CT* comp_arr) {
        std::vector<double> m;
        genMask(indices, 4, 1, m);
        Plaintext mask = m_cc->MakeCKKSPackedPlaintext(m);
        comp_arr[2] = comp1;
        leftRotate(comp1, 1 * shift, comp_arr[0]);
        leftRotate(comp1, 2 * shift, comp_arr[3]);
        leftRotate(comp1, 3 * shift, comp_arr[5]);
        comp_arr[1] = comp2;
        leftRotate(comp2, 1 * shift, comp_arr[4]);
        flipCtxt(comp_arr[0], mask);
        flipCtxt(comp_arr[3], mask);
        flipCtxt(comp_arr[5], mask);
        for (int i = 0; i < 4; i++) {
            leftRotate(ctxt, i * shift, arr[i]);
            arr[i] = m_cc->EvalMult(arr[i], mask);
        }
    }
    void slotMatching5(CT& ctxt, CT& comp1, CT& comp2, std::vector<std::vector<int>>& indices, long shift, CT* arr, CT* comp_arr) {   // :286-323
        std::vector<double> m;
        genMask(indices, 5, 1, m);
        Plaintext mask = m_cc->MakeCKKSPackedPlaintext(m);
        for (int i = 0; i < 5; i++) leftRotate(ctxt, i * shift, arr[i]);
        comp_arr[3] = comp1;
        leftRotate(comp1, shift, comp_arr[0]);
        leftRotate(comp1, 2 * shift, comp_arr[4]);
        leftRotate(comp1, 3 * shift, comp_arr[7]);
        leftRotate(comp1, 4 * shift, comp_arr[9]);
        comp_arr[2] = comp2;
        leftRotate(comp2, shift, comp_arr[6]);
        leftRotate(comp2, 2 * shift, comp_arr[1]);
        leftRotate(comp2, 3 * shift, comp_arr[5]);
        leftRotate(comp2, 4 * shift, comp_arr[8]);
        for (int i : {0, 1, 4, 5, 7, 8, 9}) flipCtxt(comp_arr[i], mask);
    }
    // role masks of a mixed stage: bit s of `sizes` selects the groups of size s (first position)
    Plaintext sizeMask(const std::vector<std::vector<int>>& indices, unsigned sizes) {
        std::vector<double> m((size_t)m_numSlots, 0.0);
        for (long i = 0; i < m_numSlots; i++)
            if (indices[1][i] == 1 && indices[0][i] >= 2 && indices[0][i] <= 5 && ((sizes >> indices[0][i]) & 1)) m[(size_t)i] = 1.0;
        return m_cc->MakeCKKSPackedPlaintext(m);
    }
    // groups of 2, 3, 4 and 5 elements in the same stage, all fed to the 5-sorter (:325-423)
    void slotMatching2345(CT& ctxt, CT& comp1, CT& comp2, std::vector<std::vector<int>>& indices, long shift, CT* arr, CT* comp_arr) {
        for (int i = 0; i < 5; i++) leftRotate(ctxt, i * shift, arr[i]);
        Plaintext m2345 = sizeMask(indices, 0x3C), m45 = sizeMask(indices, 0x30), m345 = sizeMask(indices, 0x38),
                  m3 = sizeMask(indices, 0x08), m4 = sizeMask(indices, 0x10), m5 = sizeMask(indices, 0x20);
        leftRotate(comp1, shift, comp_arr[0]);                                   // a > b
        flipCtxt(comp_arr[0], m2345);
        CT tmp = m_cc->EvalMult(comp1, m3), tmp2 = comp2->Clone();               // a > c
        leftRotate(tmp2, 2 * shift, tmp2);
        tmp2 = m_cc->EvalMult(tmp2, m45);
        flipCtxt(tmp2, m45);
        comp_arr[1] = m_cc->EvalAdd(tmp, tmp2);
        comp_arr[2] = m_cc->EvalAdd(m_cc->EvalMult(comp1, m4), m_cc->EvalMult(comp2, m5));   // a > d
        comp_arr[3] = m_cc->EvalMult(comp1, m5);                                 // a > e
        auto rotMaskFlip = [&](CT& src, long r, Plaintext& mask, bool flip, CT& out) {
            leftRotate(src, r, out);
            out = m_cc->EvalMult(out, mask);
            if (flip) flipCtxt(out, mask);
        };
        rotMaskFlip(comp1, 2 * shift, m345, true, comp_arr[4]);                  // b > c
        rotMaskFlip(comp2, 3 * shift, m45, true, comp_arr[5]);                   // b > d
        rotMaskFlip(comp2, shift, m5, false, comp_arr[6]);                       // b > e
        rotMaskFlip(comp1, 3 * shift, m45, true, comp_arr[7]);                   // c > d
        rotMaskFlip(comp2, 4 * shift, m5, true, comp_arr[8]);                    // c > e
        rotMaskFlip(comp1, 4 * shift, m5, true, comp_arr[9]);                    // d > e
    }
    void slotAssemble(CT* sorted, long num, long shift, CT& out) {   // :425-433
        out = sorted[0];
        for (long i = 1; i < num; i++) {
            CT rotated;
            rightRotate(sorted[i], i * shift, rotated);
            out = m_cc->EvalAdd(out, rotated);
        }
    }
};

// ---------------------------------------------------------------- Sorter (Sorter.cpp)
class Sorter : public SortUtils {
  protected:
    std::vector<int> m_level = {0, 1, 3, 5, 6, 7};   // levels an m-sorter consumes (Sorter.h:86-94)
    Comparison comp;

    Plaintext roleMask(const std::vector<std::vector<int>>& indices, long size, long pos) {
        std::vector<double> m;
        genMask(indices, size, pos, m);
        return m_cc->MakeCKKSPackedPlaintext(m);
    }

  public:
    Sorter(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M)
        : SortUtils(cc, enc, numSlots, k, M) {}

    // 2-sorter on the slots whose role is (2, 1); the maxima go back `shift` slots to the right (:8-38)
    void runTwoSorter(CT& ctxt, std::vector<std::vector<int>>& indices, long shift, CT& ctxt_comp, CT& ctxt_out) {
        Plaintext mask2Plain = roleMask(indices, 2, 1);
        CT slots[2], comp_slots, sorted[2], rotated;
        slotMatching2(ctxt, ctxt_comp, indices, shift, slots, comp_slots);
        twoSorter(slots, ctxt_comp, sorted);
        for (int i = 0; i < 2; i++) sorted[i] = m_cc->EvalMult(sorted[i], mask2Plain);
        rightRotate(sorted[1], shift, rotated);
        ctxt_out = m_cc->EvalAdd(sorted[0], rotated);
    }
    void runThreeSorter(CT& ctxt, std::vector<std::vector<int>>& indices, long shift, CT& ctxt_comp, CT& ctxt_out) {   // :40-70
        Plaintext mask3Plain = roleMask(indices, 3, 1);
        CT arr[3], comp_arr[3], sorted[3], r1, r2;
        slotMatching3(ctxt, ctxt_comp, indices, shift, arr, comp_arr);
        threeSorter(arr, comp_arr, sorted);
        for (int i = 0; i < 3; i++) sorted[i] = m_cc->EvalMult(sorted[i], mask3Plain);
        rightRotate(sorted[1], shift, r1);
        rightRotate(sorted[2], 2 * shift, r2);
        ctxt_out = m_cc->EvalAdd(m_cc->EvalAdd(sorted[0], r1), r2);
    }

    void runFourSorter(CT& ctxt, std::vector<std::vector<int>>& indices, long shift, CT& comp1, CT& comp2, CT& ctxt_out) {   // :71-87
        CT arr[4], comp_arr[6], sorted[4];
        slotMatching4(ctxt, comp1, comp2, indices, shift, arr, comp_arr);
        fourSorter(arr, comp_arr, sorted);
        slotAssemble(sorted, 4, shift, ctxt_out);
    }
    void runFiveSorter(CT& ctxt, std::vector<std::vector<int>>& indices, long shift, CT& comp1, CT& comp2, CT& ctxt_out) {   // :89-119
        Plaintext mask5Plain = roleMask(indices, 5, 1);
        CT arr[5], comp_arr[10], sorted[5];
        slotMatching5(ctxt, comp1, comp2, indices, shift, arr, comp_arr);
        fiveSorter(arr, comp_arr, sorted);
        for (int i = 0; i < 5; i++) sorted[i] = m_cc->EvalMult(sorted[i], mask5Plain);
        slotAssemble(sorted, 5, shift, ctxt_out);
    }
    void run2345Sorter(CT& ctxt, std::vector<std::vector<int>>& indices, long shift, CT& comp1, CT& comp2, CT& ctxt_out) {   // :121-178
        Plaintext m2345 = sizeMask(indices, 0x3C), m45 = sizeMask(indices, 0x30), m345 = sizeMask(indices, 0x38), m5 = sizeMask(indices, 0x20);
        CT arr[5], comp_arr[10], sorted[5];
        slotMatching2345(ctxt, comp1, comp2, indices, shift, arr, comp_arr);
        fiveSorter(arr, comp_arr, sorted);
        sorted[0] = m_cc->EvalMult(sorted[0], m2345);
        sorted[1] = m_cc->EvalMult(sorted[1], m2345);
        sorted[2] = m_cc->EvalMult(sorted[2], m345);
        sorted[3] = m_cc->EvalMult(sorted[3], m45);
        sorted[4] = m_cc->EvalMult(sorted[4], m5);
        slotAssemble(sorted, 5, shift, ctxt_out);
    }

    // partner alignment: ctxt_rot[i] = the element slot i is compared with; ctxt_fix = the slots that sit this stage
    // out (:180-249)
    void rightRotateForSort(CT& ctxt, std::vector<std::vector<int>>& indices, long logDist, long slope, CT& ctxt_rot, CT& ctxt_fix) {
        std::vector<double> maskLeft((size_t)m_numSlots, 0.0);
        std::vector<std::vector<double>> maskRight((size_t)m_k, std::vector<double>((size_t)m_numSlots, 0.0));
        for (long i = 0; i < m_numSlots; i++) {
            if (indices[1][i] < indices[0][i]) maskLeft[i] = 1.0;
            if (indices[0][i] > 0 && indices[0][i] == indices[1][i]) maskRight[indices[0][i] - 1][i] = 1.0;
        }
        CT left = m_cc->EvalMult(ctxt, m_cc->MakeCKKSPackedPlaintext(maskLeft));
        const long rot = getRotateDistance(m_k, logDist, slope);
        auto masked = [&](long which) { return m_cc->EvalMult(ctxt, m_cc->MakeCKKSPackedPlaintext(maskRight[(size_t)which])); };
        if (slope == 0 || slope == m_k / 2 + 1) {
            const long top = slope == 0 ? m_k - 1 : m_k - 2;
            CT right = masked(top), rotRight;
            if (slope != 0) ctxt_fix = m_cc->EvalSub(m_cc->EvalSub(ctxt, left), right);
            leftRotate(right, top * rot, rotRight);
            rightRotate(left, rot, ctxt_rot);
            ctxt_rot = m_cc->EvalAdd(ctxt_rot, rotRight);
        } else {
            std::vector<CT> right((size_t)m_k);
            for (long i = 0; i < m_k; i++) right[(size_t)i] = masked(i);
            ctxt_fix = m_cc->EvalSub(ctxt, left);
            for (long i = 0; i < m_k; i++) ctxt_fix = m_cc->EvalSub(ctxt_fix, right[(size_t)i]);
            rightRotate(left, rot, ctxt_rot);
            for (long i = 1; i < m_k; i++) {
                CT rotated;
                leftRotate(right[(size_t)i], i * rot, rotated);
                ctxt_rot = m_cc->EvalAdd(ctxt_rot, rotated);
            }
        }
    }
    void comparisonForSort(CT& ctxt, std::vector<std::vector<int>>& indices, long logDist, long slope, CT& ctxt_comp, CT& ctxt_fix,
                           SignConfig& Cfg) {   // :251-262
        CT ctxt_rot;
        rightRotateForSort(ctxt, indices, logDist, slope, ctxt_rot, ctxt_fix);
        ctxt_comp = comp.compare(m_cc, ctxt->Clone(), ctxt_rot, SignFunc::CompositeSign, Cfg);
    }

    // the partners one and two steps away (k = 5 groups need both)   (:264-282)
    void comparisonForSort2(CT& ctxt, std::vector<std::vector<int>>& indices, long logDist, long slope, CT& comp1, CT& comp2,
                            CT& ctxt_fix, SignConfig& Cfg) {
        CT rot1, rot2, dummy;
        rightRotateForSort(ctxt, indices, logDist, slope, rot1, ctxt_fix);
        rightRotateForSort(rot1, indices, logDist, slope, rot2, dummy);
        comp1 = comp.compare(m_cc, ctxt->Clone(), rot1, SignFunc::CompositeSign, Cfg);
        comp2 = comp.compare(m_cc, ctxt->Clone(), rot2, SignFunc::CompositeSign, Cfg);
    }

    int stageCount() const { return (int)(m_M + m_M * (m_M - 1) / 2 * ((m_k + 1) / 2)); }   // :298

    // one stage of the network, in place on ctxt (body of the loop at :300-397)
    void runStage(CT& ctxt, int stage, SignConfig& Cfg) {
        const int multDepth = Cfg.multDepth;
        long m, logDist, slope;
        { auto t = sortType((int)m_k, (int)m_M, stage); m = std::get<0>(t); logDist = std::get<1>(t); slope = std::get<2>(t); }
        const long shift = getRotateDistance(m_k, logDist, slope);
        auto indices = genIndices(m_numSlots, m_k, m_M, m, logDist, slope);
        CT ctxt_fix, ctxt_comp1, ctxt_comp2;
        if (m_k != 2 && m_k != 3 && m_k != 5) throw std::runtime_error("kwaySort::Sorter: k must be 2, 3 or 5");
        if (m_k == 5 && slope != 2) {               // k = 5: two partner comparisons feed a 5-, 4- or mixed sorter (:326-366)
            const size_t lv = slope == 5 / 2 + 1 ? 4 : 5;
            checkLevelAndBoot(ctxt, m_level[lv], multDepth);
            comparisonForSort2(ctxt, indices, logDist, slope, ctxt_comp1, ctxt_comp2, ctxt_fix, Cfg);
            if (slope == 5 / 2 + 1) checkLevelAndBoot(ctxt, m_level[lv], multDepth);
            checkLevelAndBoot2(ctxt_comp1, ctxt_comp2, m_level[lv], multDepth);
            if (slope == 0) { runFiveSorter(ctxt, indices, shift, ctxt_comp1, ctxt_comp2, ctxt); return; }
            if (slope == 5 / 2 + 1) runFourSorter(ctxt, indices, shift, ctxt_comp1, ctxt_comp2, ctxt);
            else run2345Sorter(ctxt, indices, shift, ctxt_comp1, ctxt_comp2, ctxt);   // slope 1
            ctxt = m_cc->EvalAdd(ctxt, ctxt_fix);
            return;
        }
        if (slope == 0) {
            checkLevelAndBoot(ctxt, m_level[(size_t)m_k], multDepth);
            comparisonForSort(ctxt, indices, logDist, slope, ctxt_comp1, ctxt_fix, Cfg);
            checkLevelAndBoot(ctxt_comp1, m_level[(size_t)m_k], multDepth);
            if (m_k == 2) runTwoSorter(ctxt, indices, shift, ctxt_comp1, ctxt);
            else runThreeSorter(ctxt, indices, shift, ctxt_comp1, ctxt);
        } else if (slope == m_k / 2 + 1) {          // k = 3, slope 2: pairs only
            checkLevelAndBoot(ctxt, m_level[(size_t)m_k - 1], multDepth);
            comparisonForSort(ctxt, indices, logDist, slope, ctxt_comp1, ctxt_fix, Cfg);
            checkLevelAndBoot(ctxt_comp1, m_level[(size_t)m_k - 1], multDepth);
            runTwoSorter(ctxt, indices, shift, ctxt_comp1, ctxt);
            ctxt = m_cc->EvalAdd(ctxt, ctxt_fix);
        } else if ((m_k == 3 && slope == 1) || (m_k == 5 && slope == 2)) {   // diagonals of 2 and of 3 in the same stage
            CT ctxt2, ctxt3;
            checkLevelAndBoot(ctxt, m_level[3], multDepth);
            comparisonForSort(ctxt, indices, logDist, slope, ctxt_comp1, ctxt_fix, Cfg);
            checkLevelAndBoot(ctxt_comp1, m_level[2], multDepth);
            runTwoSorter(ctxt, indices, shift, ctxt_comp1, ctxt2);
            checkLevelAndBoot(ctxt_comp1, m_level[3], multDepth);
            runThreeSorter(ctxt, indices, shift, ctxt_comp1, ctxt3);
            ctxt = m_cc->EvalAdd(m_cc->EvalAdd(ctxt2, ctxt_fix), ctxt3);
        } else {                                    // k = 2, slope 1
            CT ctxt2;
            checkLevelAndBoot(ctxt, m_level[2], multDepth);
            comparisonForSort(ctxt, indices, logDist, slope, ctxt_comp1, ctxt_fix, Cfg);
            checkLevelAndBoot(ctxt_comp1, m_level[2], multDepth);
            runTwoSorter(ctxt, indices, shift, ctxt_comp1, ctxt2);
            ctxt = m_cc->EvalAdd(ctxt2, ctxt_fix);
        }
    }
    // Sorter::sorter (:284-404); last_stage < 0 runs the whole network
    void sorter(CT& ctxt, CT& ctxt_out, SignConfig& Cfg, int first_stage = 0, int last_stage = -1) {
        const int end = last_stage < 0 ? stageCount() : last_stage;
        for (int stage = first_stage; stage < end; stage++) runStage(ctxt, stage, Cfg);
        ctxt_out = ctxt;
    }
};

// KWayAdapter<N>::getSizeParameters (kway_adapter.h:41-64): batch next_pow2(N), first mod 60, scaling 59, depth 40,
// rotations +-2^i
inline void adapterParameters(int n, CCParams<CryptoContextCKKSRNS>& parameters, std::vector<int>& rotations) {
    int batch = 1, lg = 0;
    while (batch < n) { batch <<= 1; lg++; }
    parameters.SetBatchSize((uint32_t)batch);
    parameters.SetFirstModSize(60);
    parameters.SetScalingModSize(59);
    parameters.SetMultiplicativeDepth(40);
    rotations.clear();
    for (int i = 0; i < lg; i++) { rotations.push_back(1 << i); rotations.push_back(-(1 << i)); }
}

}  // namespace kwaySort
