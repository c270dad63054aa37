// kway.h -- the k-way sorting network of HKC+21 with the reference's names (namespace kwaySort): Masking
// (src/k-way/Masking.cpp:26-167), EvalUtils (src/k-way/EvalUtils.cpp:18-147), SortUtils (src/k-way/SortUtils.cpp:5-82,
// 202-236), Sorter (src/k-way/Sorter.cpp:8-70,180-262,284-404) and KWayAdapter's parameter card
// (src/kway_adapter.h:41-64).  k = 2 and k = 3 are implemented (BASELINE config 4 is k = 2); the 4-/5-sorters of
// k = 5 are not.  checkLevelAndBoot calls cc->EvalBootstrap exactly where the reference does -- the B200 engine has
// no bootstrapping yet, so a network runs as far as the multiplicative depth allows and the call
// throws beyond that.  Orchestration only: every cc-> call is a CUDA kernel schedule inside libckks_b200.so.
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

    int stageCount() const { return (int)(m_M + m_M * (m_M - 1) / 2 * ((m_k + 1) / 2)); }   // :298

    // one stage of the network, in place on ctxt (body of the loop at :300-397)
    void runStage(CT& ctxt, int stage, SignConfig& Cfg) {
        const int multDepth = Cfg.multDepth;
        long m, logDist, slope;
        { auto t = sortType((int)m_k, (int)m_M, stage); m = std::get<0>(t); logDist = std::get<1>(t); slope = std::get<2>(t); }
        const long shift = getRotateDistance(m_k, logDist, slope);
        auto indices = genIndices(m_numSlots, m_k, m_M, m, logDist, slope);
        CT ctxt_fix, ctxt_comp1;
        if (m_k != 2 && m_k != 3) throw std::runtime_error("kwaySort::Sorter: only k = 2 and k = 3 are implemented");
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
        } else if (m_k == 3 && slope == 1) {        // diagonals of 2 and of 3 in the same stage
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
