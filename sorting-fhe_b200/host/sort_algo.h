// sort_algo.h -- SortBase<N> and DirectSort<N> (Optimized Rank Sort) with the reference's interface
// (reference src/sort_algo.h:36-59, 61-774).  Orchestration only: every cc-> call below is a CUDA kernel
// schedule inside libckks_b200.so.  The reference's OpenMP loops (sort_algo.h:334,349,425,713) are plain loops
// here: all work is queued on the engine's CUDA stream, where the limb dimension provides the parallelism.
#pragma once
#include <algorithm>
#include <cmath>
#include <set>
#include <vector>
#include "cheb_coeffs.h"
#include "comparison.h"
#include <map>
#include <mutex>
#include "mehp24.h"
#include "rotation.h"

enum class SortAlgo { DirectSort, BitonicSort };

template <int N> class SortBase {
  protected:
    std::shared_ptr<Encryption> m_enc;
    const Ciphertext<DCRTPoly> m_zeroCache;   // encryption of N zeros, cloned as accumulator seed (:42-45)
    virtual Ciphertext<DCRTPoly> createZeroCache() { return m_enc->encryptInput(std::vector<double>(N, 0.0)); }
  public:
    SortBase(std::shared_ptr<Encryption> enc) : m_enc(enc), m_zeroCache(createZeroCache()) {}
    virtual ~SortBase() = default;
    virtual Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly>& input_array, SignFunc SignFunc, SignConfig& Cfg) = 0;
    virtual const Ciphertext<DCRTPoly>& getZero() const { return m_zeroCache; }
    // `getZero()->Clone()` with SetSlots(slots) applied (reference src/sort_algo.h:330-331,433-434): one handle per slot count, made
    // once.  Every evaluator call is out of place, so the seed itself is never written; handing out the same handle lets the engine
    // reuse the level adjustments it caches on it (evaluator.cu adjusted_twin) instead of copying and re-adjusting 2 x 32 limbs for
    // each of the 16 baby-step blocks of every comparison batch.
    Ciphertext<DCRTPoly> zeroSeed(uint32_t slots) const {
        std::lock_guard<std::mutex> g(m_seedMu);
        auto it = m_zeroSeeds.find(slots);
        if (it == m_zeroSeeds.end()) {
            Ciphertext<DCRTPoly> z = getZero()->Clone();
            z->SetSlots(slots);
            it = m_zeroSeeds.emplace(slots, z).first;
        }
        return it->second;
    }
  private:
    mutable std::mutex m_seedMu;
    mutable std::map<uint32_t, Ciphertext<DCRTPoly>> m_zeroSeeds;
  public:
    constexpr size_t getArraySize() const { return N; }
};

namespace directsort_detail {
inline int log2i(int v) { int l = 0; while ((1 << (l + 1)) <= v) l++; return l; }
// np of constructRank (:383-416) and of rotationIndexCheckN (:670-703)
inline int rankBabySteps(int n) { return n <= 8 ? 2 : n <= 32 ? 4 : n <= 128 ? 8 : n <= 512 ? 16 : 32; }
inline int indexBabySteps(int n) { return n <= 8 ? 2 : n <= 32 ? 4 : n <= 128 ? 8 : n == 256 ? 16 : n <= 1024 ? 8 : 4; }
}  // namespace directsort_detail

template <int N> class DirectSort : public SortBase<N> {
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
    Comparison comp;
    RotationComposer<N> rot;
    int max_batch;   // ringDim / 2
    typedef Ciphertext<DCRTPoly> CT;

  public:
    std::shared_ptr<Encryption> m_enc;

    DirectSort(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> publicKey, std::vector<int> rotIndices,
               std::shared_ptr<Encryption> enc)
        : SortBase<N>(enc), m_cc(cc), m_PublicKey(publicKey), comp(enc), rot(cc, enc, rotIndices), m_enc(enc) {
        max_batch = (int)m_cc->GetRingDimension() / 2;
    }
    const std::set<int>& getRotationCalls() const { return rot.getRotationCalls(); }

    // every rotation amount sort() asks RotationComposer for at a given ring dimension; equals the reference's
    // literal key tables (:94-198) for the ring 2^17 fixtures
    static std::vector<int> requiredRotations(int ringDim) {
        using namespace directsort_detail;
        const int P = std::min(N, ringDim / 2 / N), B = N / P, S = N * P;
        std::set<int> r;
        const int np = std::min(rankBabySteps(N), P), npi = indexBabySteps(N);
        for (int i = 1; i < np; i++) r.insert(i);
        for (int b = 0; b < B; b++) for (int j = 0; j < P / np; j++) r.insert(b * P + j * np);
        for (int i = 1; i <= log2i(P); i++) r.insert(S >> i);
        if (S < ringDim / 2) r.insert(S);   // the reference also keys the full-wrap amount (:95,99,104,...)
        for (int i = 1; i < npi; i++) r.insert(i);
        for (int b = 0; b < B; b++) for (int i = 0; i < (S / N) / npi; i++) r.insert(b * P + i * npi);
        r.erase(0);
        return std::vector<int>(r.begin(), r.end());
    }

    // DirectSort<N>::getSizeParameters (:87-201): batch N, 40-bit scaling, depth table, rotation key set
    static void getSizeParameters(CCParams<CryptoContextCKKSRNS>& parameters, std::vector<int>& rotations) {
        static const int depth[][2] = {{4, 23}, {8, 24}, {16, 25}, {32, 28}, {64, 29}, {128, 30},
                                       {256, 34}, {512, 35}, {1024, 39}, {2048, 52}};
        parameters.SetBatchSize(N);
        parameters.SetScalingModSize(40);
        int d = 0;
        for (auto& e : depth) if (e[0] == N) d = e[1];
        if (!d) throw std::runtime_error("DirectSort: unsupported array size");
        parameters.SetMultiplicativeDepth(d);
        const int ringDim = parameters.GetRingDim() ? (int)parameters.GetRingDim() : (1 << 17);
        rotations = (ringDim == (1 << 17) && N >= 256) ? referenceKeySet() : requiredRotations(ringDim);
    }

    // parameters of the hybrid sorters: depth tables of tests/DirectSortHTest.cpp:28-92, DirectSortH1Test.cpp:37-101,
    // DirectSortH2Test.cpp:37-101; key set = the rank phase's amounts plus the fold steps of
    // sumColumnsToTarget / transposeColumnTarget for every block (the tests' literal tables, as their pattern)
    static void getHybridParameters(int variant, CCParams<CryptoContextCKKSRNS>& parameters, std::vector<int>& rotations) {
        static const int sizes[9] = {4, 8, 16, 32, 64, 128, 256, 512, 1024};
        static const int depth[3][9] = {{24, 25, 25, 29, 30, 31, 44, 47, 50},     // sort_hybrid
                                        {31, 35, 35, 42, 42, 46, 49, 53, 56},     // sort_hybrid1
                                        {24, 25, 25, 29, 30, 31, 35, 36, 40}};    // sort_hybrid2
        if (variant < 0 || variant > 2) throw std::runtime_error("hybrid sort: variant must be 0, 1 or 2");
        int d = 0;
        for (int i = 0; i < 9; i++) if (sizes[i] == N) d = depth[variant][i];
        if (!d) throw std::runtime_error("hybrid sort: unsupported array size");
        parameters.SetBatchSize(N);
        parameters.SetScalingModSize(40);
        parameters.SetMultiplicativeDepth(d);
        using namespace directsort_detail;
        const int ringDim = parameters.GetRingDim() ? (int)parameters.GetRingDim() : (1 << 17);
        const int P = std::min(N, ringDim / 2 / N), B = N / P, S = N * P;
        std::set<int> r;
        const int np = std::min(rankBabySteps(N), P);
        for (int i = 1; i < np; i++) r.insert(i);
        for (int b = 0; b < B; b++) for (int j = 0; j < P / np; j++) r.insert(b * P + j * np);
        for (int i = 1; i <= log2i(P); i++) r.insert(S >> i);
        const int blocks = N > 256 ? N / 256 : 1, side = N / blocks, lg = log2i(side);
        for (int b = 0; b < blocks; b++) {
            r.insert(b * 256);
            for (int i = 0; i < lg; i++) {
                const int sgn = ((b >> (lg - 1 - i)) & 1) ? -1 : 1;
                r.insert(sgn * ((side >> 1) >> i));
                r.insert(sgn * ((side * (side - 1) / 2) >> i));
            }
        }
        r.erase(0);
        rotations.assign(r.begin(), r.end());
    }

    // The reference's literal key sets for N >= 256 at ring 2^17 (:125-148) are not the set of rotation amounts
    // sort() asks for: some amounts are left to RotationComposer's decomposition, some keys are never used.  They
    // are reproduced here (as their generating pattern) so that a caller gets the same keys, hence the same op
    // sequence, as with the reference.
    static std::vector<int> referenceKeySet() {
        std::set<int> r;
        auto range = [&](int lo, int hi, int step) { for (int v = lo; v <= hi; v += step) r.insert(v); };
        if (N == 256) { range(1, 16, 1); range(24, 128, 8); range(129, 135, 1); range(144, 256, 16); }
        else if (N == 512) { range(1, 16, 1); range(24, 64, 8); range(80, 512, 16); for (int b = 1; b < 8; b++) range(64 * b + 1, 64 * b + 7, 1); }
        else if (N == 1024) { range(1, 35, 1); for (int b = 2; b < 32; b++) range(32 * b, 32 * b + 3, 1); r.insert(1024); }
        for (int v = 512; v <= 32768; v *= 2) r.insert(v);
        return std::vector<int>(r.begin(), r.end());
    }

    // ---- plaintext pattern generators (:206-306)
    std::vector<double> generateMaskVector(int num_slots, int k) {    // block k of length N set to 1
        std::vector<double> v(num_slots, 0.0);
        std::fill(v.begin() + (size_t)k * N, v.begin() + (size_t)(k + 1) * N, 1.0);
        return v;
    }
    std::vector<double> generateMaskVectorN(int num_slots, int k) { return generateMaskVector(num_slots, k); }
    std::vector<double> generateIndexVector() {
        std::vector<double> v(N);
        for (int i = 0; i < N; i++) v[i] = (double)i;
        return v;
    }
    std::vector<double> generateCheckingVectorN(int num_slots, int k) {   // block j holds (k + j) mod N
        std::vector<double> v(num_slots);
        for (int i = 0; i < num_slots; i++) v[i] = (double)((k + i / N) % N);
        return v;
    }
    std::vector<double> vectorRotate(const std::vector<double>& vec, int rotateIndex) {   // >0 left, <0 right
        std::vector<double> r = vec;
        if (r.empty()) return r;
        int n = (int)r.size(), k = ((rotateIndex % n) + n) % n;
        std::rotate(r.begin(), r.begin() + k, r.end());
        return r;
    }

    // vecRotsOpt (:326-366): sum_j Rot_{is*P + j*np}( sum_i mask_{np*j+i} (shifted right) * babyStep_i )
    CT vecRotsOpt(const std::vector<CT>& preRotatedArrays, int num_partition, int num_slots, int np, int is) {
        const int nb = num_partition / np;
        std::vector<CT> outer(nb);
        CT seed = this->zeroSeed((uint32_t)num_slots);      // `getZero()->Clone()` + SetSlots of the reference (:330-331)
        std::vector<std::vector<Plaintext>> pmsk(nb, std::vector<Plaintext>(np));
        for (int j = 0; j < nb; j++)
            for (int i = 0; i < np; i++) {
                // the masks depend on (N, slots, block, shift) only: encoded once per context, same handle every sort
                const int k = np * j + i, shift = -is * num_partition - j * np;
                pmsk[j][i] = m_cc->MakeCKKSPackedPlaintextKeyed({1, N, num_slots, k, shift},
                    [&] { return vectorRotate(generateMaskVector(num_slots, k), shift); }, 1, 0, (uint32_t)num_slots);
            }
        // T_j = EvalAdd(T_j, EvalMult(preRotatedArrays[i], pmsk[j][i])) for every i (:343-344), all blocks j in one engine call
        std::vector<CT> T = m_cc->EvalInnerProductPlainMulti(seed, std::vector<CT>(preRotatedArrays.begin(), preRotatedArrays.begin() + np), pmsk);
        for (int j = 0; j < nb; j++) outer[j] = rot.rotate(T[j], is * num_partition + j * np);
        return m_cc->EvalAddMany(outer);
    }

    // constructRank (:368-506): rank_i = #{j : a_j < a_i}
    CT constructRank(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg) {
        return finishRank(constructRankPartial(input_array, SignFunc, Cfg, 0, numBatches(), true));
    }
    int numBatches() const { return N / std::min(N, max_batch / N); }
    // The comparison batches i in [b_lo, b_hi) of constructRank's loop (:438-492), summed.  The batches are independent
    // (the reference runs them under OpenMP for N >= 512): this is the unit that shards across GPUs, the partial sums
    // being combined by an all-reduce of the ciphertext limbs.  `seed` adds the cached encryption of zeros the
    // reference starts its accumulator from (:433) -- exactly one shard must pass true.
    CT constructRankPartial(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg, int b_lo, int b_hi, bool seed) {
        const int num_partition = std::min(N, max_batch / N);
        const int num_slots = N * num_partition;
        const int np = std::min(directsort_detail::rankBabySteps(N), num_partition);
        if (b_lo < 0 || b_hi > N / num_partition || b_lo >= b_hi) throw std::runtime_error("constructRankPartial: empty or invalid batch range");
        std::vector<CT> babySteps(np);
        for (int i = 0; i < np; i++) { babySteps[i] = rot.rotate(input_array, i); babySteps[i]->SetSlots(num_slots); }
        CT rank;
        if (seed) rank = this->zeroSeed((uint32_t)num_slots);
        for (int i = b_lo; i < b_hi; i++) {
            CT shifted = vecRotsOpt(babySteps, num_partition, num_slots, np, i);
            CT dup = input_array->Clone();
            dup->SetSlots(num_slots);
            CT c = comp.compare(m_cc, dup, shifted, SignFunc, Cfg);
            if (rank) m_cc->EvalAddInPlace(rank, c); else rank = c;
        }
        return rank;
    }
    // rotate-and-sum over the partitions, then remove the self comparison's 1/2 (:494-504)
    CT finishRank(CT rank) {
        const int num_partition = std::min(N, max_batch / N), num_slots = N * num_partition;
        for (int i = 1; i < std::log2(num_partition) + 1; i++)
            m_cc->EvalAddInPlace(rank, rot.rotate(rank, num_slots / (1 << i)));
        rank->SetSlots(N);
        return m_cc->EvalSub(rank, 0.5);
    }

    // blindRotationOptN (:561-584)
    CT blindRotationOptN(const std::vector<CT>& masked_inputs, int num_slots, int np, int ib, int num_partition) {
        CT result = this->getZero()->Clone();
        const int nb = (num_slots / N) / np;
        std::vector<std::vector<Plaintext>> pmsk(nb, std::vector<Plaintext>(np));
        for (int i = 0; i < nb; i++)
            for (int j = 0; j < np; j++) {
                const int k = np * i + j;
                pmsk[i][j] = m_cc->MakeCKKSPackedPlaintextKeyed({2, N, num_slots, k, j},
                    [&] { return vectorRotate(generateMaskVectorN(num_slots, k), j); }, 1, (uint32_t)masked_inputs[j]->GetLevel(), (uint32_t)num_slots);
            }
        // tmp_i = zero + sum_j masked_inputs[j] * pmsk[i][j] (:575-576) for every block i in one engine call
        std::vector<CT> tmp = m_cc->EvalInnerProductPlainMulti(this->getZero(), std::vector<CT>(masked_inputs.begin(), masked_inputs.begin() + np), pmsk);
        for (int i = 0; i < nb; i++)
            m_cc->EvalAddInPlace(result, rot.rotate(tmp[i], ib * num_partition + i * np));
        return result;
    }

    // rotationIndexCheckN (:658-750): out_j = sum_i a_i [rank_i = j] through the doubled-sinc indicator
    CT rotationIndexCheckN(const CT& ctx_Rank, const CT& input_array) {
        return finishIndexCheck(rotationIndexCheckPartial(ctx_Rank, input_array, 0, numBatches(), true));
    }
    // batches b in [b_lo, b_hi) of the index-check loop (:713-742), summed; same sharding contract as constructRankPartial
    CT rotationIndexCheckPartial(const CT& ctx_Rank, const CT& input_array, int b_lo, int b_hi, bool seed) {
        const int num_partition = std::min(N, max_batch / N);
        const int num_slots = num_partition * N, np = directsort_detail::indexBabySteps(N);
        if (b_lo < 0 || b_hi > N / num_partition || b_lo >= b_hi) throw std::runtime_error("rotationIndexCheckPartial: empty or invalid batch range");
        CT output;
        if (seed) output = this->getZero()->Clone();
        Plaintext index_vector = m_cc->MakeCKKSPackedPlaintextKeyed({3, N}, [&] { return generateIndexVector(); }, 1, (uint32_t)ctx_Rank->GetLevel(), N);
        CT index_minus_rank = m_cc->EvalSub(index_vector, ctx_Rank);
        index_minus_rank->SetSlots(num_slots);
        input_array->SetSlots(num_slots);   // mutates the caller's ciphertext metadata, as the reference does (:711)
        for (int b = b_lo; b < b_hi; b++) {
            Plaintext chk = m_cc->MakeCKKSPackedPlaintextKeyed({4, N, num_slots, b * num_partition},
                [&] { return generateCheckingVectorN(num_slots, b * num_partition); }, 1, (uint32_t)index_minus_rank->GetLevel(), (uint32_t)num_slots);
            CT rotIndex = m_cc->EvalSub(index_minus_rank, chk);
            m_cc->EvalMultInPlace(rotIndex, 1.0 / N / 2);   // (-2N, N) -> (-1, 1/2)
            rotIndex = m_cc->EvalChebyshevSeriesPS(rotIndex, selectDoubledSincCoefficients<N>(), -1, 1);
            CT masked_input = m_cc->EvalMult(rotIndex, input_array);
            std::vector<CT> masked_inputs(np);
            for (int i = 0; i < np; i++) masked_inputs[i] = rot.rotate(masked_input, i);
            CT r = blindRotationOptN(masked_inputs, num_slots, np, b, num_partition);
            if (output) m_cc->EvalAddInPlace(output, r); else output = r;
        }
        return output;
    }
    CT finishIndexCheck(CT output) {   // (:744-748)
        const int num_partition = std::min(N, max_batch / N), num_slots = num_partition * N;
        for (int i = 1; i < std::log2(num_partition) + 1; i++)
            m_cc->EvalAddInPlace(output, rot.rotate(output, num_slots / (1 << i)));
        output->SetSlots(N);
        return output;
    }

    // ---- hybrid index checks (:776-1391): the N x N "rank_j == i" matrix of MEHP24 in one ciphertext of N*N slots
    // (256 x 256 blocks above N = 256), folded along columns and transposed back to a row -- no baby/giant rotations.
    // The three published variants differ only in the indicator they put on (i - rank_j):
    enum class HybridIndicator {
        SincOrCompare,   // sort_hybrid  (:893-1049): sinc below N = 256, composite-sign indicator from there on
        Mehp24,          // sort_hybrid1 (:1067-1211): mehp24::utils::indicatorAdv on the unscaled difference
        Sinc             // sort_hybrid2 (:1232-1372): sinc for every N
    };

    // fold by log2(matrixSize) rotations towards column `columnIndex` (:825-858) / row `rowIndex` (:860-890); the sign of
    // each step follows the bits of the target, most significant first (getBinaryPath, :815-821)
    CT foldToTarget(CT c, size_t matrixSize, size_t target, size_t firstStep, bool rowMask) {
        const size_t lg = (size_t)std::log2((double)matrixSize);
        c->SetSlots((uint32_t)(matrixSize * matrixSize));
        size_t step = firstStep;
        for (size_t i = 0; i < lg; i++, step >>= 1) {
            const bool right = (target >> (lg - 1 - i)) & 1;
            c = m_cc->EvalAdd(c, rot.rotate(c, right ? -(int)step : (int)step));
        }
        Plaintext pmsk = m_cc->MakeCKKSPackedPlaintextKeyed({5, (int64_t)matrixSize, (int64_t)target, rowMask ? 1 : 0}, [&] {
            std::vector<double> msk(matrixSize * matrixSize, 0.0);
            for (size_t i = 0; i < matrixSize; i++) msk[rowMask ? matrixSize * target + i : matrixSize * i + target] = 1.0;
            return msk; }, 1, (uint32_t)c->GetLevel(), (uint32_t)(matrixSize * matrixSize));
        return m_cc->EvalMult(c, pmsk);
    }
    CT sumColumnsToTarget(CT c, size_t matrixSize, size_t columnIndex) { return foldToTarget(c, matrixSize, columnIndex, matrixSize >> 1, false); }
    CT transposeColumnTarget(CT c, size_t matrixSize, size_t rowIndex) { return foldToTarget(c, matrixSize, rowIndex, matrixSize * (matrixSize - 1) / 2, true); }

    CT rotationIndexCheckHybrid(const CT& ctx_Rank, const CT& input_array, HybridIndicator kind) {
        const size_t maxArraySize = 256;
        const size_t num_slots = (size_t)N > maxArraySize ? (size_t)max_batch : (size_t)N * N;
        const size_t num_batch = (size_t)N > maxArraySize ? N / maxArraySize : 1;
        const size_t array_size = std::min((size_t)N, maxArraySize);
        const bool scaled = kind != HybridIndicator::Mehp24;          // ranks and row indices divided by N (:912, 1250)
        ctx_Rank->SetSlots((uint32_t)num_slots);
        CT r = scaled ? m_cc->EvalMult(ctx_Rank, 1.0 / N) : ctx_Rank;
        input_array->SetSlots((uint32_t)num_slots);

        std::vector<CT> rots_Rank(num_batch), rots_Input(num_batch);
        for (size_t b = 0; b < num_batch; b++) rots_Rank[b] = rot.rotate(r, (int)(b * maxArraySize));
        for (size_t b = 0; b < num_batch; b++) rots_Input[b] = rot.rotate(input_array, (int)(b * maxArraySize));
        const uint32_t dg_i = (uint32_t)((std::log2((double)N) + 1) / 2), df_i = 2;   // (:1126-1127)

        std::vector<CT> Masked(num_batch);
        for (size_t b = 0; b < num_batch; b++) {
            Plaintext subMaskPtx = m_cc->MakeCKKSPackedPlaintextKeyed({6, N, (int64_t)b, scaled ? 1 : 0, (int64_t)array_size, (int64_t)num_slots}, [&] {
                std::vector<double> subMask(num_slots, 0.0);
                for (size_t i = 0; i < array_size; i++)
                    for (size_t j = 0; j < array_size; j++) {
                        const double row = (double)(b * array_size + i);
                        subMask[i * array_size + j] = scaled ? row / (double)N : row;
                    }
                return subMask; }, 1, (uint32_t)ctx_Rank->GetLevel(), (uint32_t)num_slots);
            CT subMasked = this->zeroSeed((uint32_t)num_slots);
            for (size_t k = 0; k < num_batch; k++) {
                CT rotationMask = m_cc->EvalSub(subMaskPtx, rots_Rank[k]);
                if (kind == HybridIndicator::Mehp24)
                    rotationMask = mehp24::utils::indicatorAdv(m_cc, rotationMask, (double)N, dg_i, df_i);
                else if (kind == HybridIndicator::Sinc || N < 256)
                    rotationMask = m_cc->EvalChebyshevSeriesPS(rotationMask, selectCoefficients<N>(), -1, 1);
                else {
                    SignConfig Cfg = SignConfig(CompositeSignConfig(3, N < 512 ? 4 : 5, 2));
                    rotationMask = comp.indicator(m_cc, rotationMask, 0.5 / N, SignFunc::CompositeSign, Cfg);
                }
                subMasked = m_cc->EvalAdd(subMasked, m_cc->EvalMult(rots_Input[k], rotationMask));
            }
            subMasked = sumColumnsToTarget(subMasked, N / num_batch, b);
            Masked[b] = transposeColumnTarget(subMasked, N / num_batch, b);
        }
        return m_cc->EvalAddMany(Masked);
    }
    CT sort_hybrid(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg, HybridIndicator kind = HybridIndicator::SincOrCompare) {
        CT rank = constructRank(input_array, SignFunc, Cfg);
        return rotationIndexCheckHybrid(rank, input_array, kind);
    }
    CT sort_hybrid1(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg) { return sort_hybrid(input_array, SignFunc, Cfg, HybridIndicator::Mehp24); }
    CT sort_hybrid2(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg) { return sort_hybrid(input_array, SignFunc, Cfg, HybridIndicator::Sinc); }

    CT sort(const CT& input_array, SignFunc SignFunc, SignConfig& Cfg) override {   // (:752-774)
        CT rank = constructRank(input_array, SignFunc, Cfg);
        return rotationIndexCheckN(rank, input_array);
    }
};
