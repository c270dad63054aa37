// rotation.h -- rotation by arbitrary amounts from a fixed key set (reference src/rotation.h:12-238):
// Decomposer<N> splits a rotation into available steps (binary / NAF / balanced-NAF), RotationComposer<N>
// applies it with one EvalRotate per step (direct key when present, rotation.h:215-233).
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <vector>
#include "encryption.h"

struct Step {
    int8_t value;   // -1 / 0 / +1 (NAF digits); +1 for binary
    int stepSize;   // rotation amount of this step
    Step(int8_t v, int s) : value(v), stepSize(s) {}
    Step(int s) : value(1), stepSize(s) {}
};
enum class DecomposeAlgo { NAF, BNAF, BINARY };

template <int N> class Decomposer {
    std::vector<int> rotIndices;
    int maxDecomposed;
    // sum of the key indices that continue a doubling chain (rotation.h:36-45)
    int calculateMax() const {
        int total = 0, prev = 1;
        for (int idx : rotIndices) { if (prev == idx / 2) total += idx; prev = idx; }
        return total;
    }
    std::vector<Step> decomposeBinary(int r) const {
        std::vector<Step> out;
        for (int i = 31; i >= 0; --i) { int s = 1 << i; if (s < N && (r & s)) out.emplace_back(1, s); }
        return out;
    }
    // non-adjacent form, most significant step first; a step of -N/2 is replaced by the equivalent +N/2
    // (rotation.h:116-136)
    std::vector<Step> decomposeNAF(int r) const {
        std::vector<Step> out;
        for (int i = 0; r != 0; ++i, r >>= 1) {
            if (!(r & 1)) continue;
            int z = (r & 2) ? -1 : 1, s = z * (1 << i);
            if (s == -N / 2) out.emplace_back((int8_t)-z, -s); else out.emplace_back((int8_t)z, s);
            r -= z;
        }
        std::reverse(out.begin(), out.end());
        return out;
    }
    // balanced binary digits: a 1 followed by another 1 becomes -1 with a carry (rotation.h:138-165)
    std::vector<Step> decomposeBNAF(int k) const {
        std::vector<Step> out;
        for (int i = 0; k != 0; ++i) {
            int d = k % 2;
            k = (k - d) / 2;
            if (d == 1 && (k % 2) >= 1) { d = -1; k += 1; }
            if (d != 0) out.emplace_back((int8_t)d, d * (1 << i));
        }
        std::reverse(out.begin(), out.end());
        return out;
    }
  public:
    Decomposer(std::vector<int> rot) : rotIndices(rot) {
        std::sort(rotIndices.begin(), rotIndices.end());
        maxDecomposed = calculateMax();
    }
    std::vector<int> getRotIndices() { return rotIndices; }
    std::vector<Step> decompose(int rotation, int wrapN, DecomposeAlgo algo) {
        std::vector<Step> steps;
        const int largest = rotIndices.back();
        while (rotation >= largest) { steps.emplace_back(largest); rotation -= largest; }
        while (rotation > maxDecomposed) {   // peel the largest key below the remainder
            int legal = *(std::lower_bound(rotIndices.begin(), rotIndices.end(), rotation) - 1);
            steps.emplace_back(legal); rotation -= legal;
        }
        if (rotation) {
            std::vector<Step> rest = algo == DecomposeAlgo::NAF ? decomposeNAF(rotation)
                                   : algo == DecomposeAlgo::BNAF ? decomposeBNAF(rotation) : decomposeBinary(rotation);
            steps.insert(steps.end(), rest.begin(), rest.end());
        }
        steps.erase(std::remove_if(steps.begin(), steps.end(), [wrapN](const Step& s) { return s.stepSize % wrapN == 0; }),
                    steps.end());
        return steps;
    }
};

template <int N> class RotationComposer {
    CryptoContext<DCRTPoly> m_cc;
    std::shared_ptr<Encryption> m_enc;
    Decomposer<N> m_decomposer;
    DecomposeAlgo m_algo;
    std::set<int> rotation_calls, available_indices;
  public:
    RotationComposer(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, std::vector<int> rotIndices,
                     DecomposeAlgo algo = DecomposeAlgo::BINARY)
        : m_cc(cc), m_enc(enc), m_decomposer(rotIndices), m_algo(algo),
          available_indices(rotIndices.begin(), rotIndices.end()) {}
    Ciphertext<DCRTPoly> rotate(const Ciphertext<DCRTPoly>& input, int rotation) {
        if (rotation % (int)input->GetSlots() == 0) return input->Clone();
        if (available_indices.count(rotation)) return m_cc->EvalRotate(input, rotation);
        Ciphertext<DCRTPoly> result = input->Clone();
        for (const Step& s : m_decomposer.decompose(rotation, (int)input->GetSlots(), m_algo))
            result = m_cc->EvalRotate(result, s.stepSize);
        return result;
    }
    const std::set<int>& getRotationCalls() const { return rotation_calls; }
    void clearRotationCalls() { rotation_calls.clear(); }
};

// RotationTree<N> (reference src/rotation.h:239-358): a trie over the step sequences of the rotations start..end.
// Rotating through it memoises every intermediate ciphertext on its node, and the children of one node are produced
// from ONE set of hoisted ModUp digits (EvalFastRotationPrecompute / EvalFastRotation) -- the reference's only use
// of hoisting.  As in the reference, the cache belongs to the first input the tree is used with.
template <int N> class RotationTree {
  public:
    struct RotationNode {
        int stepSize;
        RotationNode* parent;
        std::map<int, std::unique_ptr<RotationNode>> children;
        std::vector<int> finalValues;                 // rotations that end at this node
        Ciphertext<DCRTPoly> rotatedCiphertext;       // memoised result of the path root -> this node
        Ciphertext<DCRTPoly> cPrecomp;                // hoisted digits of the ciphertext at this node
        RotationNode(int step, RotationNode* p = nullptr) : stepSize(step), parent(p) {}
    };

  private:
    std::unique_ptr<RotationNode> root;
    Decomposer<N> m_decomposer;
    DecomposeAlgo m_algo;
    CryptoContext<DCRTPoly> m_cc;
    uint32_t M;

    RotationNode* descend(RotationNode* node, int step, bool create) {
        auto it = node->children.find(step);
        if (it != node->children.end()) return it->second.get();
        if (!create) return nullptr;
        auto& slot = node->children[step];
        slot.reset(new RotationNode(step, node));
        return slot.get();
    }

  public:
    RotationTree(CryptoContext<DCRTPoly> cc, const std::vector<int>& rotIndices, DecomposeAlgo algo = DecomposeAlgo::NAF)
        : root(new RotationNode(0)), m_decomposer(rotIndices), m_algo(algo), m_cc(cc), M(cc->GetCyclotomicOrder()) {}

    void buildTree(int start, int end) {              // (:271-279)
        for (int i = start; i <= end; ++i) {
            RotationNode* node = root.get();
            for (const Step& s : m_decomposer.decompose(i, end, m_algo))
                if (s.value != 0) node = descend(node, s.stepSize, true);
            node->finalValues.push_back(i);
        }
    }

    Ciphertext<DCRTPoly> treeRotate(const Ciphertext<DCRTPoly>& input, int rotation) {   // (:281-291, 315-357)
        if (!root->cPrecomp) root->cPrecomp = m_cc->EvalFastRotationPrecompute(input);
        Ciphertext<DCRTPoly> cur = input;
        RotationNode* node = root.get();
        for (const Step& s : m_decomposer.decompose(rotation, (int)input->GetSlots(), m_algo)) {
            if (s.value == 0) continue;
            RotationNode* child = descend(node, s.stepSize, false);
            if (!child) return cur;                    // not in the tree: the reference reports and returns what it has
            if (!child->rotatedCiphertext) {
                if (std::abs(node->stepSize) > 2 || node->stepSize == 0) {   // hoisted below the root and below wide steps
                    if (!node->cPrecomp) node->cPrecomp = m_cc->EvalFastRotationPrecompute(cur);
                    child->rotatedCiphertext = m_cc->EvalFastRotation(cur, child->stepSize, M, node->cPrecomp);
                } else {
                    child->rotatedCiphertext = m_cc->EvalRotate(cur, child->stepSize);
                }
            }
            cur = child->rotatedCiphertext;
            node = child;
        }
        return cur;
    }
    const RotationNode* getRoot() const { return root.get(); }
};
