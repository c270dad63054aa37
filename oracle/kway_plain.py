"""oracle/kway_plain.py -- TEST INFRASTRUCTURE ONLY (plaintext model of the k-way sorting network blocks).

Plain-double restatement of the HKC+21 k-way building blocks the reference implements over ciphertexts:
  * fcnL / twoSorter / threeSorter         -- /root/reference/src/k-way/SortUtils.cpp:5-82
  * sortType / genIndices / genMask / getRotateDistance -- /root/reference/src/k-way/Masking.cpp:26-167
Used by tests/ to check the reference's literal test vectors (tests/k-way/SortUtilsTest.cpp, MaskingTest.cpp) and as
the expected-value model for the GPU k-way stage.  The product path never imports it.
"""
import numpy as np


def fcn_l(a, b, comp):
    """(a > b) * a + (a < b) * b = comp * (a - b) + b   (SortUtils.cpp:5-16)"""
    return comp * (a - b) + b


def two_sorter(a, b, comp):
    """[min, max] from comp = (a > b)   (SortUtils.cpp:32-57)"""
    hi = fcn_l(a, b, comp)
    return (a + b) - hi, hi


def three_sorter(x, comp):
    """x = [a, b, c], comp = [(a>b), (a>c), (b>c)] -> [min, mid, max]   (SortUtils.cpp:59-82)"""
    lo_ab, hi_ab = two_sorter(x[0], x[1], comp[0])
    # the comparison bits of min(a,b) / max(a,b) against c are selected by the same (a > b) bit
    lo_vs_c, hi_vs_c = two_sorter(comp[1], comp[2], comp[0])
    mx = fcn_l(hi_ab, x[2], hi_vs_c)          # compareMax
    mn = fcn_l(x[2], lo_ab, lo_vs_c)          # compareMin = fcnL with swapped operands
    return mn, (x[0] + x[1] + x[2]) - mn - mx, mx


def sort_type(k, M, stage):
    """stage -> (m, logDist, slope)   (Masking.cpp:26-47)"""
    up = (k + 1) // 2
    r = 0
    while stage >= (r + 1 + r * (r + 1) // 2 * up):
        r += 1
    n = stage - (r + r * (r - 1) // 2 * up)
    m = (n + up - 1) // up
    slope = 0 if n == 0 else ((n - 1) % up) + 1
    return m, r - m, slope


def get_rotate_distance(k, log_dist, slope):
    """Masking.cpp:157-166"""
    dist = k ** log_dist
    if slope == 0 or slope == k // 2 + 1:
        return dist
    return dist * (k - slope)


def gen_indices(num_slots, k, M, m, log_dist, slope):
    """role table [2][numSlots]: row 0 = size of the group a slot is sorted in, row 1 = its position in the group
    (Masking.cpp:49-143)"""
    res = np.zeros((2, num_slots), dtype=np.int64)
    km, dist, nxt = k ** m, k ** log_dist, k ** (m + 1)

    def diagonal(start, row, col):
        loc = 1
        while row < km and col >= 0:
            for d in range(dist):
                here = start + dist * (col + k * row) + d
                res[0][here] = loc
                if row == km - 1 or col - slope < 0:
                    for i in range(loc):
                        h2 = start + dist * ((col + i * slope) + k * (row - i)) + d
                        res[1][h2] = loc - i
                        res[0][h2] += i
            loc, row, col = loc + 1, row + 1, col - slope

    for start in range(0, k ** M, dist * nxt):
        if slope == 0:
            for s in range(km):
                for col in range(k):
                    for d in range(dist):
                        here = start + dist * (s + km * col) + d
                        res[0][here], res[1][here] = k, col + 1
        elif slope > k // 2:
            for t in range(km - 1):
                col = k - k // 2
                for loc in range(1, k):
                    for d in range(dist):
                        here = start + dist * (col + k * t + loc - 1) + d
                        res[0][here], res[1][here] = k - 1, loc
        else:
            for t in range(slope, k):
                diagonal(start, 0, t)
            for s in range(1, km - 1):
                for t in range(k - slope, k):
                    diagonal(start, s, t)
    return res


def gen_mask(indices, index0, index1):
    """Masking.cpp:145-155"""
    return ((indices[0] == index0) & (indices[1] == index1)).astype(np.float64)


def network_plain(x, k, M, compare=None, first_stage=0, last_stage=None):
    """Sorter::sorter (Sorter.cpp:284-404) on plain doubles for k in {2, 3}: `x` has next_pow2(k^M) entries (the
    ciphertext's slot count, kway_adapter.h:44), zero beyond k^M; rotations wrap at len(x).  `compare(a, b)` defaults
    to the exact (a > b) with 1/2 on ties -- the function Comparison::compare approximates."""
    n = k ** M
    x = np.array(x, dtype=np.float64)
    B = len(x)
    if compare is None:
        compare = lambda a, b: (np.sign(a - b) + 1.0) / 2.0
    left = lambda v, r: np.roll(v, -r)
    right = lambda v, r: np.roll(v, r)
    pad = lambda m: np.concatenate([m, np.zeros(B - n)])

    def partner(v, idx, log_dist, slope):  # rightRotateForSort (:180-249)
        mask_left = pad((idx[1] < idx[0]).astype(float))
        mask_right = [pad(((idx[0] > 0) & (idx[0] == idx[1]) & (idx[0] - 1 == i)).astype(float)) for i in range(k)]
        rot = get_rotate_distance(k, log_dist, slope)
        lft = v * mask_left
        fix = None
        if slope == 0 or slope == k // 2 + 1:
            top = k - 1 if slope == 0 else k - 2
            rgt = v * mask_right[top]
            if slope != 0:
                fix = v - lft - rgt
            out = right(lft, rot) + left(rgt, top * rot)
        else:
            rgt = [v * mask_right[i] for i in range(k)]
            fix = v - lft - sum(rgt)
            out = right(lft, rot)
            for i in range(1, k):
                out = out + left(rgt[i], i * rot)
        return out, fix

    def run_two(v, idx, shift, comp):  # runTwoSorter (:8-38)
        m2 = pad(gen_mask(idx, 2, 1))
        lo, hi = two_sorter(v, left(v, shift), comp)
        return lo * m2 + right(hi * m2, shift)

    def run_three(v, idx, shift, comp):  # runThreeSorter (:40-70) with slotMatching3 (SortUtils.cpp:212-236)
        m3 = pad(gen_mask(idx, 3, 1))
        arr = [left(v, i * shift) for i in range(3)]
        comps = [m3 - left(comp, shift), comp, m3 - left(comp, 2 * shift)]
        lo, mid, hi = three_sorter(arr, comps)
        return lo * m3 + right(mid * m3, shift) + right(hi * m3, 2 * shift)

    stages = M + M * (M - 1) // 2 * ((k + 1) // 2)
    for stage in range(first_stage, stages if last_stage is None else last_stage):
        m, log_dist, slope = sort_type(k, M, stage)
        shift = get_rotate_distance(k, log_dist, slope)
        idx = gen_indices(n, k, M, m, log_dist, slope)
        rot, fix = partner(x, idx, log_dist, slope)
        comp = compare(x, rot)
        if slope == 0:
            x = run_two(x, idx, shift, comp) if k == 2 else run_three(x, idx, shift, comp)
        elif slope == k // 2 + 1:
            x = run_two(x, idx, shift, comp) + fix
        elif k == 3 and slope == 1:
            x = run_two(x, idx, shift, comp) + fix + run_three(x, idx, shift, comp)
        else:
            x = run_two(x, idx, shift, comp) + fix
    return x
