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


def four_sorter(x, comp):
    """x = [a, b, c, d], comp = [(a>b), (a>c), (a>d), (b>c), (b>d), (c>d)] -> ascending   (SortUtils.cpp:80-129)"""
    lo1, hi1 = two_sorter(x[0], x[1], comp[0])
    lo2, hi2 = two_sorter(x[2], x[3], comp[5])
    lo1_c, hi1_c = two_sorter(comp[1], comp[3], comp[0])      # min/max(a,b) against c
    lo1_d, hi1_d = two_sorter(comp[2], comp[4], comp[0])      # ... against d
    hi1_lo2, hi1_hi2 = two_sorter(hi1_c, hi1_d, comp[5])      # max(a,b) against min/max(c,d)
    lo1_lo2, lo1_hi2 = two_sorter(lo1_c, lo1_d, comp[5])
    mx = fcn_l(hi1, hi2, hi1_hi2)
    left = fcn_l(lo1, hi2, lo1_hi2)
    right = fcn_l(hi1, lo2, hi1_lo2)
    second_max = fcn_l(left, right, hi1_hi2)
    mn = fcn_l(lo2, lo1, lo1_lo2)                             # compareMin
    second_min = (x[0] + x[1] + x[2] + x[3]) - mn - second_max - mx
    return mn, second_min, second_max, mx


def five_sorter(x, comp):
    """x = [a..e], comp = [a>b a>c a>d a>e b>c b>d b>e c>d c>e d>e] -> ascending   (SortUtils.cpp:131-200)"""
    abc_comp = [comp[0], comp[1], comp[4]]
    abc = three_sorter([x[0], x[1], x[2]], abc_comp)
    de = two_sorter(x[3], x[4], comp[9])
    vs_d = three_sorter([comp[2], comp[5], comp[7]], abc_comp)
    vs_e = three_sorter([comp[3], comp[6], comp[8]], abc_comp)
    M = two_sorter(vs_d[2], vs_e[2], comp[9])                 # max(abc) against min/max(d,e)
    D = two_sorter(vs_d[1], vs_e[1], comp[9])                 # mid(abc) ...
    m = two_sorter(vs_d[0], vs_e[0], comp[9])                 # min(abc) ...
    s4 = fcn_l(abc[2], de[1], M[1])
    s0 = fcn_l(de[0], abc[0], m[0])
    left = fcn_l(abc[1], de[1], D[1])
    right = fcn_l(abc[2], de[0], M[0])
    s3 = fcn_l(left, right, M[1])
    left = fcn_l(de[0], abc[1], D[0])
    right = fcn_l(de[1], abc[0], m[1])
    s1 = fcn_l(right, left, m[0])
    s2 = sum(x) - s0 - s1 - s3 - s4
    return s0, s1, s2, s3, s4


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
    """Sorter::sorter (Sorter.cpp:284-404) on plain doubles for k in {2, 3, 5}: `x` has next_pow2(k^M) entries (the
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

    def size_mask(idx, sizes):
        return pad(((idx[1] == 1) & np.isin(idx[0], sizes)).astype(float))

    def assemble(parts, shift):  # slotAssemble (SortUtils.cpp:425-433)
        out = parts[0]
        for i in range(1, len(parts)):
            out = out + right(parts[i], i * shift)
        return out

    def run_four(v, idx, shift, c1, c2):  # runFourSorter (:71-87) with slotMatching4 (SortUtils.cpp:239-284)
        m4 = pad(gen_mask(idx, 4, 1))
        comps = [m4 - left(c1, shift), c2, c1, m4 - left(c1, 2 * shift), left(c2, shift), m4 - left(c1, 3 * shift)]
        arr = [left(v, i * shift) * m4 for i in range(4)]
        return assemble(list(four_sorter(arr, comps)), shift)

    def run_five(v, idx, shift, c1, c2):  # runFiveSorter (:89-119) with slotMatching5 (SortUtils.cpp:286-323)
        m5 = pad(gen_mask(idx, 5, 1))
        comps = [None] * 10
        comps[3] = c1; comps[0] = left(c1, shift); comps[4] = left(c1, 2 * shift); comps[7] = left(c1, 3 * shift); comps[9] = left(c1, 4 * shift)
        comps[2] = c2; comps[6] = left(c2, shift); comps[1] = left(c2, 2 * shift); comps[5] = left(c2, 3 * shift); comps[8] = left(c2, 4 * shift)
        for i in (0, 1, 4, 5, 7, 8, 9):
            comps[i] = m5 - comps[i]
        arr = [left(v, i * shift) for i in range(5)]
        return assemble([p * m5 for p in five_sorter(arr, comps)], shift)

    def run_2345(v, idx, shift, c1, c2):  # run2345Sorter (:121-178) with slotMatching2345 (SortUtils.cpp:325-423)
        m2345, m45, m345 = size_mask(idx, [2, 3, 4, 5]), size_mask(idx, [4, 5]), size_mask(idx, [3, 4, 5])
        m3, m4, m5 = size_mask(idx, [3]), size_mask(idx, [4]), size_mask(idx, [5])
        arr = [left(v, i * shift) for i in range(5)]
        comps = [m2345 - left(c1, shift),
                 c1 * m3 + (m45 - left(c2, 2 * shift) * m45),
                 c1 * m4 + c2 * m5,
                 c1 * m5,
                 m345 - left(c1, 2 * shift) * m345,
                 m45 - left(c2, 3 * shift) * m45,
                 left(c2, shift) * m5,
                 m45 - left(c1, 3 * shift) * m45,
                 m5 - left(c2, 4 * shift) * m5,
                 m5 - left(c1, 4 * shift) * m5]
        s0, s1, s2, s3, s4 = five_sorter(arr, comps)
        return assemble([s0 * m2345, s1 * m2345, s2 * m345, s3 * m45, s4 * m5], shift)

    stages = M + M * (M - 1) // 2 * ((k + 1) // 2)
    for stage in range(first_stage, stages if last_stage is None else last_stage):
        m, log_dist, slope = sort_type(k, M, stage)
        shift = get_rotate_distance(k, log_dist, slope)
        idx = gen_indices(n, k, M, m, log_dist, slope)
        rot, fix = partner(x, idx, log_dist, slope)
        comp = compare(x, rot)
        if k == 5 and slope != 2:  # two partner comparisons (comparisonForSort2, :264-282)
            rot2, _ = partner(rot, idx, log_dist, slope)
            comp2 = compare(x, rot2)
            if slope == 0:
                x = run_five(x, idx, shift, comp, comp2)
            elif slope == 3:
                x = run_four(x, idx, shift, comp, comp2) + fix
            else:
                x = run_2345(x, idx, shift, comp, comp2) + fix
        elif slope == 0:
            x = run_two(x, idx, shift, comp) if k == 2 else run_three(x, idx, shift, comp)
        elif slope == k // 2 + 1:
            x = run_two(x, idx, shift, comp) + fix
        elif (k == 3 and slope == 1) or (k == 5 and slope == 2):
            x = run_two(x, idx, shift, comp) + fix + run_three(x, idx, shift, comp)
        else:
            x = run_two(x, idx, shift, comp) + fix
    return x
