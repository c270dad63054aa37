"""oracle/algos.py -- TEST INFRASTRUCTURE ONLY (CPU oracle).

Restatement of the reference's host-side algorithms (the op sequences the hot path executes),
written against an OpenFHE-like `cc` (oracle.ckks_ops.OracleCC).  Every function cites the
reference lines it follows (/root/reference/src/...).  Used to (1) pin the oracle on the reference's
known-answer vectors and level budgets, (2) produce expected outputs / op traces for the CUDA path.
"""
import math

import numpy as np

from .cheby import truncated_doubled_sinc_coeffs, truncated_sinc_coeffs

G4_COEFFS = [
    0.0, 1.077117252745569, 0.0, -0.36166113998402755, 0.0, 0.2137420717859748, 0.0, -0.15635204788780485,
    0.0, 0.11749645501187332, 0.0, -0.10074154666447852, 0.0, 0.08002086947825496, 0.0, -0.07533558758484624,
    0.0, 0.059514472116534836, 0.0, -0.06146663712787884, 0.0, 0.04570084927999001, 0.0, -0.05403683682999072,
    0.0, 0.03364293851188723, 0.0, -0.054459493266273494]  # src/sign.cpp:68-75
G3 = [4589.0 / 1024.0, -16577.0 / 1024.0, 25614.0 / 1024.0, -12860.0 / 1024.0]  # sign.cpp:18-21
F3 = [35.0 / 16.0, -35.0 / 16.0, 21.0 / 16.0, -5.0 / 16.0]  # sign.cpp:41-44
F4 = [3.14208984375, -7.33154296875, 13.19677734375, -15.71044921875, 12.21923828125, -5.99853515625,
      1.69189453125, -0.20947265625]  # sign.cpp:81-88


def _odd7(cc, x, c):
    """CompositeSign<3>::g_n / f_n evaluation order (sign.cpp:23-36, 46-59)"""
    x2 = cc.EvalSquare(x)
    x4 = cc.EvalSquare(x2)
    y = cc.EvalMult(x, c[0])
    y = cc.EvalAdd(y, cc.EvalMultAndRelinearize(cc.EvalMult(x, c[1]), x2))
    c5x = cc.EvalMult(x, c[2])
    c7x = cc.EvalMult(x, c[3])
    c7x3 = cc.EvalMultAndRelinearize(c7x, x2)
    return cc.EvalAdd(y, cc.EvalMultAndRelinearize(cc.EvalAdd(c5x, c7x3), x4))


def _f4(cc, x):
    """CompositeSign<4>::f_n (sign.cpp:79-157)"""
    c = F4
    x2 = cc.EvalSquare(x)
    cx = [cc.EvalMult(x, v) for v in c]
    x4 = cc.EvalSquare(x2)
    c3x3 = cc.EvalMultAndRelinearize(cx[1], x2)
    c7x3 = cc.EvalMultAndRelinearize(cx[3], x2)
    c11x3 = cc.EvalMultAndRelinearize(cx[5], x2)
    c15x3 = cc.EvalMultAndRelinearize(cx[7], x2)
    x8 = cc.EvalSquare(x4)
    y = cc.EvalAdd(cx[0], c3x3)
    y = cc.EvalAdd(y, cc.EvalMultAndRelinearize(cc.EvalAdd(cx[2], c7x3), x4))
    tmp1 = cc.EvalAdd(cx[4], c11x3)
    tmp2 = cc.EvalAdd(cx[6], c15x3)
    tmp1 = cc.EvalAdd(tmp1, cc.EvalMultAndRelinearize(tmp2, x4))
    return cc.EvalAdd(y, cc.EvalMultAndRelinearize(tmp1, x8))


def composite_sign(cc, x, n, dg, df):
    """compositeSign<n> (sign.cpp:160-185): g_n applied max(dg,1) times, then f_n df times
    (the first g_n call is unconditional, :174).  No bootstrap (multDepth sentinel 100)."""
    if n == 3:
        g = lambda v: _odd7(cc, v, G3)
        f = lambda v: _odd7(cc, v, F3)
    elif n == 4:
        g = lambda v: cc.EvalChebyshevSeriesPS(v, G4_COEFFS, -1, 1)
        f = lambda v: _f4(cc, v)
    else:
        raise ValueError("composite sign n must be 3 or 4")
    y = g(x)
    for _ in range(1, dg):
        y = g(y)
    for _ in range(df):
        y = f(y)
    return y


def compare(cc, a, b, n, dg, df):
    """Comparison::compare (comparison.cpp:4-22): (sign(a-b)+1)*0.5"""
    s = composite_sign(cc, cc.EvalSub(a, b), n, dg, df)
    return cc.EvalMult(cc.EvalAdd(s, 1.0), 0.5)


def indicator(cc, x, c, n, dg, df):
    """Comparison::indicator (comparison.cpp:24-40): 1 where |x| < c"""
    s1 = composite_sign(cc, cc.EvalAdd(x, c), n, dg, df)
    s2 = composite_sign(cc, cc.EvalSub(x, c), n, dg, df)
    c1 = cc.EvalMult(cc.EvalAdd(s1, 1.0), 0.5)
    c2 = cc.EvalMult(cc.EvalAdd(s2, 1.0), 0.5)
    return cc.EvalMult(c1, cc.EvalSub(1.0, c2))


class Rotator:
    """RotationComposer<N>::rotate (rotation.h:215-233) with BINARY decomposition (:105-114)"""

    def __init__(self, cc, N, indices):
        self.cc, self.N, self.avail = cc, N, sorted(set(indices))

    def rotate(self, ct, r):
        if r % ct.slots == 0:
            return ct.clone()
        if r in self.avail:
            return self.cc.EvalRotate(ct, r)
        out = ct.clone()
        for s in decompose_binary(r, self.avail, self.N, ct.slots):
            out = self.cc.EvalRotate(out, s)
        return out


def decompose_binary(rotation, rot_indices, N, wrap):
    """Decomposer<N>::decompose(..., BINARY) (rotation.h:53-114): returns step sizes"""
    idx = sorted(rot_indices)
    max_dec, step = 0, 1
    for i in idx:  # calculateMax (:36-45)
        if step == i // 2:
            max_dec += i
        step = i
    steps = []
    largest = idx[-1]
    while rotation >= largest:
        steps.append(largest)
        rotation -= largest
    if rotation:
        while rotation > max_dec:
            import bisect
            legal = idx[bisect.bisect_left(idx, rotation) - 1]
            steps.append(legal)
            rotation -= legal
        if rotation:
            for i in range(31, -1, -1):
                s = 1 << i
                if s < N and rotation & s:
                    steps.append(s)
    return [s for s in steps if s % wrap != 0]


DIRECT_SORT_PARAMS = {  # DirectSort<N>::getSizeParameters (sort_algo.h:87-201): multDepth, rotations
    4: (23, [1, 2, 4, 8, 16]),
    8: (24, [1, 2, 4, 6, 8, 16, 32, 64]),
    16: (25, [1, 2, 3, 4, 8, 12, 16, 32, 64, 128, 256]),
    32: (28, [1, 2, 3, 4, 8, 12, 16, 20, 24, 28, 32, 64, 128, 256, 512, 1024]),
    64: (29, [1, 2, 3, 4, 5, 6, 7, 8, 16, 24, 32, 40, 48, 56, 64, 128, 256, 512, 1024, 2048, 4096]),
    128: (30, [1, 2, 3, 4, 5, 6, 7, 8, 16, 24, 32, 40, 48, 56, 64, 72, 80, 88, 96, 104, 112, 120, 128, 256,
               512, 1024, 2048, 4096, 8192, 16384]),
}
RANK_NP = {4: 2, 8: 2, 16: 4, 32: 4, 64: 8, 128: 8, 256: 16, 512: 16, 1024: 32, 2048: 32}  # sort_algo.h:383-416
INDEX_NP = {4: 2, 8: 2, 16: 4, 32: 4, 64: 8, 128: 8, 256: 16, 512: 8, 1024: 8}  # sort_algo.h:670-703


class DirectSort:
    """DirectSort<N> (sort_algo.h:61-774): constructRank + rotationIndexCheckN"""

    def __init__(self, cc, N, rot_indices, sinc_coeffs=None):
        self.cc, self.N = cc, N
        self.rot = Rotator(cc, N, rot_indices)
        self.max_batch = cc.ctx.n // 2
        self.zero = cc.encrypt(cc.encode([0.0] * N, slots=N))  # SortBase::createZeroCache (:42-45)
        self.sinc = truncated_doubled_sinc_coeffs(N) if sinc_coeffs is None else sinc_coeffs

    def mask(self, num_slots, k):  # generateMaskVector (:206-214)
        v = np.zeros(num_slots)
        v[k * self.N:(k + 1) * self.N] = 1.0
        return v

    def vec_rots_opt(self, baby, num_partition, num_slots, np_, is_):
        """vecRotsOpt (:326-366)"""
        cc = self.cc
        outer = []
        for j in range(num_partition // np_):
            T = self.zero.clone()
            T.slots = num_slots
            for i in range(np_):
                msk = np.roll(self.mask(num_slots, np_ * j + i), is_ * num_partition + j * np_)  # vectorRotate(-r)
                pm = cc.MakeCKKSPackedPlaintext(msk, 1, 0, None, num_slots)
                T = cc.EvalAdd(T, cc.EvalMult(baby[i], pm))
            outer.append(self.rot.rotate(T, is_ * num_partition + j * np_))
        return cc.EvalAddMany(outer)

    def construct_rank(self, ct_in, sign_cfg):
        """constructRank (:368-506)"""
        cc, N = self.cc, self.N
        num_partition = min(N, self.max_batch // N)
        num_batch = N // num_partition
        num_slots = N * num_partition
        np_ = min(RANK_NP[N], num_partition)
        baby = []
        for i in range(np_):
            t = self.rot.rotate(ct_in, i)
            t.slots = num_slots
            baby.append(t)
        rank = self.zero.clone()
        rank.slots = num_slots
        for i in range(num_batch):
            shifted = self.vec_rots_opt(baby, num_partition, num_slots, np_, i)
            dup = ct_in.clone()
            dup.slots = num_slots
            rank = cc.EvalAdd(rank, compare(cc, dup, shifted, *sign_cfg))
        for i in range(1, int(math.log2(num_partition)) + 1):
            rank = cc.EvalAdd(rank, self.rot.rotate(rank, num_slots // (1 << i)))
        rank.slots = N
        return cc.EvalSub(rank, 0.5)

    def blind_rotation_opt_n(self, masked, num_slots, np_, ib, num_partition):
        """blindRotationOptN (:561-584)"""
        cc = self.cc
        result = self.zero.clone()
        for i in range((num_slots // self.N) // np_):
            tmp = self.zero.clone()
            for j in range(np_):
                msk = np.roll(self.mask(num_slots, np_ * i + j), -j)  # vectorRotate(msk, j): left by j
                pm = cc.MakeCKKSPackedPlaintext(msk, 1, masked[j].level, None, num_slots)
                tmp = cc.EvalAdd(tmp, cc.EvalMult(masked[j], pm))
            tmp = self.rot.rotate(tmp, ib * num_partition + i * np_)
            result = cc.EvalAdd(result, tmp)
        return result

    def rotation_index_check_n(self, rank, ct_in):
        """rotationIndexCheckN (:658-750)"""
        cc, N = self.cc, self.N
        out = self.zero.clone()
        num_partition = min(N, self.max_batch // N)
        num_batch = N // num_partition
        num_slots = num_partition * N
        np_ = INDEX_NP.get(N, 4)
        idx_pt = cc.MakeCKKSPackedPlaintext(np.arange(N, dtype=np.float64), 1, rank.level, None, N)
        d = cc.EvalSub(idx_pt, rank)
        d.slots = num_slots
        ct_in = ct_in.clone()
        ct_in.slots = num_slots
        for b in range(num_batch):
            chk = np.array([(b * num_partition + (i // N)) % N for i in range(num_slots)], dtype=np.float64)  # :272-286
            pchk = cc.MakeCKKSPackedPlaintext(chk, 1, d.level, None, num_slots)
            r = cc.EvalSub(d, pchk)
            r = cc.EvalMult(r, 1.0 / N / 2)
            r = cc.EvalChebyshevSeriesPS(r, self.sinc, -1, 1)
            masked_input = cc.EvalMult(r, ct_in)
            masked = [self.rot.rotate(masked_input, i) for i in range(np_)]
            out = cc.EvalAdd(out, self.blind_rotation_opt_n(masked, num_slots, np_, b, num_partition))
        for i in range(1, int(math.log2(num_partition)) + 1):
            out = cc.EvalAdd(out, self.rot.rotate(out, num_slots // (1 << i)))
        out.slots = N
        return out

    def _fold_to_target(self, c, size, target, first_step, row_mask):
        """sumColumnsToTarget (:825-858) / transposeColumnTarget (:860-890)"""
        cc, lg = self.cc, int(math.log2(size))
        c = c.clone(); c.slots = size * size
        step = first_step
        for i in range(lg):
            right = (target >> (lg - 1 - i)) & 1
            c = cc.EvalAdd(c, self.rot.rotate(c, -step if right else step))
            step >>= 1
        m = np.zeros(size * size)
        if row_mask:
            m[size * target:size * (target + 1)] = 1.0
        else:
            m[target::size] = 1.0
        return cc.EvalMult(c, cc.MakeCKKSPackedPlaintext(m, 1, c.level, None, size * size))

    def rotation_index_check_hybrid(self, rank, ct_in, variant):
        """rotationIndexCheckHybrid (:893-1049, variant 0), Hybrid1 (:1067-1211, variant 1), Hybrid2 (:1232-1372, 2)"""
        cc, N = self.cc, self.N
        num_slots, num_batch = (self.max_batch, N // 256) if N > 256 else (N * N, 1)
        size = min(N, 256)
        scaled = variant != 1
        rank = rank.clone(); rank.slots = num_slots
        r = cc.EvalMult(rank, 1.0 / N) if scaled else rank
        ct_in = ct_in.clone(); ct_in.slots = num_slots
        rots_rank = [self.rot.rotate(r, b * 256) for b in range(num_batch)]
        rots_in = [self.rot.rotate(ct_in, b * 256) for b in range(num_batch)]
        dg_i, df_i = int((math.log2(N) + 1) / 2), 2
        masked = []
        for b in range(num_batch):
            sub = np.zeros(num_slots)
            rows = np.repeat(np.arange(size, dtype=np.float64) + b * size, size)
            sub[:size * size] = rows / float(N) if scaled else rows
            psub = cc.MakeCKKSPackedPlaintext(sub, 1, rank.level, None, num_slots)
            acc = self.zero.clone(); acc.slots = num_slots
            for k in range(num_batch):
                m = cc.EvalSub(psub, rots_rank[k])
                if variant == 1:
                    m = Mehp24(cc, size).indicator_adv(m, float(N), dg_i, df_i)
                elif variant == 2 or N < 256:
                    if not hasattr(self, "_sinc1"):
                        self._sinc1 = truncated_sinc_coeffs(N)
                    m = cc.EvalChebyshevSeriesPS(m, self._sinc1, -1, 1)
                else:
                    m = indicator(cc, m, 0.5 / N, 3, 4 if N < 512 else 5, 2)
                acc = cc.EvalAdd(acc, cc.EvalMult(rots_in[k], m))
            acc = self._fold_to_target(acc, N // num_batch, b, (N // num_batch) >> 1, False)
            masked.append(self._fold_to_target(acc, N // num_batch, b, (N // num_batch) * (N // num_batch - 1) // 2, True))
        return cc.EvalAddMany(masked)

    def sort_hybrid(self, ct_in, sign_cfg, variant):
        """sort_hybrid / sort_hybrid1 / sort_hybrid2 (:1051-1065, 1214-1230, 1375-1391)"""
        return self.rotation_index_check_hybrid(self.construct_rank(ct_in, sign_cfg), ct_in, variant)

    def sort(self, ct_in, sign_cfg):
        """DirectSort::sort (:752-774)"""
        rank = self.construct_rank(ct_in, sign_cfg)
        return self.rotation_index_check_n(rank, ct_in)


# ---------------------------------------------------------------------------------------------------------------
# MEHP24 single-ciphertext sortFG (reference src/mehp24/mehp24_sort.cpp:251-282, mehp24_utils.cpp:21-261)
F3_POLY = [0, 35.0 / 16.0, 0, -35.0 / 16.0, 0, 21.0 / 16.0, 0, -5.0 / 16.0]
F3_FINAL_POLY = [0.5, 35.0 / 32.0, 0, -35.0 / 32.0, 0, 21.0 / 32.0, 0, -5.0 / 32.0]
G3_POLY = [0, 4589.0 / 1024.0, 0, -16577.0 / 1024.0, 0, 25614.0 / 1024.0, 0, -12860.0 / 1024.0]


def _lg(n):
    return int(np.ceil(np.log2(n)))


class Mehp24:
    """matrix-in-slots utilities on an n x n slot matrix; batch = n*n slots (tests/mehp24/Mehp24SortTest.cpp:36-37)"""

    def __init__(self, cc, n):
        self.cc, self.n, self.batch = cc, n, n * n

    def _pt(self, v):  # one-argument MakeCKKSPackedPlaintext: level 0, batch-size slots (mehp24_utils.h:66-72)
        return self.cc.MakeCKKSPackedPlaintext(v, 1, 0, None, self.batch)

    def mask_row(self, c, row):  # mehp24_utils.cpp:21-30
        m = np.zeros(self.batch); m[self.n * row:self.n * (row + 1)] = 1.0
        return self.cc.EvalMult(c, self._pt(m))

    def mask_column(self, c, col):  # :32-42
        m = np.zeros(self.batch); m[col::self.n] = 1.0
        return self.cc.EvalMult(c, self._pt(m))

    def _fold(self, c, amounts):
        for r in amounts:
            c = self.cc.EvalAdd(c, self.cc.EvalRotate(c, r))
        return c

    def replicate_row(self, c):  # :44-50
        return self._fold(c, [-(1 << (_lg(self.n) + i)) for i in range(_lg(self.n))])

    def replicate_column(self, c):  # :52-58
        return self._fold(c, [-(1 << i) for i in range(_lg(self.n))])

    def sum_rows(self, c, mask=False, row=0):  # :60-69
        c = self._fold(c, [-(1 << (_lg(self.n) + i)) for i in range(_lg(self.n))])
        return self.mask_row(c, row) if mask else c

    def sum_columns(self, c, mask=False):  # :71-80
        c = self._fold(c, [1 << i for i in range(_lg(self.n))])
        return self.mask_column(c, 0) if mask else c

    def transpose_row(self, c, mask=False):  # :82-91
        n = self.n
        c = self._fold(c, [-(n * (n - 1) // (1 << i)) for i in range(1, _lg(n) + 1)])
        return self.mask_column(c, 0) if mask else c

    def transpose_column(self, c, mask=False):  # :93-103
        n = self.n
        c = self._fold(c, [n * (n - 1) // (1 << i) for i in range(1, _lg(n) + 1)])
        return self.mask_row(c, 0) if mask else c

    def sign_adv(self, c, dg, df):  # :246-261
        for _ in range(dg):
            c = self.cc.EvalPolyLinear(c, G3_POLY)
        for _ in range(df - 1):
            c = self.cc.EvalPolyLinear(c, F3_POLY)
        return self.cc.EvalPolyLinear(c, F3_FINAL_POLY)

    def indicator_adv(self, c, b, dg, df):  # :166-174
        cc = self.cc
        tmp = cc.EvalMult(c, 1.0 / b)
        c1 = self.sign_adv(cc.EvalAdd(tmp, 0.5 / b), dg, df)
        c2 = self.sign_adv(cc.EvalSub(tmp, 0.5 / b), dg, df)
        return cc.EvalMult(c1, cc.EvalSub(1.0, c2))

    def sort_fg(self, c, sign_cfg, dg_i, df_i):  # mehp24_sort.cpp:251-282
        cc, n = self.cc, self.n
        VR = self.replicate_row(c)
        VC = self.replicate_column(self.transpose_row(c, True))
        C = compare(cc, VR, VC, *sign_cfg)
        R = self.sum_rows(C)
        sub = np.repeat(-1.0 * np.arange(n) - 0.5, n)
        M = self.indicator_adv(cc.EvalAdd(R, self._pt(sub)), float(n), dg_i, df_i)
        cc.EvalMult(M, VR)  # computed twice in the reference (:275-277)
        S = self.sum_columns(cc.EvalMult(M, VR), True)
        return self.transpose_column(S, True)


    def sort_fg_adv(self, c, dg_c, df_c, dg_i, df_i):  # mehp24_sort.cpp:219-249 (compareAdv overload)
        cc, n = self.cc, self.n
        VR = self.replicate_row(c)
        VC = self.replicate_column(self.transpose_row(c, True))
        C = self.sign_adv(cc.EvalSub(VR, VC), dg_c, df_c)          # compareAdv, mehp24_utils.cpp:139-145
        R = self.sum_rows(C)
        sub = np.repeat(-1.0 * np.arange(n) - 0.5, n)
        M = self.indicator_adv(cc.EvalAdd(R, self._pt(sub)), float(n), dg_i, df_i)
        cc.EvalMult(M, VR)  # computed twice in the reference (:243-244)
        S = self.sum_columns(cc.EvalMult(M, VR), True)
        return self.transpose_column(S, True)


def mehp24_sort_large_adv(cc, c, total_len, sub_len, dg_c, df_c, dg_i, df_i):
    """sortLargeArrayFG, compareAdv overload (mehp24_sort.cpp:607-623 over the multi-ciphertext sortFG :285-443)"""
    mh = Mehp24(cc, sub_len)
    return mehp24_sort_large(cc, c, total_len, sub_len, None, dg_i, df_i,
                             compare_blocks=lambda a, b: mh.sign_adv(cc.EvalSub(a, b), dg_c, df_c))


def mehp24_sort_large(cc, c, total_len, sub_len, sign_cfg, dg_i, df_i, compare_blocks=None):
    """sortLargeArrayFG (mehp24_sort.cpp:625-645) = splitCiphertext (mehp24_utils.cpp:266-289) -> multi-ciphertext sortFG,
    Comparison-based overload (mehp24_sort.cpp:445-605) -> combineCiphertext (mehp24_utils.cpp:291-302)"""
    mh, n = Mehp24(cc, sub_len), sub_len
    nc = total_len // sub_len
    parts = []
    for i in range(nc):
        mask = np.zeros(total_len); mask[i * n:(i + 1) * n] = 1.0
        p = cc.EvalMult(c, mh._pt(mask))
        parts.append(cc.EvalRotate(p, i * n) if i else p)
    repl_r = [mh.replicate_row(p) for p in parts]
    repl_c = [mh.replicate_column(mh.transpose_row(p, True)) for p in parts]
    acc = lambda a, x: x if a is None else cc.EvalAdd(a, x)
    cv, ch, s = [None] * nc, [None] * nc, [None] * nc
    for j in range(nc):
        for k in range(j, nc):
            cjk = compare_blocks(repl_r[j], repl_c[k]) if compare_blocks else compare(cc, repl_r[j], repl_c[k], *sign_cfg)
            cv[j] = acc(cv[j], cjk)
            if j != k:
                ch[k] = acc(ch[k], cc.EvalSub(1.0, cjk))
    for j in range(nc):
        s[j] = acc(s[j], mh.sum_rows(cv[j]))
    for j in range(1, nc):
        s[j] = acc(s[j], mh.replicate_row(mh.transpose_column(mh.sum_columns(ch[j], True), True)))
    result = []
    for j in range(nc):
        sub = np.repeat(-1.0 * (np.arange(n) + j * n) - 0.5, n)
        srt = None
        for k in range(nc):
            ind = mh.indicator_adv(cc.EvalAdd(s[k], mh._pt(sub)), float(total_len), dg_i, df_i)
            srt = acc(srt, cc.EvalMult(ind, repl_r[k]))
        result.append(mh.sum_columns(srt, True))
    result = [mh.transpose_column(r, True) for r in result]
    out = result[0]
    for i in range(1, nc):
        out = cc.EvalAdd(out, cc.EvalRotate(result[i], -i * n))
    return out


def mehp24_rotation_indices(n):  # getRotationIndices (mehp24_utils.cpp:186-213)
    sz, idx = n, []
    if n > 256:
        for i in range(n // 256):
            idx += [i * 256, -i * 256]
        sz = 256
    lg = _lg(sz)
    for i in range(lg):
        idx += [1 << i, -(1 << i), -(1 << (lg + i)), sz * (sz - 1) // (1 << (i + 1)), -(sz * (sz - 1) // (1 << (i + 1)))]
    return idx
