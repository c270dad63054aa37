"""CPU models of two index / arithmetic schemes the CUDA kernels rely on (no GPU, no oracle): they restate the device code
of sorting-fhe_b200/csrc in numpy and check the invariants the kernels' comments claim."""
import numpy as np


# ---- ntt.cu: swz / slot_of -- the shared-memory exchange of the NTT rounds
def local_index(u, k, p):
    return ((u >> p) << (p + 3)) | (k << p) | (u & ((1 << p) - 1))


def swz(e):
    return e ^ ((e >> 4) & 7) ^ (((0x96 >> (e & 7)) & 1) << 3)


def test_ntt_exchange_swizzle_is_a_conflict_free_permutation():
    """every tile size 2^4..2^10: swz permutes the tile, and for every round window p the 16 elements a half-warp touches
    (one 8-byte word each) fall into 16 different bank pairs -- one wavefront per half-warp, two per warp"""
    for T in range(4, 11):
        n = 1 << T
        assert sorted(swz(e) for e in range(n)) == list(range(n))
        threads = n // 8
        for p in range(0, T - 2):
            for k in range(8):
                for base in range(0, threads, 16):
                    lanes = range(base, min(base + 16, threads))
                    banks = [swz(local_index(u, k, p)) % 16 for u in lanes]
                    assert len(set(banks)) == len(banks), (T, p, k, base)
    # slot_of() splits the index into the thread part and the register part: swz must be linear over GF(2)
    rng = np.random.default_rng(0)
    for a, b in rng.integers(0, 1024, (200, 2)):
        assert swz(int(a) ^ int(b)) == swz(int(a)) ^ swz(int(b))


# ---- keyswitch.cu: k_conv_fp -- base conversion with exact FP64 column sums
def conv_fp_model(y, h, wide_in, wide_out):
    """one output residue before the Montgomery reduction: sum_i y_i * h_i as k_conv_fp accumulates it (21-bit pieces,
    float64 fused multiply-adds modelled by exact products below 2^53), returned as a Python integer"""
    M = (1 << 21) - 1
    s = [np.float64(0)] * 5
    for yi, hi in zip(y, h):
        yp = [np.float64(yi & M), np.float64((yi >> 21) & M)] + ([np.float64(yi >> 42)] if wide_in else [])
        hp = [np.float64(hi & M), np.float64((hi >> 21) & M)] + ([np.float64(hi >> 42)] if wide_out else [])
        for a, ya in enumerate(yp):
            for b, hb in enumerate(hp):
                prod = ya * hb                         # < 2^42: exact in float64, so fma(ya, hb, s) == s + prod exactly
                assert prod < 2.0 ** 42
                s[a + b] = s[a + b] + prod
                assert s[a + b] < 2.0 ** 48            # the kernel comment's bound; far below 2^53
    return sum(int(c) << (21 * w) for w, c in enumerate(s))


def test_conv_fp_column_sums_are_exact():
    rng = np.random.default_rng(1)
    for wide_in, wide_out in ((False, False), (True, False), (False, True), (True, True)):
        ybits, hbits = (60 if wide_in else 42), (60 if wide_out else 42)
        for _ in range(50):
            # worst case first: all-ones residues and constants, 12 source limbs (the kernel's maximum)
            y = [(1 << ybits) - 1] * 12 if _ == 0 else [int(v) for v in rng.integers(0, 1 << ybits, 12, dtype=np.uint64)]
            h = [(1 << hbits) - 1] * 12 if _ == 0 else [int(v) for v in rng.integers(0, 1 << hbits, 12, dtype=np.uint64)]
            assert conv_fp_model(y, h, wide_in, wide_out) == sum(a * b for a, b in zip(y, h))
