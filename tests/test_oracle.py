"""CPU tests of the oracle itself: the fast twins (oracle/ckks_fast.c) against the slow reference loops
(oracle/ckks_prims.c), the known-answer material the reference holds for this path (tests/SignTest.cpp,
tests/CompareTest.cpp, level budgets of src/sort_algo.h:87-201), and a full DirectSort on a small ring."""
import math

import numpy as np
import pytest

from oracle import algos
from oracle.cheby import compute_degrees_ps, eval_chebyshev_plain, truncated_doubled_sinc_coeffs
from oracle.ckks_oracle import P, Context, Params, lib
from oracle.ckks_ops import OracleCC


@pytest.fixture(scope="module")
def ctx11():
    return Context(Params(11, 6, 40))


def test_fast_ntt_matches_slow(ctx11):
    c, L = ctx11, lib()
    rng = np.random.default_rng(0)
    ids = list(range(len(c.moduli)))
    a = np.stack([rng.integers(0, c.moduli[i], c.n, dtype=np.uint64) for i in ids])
    fast = a.copy()
    c.ntt_rows(fast, ids)
    slow = a.copy()
    for r, i in enumerate(ids):
        L.orc_ntt_fwd(P(slow[r]), c.log_n, c.moduli[i], P(c.psi_br[i]))
    assert np.array_equal(fast, slow)
    c.ntt_rows(fast, ids, inverse=True)
    for r, i in enumerate(ids):
        L.orc_ntt_inv(P(slow[r]), c.log_n, c.moduli[i], P(c.psi_inv_br[i]))
    assert np.array_equal(fast, a) and np.array_equal(slow, a)


def test_ntt_is_negacyclic_convolution(ctx11):
    """pins the transform itself: pointwise product in EVALUATION format == schoolbook product mod X^n + 1"""
    c = Context(Params(8, 1, 40))
    q, n = c.moduli[0], c.n
    rng = np.random.default_rng(1)
    a = rng.integers(0, q, n, dtype=np.uint64)
    b = rng.integers(0, q, n, dtype=np.uint64)
    want = [0] * n
    for i in range(n):
        for j in range(n):
            k, s = (i + j) % n, (1 if i + j < n else -1)
            want[k] = (want[k] + s * int(a[i]) * int(b[j])) % q
    fa, fb = a.copy(), b.copy()
    c.ntt(fa, 0); c.ntt(fb, 0)
    prod = np.array([(int(x) * int(y)) % q for x, y in zip(fa, fb)], dtype=np.uint64)
    c.intt(prod, 0)
    assert [int(x) for x in prod] == want


@pytest.mark.parametrize("name", ["vadd", "vsub", "vmul", "vmuladd"])
def test_fast_vector_ops_match_slow(ctx11, name):
    c, L = ctx11, lib()
    rng = np.random.default_rng(2)
    for q in (c.moduli[0], c.moduli[1], c.moduli[-1]):
        a = rng.integers(0, q, c.n, dtype=np.uint64)
        b = rng.integers(0, q, c.n, dtype=np.uint64)
        o1, o2 = a.copy(), a.copy()
        getattr(L, "orf_" + name)(P(o1), P(a), P(b), c.n, q)
        getattr(L, "slow_" + name)(P(o2), P(a), P(b), c.n, q)
        assert np.array_equal(o1, o2)


def test_fast_scalar_automorph_conv_match_slow(ctx11):
    c, L = ctx11, lib()
    rng = np.random.default_rng(3)
    q = c.moduli[0]
    a = rng.integers(0, q, c.n, dtype=np.uint64)
    for nm, s in (("vmuls", q - 5), ("vadds", 12345), ("vmuls", 3)):
        o1, o2 = np.empty_like(a), np.empty_like(a)
        getattr(L, "orf_" + nm)(P(o1), P(a), s, c.n, q)
        getattr(L, "slow_" + nm)(P(o2), P(a), s, c.n, q)
        assert np.array_equal(o1, o2)
    for k in (5, 25, 2 * c.n - 1, pow(5, -1, 2 * c.n)):
        o1, o2 = np.empty_like(a), np.empty_like(a)
        L.orf_automorph_eval(P(o1), P(a), c.log_n, k)
        L.slow_automorph_eval(P(o2), P(a), c.log_n, k)
        assert np.array_equal(o1, o2)
    parts, md = c.ks_tables(c.size_q)
    pt = parts[0]
    x = np.stack([rng.integers(0, int(qq), c.n, dtype=np.uint64) for qq in pt["qin"]])
    o1 = np.empty((len(pt["compl"]), c.n), dtype=np.uint64)
    o2 = np.empty_like(o1)
    args = (c.n, len(pt["dig"]), len(pt["compl"]), P(pt["qin"]), P(pt["qout"]), P(pt["hatinv"]), P(pt["hatmod"]))
    L.orf_approx_switch_crt_basis(P(o1), P(x), *args)
    L.slow_approx_switch_crt_basis(P(o2), P(x), *args)
    assert np.array_equal(o1, o2)
    o1, o2 = np.empty_like(a), np.empty_like(a)
    L.orf_switch_modulus(P(o1), P(a), c.n, q, c.moduli[1])
    L.slow_switch_modulus(P(o2), P(a), c.n, q, c.moduli[1])
    assert np.array_equal(o1, o2)


def test_moduli_chain_shape():
    """DirectSort N=128 parameter card: depth 30, 40-bit scaling, 60-bit first, EXT -> 32 Q limbs,
    dnum 3 -> alpha 11, K = 8 special primes; all primes = 1 mod 2n"""
    c = Context(Params(12, 30, 40))
    assert (c.size_q, c.alpha, c.num_part_q, c.size_p) == (32, 11, 3, 8)
    assert all(q % (2 * c.n) == 1 for q in c.moduli) and len(set(c.moduli)) == len(c.moduli)
    assert c.moduli[0].bit_length() == 60 and all(39 <= q.bit_length() <= 41 for q in c.q[1:31])
    assert all(abs(math.log2(s) - 40) < 0.01 for s in c.sf_real[1:])


@pytest.fixture(scope="module")
def cc10():
    cc = OracleCC(Context(Params(10, 12, 50)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen([1, -1, 3])
    return cc


def test_encrypt_ops_decrypt(cc10):
    cc = cc10
    x = np.array([0.5, -0.3, 0.1, -0.7, 0.0, 0.8, -0.9, 0.2])
    ct = cc.encrypt(cc.encode(x, slots=8))
    assert np.allclose(cc.decrypt(ct, 8), x, atol=1e-9)
    assert np.allclose(cc.decrypt(cc.EvalMult(ct, ct), 8), x * x, atol=1e-8)
    assert np.allclose(cc.decrypt(cc.EvalRotate(ct, 1), 8), np.roll(x, -1), atol=1e-8)
    assert np.allclose(cc.decrypt(cc.EvalRotate(ct, -1), 8), np.roll(x, 1), atol=1e-8)
    assert np.allclose(cc.decrypt(cc.EvalSub(1.0, cc.EvalMult(ct, 0.5)), 8), 1 - x / 2, atol=1e-8)
    pt = cc.MakeCKKSPackedPlaintext(x, 1, 0, None, 8)
    assert np.allclose(cc.decrypt(cc.EvalMult(cc.EvalMult(ct, ct), pt), 8), x ** 3, atol=1e-8)


def test_sign_kat(cc10):
    """tests/SignTest.cpp:41-80 input vector through compositeSign<3>(dg=0, df=1): equals f_3(g_3(x)) and has the
    signs the reference checks; |x| >= 0.2 lands within 0.1 of +-1"""
    cc = cc10
    x = np.array([0.5, -0.3, 0.1, -0.7, 0.0, 0.8, -0.9, 0.2])
    got = cc.decrypt(algos.composite_sign(cc, cc.encrypt(cc.encode(x, slots=8)), 3, 0, 1), 8)
    odd7 = lambda v, c: c[0] * v + c[1] * v ** 3 + c[2] * v ** 5 + c[3] * v ** 7
    assert np.max(np.abs(got - odd7(odd7(x, algos.G3), algos.F3))) < 1e-6
    assert np.all(np.sign(got[x != 0]) == np.sign(x[x != 0]))
    assert np.all(np.abs(got - np.sign(x))[np.abs(x) >= 0.2] < 0.1)


def test_compare_kat():
    """tests/CompareTest.cpp:44-62: a={1,5,3,4}, b={2,4,3,3} -> {0,1,0.5,1} +-0.1 with Sign(4,3,3)
    (ring shrunk to 2^10 for CPU time; values normalised by 1/8 into sign's domain)"""
    cc = OracleCC(Context(Params(10, 29, 59)))
    cc.keygen(); cc.relin_keygen()
    a, b = np.array([1, 5, 3, 4]) / 8.0, np.array([2, 4, 3, 3]) / 8.0
    r = algos.compare(cc, cc.encrypt(cc.encode(a, slots=4)), cc.encrypt(cc.encode(b, slots=4)), 4, 3, 3)
    got = cc.decrypt(r, 4)
    assert np.all(np.abs(got - np.array([0, 1, 0.5, 1])) < 0.1), got
    assert r.level == 28   # 3 x g_4 (PS degree 27: 5 levels) + 3 x f_4 (4) + 1 (x 0.5)


def test_doubled_sinc_table():
    """degree of the generated table and its accuracy as an indicator (tests/SincTest.cpp:175-186 prints the
    degree per N; SURVEY 8a derives 126 for N=16)"""
    c = truncated_doubled_sinc_coeffs(16)
    assert len(c) - 1 == 126
    k, m = compute_degrees_ps(126)
    assert 126 < k * ((1 << m) - 1)
    xs = np.arange(-2 * 16 + 1, 16) / 32.0      # (i - rank - k) / 2N over the reachable range
    want = np.array([1.0 if (int(round(v * 32)) % 16 == 0) else 0.0 for v in xs])
    assert np.max(np.abs(eval_chebyshev_plain(c, xs) - want)) < 1e-4


@pytest.mark.parametrize("n,cfg", [(4, (3, 2, 2)), (8, (3, 2, 2))])
def test_direct_sort_small_ring(n, cfg):
    """DirectSort<N>::sort on ring 2^10: sorted output within the reference gate 0.01 and final level ==
    multDepth (tests/DirectSortTest.cpp:140,194)"""
    depth, rots = algos.DIRECT_SORT_PARAMS[n]
    cc = OracleCC(Context(Params(10, depth, 40)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen(rots)
    x = np.random.default_rng(n).permutation(n) / n
    out = algos.DirectSort(cc, n, rots).sort(cc.encrypt(cc.encode(x, slots=n)), cfg)
    assert out.level == depth
    assert np.max(np.abs(cc.decrypt(out, n) - np.sort(x))) < 0.01
