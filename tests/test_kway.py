"""k-way sorting network of HKC+21 (reference src/k-way/, tests/k-way/*Test.cpp, tests/KWaySortTest.cpp): index logic
and plaintext model on CPU; compare-swap stages on the GPU against the plaintext model (no bootstrapping: a network
runs as far as the multiplicative depth allows)."""
import importlib

import numpy as np
import pytest

from oracle import kway_plain as kp

sf = importlib.import_module("sorting-fhe_b200")


def next_pow2(n):
    b = 1
    while b < n:
        b *= 2
    return b


@pytest.mark.parametrize("k,M", [(2, 3), (2, 7), (3, 3), (5, 2), (5, 3)])
def test_index_logic_matches_plain_model(k, M):
    """the product's sortType / getRotateDistance / genIndices (host/kway.h) == the oracle restatement of Masking.cpp"""
    stages = sf.kway_sort_type(k, M, 0)[3]
    assert stages == M + M * (M - 1) // 2 * ((k + 1) // 2)      # Sorter.cpp:298 (k=2, M=7 -> 28 stages)
    for st in range(stages):
        m, ld, sl, _, shift = sf.kway_sort_type(k, M, st)
        assert (m, ld, sl) == kp.sort_type(k, M, st) and shift == kp.get_rotate_distance(k, ld, sl)
        assert np.array_equal(sf.kway_gen_indices(k ** M, k, M, m, ld, sl), kp.gen_indices(k ** M, k, M, m, ld, sl))


def test_adapter_parameters():
    """KWayAdapter<N>::getSizeParameters (kway_adapter.h:41-64)"""
    depth, scale, first, rots = sf.kway_params(128)
    assert (depth, scale, first) == (40, 59, 60)
    assert rots == [s * (1 << i) for i in range(7) for s in (1, -1)]


@pytest.mark.parametrize("k,M", [(2, 1), (2, 4), (2, 7), (3, 2), (3, 3), (5, 1), (5, 2), (5, 3)])
def test_plain_network_sorts(k, M):
    """the network with an exact comparator sorts (what tests/k-way/SorterTest.cpp:313-381 checks on 8 / 3 / 5 elements)"""
    n = k ** M
    rng = np.random.default_rng(k * 10 + M)
    for _ in range(4):
        v = rng.permutation(n) / n + 0.5 / n
        y = kp.network_plain(np.concatenate([v, np.zeros(next_pow2(n) - n)]), k, M)
        assert np.allclose(y[:n], np.sort(v))


def test_stage_levels_dry_run():
    """one k=2 stage at the adapter's card (depth 40, scale 59): mask (1) + Sign(3,2,2) compare (13) + fcnL (1) +
    output mask (1) = 16 levels; the third stage would need the bootstrap the reference performs there"""
    n, k, M = 128, 2, 7
    depth, scale, first, rots = sf.kway_params(n)
    eng = sf.Engine(17, depth, scale, first, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    ct = eng.encrypt(eng.encode(np.arange(n) / n, slots=n))
    s1 = sf.kway_sort(eng, ct, n, k, M, (3, 2, 2), depth, 0, 1)
    assert s1.level == 16
    s2 = sf.kway_sort(eng, s1, n, k, M, (3, 2, 2), depth, 1, 2)
    assert s2.level == 32
    with pytest.raises(sf.CkksError, match="EvalBootstrap"):
        sf.kway_sort(eng, s2, n, k, M, (3, 2, 2), depth, 2, 3)
    eng.close()


def test_four_five_sorter_exhaustive():
    """SortUtils::fourSorter / fiveSorter (src/k-way/SortUtils.cpp:80-200) on exact comparison bits: every order of 4 and
    of 5 distinct values comes out ascending (the literal vectors of tests/k-way/SortUtilsTest.cpp:157-255 are checked
    in tests/test_golden.py)"""
    import itertools
    for n, f in ((4, kp.four_sorter), (5, kp.five_sorter)):
        for perm in itertools.permutations(range(n)):
            x = [np.array([float(v)]) for v in perm]
            comp = [(x[i] > x[j]).astype(float) for i in range(n) for j in range(i + 1, n)]
            assert [float(v[0]) for v in f(x, comp)] == list(range(n)), perm


def test_k5_stage_levels_dry_run():
    """k = 5, M = 2 (25 elements): the five stages are a 5-sorter, a 5-sorter, the mixed 2/3/4/5 stage, the 2+3 stage and a
    4-sorter (Sorter.cpp:326-380); level use = mask + Sign(3,3,2) compare (16) + the sorter's own levels"""
    k, M, n, B = 5, 2, 25, 32
    rots = [s * (1 << i) for i in range(5) for s in (1, -1)]
    eng = sf.Engine(11, 34, 50, 60, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    assert sf.kway_sort_type(k, M, 0)[3] == 5
    levels = []
    for st in range(5):
        ct = eng.encrypt(eng.encode(np.arange(B) / B, slots=B))
        levels.append(sf.kway_sort(eng, ct, n, k, M, (3, 3, 2), 34, st, st + 1).level)
    assert [sf.kway_sort_type(k, M, st)[2] for st in range(5)] == [0, 0, 1, 2, 3]
    assert levels == [24, 24, 25, 20, 22]
    eng.close()


@pytest.mark.gpu
def test_k5_stages_match_plain_network_on_gpu():
    """every stage of the k = 5, M = 2 network on the GPU, started from the plaintext network's state before that stage,
    equals the plaintext network's state after it within 0.01 (4-, 5- and mixed sorters, slot matching, assembly)"""
    k, M, n, B = 5, 2, 25, 32
    rots = [s * (1 << i) for i in range(5) for s in (1, -1)]
    eng = sf.Engine(11, 34, 50, 60, seed=9)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    v = np.random.default_rng(25).permutation(n) / n + 0.5 / n
    state = np.concatenate([v, np.zeros(B - n)])
    for st in range(5):
        want = kp.network_plain(state, k, M, first_stage=st, last_stage=st + 1)
        out = sf.kway_sort(eng, eng.encrypt(eng.encode(state, slots=B)), n, k, M, (3, 3, 2), 34, st, st + 1)
        got = eng.decrypt(out, B)
        assert np.max(np.abs(got - want)) < 0.01, (st, got, want)
        state = want
    assert np.allclose(state[:n], np.sort(v))
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("k,M", [(2, 2), (3, 1), (5, 1)])
def test_small_network_sorts_on_gpu(k, M):
    """whole network on k^M elements with enough depth to need no bootstrap (ring 2^11, 50-bit scaling); decrypted
    output = sorted input within the reference's 0.01 gate (tests/k-way/KWaySort2Test.cpp:214)"""
    n = k ** M
    B = next_pow2(n)
    stages = sf.kway_sort_type(k, M, 0)[3]
    depth = 20 * stages + (8 if k == 5 else 4)
    rots = [s * (1 << i) for i in range(max(1, int(np.log2(B)))) for s in (1, -1)]
    eng = sf.Engine(11, depth, 50, 60, seed=5)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    v = np.random.default_rng(7).permutation(n) / n + 0.5 / n
    ct = eng.encrypt(eng.encode(np.concatenate([v, np.zeros(B - n)]), slots=B))
    out = sf.kway_sort(eng, ct, n, k, M, (3, 3, 2), depth)
    got = eng.decrypt(out, B)[:n]
    assert np.max(np.abs(got - np.sort(v))) < 0.01, got
    eng.close()


@pytest.mark.gpu
def test_n128_compare_swap_stages_ring17():
    """BASELINE config 4 shape: k=2, M=7, N=128 at the adapter's card (ring 2^17, depth 40, first 60 / scaling 59).
    The first two compare-swap stages (all that fit before the reference's first bootstrap) equal the plaintext
    network with an exact comparator within 0.01."""
    n, k, M = 128, 2, 7
    depth, scale, first, rots = sf.kway_params(n)
    eng = sf.Engine(17, depth, scale, first, seed=2)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    v = np.random.default_rng(3).permutation(n) / n
    out = sf.kway_sort(eng, eng.encrypt(eng.encode(v, slots=n)), n, k, M, (3, 3, 2), depth, 0, 2)
    want = kp.network_plain(v, k, M, last_stage=2)
    got = eng.decrypt(out, n)
    assert out.level <= depth
    assert np.max(np.abs(got - want)) < 0.01
    eng.close()


@pytest.mark.gpu
def test_n128_full_network_with_bootstrapping_ring17():
    """BASELINE config 4 end to end: the whole k = 2, M = 7 network on 128 values (28 stages, src/k-way/Sorter.cpp:284-404) at
    KWayAdapter's card (ring 2^17, depth 40, first 60 / scaling 59, level budget {4,4}: src/kway_adapter.h:41-64) with
    EvalBootstrap wherever checkLevelAndBoot / sign's lazy bootstrap ask for it (EvalUtils.cpp:57-86, sign.cpp:164-171);
    sign configuration of tests/k-way/KWaySort2Test.cpp:117-121,157; gate 0.01 (:214)."""
    n, k, M = 128, 2, 7
    depth, scale, first, rots = sf.kway_params(n)
    eng = sf.Engine(17, depth, scale, first, seed=2)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        eng.EvalBootstrapSetup((4, 4), (0, 0), n)
        eng.EvalBootstrapKeyGen(n)
        v = np.random.default_rng(3).permutation(n) * ((1.0 - 1e-8) / n)      # getVectorWithMinDiff(N, 0, 1, (1 - 1e-8) / N)
        eng.stats(reset=True)
        out = sf.kway_sort(eng, eng.encrypt(eng.encode(v, slots=n)), n, k, M, (3, 2, 3), depth)
        got = eng.decrypt(out, n)
        boots = sum(cnt for name, nl, cnt in eng.stats()["ops"] if name == "bootstrap")
        assert boots >= 10, boots
        assert np.max(np.abs(got - np.sort(v))) < 0.01, np.max(np.abs(got - np.sort(v)))
    finally:
        eng.close()
