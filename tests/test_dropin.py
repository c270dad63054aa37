"""Drop-in boundary: the reference's UNMODIFIED sources (src/sort_algo.h, sign.cpp, comparison.cpp, encryption.cpp,
mehp24/*.cpp, k-way/*.cpp) compiled against sorting-fhe_b200/host/compat/openfhe.h and libckks_b200.so.

* CPU (build container, where /root/reference exists): every source passes the compiler and the driver links.
* GPU: tests/dropin/_ref/dropin_ref (built here by tests/dropin/build.sh, travels with the snapshot) runs the
  reference's own orchestration -- DirectSort<N>::sort with its `#pragma omp parallel for` regions live
  (src/sort_algo.h:334,349,425,713), mehp24::sortFG, Comparison::compare -- on the engine; the limbs it writes must equal
  what the re-expressed host layer (sorting-fhe_b200/host/) produces on the same seed.  That pins the host layer to the
  reference's code rather than to the oracle's restatement of it, and exercises the thread-safe context."""
import importlib
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
BIN = os.path.join(ROOT, "tests", "dropin", "_ref", "dropin_ref")
sf = importlib.import_module("sorting-fhe_b200")


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not present")
def test_reference_sources_compile_unmodified():
    inc = ["-I", os.path.join(ROOT, "sorting-fhe_b200", "host", "compat"), "-I", REF + "/src", "-I", REF + "/src/k-way",
           "-I", REF + "/src/mehp24"]
    srcs = ["sign.cpp", "comparison.cpp", "encryption.cpp", "mehp24/mehp24_utils.cpp", "mehp24/mehp24_sort.cpp",
            "k-way/EvalUtils.cpp", "k-way/Masking.cpp", "k-way/SortUtils.cpp", "k-way/Sorter.cpp"]
    for s in srcs:
        r = subprocess.run(["g++", "-std=c++17", "-fopenmp", "-fsyntax-only", "-w"] + inc + [os.path.join(REF, "src", s)],
                           capture_output=True, text=True)
        assert r.returncode == 0, (s, r.stderr[-2000:])
    r = subprocess.run(["bash", os.path.join(ROOT, "tests", "dropin", "build.sh")], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(BIN), r.stderr[-2000:]


def _input(n, seed):
    return np.array([((i * (2 * seed + 3) + seed) % n) / n for i in range(n)])


def _load(path):
    raw = np.fromfile(path, dtype=np.uint64)
    parts, nl, level, deg, slots, ring = [int(v) for v in raw[:6].view(np.int64)]
    return raw[6:].reshape(parts, nl, ring), level, deg, slots


@pytest.fixture(scope="module")
def dropin_out(tmp_path_factory):
    """one run of the reference-built driver; every test below reads its dumps"""
    if not os.path.exists(BIN):
        pytest.skip("tests/dropin/_ref/dropin_ref not built (needs /root/reference at build time)")
    out = tmp_path_factory.mktemp("dropin")
    env = dict(os.environ, OMP_NUM_THREADS="4", OMP_NESTED="TRUE")
    r = subprocess.run([BIN, str(out)], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0 and "dropin ok" in r.stdout, (r.stdout[-1500:], r.stderr[-1500:])
    return out


@pytest.mark.gpu
def test_reference_mehp24_overloads_match_host_layer(dropin_out):
    """SURVEY 8 A19 "other overloads": the reference's own mehp24::sortFG (compareAdv, mehp24_sort.cpp:219-249),
    sortLargeArrayFG over two blocks (:285-443, 607-623) and mehp24::sort with the Chebyshev predicates (:23-47, 49-217),
    compiled unmodified and run on the engine, against this repo's host layer on the same seed: identical limbs."""
    n = 4
    rots = list(sf.mehp24_params(n)[1]) + [n, -n]
    eng = sf.Engine(10, 45, 40, seed=1)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        x4, x8 = _input(n, 2), _input(2 * n, 3)
        ct = eng.encrypt(eng.encode(x4, slots=n * n))                      # encryption counter 0, as in the driver
        adv = sf.mehp24_sort_fg_adv(eng, ct, n, 3, 2, 2, 2)
        cheb = sf.mehp24_sort_cheb(eng, ct, n, -1.0, 1.0, 59, 119)
        ct8 = eng.encrypt(eng.encode(x8, slots=n * n))                     # counter 1
        large = sf.mehp24_sort_large_adv(eng, ct8, 2 * n, n, 3, 2, 2, 2)
        multi = sf.mehp24_sort_cheb_multi(eng, sf.mehp24_split(eng, ct8, 2 * n, n), n, -1.0, 1.0, 119, 247)
        for name, got in (("mehp24_adv_4", adv), ("mehp24_cheb_4", cheb), ("mehp24_large_adv_8", large),
                          ("mehp24_cheb_multi_8_0", multi[0]), ("mehp24_cheb_multi_8_1", multi[1])):
            limbs, level, deg, slots = _load(os.path.join(str(dropin_out), name + ".bin"))
            assert (level, deg) == (got.level, got.deg), name
            assert np.array_equal(limbs, got.export()), name + ": reference orchestration and host layer differ"
        assert np.max(np.abs(eng.decrypt(adv, n * n)[:n] - np.sort(x4))) < 0.01
        assert np.max(np.abs(eng.decrypt(large, n * n)[:2 * n] - np.sort(x8))) < 0.01
        # mehp24::sort leaves the sorted values in column 0 of the matrix (no masking, no transposition)
        col0 = eng.decrypt(cheb, n * n).reshape(n, n)[:, 0]
        assert np.max(np.abs(col0 - np.sort(x4))) < 0.1, col0
        both = np.concatenate([eng.decrypt(m, n * n).reshape(n, n)[:, 0] for m in multi])
        assert np.max(np.abs(both - np.sort(x8))) < 0.1, both
    finally:
        eng.close()


@pytest.mark.gpu
def test_reference_orchestration_on_the_engine_matches_host_layer(dropin_out):
    tmp_path = dropin_out
    # ---- src/sort.h's SortContext<4> through files (host/serial.h): same values as the in-memory sort, and sorted
    rows = np.loadtxt(os.path.join(str(tmp_path), "serial.txt"))
    assert rows.shape == (4, 3)
    assert np.max(np.abs(rows[:, 1] - rows[:, 2])) < 1e-4, rows      # fresh encryption randomness in the imported context: values, not limbs
    assert np.max(np.abs(rows[:, 2] - np.sort(rows[:, 0]))) < 0.01, rows
    # ---- DirectSort<4>, <8>
    for n in (4, 8):
        depth, rots = sf.direct_sort_params(n, 1 << 10)
        eng = sf.Engine(10, depth, 40, seed=1)
        try:
            eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
            x = _input(n, 1)
            out = sf.DirectSort(eng, n, rots).sort(eng.encrypt(eng.encode(x, slots=n)), (3, 2, 2))
            limbs, level, deg, slots = _load(os.path.join(str(tmp_path), "direct_sort_%d.bin" % n))
            assert (level, deg) == (out.level, out.deg) and level == depth
            assert np.array_equal(limbs, out.export()), "DirectSort<%d>: reference orchestration and host layer differ" % n
            assert np.max(np.abs(eng.decrypt(out, n) - np.sort(x))) < 0.01
        finally:
            eng.close()
    # ---- mehp24::sortFG N = 4
    n = 4
    depth, rots = sf.mehp24_params(n)
    eng = sf.Engine(10, depth, 40, seed=1)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        x = _input(n, 2)
        out = sf.mehp24_sort_fg(eng, eng.encrypt(eng.encode(x, slots=n * n)), n, (3, 2, 2))
        limbs, level, deg, slots = _load(os.path.join(str(tmp_path), "mehp24_4.bin"))
        assert np.array_equal(limbs, out.export()), "mehp24::sortFG: reference orchestration and host layer differ"
    finally:
        eng.close()
    # ---- CompareTest
    eng = sf.Engine(12, 50, 59, seed=1)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen()
        ca = eng.encrypt(eng.encode(np.array([1.0, 5.0, 3.0, 4.0]), slots=4))
        cb = eng.encrypt(eng.encode(np.array([2.0, 4.0, 3.0, 3.0]), slots=4))
        out = sf.compare(eng, ca, cb, (4, 3, 3))
        limbs, level, deg, slots = _load(os.path.join(str(tmp_path), "compare.bin"))
        assert np.array_equal(limbs, out.export())
        assert np.all(np.abs(eng.decrypt(out, 4) - np.array([0, 1, 0.5, 1])) < 0.1)
    finally:
        eng.close()
