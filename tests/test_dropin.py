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


@pytest.mark.gpu
def test_reference_orchestration_on_the_engine_matches_host_layer(tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("tests/dropin/_ref/dropin_ref not built (needs /root/reference at build time)")
    env = dict(os.environ, OMP_NUM_THREADS="4", OMP_NESTED="TRUE")
    r = subprocess.run([BIN, str(tmp_path)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "dropin ok" in r.stdout, (r.stdout[-1500:], r.stderr[-1500:])
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
