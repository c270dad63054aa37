"""Full-size limb parity (`-m gpu`): the engine must reproduce, bit for bit, the limbs the CPU oracle produced for the
BASELINE configurations at ring 2^17 -- digests committed in tests/golden/fullsize_limbs.json by
tests/golden/make_fullsize.py (a complete DirectSort<128>::sort takes the oracle minutes, so it is run once in the build
container, not on the GPU box).  This is the only place where k_conv_fp<11,1>, <8,2>, the two-pass 2^17 NTT with
on-the-fly twiddles, the side-stream plaintext morphing and the plaintext cache all run together against the oracle.
Reference shapes: tests/DirectSortTest.cpp:29-44,101-102,116-117,194; tests/CompareTest.cpp:13-62;
tests/mehp24/Mehp24SortTest.cpp:36-60."""
import hashlib
import importlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
sf = importlib.import_module("sorting-fhe_b200")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_limbs.json")


def gold(name):
    with open(GOLD) as f:
        d = json.load(f)
    if name not in d:
        pytest.skip("no committed digest for %s (run tests/golden/make_fullsize.py %s)" % (name, name))
    return d[name]


def check(ct, g, what):
    a = ct.export()
    assert list(a.shape) == g["shape"], (what, a.shape, g["shape"])
    assert (ct.level, ct.deg, ct.slots) == (g["level"], g["deg"], g["slots"]), what
    assert float(ct.sf).hex() == g["sf_hex"], (what, float(ct.sf).hex(), g["sf_hex"])
    assert [int(v) for v in a.reshape(-1)[:4]] == g["head"], (what, "first words differ")
    assert hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest() == g["sha256"], what + ": limb digest differs from the oracle's"


@pytest.mark.parametrize("name", ["direct_sort_n128", "direct_sort_n16"])
def test_direct_sort_fullsize_limbs(name):
    g = gold(name)
    p = g["params"]
    n, cfg = p["n"], tuple(p["sign_cfg"])
    depth, rots = sf.direct_sort_params(n, 1 << p["log_n"])
    assert depth == p["depth"] and list(rots) == p["rots"]
    eng = sf.Engine(p["log_n"], depth, p["scale"], p["first"], 0, p["ext"], p["seed"])
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        x = np.random.default_rng(1).permutation(n) / n
        ct = eng.encrypt(eng.encode(x, slots=n))                  # encryption counter 0, as in the oracle run
        check(ct, g["digests"]["input"], "input")
        sorter = sf.DirectSort(eng, n, rots)
        for rep in range(2):                                      # second pass: every mask comes from the plaintext cache
            eng.set_encryption_counter(1)                         # the sorter's zero cache is encryption 1 in the oracle run
            rank = sorter.construct_rank(ct, cfg)
            check(rank, g["digests"]["rank"], "rank (pass %d)" % rep)
            eng.set_encryption_counter(1)
            out = sorter.rotation_index_check(rank, ct)
            check(out, g["digests"]["output"], "output (pass %d)" % rep)
        got = eng.decrypt(out, n)
        assert np.array_equal(got, np.array(g["decrypted"]["output"]))
        assert np.max(np.abs(got - np.sort(x))) < 1e-3
        # tests/DirectSortNTest.cpp:124: every rank within 1e-4 of the integer rank
        ranks = eng.decrypt(rank, n)
        assert np.max(np.abs(ranks - np.argsort(np.argsort(x)))) < 1e-4
        e, b, h, m = eng.plaintext_cache_stats()
        assert h >= m > 0, (e, b, h, m)
    finally:
        eng.close()


def test_compare_fullslot_limbs():
    """BASELINE config 2: Comparison::compare with Sign(4,3,3) on all 65 536 slots"""
    g = gold("compare_fullslot")
    p = g["params"]
    eng = sf.Engine(p["log_n"], p["depth"], p["scale"], p["first"], 0, p["ext"], p["seed"])
    try:
        eng.KeyGen(); eng.EvalMultKeyGen()
        slots = p["slots"]
        a = np.random.default_rng(2).permutation(slots) / slots
        b = np.roll(a, 1)
        ca, cb = eng.encrypt(eng.encode(a, slots=slots)), eng.encrypt(eng.encode(b, slots=slots))
        check(ca, g["digests"]["a"], "a"); check(cb, g["digests"]["b"], "b")
        out = sf.compare(eng, ca, cb, tuple(p["sign_cfg"]))
        check(out, g["digests"]["output"], "compare output")
        got = eng.decrypt(out, slots)
        assert np.max(np.abs(got - (a > b))) < 0.1                 # tests/CompareTest.cpp:58-61 gate
    finally:
        eng.close()


def test_mehp24_n256_limbs():
    """BASELINE config 5a: mehp24 sortFG N=256 in one 65 536-slot ciphertext, depth 49"""
    g = gold("mehp24_n256")
    p = g["params"]
    n = p["n"]
    depth, rots = sf.mehp24_params(n)
    assert depth == p["depth"] and list(rots) == p["rots"]
    eng = sf.Engine(p["log_n"], depth, p["scale"], p["first"], 0, p["ext"], p["seed"])
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        x = np.random.default_rng(3).permutation(n) / n
        ct = eng.encrypt(eng.encode(x, slots=n * n))
        check(ct, g["digests"]["input"], "input")
        out = sf.mehp24_sort_fg(eng, ct, n, tuple(p["sign_cfg"]))
        check(out, g["digests"]["output"], "sortFG output")
        assert np.max(np.abs(eng.decrypt(out, n * n)[:n] - np.sort(x))) < 0.01
    finally:
        eng.close()
