"""Golden fixtures (tests/golden/, written by tests/golden/make_golden.py from the reference's own test sources and
from seeded oracle runs).  CPU part: the oracle and the host layer against the reference's known-answer vectors.
GPU part: the engine's limbs against the committed oracle digests (through the C-ABI)."""
import hashlib
import importlib
import json
import os

import numpy as np
import pytest

from oracle import algos
from oracle.ckks_oracle import Context, Params
from oracle.ckks_ops import OracleCC

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "reference_kat.json")))
LIMBS = json.load(open(os.path.join(HERE, "golden", "oracle_limbs.json")))
sf = importlib.import_module("sorting-fhe_b200")


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


# ------------------------------------------------------------------------------------------------ CPU
def test_oracle_compare_kat():
    """tests/CompareTest.cpp:44-62 (fixture depth 50 / scale 59 shrunk to ring 2^10, values scaled by 1/8 into
    sign's domain, which leaves the expected {0, 1, 0.5, 1} unchanged)"""
    k = KAT["compare"]
    cc = OracleCC(Context(Params(10, 29, 59)))
    cc.keygen(); cc.relin_keygen()
    a, b = np.array(k["a"]) / 8.0, np.array(k["b"]) / 8.0
    r = algos.compare(cc, cc.encrypt(cc.encode(a, slots=4)), cc.encrypt(cc.encode(b, slots=4)), *k["sign_cfg"])
    assert np.all(np.abs(cc.decrypt(r, 4) - np.array(k["expected"])) < k["tol"])


def test_oracle_compare_kat_as_written():
    """tests/CompareTest.cpp:13-62 verbatim: a = {1,5,3,4}, b = {2,4,3,3} unscaled (the differences {-1,1,0,1} lie on the
    boundary of sign's domain), ring 2^12, depth 50, scale 59, Sign(4,3,3) -> {0, 1, 0.5, 1} +- 0.1"""
    k = KAT["compare"]
    cc = OracleCC(Context(Params(12, 50, 59)))
    cc.keygen(); cc.relin_keygen()
    a, b = np.array(k["a"]), np.array(k["b"])
    r = algos.compare(cc, cc.encrypt(cc.encode(a, slots=4)), cc.encrypt(cc.encode(b, slots=4)), *k["sign_cfg"])
    assert np.all(np.abs(cc.decrypt(r, 4) - np.array(k["expected"])) < k["tol"])


@pytest.mark.parametrize("name", ["sign3", "sign4"])
def test_oracle_sign_kat(name):
    """tests/SignTest.cpp:41-80 (compositeSign<3>, dg=0 df=1) and :82-122 (compositeSign<4>, dg=df=3)"""
    k = KAT[name]
    n, dg, df = k["sign_cfg"]
    depth = 3 * (dg + df) + 3 if n == 3 else 5 * dg + 4 * df + 3
    cc = OracleCC(Context(Params(10, depth, 50)))
    cc.keygen(); cc.relin_keygen()
    x = np.array(k["input"])
    got = cc.decrypt(algos.composite_sign(cc, cc.encrypt(cc.encode(x, slots=8)), n, dg, df), 8)
    exp = np.array(k["expected"])
    if name == "sign3":     # the reference asserts sign agreement for every non-zero input and +-0.1 where |x| is large
        assert np.all(np.sign(got[x != 0]) == exp[x != 0])
        assert np.all(np.abs(got - exp)[np.abs(x) >= 0.2] < k["tol"]) and abs(got[4]) < k["tol"]
    else:
        assert np.all(np.abs(got - exp) < k["tol"]), got


def test_rotation_tree_rows():
    """tests/RotationTest.cpp:77-109: expected rows are plain cyclic shifts -- pins the rotation direction convention
    (positive index = left rotation) used by the oracle and by RotationComposer"""
    k = KAT["rotation_tree"]
    x = np.array(k["input"])
    for r, row in zip(k["rotations"], k["expected"]):
        assert np.array_equal(np.roll(x, -r), np.array(row))
    cc = OracleCC(Context(Params(10, 1, 50)))
    cc.keygen(); cc.rot_keygen([1, -1, 2, -2, 4])
    ct = cc.encrypt(cc.encode(x, slots=8))
    for r in (1, -1, 2, -2, 4):
        assert np.allclose(cc.decrypt(cc.EvalRotate(ct, r), 8), k["expected"][r + 4], atol=k["tol"])


def test_decompose_kat():
    """tests/DecomposeTest.cpp:64-74 through the product's C entry point"""
    k = KAT["decompose"]
    keys = [1, 2, 4, 8, 16, 32, 64]
    for r in k["numbers"]:
        for algo in k["algos"]:
            assert sum(sf.decompose(128, r, 128, keys, algo)) == r


def test_direct_sort_params_kat():
    """src/sort_algo.h:87-201 literal tables == the product's requiredRotations() / depth table (ring 2^17)"""
    for n, e in KAT["direct_sort_params"]["table"].items():
        if int(n) > 1024:
            continue
        depth, rots = sf.direct_sort_params(int(n), 1 << 17)
        assert depth == e["multDepth"], n
        assert sorted(set(rots)) == sorted(set(e["rotations"])), n
        if int(n) in algos.DIRECT_SORT_PARAMS:
            assert algos.DIRECT_SORT_PARAMS[int(n)] == (e["multDepth"], e["rotations"])


def test_kway_kat_plain():
    """tests/k-way/SortUtilsTest.cpp fcnL / two- / three- / four- / five-sorter literal vectors and MaskingTest getRotateDistance:
    the plaintext model of the k-way building blocks reproduces them"""
    from oracle import kway_plain
    for c in KAT["kway_sortutils"]["cases"]:
        v = {k: np.array(x) for k, x in c["vectors"].items()}
        if c["test"] == "FcnL":
            assert np.allclose(kway_plain.fcn_l(v["input1"], v["input2"], v["compValues"]), v["expected"])
        elif c["test"] == "TwoSorter":
            lo, hi = kway_plain.two_sorter(v["input1"], v["input2"], v["compValues"])
            assert np.allclose(lo, v["expectedMin"]) and np.allclose(hi, v["expectedMax"])
        elif c["test"] == "ThreeSorter":
            lo, mid, hi = kway_plain.three_sorter([v["input1"], v["input2"], v["input3"]], [v["comp1"], v["comp2"], v["comp3"]])
            assert np.allclose(lo, v["expectedMin"]) and np.allclose(mid, v["expectedMid"]) and np.allclose(hi, v["expectedMax"])
        elif c["test"] == "FourSorter":      # tests/k-way/SortUtilsTest.cpp:157-202
            got = kway_plain.four_sorter([v["input%d" % i] for i in range(1, 5)], [v["comp%d" % i] for i in range(1, 7)])
            assert np.allclose(np.array(got), v["expected"])
        elif c["test"] == "FiveSorter":      # tests/k-way/SortUtilsTest.cpp:204-255
            got = kway_plain.five_sorter(list(v["inputs"]), list(v["comps"]))
            assert np.allclose(np.array(got), v["expected"])
    assert {c["test"] for c in KAT["kway_sortutils"]["cases"]} >= {"FcnL", "TwoSorter", "ThreeSorter", "FourSorter", "FiveSorter"}
    for d in KAT["kway_masking"]["rotate_distance"]:
        assert kway_plain.get_rotate_distance(d["k"], d["logDist"], d["slope"]) == d["expected"]


def test_oracle_digests_unchanged():
    """the oracle still produces the committed limbs (guards the checker itself)"""
    g = LIMBS["ring12_depth6_seed1"]
    p = g["params"]
    cc = OracleCC(Context(Params(p["log_n"], p["depth"], p["scale"], p["first"], 0, p["ext"], p["seed"])))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen(p["rots"])
    x = np.array(g["input"])
    ct = cc.encrypt(cc.encode(x, slots=8))
    assert digest(ct.d) == g["digests"]["encrypt"]["sha256"]
    assert digest(cc.EvalMult(ct, ct).d) == g["digests"]["mult"]["sha256"]
    assert digest(cc.EvalRotate(ct, 1).d) == g["digests"]["rotate1"]["sha256"]


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_engine_matches_committed_digests():
    g = LIMBS["ring12_depth6_seed1"]
    p = g["params"]
    eng = sf.Engine(p["log_n"], p["depth"], p["scale"], p["first"], 0, p["ext"], p["seed"])
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(p["rots"])
    x = np.array(g["input"])
    ct = eng.encrypt(eng.encode(x, slots=8))
    seq = {"encrypt": ct}
    seq["mult"] = eng.EvalMult(ct, ct)
    seq["rotate1"] = eng.EvalRotate(ct, 1)
    seq["rotate_m1_of_mult"] = eng.EvalRotate(seq["mult"], -1)
    seq["mult_pt"] = eng.EvalMult(seq["mult"], eng.MakeCKKSPackedPlaintext(x, 1, 0, None, 8))
    seq["sub_const"] = eng.EvalSub(1.0, eng.EvalMult(ct, 0.5))
    seq["cube"] = eng.EvalMult(seq["mult"], ct)
    seq["sign3_0_1"] = sf.sign(eng, ct, (3, 0, 1))
    for k, v in seq.items():
        e = g["digests"][k]
        assert (v.level, v.deg) == (e["level"], e["deg"]), k
        assert digest(v.export()) == e["sha256"], k
    eng.close()


@pytest.mark.gpu
def test_engine_direct_sort_n4_matches_committed_digest():
    g = LIMBS["direct_sort_n4_ring10_seed1"]
    p = g["params"]
    eng = sf.Engine(p["log_n"], p["depth"], p["scale"], seed=p["seed"])
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(p["rots"])
    arr = np.array(g["input"])
    ct_in = eng.encrypt(eng.encode(arr, slots=4))     # encryption counter 0, as in make_golden.py
    out = sf.DirectSort(eng, 4, p["rots"]).sort(ct_in, tuple(p["sign_cfg"]))
    assert out.level == g["level"] and digest(out.export()) == g["sha256"]
    assert np.allclose(eng.decrypt(out, 4), g["decrypted"], atol=1e-12)
    assert np.max(np.abs(eng.decrypt(out, 4) - np.sort(arr))) < 0.01      # tests/DirectSortTest.cpp:194
    eng.close()


@pytest.mark.gpu
def test_engine_rotation_tree_kat():
    """tests/RotationTest.cpp:77-109 through RotationTree<8> on the GPU: the nine literal rows (1e-6), and every
    result bit-identical to the oracle's un-hoisted EvalRotate chain over the same NAF steps"""
    sf = importlib.import_module("sorting-fhe_b200")
    k = KAT["rotation_tree"]
    rots = [-1, -2, -4, 1, 2, 4]
    ctx = Context(Params(12, 2, 50))
    orc = OracleCC(ctx); orc.keygen(); orc.rot_keygen(rots)
    eng = sf.Engine(12, 2, 50); eng.KeyGen(); eng.EvalRotateKeyGen(rots)
    try:
        co = orc.encrypt(orc.encode(np.array(k["input"]), slots=8))
        ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
        # buildTree(start, end) decomposes with `end` as the wrap length (src/rotation.h:271-279) while treeRotate wraps at
        # the slot count: build up to 8 = slots so that both agree, and check the KAT's nine rows -4..4
        outs = sf.tree_rotate(eng, ce, 8, -4, 8, rots, "NAF")
        assert len(outs) == 13
        for r, out in zip(range(-4, 9), outs):
            row = k["expected"][r + 4] if r <= 4 else np.roll(k["input"], -r)
            assert np.allclose(eng.decrypt(out, 8), row, atol=k["tol"]), r
            want = co
            for step in sf.decompose(8, r, 8, rots, "NAF"):
                want = orc.EvalRotate(want, step)
            assert np.array_equal(out.export(), want.d), r
        # the reference's own call, buildTree(-4, 4): steps that are multiples of 4 were dropped at build time, so -4 / 4
        # are not in the tree and treeRotate hands back the input (:297-300, 323-327)
        quirk = sf.tree_rotate(eng, ce, 8, -4, 4, rots, "NAF")
        assert np.array_equal(quirk[0].export(), co.d) and np.array_equal(quirk[5].export(), outs[5].export())
    finally:
        eng.close()
