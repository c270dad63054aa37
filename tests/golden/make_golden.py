#!/usr/bin/env python
"""Regenerates the committed fixtures under tests/golden/ (run in the build container, where /root/reference exists).

  python tests/golden/make_golden.py

1. reference_kat.json -- the literal known-answer vectors of the reference's own tests, read out of the C++
   sources with their file:line (nothing is typed by hand): CompareTest, SignTest, RotationTest, DecomposeTest,
   k-way Masking / SortUtils tests, and DirectSort<N>::getSizeParameters' depth / rotation-key tables.
2. oracle_limbs.json -- SHA-256 of the oracle's ciphertext limbs for a handful of seeded op sequences on small rings.
   The GPU parity tests compare the engine's limbs both with the live oracle and with these committed digests, so an
   accidental change of the oracle itself is caught too.

The reference cannot be executed (it needs OpenFHE 1.1.4, absent here: SURVEY 8c), so (1) is what its tests pin.
"""
import hashlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def src(path):
    with open(os.path.join(REF, path)) as f:
        return f.read().split("\n")


def brace_list(lines, start, name):
    """parse `name = {...};` (possibly multi-line, possibly nested one level) starting at 1-based line `start`"""
    text, i = "", start - 1
    assert name in lines[i], (name, lines[i])
    while True:
        text += re.sub(r"//.*", "", lines[i]) + " "
        if ";" in lines[i]:
            break
        i += 1
    body = text[text.index("{"):text.rindex("}") + 1]
    py = body.replace("{", "[").replace("}", "]")
    py = re.sub(r",\s*\]", "]", py)   # trailing commas
    return json.loads(py), "%d-%d" % (start, i + 1)


def find(lines, pattern, after=1):
    for i in range(after - 1, len(lines)):
        if re.search(pattern, lines[i]):
            return i + 1
    raise KeyError(pattern)


def reference_kat():
    out = {}
    # ---- tests/CompareTest.cpp
    L = src("tests/CompareTest.cpp")
    a, la = brace_list(L, find(L, r"std::vector<double> a = "), "a =")
    b, lb = brace_list(L, find(L, r"std::vector<double> b = "), "b =")
    e, le = brace_list(L, find(L, r"std::vector<double> expected = "), "expected")
    cfg = re.search(r"CompositeSignConfig\((\d+), (\d+), (\d+)\)", "\n".join(L)).groups()
    tol = float(re.search(r"EXPECT_NEAR\(resultValues\[i\], expected\[i\], ([0-9.]+)\)", "\n".join(L)).group(1))
    out["compare"] = {"cite": "tests/CompareTest.cpp:%s,%s,%s" % (la, lb, le), "a": a, "b": b, "expected": e,
                      "sign_cfg": [int(x) for x in cfg], "tol": tol}
    # ---- tests/SignTest.cpp
    L = src("tests/SignTest.cpp")
    t1 = find(L, r"TEST_F\(ArraySortTest, CompositeSignTest\)")
    i1, c1 = brace_list(L, find(L, r"std::vector<double> input = ", t1), "input")
    e1, c1e = brace_list(L, find(L, r"std::vector<double> expected = ", t1), "expected")
    t2 = find(L, r"TEST_F\(ArraySortTest, VerySmallElementsTest\)")
    i2, c2 = brace_list(L, find(L, r"std::vector<double> input = ", t2), "input")
    e2, c2e = brace_list(L, find(L, r"std::vector<double> expected = ", t2), "expected")
    out["sign3"] = {"cite": "tests/SignTest.cpp:%s,%s" % (c1, c1e), "input": i1, "expected": e1, "sign_cfg": [3, 0, 1], "tol": 0.1}
    out["sign4"] = {"cite": "tests/SignTest.cpp:%s,%s" % (c2, c2e), "input": i2, "expected": e2, "sign_cfg": [4, 3, 3], "tol": 0.1}
    # ---- tests/RotationTest.cpp (RotationTree over -4..4)
    L = src("tests/RotationTest.cpp")
    t = find(L, r"RotateTreeVector")
    inp, ci = brace_list(L, find(L, r"std::vector<double> input = ", t), "input")
    rows, cr = brace_list(L, find(L, r"expectedResults = ", t), "expectedResults")
    out["rotation_tree"] = {"cite": "tests/RotationTest.cpp:%s,%s" % (ci, cr), "input": inp, "rotations": list(range(-4, 5)),
                            "expected": rows, "tol": 1e-6}
    # ---- tests/DecomposeTest.cpp
    L = src("tests/DecomposeTest.cpp")
    nums, cn = brace_list(L, find(L, r"std::vector<int> testNumbers = "), "testNumbers")
    out["decompose"] = {"cite": "tests/DecomposeTest.cpp:%s" % cn, "numbers": nums, "algos": ["NAF", "BNAF", "BINARY"]}
    # ---- tests/k-way/MaskingTest.cpp
    L = src("tests/k-way/MaskingTest.cpp")
    dist = []
    for i, ln in enumerate(L):
        m = re.search(r"EXPECT_EQ\(getRotateDistance\((\d+), (\d+), (\d+)\), (\d+)\)", ln)
        if m:
            dist.append({"line": i + 1, "k": int(m.group(1)), "logDist": int(m.group(2)), "slope": int(m.group(3)),
                         "expected": int(m.group(4))})
    t = find(L, r"TEST_F\(MaskingTest, GenMask\)")
    idx, cidx = brace_list(L, find(L, r"indices = ", t), "indices")
    out["kway_masking"] = {"cite": "tests/k-way/MaskingTest.cpp:%s + getRotateDistance lines" % cidx, "rotate_distance": dist,
                           "gen_mask": {"indices": idx, "args": [1, 1], "ones_at": [0, 4]}}
    # ---- tests/k-way/SortUtilsTest.cpp (fcnL, two/three-sorter literal vectors)
    L = src("tests/k-way/SortUtilsTest.cpp")
    cases, pos = [], 1
    for name in ("FcnL", "TwoSorter", "ThreeSorter", "FourSorter", "FiveSorter"):
        try:
            t = find(L, r"TEST_F\(\w+, \w*%s\w*\)" % name)
        except KeyError:
            continue
        end = find(L, r"^}", t)
        vecs = {}
        for i in range(t, end):
            m = re.search(r"std::vector<(?:std::vector<)?double>+ (\w+) = \{", L[i - 1])
            if m:
                try:
                    vecs[m.group(1)], _ = brace_list(L, i, m.group(1))
                except Exception:
                    pass                      # lists of identifiers (inputs = {input1, ...}) are not literal data
        cases.append({"test": name, "lines": "%d-%d" % (t, end), "vectors": vecs})
    out["kway_sortutils"] = {"cite": "tests/k-way/SortUtilsTest.cpp", "cases": cases, "tol": 0.1}
    # ---- src/sort_algo.h getSizeParameters
    L = src("src/sort_algo.h")
    params, i = {}, find(L, r"static void getSizeParameters")
    end = find(L, r"parameters\.SetMultiplicativeDepth", i)
    text = "\n".join(L[i - 1:end])
    for m in re.finditer(r"case (\d+):(.*?)break;", text, re.S):
        depth = int(re.search(r"multDepth = (\d+)", m.group(2)).group(1))
        rots = [int(x) for x in re.findall(r"-?\d+", re.search(r"rotations = \{(.*?)\}", m.group(2), re.S).group(1))]
        params[m.group(1)] = {"multDepth": depth, "rotations": rots}
    out["direct_sort_params"] = {"cite": "src/sort_algo.h:%d-%d" % (i, end), "table": params}
    return out


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint64).tobytes()).hexdigest()


def oracle_limbs():
    from oracle import algos
    from oracle.ckks_oracle import Context, Params
    from oracle.ckks_ops import OracleCC
    out = {}
    cc = OracleCC(Context(Params(12, 6, 40, 60, 0, True, 1)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen([1, -1, 3, 4])
    x = np.array([0.5, -0.3, 0.1, -0.7, 0.0, 0.8, -0.9, 0.2])
    ct = cc.encrypt(cc.encode(x, slots=8))
    seq = {"encrypt": ct}
    seq["mult"] = cc.EvalMult(ct, ct)
    seq["rotate1"] = cc.EvalRotate(ct, 1)
    seq["rotate_m1_of_mult"] = cc.EvalRotate(seq["mult"], -1)
    seq["mult_pt"] = cc.EvalMult(seq["mult"], cc.MakeCKKSPackedPlaintext(x, 1, 0, None, 8))
    seq["sub_const"] = cc.EvalSub(1.0, cc.EvalMult(ct, 0.5))
    seq["cube"] = cc.EvalMult(seq["mult"], ct)
    seq["sign3_0_1"] = algos.composite_sign(cc, ct, 3, 0, 1)
    out["ring12_depth6_seed1"] = {
        "params": {"log_n": 12, "depth": 6, "scale": 40, "first": 60, "ext": True, "seed": 1, "rots": [1, -1, 3, 4]},
        "input": x.tolist(),
        "digests": {k: {"sha256": digest(v.d), "level": int(v.level), "deg": int(v.deg), "limbs": int(v.d.shape[1])}
                    for k, v in seq.items()}}
    # DirectSort N=4 on ring 2^10, whole pipeline
    n = 4
    depth, rots = algos.DIRECT_SORT_PARAMS[n]
    cc = OracleCC(Context(Params(10, depth, 40, 60, 0, True, 1)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen(rots)
    arr = np.random.default_rng(n).permutation(n) / n
    ct_in = cc.encrypt(cc.encode(arr, slots=n))       # encryption counter 0; the sorter's zero cache takes counter 1
    res = algos.DirectSort(cc, n, rots).sort(ct_in, (3, 2, 2))
    out["direct_sort_n4_ring10_seed1"] = {
        "params": {"log_n": 10, "depth": depth, "scale": 40, "seed": 1, "rots": rots, "sign_cfg": [3, 2, 2]},
        "input": arr.tolist(), "sha256": digest(res.d), "level": int(res.level),
        "decrypted": [float(v) for v in cc.decrypt(res, n)]}
    return out


if __name__ == "__main__":
    with open(os.path.join(HERE, "reference_kat.json"), "w") as f:
        json.dump(reference_kat(), f, indent=1)
    with open(os.path.join(HERE, "oracle_limbs.json"), "w") as f:
        json.dump(oracle_limbs(), f, indent=1)
    print("wrote reference_kat.json, oracle_limbs.json")
