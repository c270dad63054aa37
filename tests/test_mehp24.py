"""MEHP24 ranking sort (reference src/mehp24/, tests/mehp24/Mehp24SortTest.cpp): oracle restatement on CPU, level
budgets of the host layer in dry-run mode, and the CUDA path bit-exact against the oracle / within the reference's
0.01 gate at the BASELINE size N=256."""
import importlib

import numpy as np
import pytest

from oracle import algos
from oracle.ckks_oracle import Context, Params
from oracle.ckks_ops import OracleCC
from tests.common import make_pair, same

sf = importlib.import_module("sorting-fhe_b200")
SIGN = {4: (3, 2, 2), 8: (3, 2, 2), 16: (3, 2, 2), 32: (3, 3, 2), 64: (3, 3, 2), 128: (3, 3, 2), 256: (3, 4, 2)}


def dg_i(n):
    return (int(np.log2(n)) + 1) // 2      # tests/mehp24/Mehp24SortTest.cpp:137


def test_rotation_indices_match_reference_formula():
    """getRotationIndices (mehp24_utils.cpp:186-213): product == oracle restatement, and every rotation sortFG issues
    has a key"""
    for n in (4, 8, 64, 256):
        depth, rots = sf.mehp24_params(n)
        assert rots == algos.mehp24_rotation_indices(n)
        lg = int(np.log2(n))
        need = {-(1 << (lg + i)) for i in range(lg)} | {-(1 << i) for i in range(lg)} | {1 << i for i in range(lg)}
        need |= {-(n * (n - 1) // (1 << i)) for i in range(1, lg + 1)} | {n * (n - 1) // (1 << i) for i in range(1, lg + 1)}
        assert need <= set(rots)
    assert [sf.mehp24_params(n)[0] for n in (4, 8, 16, 32, 64, 128, 256)] == [31, 35, 35, 42, 42, 46, 49]


def test_oracle_sort_fg_n4():
    """sortFG N=4 on ring 2^10 ends exactly at the fixture's multDepth 31 and sorts within 0.01"""
    n = 4
    cc = OracleCC(Context(Params(10, 31, 40)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen(algos.mehp24_rotation_indices(n))
    x = np.random.default_rng(5).permutation(n) / n
    out = algos.Mehp24(cc, n).sort_fg(cc.encrypt(cc.encode(x, slots=n * n)), SIGN[n], dg_i(n), 2)
    assert out.level == 31
    assert np.max(np.abs(cc.decrypt(out, n * n)[:n] - np.sort(x))) < 0.01


@pytest.mark.parametrize("n", [4, 16, 64, 256])
def test_level_budget_dry_run(n):
    """the fixture's depth table leaves no slack: sortFG must end at level == multDepth (metadata-only context)"""
    depth, rots = sf.mehp24_params(n)
    eng = sf.Engine(17, depth, 40, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    ct = eng.encrypt(eng.encode(np.arange(n) / n, slots=n * n))
    out = sf.mehp24_sort_fg(eng, ct, n, SIGN[n])
    assert out.level == depth
    ops = {}
    for name, nl, cnt in eng.stats()["ops"]:
        ops[name] = ops.get(name, 0) + cnt
    assert ops["rotate"] == 6 * int(np.log2(n))          # 6 log-step folds (~48 rotations at N=256)
    eng.close()


@pytest.mark.gpu
def test_matrix_ops_bit_exact():
    n = 4
    rots = algos.mehp24_rotation_indices(n)
    eng, orc = make_pair(10, 6, rots=rots)
    x = np.arange(1, n + 1) / 8.0
    co = orc.encrypt(orc.encode(x, slots=n * n))
    ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
    m = algos.Mehp24(orc, n)
    same(sf.mehp24_matrix_op(eng, ce, n, "replicateRow"), m.replicate_row(co))
    same(sf.mehp24_matrix_op(eng, ce, n, "transposeRow", True), m.transpose_row(co, True))
    vr = sf.mehp24_matrix_op(eng, ce, n, "replicateRow")
    same(sf.mehp24_matrix_op(eng, vr, n, "sumColumns", True), m.sum_columns(m.replicate_row(co), True))
    same(sf.mehp24_matrix_op(eng, vr, n, "sumRows"), m.sum_rows(m.replicate_row(co)))
    got = eng.decrypt(sf.mehp24_matrix_op(eng, ce, n, "replicateRow"), n * n)
    assert np.allclose(got, np.tile(x, n), atol=1e-7)
    eng.close()


@pytest.mark.gpu
def test_sort_fg_n4_bit_exact():
    """whole sortFG pipeline, every limb identical to the oracle (ring 2^10, fixture depth 31)"""
    n = 4
    depth, rots = sf.mehp24_params(n)
    eng, orc = make_pair(10, depth, rots=rots)
    x = np.random.default_rng(5).permutation(n) / n
    co = orc.encrypt(orc.encode(x, slots=n * n))
    ce = eng.encrypt(eng.encode(x, slots=n * n))
    same(ce, co)
    ro = algos.Mehp24(orc, n).sort_fg(co, SIGN[n], dg_i(n), 2)
    re = sf.mehp24_sort_fg(eng, ce, n, SIGN[n])
    same(re, ro)
    assert re.level == depth
    assert np.max(np.abs(eng.decrypt(re, n * n)[:n] - np.sort(x))) < 0.01
    eng.close()


@pytest.mark.gpu
def test_sort_fg_n256_ring17_properties():
    """BASELINE config 5 shape: MEHP24 sortFG N=256 (65 536 slots, ring 2^17, depth 49, Sign(3,4,2)); reference gate
    max error < 0.01 (tests/mehp24/Mehp24SortTest.cpp:207), output is the sorted permutation of the input"""
    n = 256
    depth, rots = sf.mehp24_params(n)
    eng = sf.Engine(17, depth, 40, seed=3)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    x = np.random.default_rng(11).permutation(n) / n
    out = sf.mehp24_sort_fg(eng, eng.encrypt(eng.encode(x, slots=n * n)), n, SIGN[n])
    got = eng.decrypt(out, n * n)[:n]
    assert out.level == depth
    assert np.max(np.abs(got - np.sort(x))) < 0.01
    assert np.array_equal(np.round(got * n).astype(int), np.arange(n))
    eng.close()


def test_large_array_level_budget_dry_run():
    """sortLargeArrayFG at the fixture's N = 512 card (depth 57, Sign(3,4,2), two 256 x 256 blocks, ring 2^17,
    tests/mehp24/Mehp24SortTest.cpp:62-64,150-157) stays inside the depth; metadata-only run"""
    depth, rots = sf.mehp24_params(512)
    assert depth == 57 and 256 in rots and -256 in rots
    eng = sf.Engine(17, depth, 40, 60, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    ct = eng.encrypt(eng.encode(np.arange(512) / 512, slots=65536))
    out = sf.mehp24_sort_large(eng, ct, 512, 256, (3, 4, 2))
    assert out.level <= depth
    eng.close()


@pytest.mark.gpu
def test_sort_large_array_bit_exact_small():
    """mehp24::sortLargeArrayFG (split -> multi-ciphertext sortFG -> combine, src/mehp24/mehp24_sort.cpp:445-645) on 8 values
    as two 4 x 4 blocks: every limb identical to the oracle restatement, decrypted output sorted within 0.01"""
    total, sub = 8, 4
    rots = sorted(set(sf.mehp24_params(sub)[1] + [sub, -sub]))
    eng, orc = make_pair(10, 40, rots=rots)
    x = np.random.default_rng(8).permutation(total) / total
    co = orc.encrypt(orc.encode(x, slots=sub * sub))
    ce = eng.encrypt(eng.encode(x, slots=sub * sub))
    d = int((np.log2(total) + 1) / 2)
    ro = algos.mehp24_sort_large(orc, co, total, sub, (3, 2, 2), d, 2)
    re = sf.mehp24_sort_large(eng, ce, total, sub, (3, 2, 2), d, 2)
    same(re, ro)
    assert np.max(np.abs(eng.decrypt(re, sub * sub)[:total] - np.sort(x))) < 0.01
    eng.close()
