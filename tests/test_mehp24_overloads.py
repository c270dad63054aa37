"""MEHP24 overloads that do not go through Comparison (SURVEY 8 A19 "other overloads"; reference
src/mehp24/mehp24_sort.cpp:23-47, 49-217, 219-249, 285-443, 607-623).  No reference fixture drives them
(tests/mehp24/Mehp24SortTest.cpp uses the Comparison-based ones), so the parameters below are this repo's:
  * compareAdv / indicatorAdv pipelines: bit-exact against the oracle twin (oracle/algos.py Mehp24.sort_fg_adv,
    mehp24_sort_large_adv) -- fixed polynomial constants, no interpolation involved;
  * Chebyshev-predicate `sort`: the interpolation coefficients are computed on the host (shim), so limbs are pinned
    against the reference's own code in tests/test_dropin.py; here: level budget (CPU, dry run) and sortedness."""
import importlib

import numpy as np
import pytest

from oracle import algos
from tests.common import make_pair, same

sf = importlib.import_module("sorting-fhe_b200")


def test_level_budgets_dry_run():
    """levels consumed on a metadata-only context: compareAdv (3, 2) + indicatorAdv (2, 2) end at 42, two blocks at 45,
    `sort` with degrees (59, 119) well below"""
    n = 4
    rots = list(sf.mehp24_params(n)[1]) + [n, -n]
    eng = sf.Engine(10, 50, 40, dry_run=True)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        ct = eng.encrypt(eng.encode(np.arange(n) / n, slots=n * n))
        assert sf.mehp24_sort_fg_adv(eng, ct, n, 3, 2, 2, 2).level == 42
        ct8 = eng.encrypt(eng.encode(np.arange(2 * n) / (2 * n), slots=n * n))
        assert sf.mehp24_sort_large_adv(eng, ct8, 2 * n, n, 3, 2, 2, 2).level == 45
        single = sf.mehp24_sort_cheb(eng, ct, n, -1.0, 1.0, 59, 119)
        multi = sf.mehp24_sort_cheb_multi(eng, sf.mehp24_split(eng, ct8, 2 * n, n), n, -1.0, 1.0, 119, 247)
        assert len(multi) == 2 and single.level <= 30 and multi[0].level == multi[1].level <= 32
        with pytest.raises(RuntimeError):
            sf.mehp24_sort_fg_adv(eng, ct, 6, 3, 2, 2, 2)              # not a power of two
    finally:
        eng.close()


@pytest.mark.gpu
def test_sort_fg_adv_bit_exact_vs_oracle():
    n = 4
    rots = algos.mehp24_rotation_indices(n)
    assert list(rots) == list(sf.mehp24_params(n)[1])
    eng, orc = make_pair(10, 42, rots=rots)
    try:
        x = np.array([0.5, 0.0, 0.75, 0.25])
        co = orc.encrypt(orc.encode(x, slots=n * n))
        ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
        got = sf.mehp24_sort_fg_adv(eng, ce, n, 3, 2, 2, 2)
        want = algos.Mehp24(orc, n).sort_fg_adv(co, 3, 2, 2, 2)
        same(got, want)
        assert got.level == 42
        assert np.max(np.abs(eng.decrypt(got, n * n)[:n] - np.sort(x))) < 1e-3
    finally:
        eng.close()


@pytest.mark.gpu
def test_sort_large_adv_two_blocks_bit_exact_vs_oracle():
    n = 4
    rots = sorted(set(list(algos.mehp24_rotation_indices(n)) + [n, -n]))
    eng, orc = make_pair(10, 45, rots=rots)
    try:
        x = np.random.default_rng(4).permutation(2 * n) / (2 * n)
        co = orc.encrypt(orc.encode(x, slots=n * n))
        ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
        got = sf.mehp24_sort_large_adv(eng, ce, 2 * n, n, 3, 2, 2, 2)
        want = algos.mehp24_sort_large_adv(orc, co, 2 * n, n, 3, 2, 2, 2)
        same(got, want)
        assert np.max(np.abs(eng.decrypt(got, n * n)[:2 * n] - np.sort(x))) < 1e-3
    finally:
        eng.close()


@pytest.mark.gpu
def test_sort_cheb_sorts_into_column_zero():
    n = 4
    rots = list(sf.mehp24_params(n)[1]) + [n, -n]
    eng = sf.Engine(10, 32, 50, seed=2)
    try:
        eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
        x = np.array([0.75, 0.25, 0.0, 0.5])
        out = sf.mehp24_sort_cheb(eng, eng.encrypt(eng.encode(x, slots=n * n)), n, -1.0, 1.0, 59, 119)
        col0 = eng.decrypt(out, n * n).reshape(n, n)[:, 0]
        assert np.max(np.abs(col0 - np.sort(x))) < 0.1, col0
        x8 = np.random.default_rng(6).permutation(2 * n) / (2 * n)
        parts = sf.mehp24_split(eng, eng.encrypt(eng.encode(x8, slots=n * n)), 2 * n, n)
        outs = sf.mehp24_sort_cheb_multi(eng, parts, n, -1.0, 1.0, 119, 247)      # degrees (59, 119) leave 0.14 at spacing 1/8
        both = np.concatenate([eng.decrypt(o, n * n).reshape(n, n)[:, 0] for o in outs])
        assert np.max(np.abs(both - np.sort(x8))) < 0.1, both
    finally:
        eng.close()
