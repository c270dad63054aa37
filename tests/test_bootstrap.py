"""CKKS bootstrapping: EvalBootstrapSetup / EvalBootstrapKeyGen / EvalBootstrap.
CPU: the oracle's restatement refreshes a ciphertext (tests/k-way/EvalUtilsTest.cpp:132-158 shape).  GPU: the three new
primitives (complex encode, conjugation, ModRaise) and the whole bootstrap are bit-exact against the oracle; the
reference's round-trip KAT holds; the precomputed constants of the two sides are identical doubles."""
import importlib

import numpy as np
import pytest

from oracle.ckks_oracle import Context, Params
from oracle.ckks_ops import OracleCC
from tests.common import make_pair, same

sf = importlib.import_module("sorting-fhe_b200")


def test_oracle_bootstrap_round_trip():
    """EvalUtilsTest.Bootstrapping: {0.1,0.2,0.3,0.4} encoded at level 40 of depth 50 (scale 59, ring 2^12, 16 slots,
    level budget {3,3}) comes back within 0.2 at a lower level"""
    cc = OracleCC(Context(Params(12, 50, 59, 60, 0, True, 1)))
    cc.keygen(); cc.relin_keygen()
    cc.EvalBootstrapSetup((3, 3), (0, 0), 16)
    cc.EvalBootstrapKeyGen(16)
    x = np.array([0.1, 0.2, 0.3, 0.4])
    ct = cc.encrypt(cc.encode(x, 1, 40, 16))
    out = cc.EvalBootstrap(ct)
    assert out.level < 40
    got = cc.decrypt(out, 4)
    assert np.max(np.abs(got - x)) < 1e-5, got
    assert np.max(np.abs(cc.decrypt(out, 16)[4:])) < 1e-5


@pytest.mark.gpu
def test_bootstrap_primitives_bit_exact():
    eng, orc = make_pair(12, 6)
    try:
        z = np.array([0.25 - 0.5j, 1.0 + 0.125j, -0.75j, 0.3])
        same(eng.encode_complex(z, 1, 2, 8), orc.encode(z, 1, 2, 8))
        x = np.array([0.1, -0.2, 0.3, 0.4])
        co = orc.encrypt(orc.encode(x, slots=4))
        ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
        # conjugation key + automorphism: conj of a product with a complex plaintext
        orc.conj_keygen(); eng.EvalBootstrapSetup((1, 1), (0, 0), 4); eng.EvalBootstrapKeyGen(4)
        po, pe = orc.encode(z, 1, 0, 4), eng.encode_complex(z, 1, 0, 4)
        mo, me = orc.EvalMult(co, po), eng.EvalMult(ce, pe)
        same(me, mo)
        same(eng.EvalConjugate(me), orc.EvalConjugate(mo))
        # ModRaise: reduce to the last limb, raise to level 1
        lo = co.clone(); orc.rescale(lo); orc.level_reduce(lo, lo.nlimbs - 1)
        le = eng.import_ct(lo.d, lo.level, lo.deg, lo.sf, lo.slots)
        same(eng.mod_raise(le, 1), orc.mod_raise(lo, 1))
    finally:
        eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("log_n,depth,slots,budget,level", [(12, 50, 16, (3, 3), 40), (12, 40, 128, (4, 4), 37)])
def test_bootstrap_bit_exact_and_kat(log_n, depth, slots, budget, level):
    eng, orc = make_pair(log_n, depth, scale=59)
    try:
        for cc in (eng, orc):
            cc.EvalBootstrapSetup(budget, (0, 0), slots)
            cc.EvalBootstrapKeyGen(slots)
        x = np.array([0.1, 0.2, 0.3, 0.4, -0.5, 0.9, -1.0, 0.0])
        co = orc.encrypt(orc.encode(x, 1, level, slots))
        ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
        ro, re = orc.EvalBootstrap(co), eng.EvalBootstrap(ce)
        same(re, ro)
        got = eng.decrypt(re, slots)
        assert re.level < level
        assert np.max(np.abs(got[:8] - x)) < 1e-5 and np.max(np.abs(got[8:])) < 1e-5, got[:8]
        # a refreshed ciphertext keeps working: square it and bootstrap again
        sq = eng.EvalMult(re, re)
        again = eng.EvalBootstrap(sq)
        assert np.max(np.abs(eng.decrypt(again, 8) - x * x)) < 1e-4
    finally:
        eng.close()
