"""shared helpers for the parity tests: matching (oracle, engine) pairs on the same seed"""
import importlib

import numpy as np

from oracle.ckks_oracle import Context, Params
from oracle.ckks_ops import OracleCC


def engine_module():
    return importlib.import_module("sorting-fhe_b200")


def make_pair(log_n, depth, scale=40, first=60, seed=1, rots=(), dry=False, ext=True):
    sf = engine_module()
    eng = sf.Engine(log_n, depth, scale, first, 0, ext, seed, dry_run=dry)
    orc = OracleCC(Context(Params(log_n, depth, scale, first, 0, ext, seed)))
    for cc in (eng, orc):
        if cc is eng:
            cc.KeyGen(); cc.EvalMultKeyGen()
            if rots:
                cc.EvalRotateKeyGen(list(rots))
        else:
            cc.keygen(); cc.relin_keygen(); cc.rot_keygen(list(rots))
    return eng, orc


def same(ct_eng, ct_orc):
    """bit-exact comparison of an engine ciphertext with an oracle ciphertext (limbs + metadata)"""
    a = ct_eng.export()
    assert a.shape == ct_orc.d.shape, (a.shape, ct_orc.d.shape)
    assert ct_eng.level == ct_orc.level and ct_eng.deg == ct_orc.deg and ct_eng.slots == ct_orc.slots
    assert ct_eng.sf == ct_orc.sf, (ct_eng.sf, ct_orc.sf)
    bad = np.argwhere(a != ct_orc.d)
    assert bad.size == 0, "first mismatches (part, limb, idx): %s of %d" % (bad[:5].tolist(), len(bad))
    return True
