"""GPU parity tests proper: every CUDA primitive, called through the C-ABI, bit-exact against the
CPU oracle on the same seeded inputs (north_star level 1)."""
import numpy as np
import pytest

from oracle.ckks_oracle import P, lib
from tests.common import make_pair, same

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pair12():
    eng, orc = make_pair(12, 8, rots=(1, -1, 2, 5, 64))
    yield eng, orc
    eng.close()


@pytest.mark.parametrize("log_n", [8, 10, 11, 12, 13, 14, 16, 17])
def test_ntt_bit_exact(log_n):
    eng, orc = None, None
    sf = __import__("importlib").import_module("sorting-fhe_b200")
    eng = sf.Engine(log_n, 3, 40)
    from oracle.ckks_oracle import Context, Params
    ctx = Context(Params(log_n, 3, 40))
    M = ctx.size_q + ctx.size_p
    rng = np.random.default_rng(log_n)
    ids = list(range(M))
    data = np.stack([rng.integers(0, ctx.moduli[i], ctx.n, dtype=np.uint64) for i in ids])
    ref = data.copy()
    for r, i in enumerate(ids):
        ctx.ntt(ref[r], i)
    got = eng.ntt_host(data.copy(), ids, inverse=False)
    bad = np.argwhere(got != ref)
    assert bad.size == 0, "forward NTT mismatches at %s" % bad[:5].tolist()
    back = eng.ntt_host(got.copy(), ids, inverse=True)
    assert np.array_equal(back, data), "inverse NTT does not round-trip"
    ref2 = ref.copy()
    for r, i in enumerate(ids):
        ctx.intt(ref2[r], i)
    assert np.array_equal(ref2, data)
    eng.close()


def test_keys_bit_exact(pair12):
    eng, orc = pair12
    assert np.array_equal(eng.export_secret(), orc.secret)
    pk = eng.export_public_key()
    assert np.array_equal(pk[0], orc.pk[0]) and np.array_equal(pk[1], orc.pk[1])
    rk = eng.export_eval_key()
    for j in range(eng.num_part_q):
        assert np.array_equal(rk[j, 0], orc.relin[1][j]), "relin b digit %d" % j
        assert np.array_equal(rk[j, 1], orc.relin[0][j]), "relin a digit %d" % j
    for r in (1, -1, 5):
        k = eng.export_eval_key(rotation=r)
        av, bv = orc.rot[orc.auto_index(r)]
        for j in range(eng.num_part_q):
            assert np.array_equal(k[j, 0], bv[j]) and np.array_equal(k[j, 1], av[j])


def test_encode_encrypt_decrypt(pair12):
    eng, orc = pair12
    x = np.linspace(-1, 1, 16)
    for level, slots in [(0, 16), (0, 2048), (3, 64), (5, 16)]:
        pe, po = eng.encode(x, 1, level, slots), orc.encode(x, 1, level, slots)
        same(pe, po)
        assert np.allclose(eng.decrypt(pe, 16), x, atol=1e-8)
    orc.ctx.enc_counter = 0
    ce, co = eng.encrypt(eng.encode(x, slots=16)), orc.encrypt(orc.encode(x, slots=16))
    same(ce, co)
    de, do = eng.decrypt(ce, 16), orc.decrypt(co, 16)
    assert np.array_equal(de, do), "decrypt/decode must be bit-identical to the oracle"
    assert np.allclose(de, x, atol=1e-7)


def test_eval_ops_bit_exact(pair12):
    eng, orc = pair12
    x = np.array([0.1, 0.2, -0.3, 0.4, 0.5, -0.6, 0.7, 0.8])
    orc.ctx.enc_counter = 100
    # engine counter: encrypt consumed 1 in the previous test at most; import the oracle ct instead
    co = orc.encrypt(orc.encode(x, slots=8))
    ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
    same(ce, co)
    # EvalMult + relinearisation (+ implicit rescale of both fresh operands)
    me, mo = eng.EvalMult(ce, ce), orc.EvalMult(co, co)
    same(me, mo)
    # second multiplication: lazy rescale then tensor + relin
    m2e, m2o = eng.EvalMult(me, ce), orc.EvalMult(mo, co)
    same(m2e, m2o)
    assert np.allclose(eng.decrypt(m2e, 8), x ** 3, atol=1e-6)
    # rotations (key switch + automorphism), positive / negative / composite index
    for r in (1, -1, 5, 64):
        same(eng.EvalRotate(m2e, r), orc.EvalRotate(m2o, r))
    assert np.allclose(eng.decrypt(eng.EvalRotate(ce, 1), 8), np.roll(x, -1), atol=1e-7)
    # level/scale alignment paths of EvalAdd/EvalSub
    same(eng.EvalAdd(m2e, ce), orc.EvalAdd(m2o, co))
    same(eng.EvalSub(ce, m2e), orc.EvalSub(co, m2o))
    same(eng.EvalAdd(m2e, 0.5), orc.EvalAdd(m2o, 0.5))
    same(eng.EvalSub(m2e, 0.25), orc.EvalSub(m2o, 0.25))
    same(eng.EvalSub(1.0, m2e), orc.EvalSub(1.0, m2o))
    same(eng.EvalMult(m2e, -0.25), orc.EvalMult(m2o, -0.25))
    same(eng.EvalNegate(m2e), orc.EvalNegate(m2o))
    # ct x pt and ct - pt with plaintexts at a different level
    pe, po = eng.encode(x, 1, 0, 8), orc.encode(x, 1, 0, 8)
    same(eng.EvalMult(m2e, pe), orc.EvalMult(m2o, po))
    same(eng.EvalSub(pe, m2e), orc.EvalSub(po, m2o))
    same(eng.EvalAddMany([ce, me, m2e]), orc.EvalAddMany([co, mo, m2o]))


def test_chebyshev_ps_bit_exact(pair12):
    eng, orc = pair12
    from oracle.algos import G4_COEFFS
    x = np.linspace(-0.9, 0.9, 16)
    co = orc.encrypt(orc.encode(x, slots=16))
    ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
    re, ro = eng.EvalChebyshevSeriesPS(ce, G4_COEFFS, -1, 1), orc.EvalChebyshevSeriesPS(co, G4_COEFFS, -1, 1)
    same(re, ro)
    assert re.level == 5


def test_missing_rotation_key_fails(pair12):
    eng, _ = pair12
    sf = __import__("importlib").import_module("sorting-fhe_b200")
    ct = eng.encrypt(eng.encode([1.0, 2.0], slots=2))
    with pytest.raises(sf.CkksError):
        eng.EvalRotate(ct, 3)


def test_hoisted_rotations_bit_exact(pair12):
    """EvalFastRotationPrecompute / EvalFastRotation (src/rotation.h:286,344-347) and the per-handle ModUp cache of
    EvalRotate give the same limbs as an un-hoisted EvalRotate (the oracle's)"""
    eng, orc = pair12
    x = np.linspace(-0.9, 0.9, 8)
    co = orc.encrypt(orc.encode(x, slots=8))
    ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
    digits = eng.EvalFastRotationPrecompute(ce)
    for r in (1, -1, 5):
        want = orc.EvalRotate(co, r)
        same(eng.EvalFastRotation(ce, r, 2 * eng.n, digits), want)
        same(eng.EvalRotate(ce, r), want)        # 2nd and 3rd call reuse the digits cached on the handle
    sq = eng.EvalMult(ce, ce)                    # deeper level, noiseScaleDeg 2
    same(eng.EvalRotate(eng.EvalRotate(sq, 1), 5), orc.EvalRotate(orc.EvalRotate(orc.EvalMult(co, co), 1), 5))


def test_compact_plaintext_ops_bit_exact(pair12):
    """sparsely packed plaintexts are kept compact on the device (one value per replication group); every op that
    consumes them must give the limbs of the full-length form"""
    eng, orc = pair12
    x = np.array([0.25, -0.5, 0.75, 0.1])
    co = orc.encrypt(orc.encode(x, slots=4))
    ce = eng.import_ct(co.d, co.level, co.deg, co.sf, co.slots)
    for slots in (4, 64, eng.n // 2):
        v = np.arange(slots) / slots - 0.3
        for level in (0, 2):
            po, pe = orc.MakeCKKSPackedPlaintext(v, 1, level, None, slots), eng.MakeCKKSPackedPlaintext(v, 1, level, None, slots)
            assert np.array_equal(pe.export(), po.d)
            same(eng.EvalMult(ce, pe), orc.EvalMult(co, po))
            same(eng.EvalAdd(ce, pe), orc.EvalAdd(co, po))
            same(eng.EvalSub(pe, ce), orc.EvalSub(po, co))
            assert np.allclose(eng.decrypt(eng.encrypt(pe), slots)[:4], v[:4], atol=1e-6)
            assert np.allclose(eng.decrypt(pe, slots)[:4], v[:4], atol=1e-6)


def test_inner_product_plain_bit_exact(pair12):
    """ckks_eval_inner_product_plain (the masked sums of src/sort_algo.h:336-345, 567-577 as one call): the fused pass
    and the fallback chain both equal the oracle's EvalAdd(T, EvalMult(ct, pt)) chain limb for limb"""
    eng, orc = pair12
    rng = np.random.default_rng(7)
    xs = [rng.uniform(-1, 1, 8) for _ in range(3)]
    cos = [orc.encrypt(orc.encode(x, slots=8)) for x in xs]
    ces = [eng.import_ct(c.d, c.level, c.deg, c.sf, c.slots) for c in cos]
    zo = orc.encrypt(orc.encode(np.zeros(8), slots=8))
    ze = eng.import_ct(zo.d, zo.level, zo.deg, zo.sf, zo.slots)
    for slots in (8, 64, eng.n // 2):                       # compact and full-length plaintexts
        ms = [rng.integers(0, 2, slots).astype(np.float64) for _ in range(3)]
        pos = [orc.MakeCKKSPackedPlaintext(m, 1, 0, None, slots) for m in ms]
        pes = [eng.MakeCKKSPackedPlaintext(m, 1, 0, None, slots) for m in ms]
        want = zo
        for c, p in zip(cos, pos):
            want = orc.EvalAdd(want, orc.EvalMult(c, p))
        same(eng.EvalInnerProductPlain(ze, ces, pes), want)                 # aligned: fused pass (lifts the deg-1 accumulator)
        same(eng.EvalInnerProductPlain(eng.EvalInnerProductPlain(ze, ces[:1], pes[:1]), ces[1:], pes[1:]), want)   # deg-2 accumulator
    # operands that need adjusting (noiseScaleDeg 2 inputs) take the chain
    sq_o = [orc.EvalMult(c, c) for c in cos]
    sq_e = [eng.EvalMult(c, c) for c in ces]
    ms = [rng.integers(0, 2, 8).astype(np.float64) for _ in range(3)]
    pos = [orc.MakeCKKSPackedPlaintext(m, 1, 0, None, 8) for m in ms]
    pes = [eng.MakeCKKSPackedPlaintext(m, 1, 0, None, 8) for m in ms]
    want = zo
    for c, p in zip(sq_o, pos):
        want = orc.EvalAdd(want, orc.EvalMult(c, p))
    same(eng.EvalInnerProductPlain(ze, sq_e, pes), want)


def test_inner_product_plain_multi_bit_exact(pair12):
    """ckks_eval_inner_product_plain_multi (all mask sets of one vecRotsOpt / blindRotationOptN call, src/sort_algo.h:331-352,
    563-580): every output equals the oracle's EvalAdd(T, EvalMult(ct, pt)) chain for its own mask set, for 1..6 sets (one to
    two kernel passes of at most four accumulators), compact and full-length plaintexts, and through the per-set fallback"""
    eng, orc = pair12
    rng = np.random.default_rng(11)
    xs = [rng.uniform(-1, 1, 8) for _ in range(3)]
    cos = [orc.encrypt(orc.encode(x, slots=8)) for x in xs]
    ces = [eng.import_ct(c.d, c.level, c.deg, c.sf, c.slots) for c in cos]
    zo = orc.encrypt(orc.encode(np.zeros(8), slots=8))
    ze = eng.import_ct(zo.d, zo.level, zo.deg, zo.sf, zo.slots)

    def chain(cts, pts):
        want = zo
        for c, p in zip(cts, pts):
            want = orc.EvalAdd(want, orc.EvalMult(c, p))
        return want

    for slots, groups in ((8, 1), (8, 2), (64, 3), (eng.n // 2, 4), (64, 6)):
        ms = [[rng.integers(0, 2, slots).astype(np.float64) for _ in range(3)] for _ in range(groups)]
        pos = [[orc.MakeCKKSPackedPlaintext(m, 1, 0, None, slots) for m in g] for g in ms]
        pes = [[eng.MakeCKKSPackedPlaintext(m, 1, 0, None, slots) for m in g] for g in ms]
        outs = eng.EvalInnerProductPlainMulti(ze, ces, pes)
        assert len(outs) == groups
        for g in range(groups):
            same(outs[g], chain(cos, pos[g]))
    # noiseScaleDeg-2 ciphertexts are rescaled through their cached twins first; still one fused pass per four sets
    sq_o = [orc.EvalMult(c, c) for c in cos]
    sq_e = [eng.EvalMult(c, c) for c in ces]
    ms = [[rng.integers(0, 2, 8).astype(np.float64) for _ in range(3)] for _ in range(2)]
    pos = [[orc.MakeCKKSPackedPlaintext(m, 1, 0, None, 8) for m in g] for g in ms]
    pes = [[eng.MakeCKKSPackedPlaintext(m, 1, 0, None, 8) for m in g] for g in ms]
    outs = eng.EvalInnerProductPlainMulti(ze, sq_e, pes)
    for g in range(2):
        same(outs[g], chain(sq_o, pos[g]))
    with pytest.raises(ValueError):
        eng.EvalInnerProductPlainMulti(ze, ces, [pes[0][:2]])


def test_clone_is_copy_on_write_and_adjusted_twins_bit_exact(pair12):
    """ckks_ct_clone shares the source's limbs until SetSlots / ckks_ct_mod_reduce / ckks_ct_device_ptr is applied to one of the
    handles (capi.cu); the level adjustment of two depth-2 operands at different levels (scaled rescale, cached as an adjusted twin
    on the shallower handle) gives the oracle's limbs on the first use, on the cached second use, through a clone, and after the
    source handle has been freed"""
    eng, orc = pair12
    rng = np.random.default_rng(21)
    xs = [rng.uniform(-1, 1, 8) for _ in range(3)]
    cos = [orc.encrypt(orc.encode(x, slots=8)) for x in xs]
    ces = [eng.import_ct(c.d, c.level, c.deg, c.sf, c.slots) for c in cos]
    # shallow = depth-2 product at the top; deep = the same kind of product three levels further down
    sh_o, sh_e = orc.EvalMult(cos[0], cos[1]), eng.EvalMult(ces[0], ces[1])
    d_o, d_e = cos[2], ces[2]
    for _ in range(3):
        d_o, d_e = orc.EvalMult(d_o, d_o), eng.EvalMult(d_e, d_e)
    assert sh_e.deg == 2 and d_e.deg == 2 and d_e.level > sh_e.level + 1
    want = orc.EvalAdd(sh_o, d_o)
    same(eng.EvalAdd(sh_e, d_e), want)                       # first use: scaled rescale + level drop in one step
    same(eng.EvalAdd(sh_e, d_e), want)                       # cached adjusted twin
    cl = sh_e.clone()
    same(eng.EvalAdd(cl, d_e), want)                         # the clone shares limbs and twins
    before = sh_e.export().copy()
    cl.slots = 4                                             # unshares the clone only
    assert sh_e.slots == 8 and cl.slots == 4
    ptr, words = eng.ct_device_ptr(cl)
    ptr0, _ = eng.ct_device_ptr(sh_e)
    assert ptr != ptr0
    cl2 = sh_e.clone()
    eng.ct_mod_reduce(cl2)                                   # in-place entry point on a shared handle: works on a private copy
    assert np.array_equal(sh_e.export(), before) and np.array_equal(cl2.export(), before)
    del sh_e                                                 # the clone keeps its own data alive
    eng.sync()
    assert np.array_equal(cl.export(), before)
    same(eng.EvalAdd(cl2, d_e), want)
