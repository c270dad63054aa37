"""ckks_key_switch_dev_noalloc (`-m gpu`; collected last on purpose): the allocation-free key switch over the library's
static 16 MiB device scratch must return, bit for bit, what ckks_key_switch_dev returns -- on two calls in a row (the
scratch keeps the previous call's contents), for the relinearisation key and a rotation key, at two limb counts -- and
must reject what it cannot serve with a clear error.  The entry point has no counterpart in the reference
(include/ckks_b200.h explains why it exists); the bar is the engine's own: integer work, bit-exact."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
sf = importlib.import_module("sorting-fhe_b200")


def _rows(eng, nl):
    """scratch rows the entry point needs (include/ckks_b200.h)"""
    k, alpha = eng.size_p, eng.alpha
    digits = [min(alpha * (j + 1), nl) - alpha * j for j in range(-(-nl // alpha))]
    return nl + sum(nl + k - d for d in digits) + 2 * (nl + k) + 2 * k + 2 * nl


@pytest.fixture(scope="module")
def engine():
    stream = torch.cuda.Stream()
    eng = sf.Engine(13, 12, 40, seed=5, stream=stream.cuda_stream)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen([1, -3])
    yield eng, stream
    eng.close()


@pytest.mark.parametrize("rotation", [None, 1, -3])
def test_noalloc_key_switch_matches_the_pooled_one_twice(engine, rotation):
    eng, stream = engine
    n, q = eng.n, eng.moduli()[0]
    rng = np.random.default_rng(17 if rotation is None else 100 + rotation)
    for nl in (eng.size_q, 5):
        assert _rows(eng, nl) * n <= 1 << 21
        want_all, got_all = [], []
        for rep in range(2):                                              # two calls in a row, different inputs
            src_h = np.stack([rng.integers(0, q[i], n, dtype=np.uint64) for i in range(nl)])
            src = torch.from_numpy(src_h.view(np.int64)).cuda()
            want = torch.zeros((2, nl, n), dtype=torch.int64, device="cuda")
            got = torch.full((2, nl, n), -1, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            eng.key_switch_dev(src.data_ptr(), nl, want.data_ptr(), rotation)      # also the set-up call: builds the tables
            eng.sync()
            free0 = torch.cuda.mem_get_info()[0]
            eng.key_switch_dev_noalloc(src.data_ptr(), nl, got.data_ptr(), rotation)
            eng.sync()
            if rep:   # the very first call on a device is the set-up: the symbol lookup loads the module that holds the array
                assert torch.cuda.mem_get_info()[0] == free0, "the allocation-free entry point changed the device's free memory"
            assert torch.equal(got, want), "rep %d, %d limbs: differs from ckks_key_switch_dev" % (rep, nl)
            assert bool((got >= 0).all())
            want_all.append(want.cpu()); got_all.append(got.cpu())
        assert not torch.equal(want_all[0], want_all[1])                    # the two calls really computed different things
        # and again on the first input after the scratch held the second call's data
        # (stale contents must not leak into the result)
        eng.key_switch_dev_noalloc(src.data_ptr(), nl, got.data_ptr(), rotation)
        eng.sync()
        assert torch.equal(got.cpu(), want_all[1])


def test_noalloc_rejects_what_it_cannot_serve(engine):
    eng, stream = engine
    n = eng.n
    src = torch.zeros((eng.size_q, n), dtype=torch.int64, device="cuda")
    out = torch.zeros((2, eng.size_q, n), dtype=torch.int64, device="cuda")
    with pytest.raises(RuntimeError, match="key not present"):
        eng.key_switch_dev_noalloc(src.data_ptr(), eng.size_q, out.data_ptr(), 7)
    with pytest.raises(RuntimeError, match="bad limb count"):
        eng.key_switch_dev_noalloc(src.data_ptr(), eng.size_q + 1, out.data_ptr())
    with pytest.raises(RuntimeError, match="allocated on first use"):
        eng.key_switch_dev_noalloc(src.data_ptr(), 3, out.data_ptr())           # no key switch ran at 3 limbs yet


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    eng = sf.Engine(15, 12, 40, seed=5)                                        # ring 2^15: 13 limbs need > 2^21 words
    try:
        eng.KeyGen(); eng.EvalMultKeyGen()
        nl, n = eng.size_q, eng.n
        assert _rows(eng, nl) * n > 1 << 21
        src = torch.zeros((nl, n), dtype=torch.int64, device="cuda")
        out = torch.zeros((2, nl, n), dtype=torch.int64, device="cuda")
        eng.key_switch_dev(src.data_ptr(), nl, out.data_ptr())                  # tables exist, so the size check is what fires
        eng.sync()
        with pytest.raises(RuntimeError, match="beyond the static scratch of 16777216 bytes"):
            eng.key_switch_dev_noalloc(src.data_ptr(), nl, out.data_ptr())
    finally:
        eng.close()
