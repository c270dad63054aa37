"""CPU tests of the product's host side: the C-ABI libraries load and export every declared symbol, and the
sorter layer's pure logic (Decomposer, parameter tables, coefficient generator, level budgets in dry-run mode)
behaves like the reference's.  No compute call needs a GPU here."""
import ctypes
import importlib
import os
import re

import numpy as np
import pytest

from oracle import algos
from oracle.cheby import truncated_doubled_sinc_coeffs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sf = importlib.import_module("sorting-fhe_b200")


@pytest.mark.parametrize("header,lib", [("ckks_b200.h", "libckks_b200.so"), ("sortfhe_host.h", "libsortfhe_host.so")])
def test_c_abi_exports_every_declared_symbol(header, lib):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b((?:ckks|sortfhe)_[a-z0-9_]+)\s*\(", text)
    assert len(names) > 8
    L = ctypes.CDLL(os.path.join(ROOT, "sorting-fhe_b200", lib))
    for n in set(names):
        assert hasattr(L, n), n


def test_compute_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(sf.CkksError, match="no CUDA device"):
        sf.Engine(10, 2)


def test_unsupported_digit_count_is_a_clean_error():
    """numLargeDigits beyond the key switch's 4-digit tables must fail at context creation, not overflow"""
    with pytest.raises(sf.CkksError, match="at most 4 digits"):
        sf.Engine(10, 9, 40, 60, 6, dry_run=True)
    sf.Engine(10, 9, 40, 60, 4, dry_run=True).close()


def test_entropy_seeded_context_and_options():
    eng = sf.Engine(10, 3, seed=0, dry_run=True)       # seed 0: the sampler key comes from getrandom()
    eng.set_option("add_const_mode", 1)
    eng.set_option("modup_cache", 0)
    with pytest.raises(sf.CkksError):
        eng.set_option("add_const_mode", 2)
    assert eng.info(8) > 0
    eng.close()


def test_decomposer_reference_vectors():
    """tests/DecomposeTest.cpp:64-74: Decomposer<128>{1..64}, sum of steps == rotation for 15 values x 3 algos"""
    keys = [1, 2, 4, 8, 16, 32, 64]
    for r in [1, 2, 3, 4, 7, 8, 15, 16, 31, 32, 63, 64, 65, 127, 128]:
        for algo in ("NAF", "BNAF", "BINARY"):
            assert sum(sf.decompose(128, r, 128, keys, algo)) == r, (r, algo)
    assert sf.decompose(128, 21, 128, keys, "BINARY") == algos.decompose_binary(21, keys, 128, 128) == [16, 4, 1]
    assert sf.decompose(128, 7, 128, keys, "NAF") == [8, -1]


def test_direct_sort_parameter_tables():
    """DirectSort<N>::getSizeParameters (src/sort_algo.h:87-201): depth and rotation key set per N"""
    for n, (depth, rots) in algos.DIRECT_SORT_PARAMS.items():
        assert sf.direct_sort_params(n, 1 << 17) == (depth, rots)


def test_sinc_coefficients_match_generator():
    """utils/generate_cheb_doubled_coeffs.cpp: the product's C++ generator and the oracle's restatement agree to
    the last printed digit"""
    for n in (4, 16):
        assert np.array_equal(sf.sinc_coefficients(n), truncated_doubled_sinc_coeffs(n))


@pytest.mark.parametrize("n,cfg", [(4, (3, 2, 2)), (16, (3, 2, 2)), (64, (3, 3, 2)), (128, (3, 3, 2)), (256, (3, 4, 2))])
def test_level_budget_dry_run(n, cfg):
    """SURVEY 8a level-budget identity: DirectSort<N>::sort ends exactly at multDepth (tests/DirectSortTest.cpp:140).
    dry_run contexts execute the whole host schedule with metadata only."""
    depth, rots = sf.direct_sort_params(n, 1 << 17)
    eng = sf.Engine(17, depth, 40, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    ct = eng.encrypt(eng.encode(np.arange(n) / n, slots=n))
    out = sf.DirectSort(eng, n, rots).sort(ct, cfg)
    assert out.level == depth and out.slots == n
    ops = {}
    for name, nl, cnt in eng.stats()["ops"]:
        ops[name] = ops.get(name, 0) + cnt
    if n == 128:   # SURVEY 3.1 op table: 58 rotations, 258 encodes (+1 zero cache)
        assert ops["rotate"] == 58 and ops["encode"] in (258, 259, 260)
    eng.close()


def test_production_sampler_block_function_known_answers():
    """contexts created with seed 0 sample through ChaCha20 (key from getrandom, counter = sample index, nonce = sampler
    stream); the host instance of that block function must reproduce RFC 8439 section 2.3.2 and, where the `cryptography`
    package is present, the library's keystream for random keys / counters / nonces"""
    import struct
    key = bytes(range(32))
    got = sf.rng_block(key, 0x4A000000, 1 | (0x09000000 << 32))       # counter 1, nonce 00:00:00:09 00:00:00:4a 00:00:00:00
    want = ("e4e7f110 15593bd1 1fdd0f50 c47120a3 c7f4d1c7 0368c033 9aaa2204 4e6cd4c3 "
            "466482d2 09aa9f07 05d7c214 a2028bd9 d19c12b5 b94e16de e883d0cb 4e3c50a2")
    assert list(struct.unpack("<16I", got)) == [int(w, 16) for w in want.split()]
    try:
        from cryptography.hazmat.primitives.ciphers import Cipher, algorithms
    except ImportError:
        return
    rng = np.random.default_rng(0)
    for _ in range(8):
        kb, stream, idx = rng.bytes(32), int(rng.integers(0, 2 ** 63)), int(rng.integers(0, 2 ** 63))
        ks = Cipher(algorithms.ChaCha20(kb, struct.pack("<QQ", idx, stream)), mode=None).encryptor().update(b"\0" * 64)
        assert ks == sf.rng_block(kb, stream, idx)


@pytest.mark.parametrize("degree", [5, 13, 27, 59, 119, 247, 495, 848])
def test_paterson_stockmeyer_plan_matches_evaluation_dry_run(degree):
    """cheby.cu forms the weighted sums of a Paterson-Stockmeyer evaluation ahead of the recursion, from a plan made on the
    coefficients alone (ps_plan / wsum_precompute); the recursion then checks every sum it asks for against the plan and fails with
    "plan and evaluation disagree" otherwise.  Dry-run contexts run the whole host schedule: every degree class the reference
    reaches (g_4's 27, the sinc tables up to 848 of src/generated_doubled_sinc_coeffs.h) plus dense and sparse coefficient vectors"""
    rng = np.random.default_rng(degree)
    eng = sf.Engine(12, 14, 40, dry_run=True)
    eng.KeyGen(); eng.EvalMultKeyGen()
    x = eng.encrypt(eng.encode(np.linspace(-0.9, 0.9, 8), slots=8))
    dense = rng.uniform(-1, 1, degree + 1)
    sparse = dense.copy(); sparse[::2] = 0.0; sparse[-1] = 0.5          # odd polynomial, like the sign / sinc series
    lead = np.zeros(degree + 1); lead[-1] = 1.0; lead[1] = 0.25           # almost empty: single-term leaves and short sums
    for coeffs in (dense, sparse, lead):
        y = eng.EvalChebyshevSeriesPS(x, coeffs.tolist())
        assert y.level > x.level
    eng.close()
