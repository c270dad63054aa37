"""Device times of the BASELINE.json configurations other than the headline one (bench.py), one JSON object on stdout.
  python tools/bench_configs.py [--only NAME]
Each entry: milliseconds per call (CUDA events on the engine stream, best of 3 after one warm-up) and the decrypted
max abs error where the call is a whole sort."""
import argparse
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

sf = importlib.import_module("sorting-fhe_b200")


def timed(eng, stream, fn, reps=3):
    fn(); eng.sync()
    best, out = 1e30, None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng.sync(); e0.record(stream)
        out = fn()
        e1.record(stream); eng.sync(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


def engine(log_n, depth, scale=40, first=60, rots=(), seed=1):
    stream = torch.cuda.Stream()
    eng = sf.Engine(log_n, depth, scale, first, seed=seed, stream=stream.cuda_stream)
    eng.KeyGen(); eng.EvalMultKeyGen()
    if rots:
        eng.EvalRotateKeyGen(list(rots))
    return eng, stream


def direct_sort(n, cfg, log_n=17):
    depth, rots = sf.direct_sort_params(n, 1 << log_n)
    eng, stream = engine(log_n, depth, rots=rots)
    x = np.random.default_rng(n).permutation(n) / n
    ct = eng.encrypt(eng.encode(x, slots=n))
    s = sf.DirectSort(eng, n, rots)

    def run():
        ct.slots = n
        return s.sort(ct, cfg)
    ms, out = timed(eng, stream, run)
    ct.slots = n
    ms_rank, rank = timed(eng, stream, lambda: s.construct_rank(ct, cfg))
    res = {"ms": ms, "ms_construct_rank": ms_rank, "ms_index_check": ms - ms_rank, "depth": depth, "rotation_keys": len(rots),
           "max_abs_err": float(np.max(np.abs(eng.decrypt(out, n) - np.sort(x))))}
    eng.close()
    return res


def compare_full_slot():
    """config 2: Comparison::compare with Sign(4,3,3) on a full-slot ciphertext (65 536 slots, ring 2^17)"""
    eng, stream = engine(17, 29, scale=59)
    slots = eng.n // 2
    rng = np.random.default_rng(2)
    a, b = rng.permutation(slots) / slots, rng.permutation(slots) / slots
    ca, cb = eng.encrypt(eng.encode(a, slots=slots)), eng.encrypt(eng.encode(b, slots=slots))
    ms, out = timed(eng, stream, lambda: sf.compare(eng, ca, cb, (4, 3, 3)))
    got = eng.decrypt(out, slots)
    want = (np.sign(a - b) + 1) / 2
    far = np.abs(a - b) > 0.01
    res = {"ms": ms, "slots": slots, "level": out.level, "max_abs_err_where_gap_gt_0.01": float(np.max(np.abs(got - want)[far]))}
    eng.close()
    return res


def kway_stage():
    """config 4: the k=2, M=7 network on N=128 at the adapter's card: one compare-swap stage, one EvalBootstrap, and the
    whole 28-stage network with bootstrapping (src/k-way/Sorter.cpp:284-404, level budget {4,4})"""
    n, k, M = 128, 2, 7
    depth, scale, first, rots = sf.kway_params(n)
    eng, stream = engine(17, depth, scale, first, rots)
    eng.EvalBootstrapSetup((4, 4), (0, 0), n)
    eng.EvalBootstrapKeyGen(n)
    x = np.random.default_rng(3).permutation(n) * ((1.0 - 1e-8) / n)
    ct = eng.encrypt(eng.encode(x, slots=n))
    ms, out = timed(eng, stream, lambda: sf.kway_sort(eng, ct, n, k, M, (3, 2, 3), depth, 0, 1))
    deep = eng.encrypt(eng.encode(x, 1, depth - 2, n))
    ms_boot, fresh = timed(eng, stream, lambda: eng.EvalBootstrap(deep))
    eng.stats(reset=True)
    ms_all, res_ct = timed(eng, stream, lambda: sf.kway_sort(eng, ct, n, k, M, (3, 2, 3), depth))
    boots = sum(cnt for name, nl, cnt in eng.stats()["ops"] if name == "bootstrap")
    got = eng.decrypt(res_ct, n)
    res = {"ms_first_stage": ms, "levels_first_stage": out.level, "ms_bootstrap": ms_boot, "level_after_bootstrap": fresh.level,
           "bootstrap_max_abs_err": float(np.max(np.abs(eng.decrypt(fresh, n) - x))),
           "ms_full_network_28_stages": ms_all, "bootstraps_per_run": boots / 4.0, "key_bytes": eng.info(7),
           "max_abs_err": float(np.max(np.abs(got - np.sort(x))))}
    eng.close()
    return res


def mehp24(n=256):
    depth, rots = sf.mehp24_params(n)
    eng, stream = engine(17, depth, rots=rots, seed=3)
    x = np.random.default_rng(11).permutation(n) / n
    ct = eng.encrypt(eng.encode(x, slots=n * n))
    cfg = (3, 4, 2) if n > 128 else (3, 3, 2) if n > 16 else (3, 2, 2)
    ms, out = timed(eng, stream, lambda: sf.mehp24_sort_fg(eng, ct, n, cfg))
    res = {"ms": ms, "depth": depth, "level": out.level, "key_bytes": eng.info(7),
           "max_abs_err": float(np.max(np.abs(eng.decrypt(out, n * n)[:n] - np.sort(x))))}
    eng.close()
    return res


def hybrid1(n=128):
    """sort_hybrid1 at N = 128: the reference README's only published time (74.12 s, its CPU, 2^17 ring, depth 46)"""
    depth, rots = sf.hybrid_sort_params(n, "hybrid1")
    eng, stream = engine(17, depth, rots=rots, seed=5)
    x = np.random.default_rng(n).permutation(n) / n
    ct = eng.encrypt(eng.encode(x, slots=n))
    s = sf.DirectSort(eng, n, rots)

    def run():
        ct.slots = n
        return s.sort_hybrid(ct, (3, 3, 2), "hybrid1")
    ms, out = timed(eng, stream, run)
    res = {"ms": ms, "depth": depth, "rotation_keys": len(rots), "key_bytes": eng.info(7), "reference_published_s": 74.12,
           "max_abs_err": float(np.max(np.abs(eng.decrypt(out, n * n)[:n] - np.sort(x))))}
    eng.close()
    return res


def mehp24_large(n=512, sub=256):
    """mehp24::sortLargeArrayFG N = 512 as two 256 x 256 blocks (fixture depth 57, Sign(3,4,2))"""
    depth, rots = sf.mehp24_params(n)
    eng, stream = engine(17, depth, rots=rots, seed=4)
    x = np.random.default_rng(12).permutation(n) / n
    ct = eng.encrypt(eng.encode(x, slots=sub * sub))
    ms, out = timed(eng, stream, lambda: sf.mehp24_sort_large(eng, ct, n, sub, (3, 4, 2)), reps=2)
    res = {"ms": ms, "depth": depth, "level": out.level, "rotation_keys": len(rots), "key_bytes": eng.info(7),
           "max_abs_err": float(np.max(np.abs(eng.decrypt(out, sub * sub)[:n] - np.sort(x))))}
    eng.close()
    return res


CONFIGS = {
    "mehp24_sort_large_n512": mehp24_large,
    "sort_hybrid1_n128": hybrid1,
    "direct_sort_n16": lambda: direct_sort(16, (3, 2, 2)),
    "compare_sign433_full_slot": compare_full_slot,
    "direct_sort_n128": lambda: direct_sort(128, (3, 3, 2)),
    "kway_k2_n128_stage": kway_stage,
    "mehp24_sortfg_n256": mehp24,
    "direct_sort_n64": lambda: direct_sort(64, (3, 3, 2)),
}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    out = {}
    for name, fn in CONFIGS.items():
        if a.only and a.only != name:
            continue
        out[name] = fn()
        torch.cuda.empty_cache()
    print(json.dumps(out, indent=1))
