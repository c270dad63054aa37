#!/usr/bin/env python
"""Full-size oracle runs -> tests/golden/fullsize_limbs.json (run in the build container; CPU only, minutes).

  python tests/golden/make_fullsize.py [direct_sort_n128] [compare_fullslot] [mehp24_n256] [direct_sort_n16] ...

Each case executes the oracle's restatement of the reference op sequence at the BASELINE size on seeded inputs and
records SHA-256 digests of the ciphertext limbs at the checkpoints the GPU parity tests compare (`-m gpu`,
tests/test_fullsize.py): the engine must reproduce every limb of a complete DirectSort<128>::sort at ring 2^17, not
only its decrypted values.  The wall time of each run is recorded too (it is the honest CPU-port time of one sort on
this container's cores; bench.py --impl reference times the same call on the GPU box).

Cases (BASELINE.json configs):
  direct_sort_n128   config 3: DirectSort N=128, ring 2^17, depth 30, Sign(3,3,2)   tests/DirectSortTest.cpp:29-44,116-117
  direct_sort_n16    config 1: DirectSort N=16, ring 2^17, depth 25, Sign(3,2,2)    tests/DirectSortTest.cpp:101-102
  compare_fullslot   config 2: Comparison::compare Sign(4,3,3) on a 65 536-slot ciphertext (CompareTest parameters at ring 2^17)
  mehp24_n256        config 5a: mehp24 sortFG N=256 (65 536 slots), depth 49        tests/mehp24/Mehp24SortTest.cpp:36-60
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
OUT = os.path.join(HERE, "fullsize_limbs.json")

from oracle import algos  # noqa: E402
from oracle.ckks_oracle import Context, Params  # noqa: E402
from oracle.ckks_ops import OracleCC  # noqa: E402


def digest(ct):
    return {"sha256": hashlib.sha256(np.ascontiguousarray(ct.d).tobytes()).hexdigest(), "shape": list(ct.d.shape),
            "level": int(ct.level), "deg": int(ct.deg), "sf_hex": float(ct.sf).hex(), "slots": int(ct.slots),
            "head": [int(v) for v in ct.d.reshape(-1)[:4]]}


def make_cc(log_n, depth, scale, first, seed, rots):
    t = time.perf_counter()
    cc = OracleCC(Context(Params(log_n, depth, scale, first, 0, True, seed)))
    cc.keygen(); cc.relin_keygen(); cc.rot_keygen(list(rots))
    return cc, time.perf_counter() - t


def ops_of(cc):
    return sorted([[k[0], int(k[1]), int(v)] for k, v in cc.stats.items()])


def case_direct_sort(n, cfg, log_n=17, seed=1, input_seed=1):
    depth, rots = algos.DIRECT_SORT_PARAMS[n]
    cc, t_keys = make_cc(log_n, depth, 40, 60, seed, rots)
    x = np.random.default_rng(input_seed).permutation(n) / n
    ct = cc.encrypt(cc.encode(x, slots=n))           # encryption counter 0
    sorter = algos.DirectSort(cc, n, rots)           # zero cache: encryption counter 1
    cc.stats.clear()
    t0 = time.perf_counter()
    rank = sorter.construct_rank(ct, cfg)
    t1 = time.perf_counter()
    out = sorter.rotation_index_check_n(rank, ct)
    t2 = time.perf_counter()
    got = cc.decrypt(out, n)
    ranks = cc.decrypt(rank, n)
    return {"params": {"log_n": log_n, "depth": depth, "scale": 40, "first": 60, "ext": True, "seed": seed, "rots": list(rots),
                       "n": n, "sign_cfg": list(cfg), "input_rng": "default_rng(%d).permutation(n)/n" % input_seed},
            "digests": {"input": digest(ct), "rank": digest(rank), "output": digest(out)},
            "decrypted": {"output": [float(v) for v in got], "rank": [float(v) for v in ranks]},
            "max_abs_err": float(np.max(np.abs(got - np.sort(x)))),
            "rank_max_abs_err": float(np.max(np.abs(ranks - np.argsort(np.argsort(x))))),
            "oracle_seconds": {"keygen": t_keys, "construct_rank": t1 - t0, "index_check": t2 - t1, "sort": t2 - t0,
                               "host_threads": os.cpu_count()},
            "ops": ops_of(cc)}


def case_compare_fullslot(log_n=17, seed=1):
    """config 2: compare (4,3,3) on every slot; CompareTest's parameters (depth 50, scale 59, tests/CompareTest.cpp:13-27)
    at ring 2^17; the final level is recorded"""
    slots = 1 << (log_n - 1)
    depth = 50
    cc, t_keys = make_cc(log_n, depth, 59, 60, seed, [])
    rng = np.random.default_rng(2)
    a = rng.permutation(slots) / slots
    b = np.roll(a, 1)
    ca, cb = cc.encrypt(cc.encode(a, slots=slots)), cc.encrypt(cc.encode(b, slots=slots))
    cc.stats.clear()
    t0 = time.perf_counter()
    out = algos.compare(cc, ca, cb, 4, 3, 3)
    t1 = time.perf_counter()
    got = cc.decrypt(out, slots)
    want = (a > b).astype(np.float64)
    return {"params": {"log_n": log_n, "depth": depth, "scale": 59, "first": 60, "ext": True, "seed": seed, "rots": [],
                       "slots": slots, "sign_cfg": [4, 3, 3], "input_rng": "a = default_rng(2).permutation(slots)/slots; b = roll(a, 1)"},
            "digests": {"a": digest(ca), "b": digest(cb), "output": digest(out)},
            "decrypted_head": [float(v) for v in got[:16]],
            "max_abs_err": float(np.max(np.abs(got - want))),
            "oracle_seconds": {"keygen": t_keys, "compare": t1 - t0, "host_threads": os.cpu_count()},
            "ops": ops_of(cc)}


def case_mehp24(n=256, log_n=17, seed=1):
    depth = 49                                         # tests/mehp24/Mehp24SortTest.cpp:58-60
    rots = algos.mehp24_rotation_indices(n)
    cc, t_keys = make_cc(log_n, depth, 40, 60, seed, rots)
    x = np.random.default_rng(3).permutation(n) / n
    ct = cc.encrypt(cc.encode(x, slots=n * n))
    cc.stats.clear()
    t0 = time.perf_counter()
    dg_i = (int(np.log2(n)) + 1) // 2                  # tests/mehp24/Mehp24SortTest.cpp:137
    out = algos.Mehp24(cc, n).sort_fg(ct, (3, 4, 2), dg_i, 2)
    t1 = time.perf_counter()
    got = cc.decrypt(out, n * n)[:n]
    return {"params": {"log_n": log_n, "depth": depth, "scale": 40, "first": 60, "ext": True, "seed": seed, "rots": list(rots),
                       "n": n, "sign_cfg": [3, 4, 2], "input_rng": "default_rng(3).permutation(n)/n"},
            "digests": {"input": digest(ct), "output": digest(out)},
            "decrypted": {"output": [float(v) for v in got]},
            "max_abs_err": float(np.max(np.abs(got - np.sort(x)))),
            "oracle_seconds": {"keygen": t_keys, "sort": t1 - t0, "host_threads": os.cpu_count()},
            "ops": ops_of(cc)}


CASES = {
    "direct_sort_n128": lambda: case_direct_sort(128, (3, 3, 2)),
    "direct_sort_n16": lambda: case_direct_sort(16, (3, 2, 2)),
    "compare_fullslot": case_compare_fullslot,
    "mehp24_n256": case_mehp24,
}


def main():
    names = sys.argv[1:] or list(CASES)
    for name in names:
        t = time.perf_counter()
        r = CASES[name]()
        res = json.load(open(OUT)) if os.path.exists(OUT) else {}
        res[name] = r
        print("%s: %.1f s, max_abs_err %.3g" % (name, time.perf_counter() - t, r["max_abs_err"]), flush=True)
        with open(OUT, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
