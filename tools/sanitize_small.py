"""Small end-to-end run for compute-sanitizer (memcheck / racecheck): every kernel class at a small ring -- key generation,
encrypt, DirectSort N=8 (relinearisations with the fused ModDown epilogue, rotations through the automorphism scatter,
rescales, Paterson-Stockmeyer), one bootstrapping-free k-way stage is left out (same kernels) -- plus the allocation-free
key switch.  usage: compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
sf = importlib.import_module("sorting-fhe_b200")
n = 8
depth, rots = sf.direct_sort_params(n, 1 << 10)
eng = sf.Engine(10, depth, 40, seed=3)
eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
x = np.random.default_rng(2).permutation(n) / n
out = sf.DirectSort(eng, n, rots).sort(eng.encrypt(eng.encode(x, slots=n)), (3, 2, 2))
err = float(np.max(np.abs(eng.decrypt(out, n) - np.sort(x))))
nl = eng.size_q
src = torch.randint(0, 1 << 39, (nl, eng.n), dtype=torch.int64, device="cuda")
a = torch.zeros((2, nl, eng.n), dtype=torch.int64, device="cuda"); b = torch.zeros_like(a)
eng.key_switch_dev(src.data_ptr(), nl, a.data_ptr(), rots[0]); eng.sync()
eng.key_switch_dev_noalloc(src.data_ptr(), nl, b.data_ptr(), rots[0]); eng.sync()
print("sort max_abs_err %.3g, noalloc equal %s" % (err, bool(torch.equal(a, b))))
assert err < 0.01 and torch.equal(a, b)
eng.close()
