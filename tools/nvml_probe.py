"""How much does an NVML clock query cost next to the engine's kernel launches?  (bench.py clock sampling design)
  python tools/nvml_probe.py"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import pynvml as nv
sf = importlib.import_module("sorting-fhe_b200")

n = 128
depth, rots = sf.direct_sort_params(n, 1 << 17)
stream = torch.cuda.Stream()
eng = sf.Engine(17, depth, 40, seed=1, stream=stream.cuda_stream)
eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
sorter = sf.DirectSort(eng, n, rots)
x = np.random.default_rng(0).permutation(n) / n
ct = eng.encrypt(eng.encode(x, slots=n))
for _ in range(3):
    sorter.sort(ct, (3, 3, 2)); ct.slots = n
eng.sync()
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(0)


def q():
    t = time.perf_counter()
    c = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
    return c, r, (time.perf_counter() - t) * 1e3


def region(mode, K=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    qs = []
    e0.record(stream)
    for i in range(K):
        ct.slots = n
        if mode == "before_sync" and i % 3 == 1:
            o = sorter.sort(ct, (3, 3, 2)); qs.append(q()); eng.sync()      # query while the GPU works through the sort
        else:
            o = sorter.sort(ct, (3, 3, 2)); eng.sync()
            if mode == "after_sync" and i % 3 == 1:
                qs.append(q())
    e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) / K, qs


for mode in ("none", "after_sync", "before_sync", "none"):
    ms, qs = region(mode)
    print(mode, "ms/step %.2f" % ms, [(c, r, round(t, 2)) for c, r, t in qs], flush=True)
# per-step wall + device times: where does an occasional slow region come from?
for rep in range(4):
    rows = []
    for i in range(10):
        ct.slots = n
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        o = sorter.sort(ct, (3, 3, 2))
        t1 = time.perf_counter()
        e1.record(stream)
        eng.sync()
        t2 = time.perf_counter()
        rows.append((round(e0.elapsed_time(e1), 1), round((t1 - t0) * 1e3, 1), round((t2 - t0) * 1e3, 1)))
    print("rep", rep, "(device ms, host enqueue ms, wall ms):", rows, flush=True)
eng.close()
