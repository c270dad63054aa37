"""Host enqueue time vs device time of one DirectSort N=128 sort: is the stream ever starved by the host?"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
sf = importlib.import_module("sorting-fhe_b200")
n = 128
depth, rots = sf.direct_sort_params(n, 1 << 17)
stream = torch.cuda.Stream()
eng = sf.Engine(17, depth, 40, 60, seed=1, stream=stream.cuda_stream)
eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
x = np.random.default_rng(0).permutation(n) / n
ct = eng.encrypt(eng.encode(x, slots=n))
s = sf.DirectSort(eng, n, rots)
prev = None
for it in range(4):
    ct.slots = n
    eng.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    out = s.sort(ct, (3, 3, 2))
    t1 = time.perf_counter(); e1.record(stream)
    eng.sync(); t2 = time.perf_counter()
    st = eng.stats(); cur = (st.get('host_ns'), st.get('host_calls'))
    if prev and cur[0]:
        print('   per-sort host ms [launch, alloc, free, other]:', [round((a - b) / 1e6, 2) for a, b in zip(cur[0], prev[0])], 'calls', [a - b for a, b in zip(cur[1], prev[1])])
    prev = cur
    print("iter %d: host enqueue %.1f ms, wall to completion %.1f ms, device %.1f ms" % (it, (t1 - t0) * 1e3, (t2 - t0) * 1e3, e0.elapsed_time(e1)))
def phase(name, fn):
    eng.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    r = fn()
    t1 = time.perf_counter(); e1.record(stream)
    eng.sync()
    print("%s: host enqueue %.1f ms, device %.1f ms" % (name, (t1 - t0) * 1e3, e0.elapsed_time(e1)))
    return r
for it in range(2):
    ct.slots = n
    rank = phase("construct_rank", lambda: s.construct_rank(ct, (3, 3, 2)))
    phase("rotation_index_check", lambda: s.rotation_index_check(rank, ct))
st = eng.stats()
print('host_ns', st.get('host_ns'), 'calls', st.get('host_calls'))
eng.close()
