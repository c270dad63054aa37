"""Driver for ncu captures INSIDE a sort: warms up (keys, plaintext cache, memory pool), then runs ONE DirectSort N=128 sort
between cudaProfilerStart / cudaProfilerStop, so that `ncu --profile-from-start off` sees exactly the launches of one
steady-state sort (warm L2, real neighbours), not an isolated cold-cache kernel.
  ncu --profile-from-start off --clock-control none -k regex:ntt_kernel \
      --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --page raw \
      --log-file gpurun_out/ntt_in_sort.csv python tools/prof_sort.py
usage: python tools/prof_sort.py [n=128] [log_n=17]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
sf = importlib.import_module("sorting-fhe_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
log_n = int(sys.argv[2]) if len(sys.argv) > 2 else 17
cfg = {16: (3, 2, 2), 64: (3, 3, 2), 128: (3, 3, 2)}[n]
depth, rots = sf.direct_sort_params(n, 1 << log_n)
stream = torch.cuda.Stream()
eng = sf.Engine(log_n, depth, 40, seed=1, stream=stream.cuda_stream)
eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
sorter = sf.DirectSort(eng, n, rots)
x = np.random.default_rng(100).permutation(n) / n
ct = eng.encrypt(eng.encode(x, slots=n))
for _ in range(2):
    ct.slots = n
    sorter.sort(ct, cfg)
eng.sync(); torch.cuda.synchronize()
torch.cuda.profiler.start()
ct.slots = n
out = sorter.sort(ct, cfg)
eng.sync(); torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("max_abs_err", float(np.max(np.abs(eng.decrypt(out, n) - np.sort(x)))))
eng.close()
