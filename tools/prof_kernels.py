"""Tiny driver for ncu: a handful of NTT / INTT / key-switch launches at the DirectSort N=128 parameters.
usage: python tools/prof_kernels.py [log_n] [depth] [nl] [what=ntt|ks]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sf = importlib.import_module("sorting-fhe_b200")
log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 17
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 30
nl = int(sys.argv[3]) if len(sys.argv) > 3 else 32
what = sys.argv[4] if len(sys.argv) > 4 else "ntt"
stream = torch.cuda.Stream()
eng = sf.Engine(log_n, depth, 40, stream=stream.cuda_stream)
buf = torch.randint(0, 1 << 39, (nl, eng.n), dtype=torch.int64, device="cuda")
# two warm iterations, then ONE between cudaProfilerStart / Stop (ncu --profile-from-start off sees only that one)
if what == "ks":
    eng.KeyGen(); eng.EvalMultKeyGen()
    out = torch.empty((2, nl, eng.n), dtype=torch.int64, device="cuda")
    for it in range(3):
        if it == 2:
            eng.sync(); torch.cuda.synchronize(); torch.cuda.profiler.start()
        eng.key_switch_dev(buf.data_ptr(), nl, out.data_ptr())
else:
    for it in range(3):
        if it == 2:
            eng.sync(); torch.cuda.synchronize(); torch.cuda.profiler.start()
        eng.ntt_dev(buf.data_ptr(), nl, 0, False)
        eng.ntt_dev(buf.data_ptr(), nl, 0, True)
eng.sync()
torch.cuda.profiler.stop()
print("ok")
eng.close()
