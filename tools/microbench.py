"""Kernel microbenchmarks on one B200: NTT / INTT and hybrid key switch at DirectSort N=128 parameters.
Timed with CUDA events on the engine's stream (a torch stream handed to the engine)."""
import importlib, json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sf = importlib.import_module("sorting-fhe_b200")

def timeit(stream, fn, iters=20, warm=3):
    for _ in range(warm): fn()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(iters): fn()
        e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us

def main():
    log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 17
    depth = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    stream = torch.cuda.Stream()
    eng = sf.Engine(log_n, depth, 40, stream=stream.cuda_stream)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen([1])
    n, Q, K = eng.n, eng.size_q, eng.size_p
    res = {"log_n": log_n, "size_q": Q, "size_p": K, "alpha": eng.alpha}
    for nl in (Q, 24, 16, 8):
        if nl > Q: continue
        buf = torch.randint(0, 1 << 39, (nl, n), dtype=torch.int64, device="cuda")
        out = torch.empty((2, nl, n), dtype=torch.int64, device="cuda")
        t_ntt = timeit(stream, lambda: eng.ntt_dev(buf.data_ptr(), nl, 0, False))
        t_intt = timeit(stream, lambda: eng.ntt_dev(buf.data_ptr(), nl, 0, True))
        t_ks = timeit(stream, lambda: eng.key_switch_dev(buf.data_ptr(), nl, out.data_ptr()))
        beta = -(-nl // eng.alpha)
        ks_bytes = 8 * n * (nl + 2 * beta * (nl + K) + 2 * nl)
        res[f"nl{nl}"] = {"ntt_us": t_ntt, "intt_us": t_intt, "ntt_GBs": 16 * n * nl / t_ntt / 1e3,
                          "ks_us": t_ks, "ks_alg_MB": ks_bytes / 1e6, "ks_GBs": ks_bytes / t_ks / 1e3}
    print(json.dumps(res, indent=1))
    eng.close()
main()
