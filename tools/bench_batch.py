"""BASELINE config 5b: a batch of independent DirectSort arrays sharded across the GPUs of one box (no data-path collective).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/bench_batch.py [--arrays 64] [--n 64]
Every rank generates the same keys (same seed) and sorts its share of the batch (shard.sort_batch_sharded); time = max over
ranks of the device time of the share, CUDA events on the engine stream.  Rank 0 prints one JSON line."""
import argparse, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

sf = importlib.import_module("sorting-fhe_b200")
shard = importlib.import_module("sorting-fhe_b200.shard")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arrays", type=int, default=64)
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--log-n", type=int, default=17)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = (3, 3, 2)
    depth, rots = sf.direct_sort_params(a.n, 1 << a.log_n)
    stream = torch.cuda.Stream()
    eng = sf.Engine(a.log_n, depth, 40, seed=5, stream=stream.cuda_stream)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    rng = np.random.default_rng(3)
    arrays = [rng.permutation(a.n) / a.n for _ in range(a.arrays)]
    lo, hi = shard.shard_range(a.arrays, rank, world)
    cts = [eng.encrypt(eng.encode(x, slots=a.n)) if lo <= i < hi else None for i, x in enumerate(arrays)]
    sf.DirectSort(eng, a.n, rots).sort(cts[lo], cfg)            # warm-up (memory pool, tables)
    cts[lo].slots = a.n
    eng.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    sorter = sf.DirectSort(eng, a.n, rots)
    outs = []
    for c in cts[lo:hi]:
        outs.append(sorter.sort(c, cfg))
        eng.sync()
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], device="cuda")
    err = torch.tensor([max(float(np.max(np.abs(eng.decrypt(o, a.n) - np.sort(arrays[lo + i])))) for i, o in enumerate(outs))], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(err, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"workload": "%d independent DirectSort N=%d arrays, ring 2^%d, depth %d" % (a.arrays, a.n, a.log_n, depth),
                          "n_gpus": world, "batch_s": t.item(), "arrays_per_s": a.arrays / t.item(), "ms_per_array_per_gpu": t.item() / (hi - lo) * 1e3,
                          "max_abs_err": err.item(), "collectives_on_data_path": 0}))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


main()
