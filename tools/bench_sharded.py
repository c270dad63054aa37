"""Strong-scaling measurement of the multi-GPU paths that HAVE a collective:
  * DirectSort with num_batch > 1 (reference src/sort_algo.h:438-492, 713-742): N = 512 (depth 35, 4 batches) or N = 1024
    (depth 39, 16 batches) at ring 2^17 through shard.direct_sort_sharded -- two limb all-reduces per sort;
  * MEHP24 sortLargeArrayFG (src/mehp24/mehp24_sort.cpp:484-522, 554-603): N = 512 / 1024 as blocks of 256 through
    shard.mehp24_sort_large_sharded -- 2 nc - 1 limb all-reduces and nc broadcasts.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/bench_sharded.py --what direct --size 512

Every rank holds the same keys (same seed) and the same input ciphertext; the result is identical on every rank and --
checked through its SHA-256, printed in the JSON line -- identical to the 1-GPU run.  Time = max over ranks of the
CUDA-event time on the engine stream; the all-reduce share is measured with event pairs around every collective."""
import argparse
import hashlib
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

sf = importlib.import_module("sorting-fhe_b200")
shard = importlib.import_module("sorting-fhe_b200.shard")
SIGN = {256: (3, 4, 2), 512: (3, 4, 2), 1024: (3, 5, 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="direct", choices=["direct", "mehp24"])
    ap.add_argument("--size", dest="n", type=int, default=512)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, cfg = a.n, SIGN[a.n]
    stream = torch.cuda.Stream()
    if a.what == "direct":
        depth, rots = sf.direct_sort_params(n, 1 << 17)
        slots = n
    else:
        depth, rots = sf.mehp24_params(n)
        slots = 1 << 16
    eng = sf.Engine(17, depth, 40, seed=11, stream=stream.cuda_stream)
    eng.KeyGen(); eng.EvalMultKeyGen(); eng.EvalRotateKeyGen(rots)
    x = np.random.default_rng(5).permutation(n) / n
    ct = eng.encrypt(eng.encode(x, slots=slots))
    timing = []

    def run():
        ct.slots = slots
        eng.set_encryption_counter(1)                     # every repetition builds the same sorter (zero cache) on every rank
        if a.what == "direct":
            return shard.direct_sort_sharded(eng, n, rots, ct, cfg, timing=timing, halves=True)
        return shard.mehp24_sort_large_sharded(eng, ct, n, 256, cfg)

    out = run(); eng.sync()                               # warm-up: memory pool, plaintext cache, NCCL communicator
    times = []
    for _ in range(a.reps):
        timing.clear()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = run()
        e1.record(stream)
        eng.sync(); torch.cuda.synchronize()
        times.append((e0.elapsed_time(e1), sum(p.elapsed_time(q) for p, q in timing)))
    best = min(times)
    t = torch.tensor([best[0], best[1]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    got = eng.decrypt(out, slots)[:n]
    digest = hashlib.sha256(np.ascontiguousarray(out.export()).tobytes()).hexdigest()
    if world > 1:
        digs = [None] * world
        dist.all_gather_object(digs, digest)
        assert all(d == digest for d in digs), "ranks disagree on the result limbs"
    if rank == 0:
        print(json.dumps({"what": a.what, "n": n, "n_gpus": world, "depth": depth, "rotation_keys": len(rots), "key_bytes": eng.info(7),
                          "num_batch": sf.num_batches(n, 1 << 17) if a.what == "direct" else None,
                          "ms": float(t[0]), "allreduce_ms": float(t[1]), "allreduce_share": float(t[1] / t[0]) if t[0] else None,
                          "collectives_per_sort": len(timing), "sha256": digest, "level": out.level,
                          "max_abs_err": float(np.max(np.abs(got - np.sort(x))))}))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
