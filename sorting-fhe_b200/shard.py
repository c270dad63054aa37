"""Multi-GPU partitioning of the sorting path (one process per GPU, torch.distributed).

Two shapes shard naturally:
  * a batch of independent sorts        -> `shard_range` over the array index, evaluation keys replicated per GPU,
                                           no data-path collective;
  * DirectSort with num_batch > 1       -> the comparison batches of constructRank (reference src/sort_algo.h:438-492)
                                           and the index-check batches (:713-742) are independent; each rank sums its
                                           batches and the partial ciphertexts are combined by ONE all-reduce each,
                                           replacing the `omp critical` + EvalAddInPlace of :471-472 and :740-741.
The all-reduce is a plain integer SUM of the raw uint64 limbs (NCCL over NVLink on GPUs, gloo in the CPU tests)
followed by one `mod q_limb` pass: residues are < 2^61, so up to 8 addends stay below 2^63, and modular addition is
associative -- the result is bit-identical to the single-GPU sum.  A single DirectSort with num_batch == 1 (N <= 256 at
ring 2^17) is one dependent chain and does not shard: replicas only.
  * MEHP24 multi-ciphertext sortFG      -> the nc (nc + 1) / 2 block comparisons (reference src/mehp24/mehp24_sort.cpp:
    (N > 256 as nc blocks)                 484-522) and the nc target blocks (:554-603) are independent; the comparison
                                           sums are combined by the same limb all-reduce (2 nc - 1 ciphertexts), each
                                           target block is computed by one rank and broadcast to the others.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(num_items, rank, world):
    """contiguous, balanced [lo, hi) share of `num_items` independent units for `rank` of `world`"""
    base, extra = divmod(num_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_limbs(limbs, moduli, group=None):
    """in-place SUM all-reduce of residue rows followed by `mod q_row`.  limbs: int64 tensor [rows, n] holding residues
    in [0, q_row); moduli: one modulus per row.  Device-agnostic (this is the form the gloo tests exercise); on CUDA
    the engine's own kernel does the reduction step, see allreduce_ciphertext."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world > 8:
        raise ValueError("allreduce_limbs: more than 8 addends could overflow 63 bits")
    if world > 1:
        dist.all_reduce(limbs, op=dist.ReduceOp.SUM, group=group)
    q = torch.as_tensor(np.asarray(moduli, dtype=np.int64), device=limbs.device).view(-1, 1)
    limbs.remainder_(q)
    return limbs


class _DevView:
    """zero-copy view of a ciphertext's device limbs for torch (CUDA array interface)"""

    def __init__(self, ptr, words):
        self.__cuda_array_interface__ = {"shape": (words,), "typestr": "<i8", "data": (ptr, False), "version": 2}


def engine_stream(eng):
    """the engine's CUDA stream as a torch stream: collectives are issued with it current, so NCCL orders itself after
    the kernels that produce the limbs and the engine's next kernel after the collective -- no host synchronisation"""
    s = getattr(eng, "_torch_stream", None)
    if s is None:
        s = eng._torch_stream = torch.cuda.ExternalStream(eng.info(9))
    return s


def allreduce_ciphertext(eng, ct, group=None, timing=None):
    """sum a partial ciphertext over all ranks, in place: NCCL all-reduce on the limbs where they lie, enqueued behind the
    engine's stream, then the engine's mod-reduce kernel (ckks_ct_device_ptr / ckks_ct_mod_reduce of include/ckks_b200.h).
    timing: optional list that receives a (start, end) CUDA event pair around the collective (recorded on the engine stream)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return ct
    ptr, words = eng.ct_device_ptr(ct)
    t = torch.as_tensor(_DevView(ptr, words), device="cuda")
    s = engine_stream(eng)
    with torch.cuda.stream(s):
        if timing is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)   # the wait() of the blocking form is a stream wait for NCCL
        if timing is not None:
            e1.record(s)
            timing.append((e0, e1))
    eng.ct_mod_reduce(ct)
    return ct


def direct_sort_sharded(eng, n, rots, ct, sign_cfg, group=None, timing=None, halves=False):
    """DirectSort<N>::sort with the comparison / index-check batches split over the ranks of `group` (every rank holds
    the same input ciphertext and keys and returns the same sorted ciphertext).  Falls back to the plain sort when
    the shape has a single batch."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    from . import DirectSort, num_batches
    nb = num_batches(n, eng.n)
    sorter = DirectSort(eng, n, rots)
    if nb == 1 or (world == 1 and not halves):
        return sorter.sort(ct, sign_cfg)              # halves=True: the shardable halves on one rank (same calls, hence
                                                      # the same zero-cache encryptions and limbs, as the multi-rank run)
    if nb < world:
        raise ValueError("direct_sort_sharded: %d batches cannot occupy %d ranks (use a sub-group)" % (nb, world))
    lo, hi = shard_range(nb, rank, world)                 # balanced, the first nb % world ranks take one batch more
    part = sorter.construct_rank_partial(ct, sign_cfg, lo, hi, seed=(rank == 0))
    rank_ct = sorter.finish_rank(allreduce_ciphertext(eng, part, group, timing))
    part = sorter.index_check_partial(rank_ct, ct, lo, hi, seed=(rank == 0))
    return sorter.finish_index_check(allreduce_ciphertext(eng, part, group, timing))


def sort_batch_sharded(eng, n, rots, cts, sign_cfg, group=None):
    """a batch of independent sorts: rank r sorts cts[lo:hi] of its own; no collective on the data path.
    Returns (lo, [sorted ciphertexts])."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    from . import DirectSort
    lo, hi = shard_range(len(cts), rank, world)
    sorter = DirectSort(eng, n, rots)
    return lo, [sorter.sort(c, sign_cfg) for c in cts[lo:hi]]


def broadcast_ciphertext(eng, ct, src, group=None):
    """every rank ends up with rank `src`'s ciphertext: metadata as a small tensor, limbs broadcast into a fresh handle
    of that shape where they lie on the device.  `ct` is ignored on the other ranks."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return ct
    rank = dist.get_rank(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    meta = torch.zeros(6, dtype=torch.float64, device=dev)
    if rank == src:
        meta = torch.tensor([ct.parts, ct.nlimbs, ct.level, ct.deg, ct.sf, ct.slots], dtype=torch.float64, device=dev)
    dist.broadcast(meta, src, group=group)
    parts, nlimbs, level, deg, sf_, slots = (int(meta[0]), int(meta[1]), int(meta[2]), int(meta[3]), float(meta[4]), int(meta[5]))
    if rank != src:
        ct = eng.import_ct(np.zeros((parts, nlimbs, eng.n), dtype=np.uint64), level, deg, sf_, slots)
    ptr, words = eng.ct_device_ptr(ct)
    t = torch.as_tensor(_DevView(ptr, words), device="cuda")
    with torch.cuda.stream(engine_stream(eng)):
        dist.broadcast(t, src, group=group)
    return ct


def mehp24_sort_large_sharded(eng, ct, total_len, sub_len, sign_cfg, dg_i=None, df_i=2, group=None):
    """mehp24::sortLargeArrayFG with the block comparisons and the target blocks of the multi-ciphertext sortFG split over
    the ranks of `group`.  Every rank holds the same input ciphertext and keys and returns the same sorted ciphertext,
    limb for limb the one `mehp24_sort_large` computes on a single GPU (the partial sums have one level and scale, and
    modular addition is associative)."""
    from . import (mehp24_split, mehp24_combine, mehp24_compare_units, mehp24_block_compare_partial, mehp24_block_ranks,
                   mehp24_block_sorted, mehp24_sort_large)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if dg_i is None:
        dg_i = int((np.log2(total_len) + 1) / 2)
    nc = total_len // sub_len
    if world == 1 or nc == 1:
        return mehp24_sort_large(eng, ct, total_len, sub_len, sign_cfg, dg_i, df_i)
    units = mehp24_compare_units(nc)
    if world > units:
        raise ValueError("mehp24_sort_large_sharded: %d block comparisons do not occupy %d ranks" % (units, world))
    parts = mehp24_split(eng, ct, total_len, sub_len)
    lo, hi = shard_range(units, rank, world)
    cv, ch = mehp24_block_compare_partial(eng, parts, sub_len, sign_cfg, lo, hi)
    for c in cv + ch[1:]:
        allreduce_ciphertext(eng, c, group)
    s = mehp24_block_ranks(eng, cv, ch, sub_len)
    owner = [next(r for r in range(world) if shard_range(nc, r, world)[0] <= j < shard_range(nc, r, world)[1]) for j in range(nc)]
    mine = {j: mehp24_block_sorted(eng, parts, s, sub_len, j, dg_i, df_i) for j in range(nc) if owner[j] == rank}
    blocks = [broadcast_ciphertext(eng, mine.get(j), owner[j], group) for j in range(nc)]
    return mehp24_combine(eng, blocks, sub_len)
