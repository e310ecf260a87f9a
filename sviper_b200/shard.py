"""Host-side sharding of candidate variants over the GPUs of one box.

Variants are independent (include/sviper/sviper.h:130-510 touches only its own Variant), so the reference's
OpenMP loop over VCF records (src/sviper.cpp:43-47) becomes: deal the records to the ranks, every rank polishes
its shard on its own GPU, the results are gathered on the HOST of rank 0 into input order (io.h:89-90 writes the
VCF in input order). There is no data-path collective and nothing moves between GPUs; torch.distributed (gloo) only
carries the small per-variant results to rank 0. bench.py --gpus N runs BASELINE.json configs[2] through this module.
"""


def window_work(sv_type, sv_length, flanking_region=400, ref_flank_length=500):
    """Alignment work of a record, estimated from what the candidate VCF says (no reads needed): the flanked consensus is
    2k + 2 * 500 bases long, plus SVLEN for an insertion (sviper.h:318-322, 419-422), and the number of short reads over it
    grows with the same length, so the cells of one alignment round grow with its square."""
    L = 2 * flanking_region + 2 * ref_flank_length + (sv_length if sv_type == "INS" else 0)
    return float(L) * float(L)


def deal(work, world):
    """Deals records to ranks so that the estimated work is even: records in order of decreasing work, back and forth
    over the ranks (snake). Deterministic, so every rank computes the same assignment without communication. Returns one
    ascending index list per rank. (Round-robin by index is what round 1 did; on a list that alternates DEL / INS it puts every
    insertion on the odd ranks: 20 % imbalance at 2-8 GPUs, SCALE run of round 2.)"""
    if world < 1:
        raise ValueError("bad world size")
    order = sorted(range(len(work)), key=lambda i: (-work[i], i))
    shards = [[] for _ in range(world)]
    for pos, i in enumerate(order):
        r = pos % (2 * world)
        shards[r if r < world else 2 * world - 1 - r].append(i)
    return [sorted(s) for s in shards]


def shard_indices(n_items, rank, world, work=None):
    """record indices of this rank (work = None: all records count the same)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return deal(work if work is not None else [1.0] * n_items, world)[rank]


def merge_in_input_order(n_items, per_rank, work=None):
    """per_rank[r] = results of rank r's shard in shard order -> list in input order"""
    world = len(per_rank)
    owners = deal(work if work is not None else [1.0] * n_items, world)
    out = [None] * n_items
    for r, res in enumerate(per_rank):
        if len(owners[r]) != len(res):
            raise ValueError("rank %d returned %d results for %d records" % (r, len(res), len(owners[r])))
        for i, x in zip(owners[r], res):
            out[i] = x
    return out


def polish_sharded(variants, polish_fn, rank, world, work=None, group=None, dst=0):
    """Polishes this rank's shard with polish_fn(list_of_variants) -> list_of_results and gathers all shards on rank dst.
    Returns the results in input order on rank dst, None elsewhere. work: estimated work per variant (window_work)."""
    mine = [variants[i] for i in shard_indices(len(variants), rank, world, work)]
    res = polish_fn(mine) if mine else []
    if world == 1:
        return merge_in_input_order(len(variants), [res], work)
    import torch.distributed as dist
    parts = [None] * world if rank == dst else None
    dist.gather_object(res, parts, dst=dst, group=group)
    return merge_in_input_order(len(variants), parts, work) if rank == dst else None


_gather_calls = 0


def gather_bytes(buf, rank, world, group=None, dst=0, shm_dir=None):
    """Host gather of one uint8 numpy array per rank onto rank dst (all ranks on ONE node, as bench.py --gpus N runs them).
    The payload goes through POSIX shared memory: every other rank writes its array to a file under /dev/shm, rank dst reads
    and removes the files; torch.distributed (gloo) carries only one (size, ok) pair per rank, which is also the signal that
    the files are complete. (The first version sent the arrays as gloo tensors over loopback TCP: ~8 ms for the 9 MB of
    configs[2] at 8 ranks, a fifth of a rank's device time.) If a rank cannot write its file, all ranks see that in the
    exchanged flags and the arrays go through gloo instead. Returns the list of arrays (rank order) on dst, None elsewhere."""
    global _gather_calls
    import numpy as np
    if world == 1:
        return [buf]
    import os
    import torch
    import torch.distributed as dist
    _gather_calls += 1
    shm_dir = shm_dir or os.environ.get("SVP_SHM_DIR", "/dev/shm")
    name = lambda r: os.path.join(shm_dir, "svp_gather_%s_%d_%d" % (os.environ.get("MASTER_PORT", "0"), _gather_calls, r))
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    ok = 1
    if rank != dst:
        try:
            with open(name(rank), "wb") as f:
                f.write(memoryview(buf))
        except OSError:
            ok = 0
    mine = torch.tensor([buf.size, ok], dtype=torch.int64)
    info = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(info, mine, group=group)                    # every file is complete once this returns
    sizes = [int(x[0]) for x in info]
    if all(int(x[1]) == 1 for x in info):
        if rank != dst:
            return None
        out = []
        for r in range(world):
            if r == dst:
                out.append(buf)
                continue
            a = np.fromfile(name(r), dtype=np.uint8)
            os.unlink(name(r))
            if a.size != sizes[r]:
                raise RuntimeError("rank %d wrote %d bytes, announced %d" % (r, a.size, sizes[r]))
            out.append(a)
        return out
    # fallback: the arrays as padded uint8 tensors over gloo
    if rank != dst and ok:
        try:
            os.unlink(name(rank))
        except OSError:
            pass
    cap = max(sizes)
    padded = torch.zeros(cap, dtype=torch.uint8)
    padded[:buf.size] = torch.from_numpy(buf)
    bufs = [torch.zeros(cap, dtype=torch.uint8) for _ in range(world)] if rank == dst else None
    dist.gather(padded, bufs, dst=dst, group=group)
    return [b[:n].numpy() for b, n in zip(bufs, sizes)] if rank == dst else None
