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
