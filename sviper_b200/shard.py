"""Host-side sharding of candidate variants over the GPUs of one box.

Variants are independent (include/sviper/sviper.h:130-510 touches only its own Variant), so the reference's
OpenMP loop over VCF records (src/sviper.cpp:43-47) becomes: deal record indices to ranks, every rank polishes
its shard on its own GPU, results are gathered on the host in input order (io.h:89-90 writes the VCF in input
order). There is no data-path collective; torch.distributed is used only to gather the small per-variant
results (any backend: nccl on the GPU box, gloo in the CPU tests)."""


def shard_indices(n_items, rank, world):
    """record indices of this rank: round-robin, so neighbouring (similarly sized) records spread over ranks"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_items, world))


def merge_in_input_order(n_items, per_rank):
    """per_rank[r] = results of shard_indices(n_items, r, world) in shard order -> list in input order"""
    world = len(per_rank)
    out = [None] * n_items
    for r, res in enumerate(per_rank):
        idx = shard_indices(n_items, r, world)
        if len(idx) != len(res):
            raise ValueError("rank %d returned %d results for %d records" % (r, len(res), len(idx)))
        for i, x in zip(idx, res):
            out[i] = x
    return out


def polish_sharded(variants, polish_fn, rank, world, gather_fn=None):
    """Polishes this rank's shard with polish_fn(list_of_variants) -> list_of_results and gathers all shards.
    gather_fn(obj) -> [obj_rank0, ...]; default: torch.distributed.all_gather_object when world > 1."""
    mine = [variants[i] for i in shard_indices(len(variants), rank, world)]
    res = polish_fn(mine) if mine else []
    if world == 1:
        return merge_in_input_order(len(variants), [res])
    if gather_fn is None:
        import torch.distributed as dist

        def gather_fn(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out
    return merge_in_input_order(len(variants), gather_fn(res))
