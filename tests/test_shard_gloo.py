"""N > 1 host logic on CPU: world_size 2 over gloo. Each rank 'polishes' its shard with a stand-in function
(the CUDA path needs a GPU; what is tested here is the sharding, the gather and the input-order merge)."""
import os
import socket

import pytest

from sviper_b200 import shard


def test_shard_indices_partition():
    for n in (0, 1, 7, 100):
        for world in (1, 2, 3, 8):
            seen = sorted(i for r in range(world) for i in shard.shard_indices(n, r, world))
            assert seen == list(range(n))
    with pytest.raises(ValueError):
        shard.shard_indices(5, 2, 2)


def test_merge_in_input_order():
    per_rank = [[("r0", i) for i in shard.shard_indices(7, 0, 3)], [("r1", i) for i in shard.shard_indices(7, 1, 3)],
                [("r2", i) for i in shard.shard_indices(7, 2, 3)]]
    out = shard.merge_in_input_order(7, per_rank)
    assert [x[1] for x in out] == list(range(7))
    with pytest.raises(ValueError):
        shard.merge_in_input_order(7, [per_rank[0], per_rank[1][:-1], per_rank[2]])


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    variants = [{"id": i, "payload": "v%d" % i} for i in range(11)]

    def fake_polish(vs):   # stand-in for Context.polish_batch
        return [{"id": v["id"], "polished": v["payload"].upper(), "rank": rank} for v in vs]

    out = shard.polish_sharded(variants, fake_polish, rank, world)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, out))


def test_polish_sharded_world2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank in (0, 1):
        out = got[rank]
        assert [x["id"] for x in out] == list(range(11))                # input order on every rank
        assert [x["polished"] for x in out] == ["V%d" % i for i in range(11)]
        assert [x["rank"] for x in out] == [i % 2 for i in range(11)]   # round-robin shards
