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


def test_deal_balances_estimated_work():
    """an alternating DEL / INS list (configs[2]): round-robin by index would put every insertion on the odd ranks"""
    import random
    rnd = random.Random(3)
    recs = [("DEL" if i % 2 == 0 else "INS", rnd.randint(50, 1000)) for i in range(5000)]
    work = [shard.window_work(t, n) for t, n in recs]
    for world in (2, 4, 8):
        shards = shard.deal(work, world)
        assert sorted(i for s in shards for i in s) == list(range(5000))
        load = [sum(work[i] for i in s) for s in shards]
        assert max(load) / (sum(load) / world) < 1.01
        naive = [sum(work[i] for i in range(r, 5000, world)) for r in range(world)]
        assert max(naive) / (sum(naive) / world) > 1.15
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
    assert shard.window_work("INS", 1000) > 2 * shard.window_work("DEL", 1000)


def test_merge_in_input_order():
    work = [float(1 + (i * 7) % 5) for i in range(7)]
    per_rank = [[("r%d" % r, i) for i in shard.shard_indices(7, r, 3, work)] for r in range(3)]
    out = shard.merge_in_input_order(7, per_rank, work)
    assert [x[1] for x in out] == list(range(7))
    with pytest.raises(ValueError):
        shard.merge_in_input_order(7, [per_rank[0], per_rank[1][:-1], per_rank[2]], work)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    variants = [{"id": i, "payload": "v%d" % i, "type": "INS" if i % 2 else "DEL", "svlen": 100 + 80 * i} for i in range(11)]
    work = [shard.window_work(v["type"], v["svlen"]) for v in variants]

    def fake_polish(vs):   # stand-in for Context.polish_batch
        return [{"id": v["id"], "polished": v["payload"].upper(), "rank": rank} for v in vs]

    out = shard.polish_sharded(variants, fake_polish, rank, world, work)
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
    assert got[1] is None                                               # gathered on the host of rank 0 only
    out = got[0]
    assert [x["id"] for x in out] == list(range(11))                    # input order
    assert [x["polished"] for x in out] == ["V%d" % i for i in range(11)]
    work = [shard.window_work("INS" if i % 2 else "DEL", 100 + 80 * i) for i in range(11)]
    owners = shard.deal(work, 2)
    assert [x["rank"] for x in out] == [0 if i in owners[0] else 1 for i in range(11)]
    assert {x["rank"] for x in out} == {0, 1}
