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


def _gather_worker(rank, world, port, q, shm_dir):
    import numpy as np
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    res = []
    for call in range(3):                                   # several gathers in a row: file names must not collide
        mine = (np.arange(1000 * (rank + 1) + call, dtype=np.int64) * (rank + 7) % 251).astype(np.uint8)
        got = shard.gather_bytes(mine, rank, world, shm_dir=shm_dir)
        res.append(None if got is None else [g.tolist() for g in got])
    empty = shard.gather_bytes(np.zeros(0, np.uint8), rank, world, shm_dir=shm_dir)
    res.append(None if empty is None else [len(g) for g in empty])
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, res))


@pytest.mark.parametrize("world,shm_dir", [(2, None), (3, None), (2, "/nonexistent_dir_for_the_fallback")])
def test_gather_bytes_gloo(world, shm_dir):
    """the host gather of the sharded candidate list: payload through /dev/shm, sizes and completion through gloo; with an
    unusable shared-memory directory every rank falls back to gloo tensors. Same results either way."""
    import numpy as np
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gather_worker, args=(r, world, port, q, shm_dir)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in range(1, world):
        assert got[r] == [None] * 4
    for call in range(3):
        want = [((np.arange(1000 * (r + 1) + call, dtype=np.int64) * (r + 7) % 251).astype(np.uint8)).tolist() for r in range(world)]
        assert got[0][call] == want
    assert got[0][3] == [0] * world
    assert not [f for f in os.listdir("/dev/shm") if f.startswith("svp_gather_%d_" % port)]       # nothing left behind
    assert shard.gather_bytes(np.arange(5, dtype=np.uint8), 0, 1)[0].tolist() == [0, 1, 2, 3, 4]


def _bench_gather_worker(rank, world, port, q):
    """what bench_configs.run_sharded_c3 does after the device call, with made-up result buffers"""
    import ctypes as C
    import sys
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench_configs as bc
    from sviper_b200 import api
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_total = 41
    owners = bc.shard_records(n_total, world)
    own = owners[rank]
    res = (api.Result * len(own))()
    parts = []
    for i, j in enumerate(own):
        rng = np.random.default_rng(1000 + j)                       # record j looks the same whoever owns it
        res[i].status = 0; res[i].polished_len = 50 + j; res[i].n_cigar = 1 + j % 4; res[i].rounds = 3 + j % 2
        res[i].substituted_bases = j; res[i].inserted_bases = 2 * j; res[i].deleted_bases = 7
        parts.append((None, rng.integers(0, 4, 128).astype(np.uint8), None, C.create_string_buffer(b"MIDMI" + b"\0" * 11, 16),
                      rng.integers(1, 99, 16).astype(np.uint32)))
    got = shard.gather_bytes(bc.pack_shard(dict(n=len(own), res=res, parts=parts)), rank, world)
    ok = None
    if rank == 0:
        g = bc.GatheredResults(n_total, owners, got)
        ok = len(g) == n_total
        for j in range(n_total):
            rng = np.random.default_rng(1000 + j)
            pol_w = rng.integers(0, 4, 128).astype(np.uint8)[:50 + j]; cnt_w = rng.integers(1, 99, 16).astype(np.uint32)[:1 + j % 4]
            pol, ops, cnts, k4 = g.record(j)
            ok = ok and np.array_equal(pol, pol_w) and bytes(ops) == b"MIDMI"[:1 + j % 4] and np.array_equal(cnts, cnt_w)
            ok = ok and list(k4) == [3 + j % 2, j, 2 * j, 7]
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


def test_bench_gather_path_world3_gloo():
    """pack_shard -> shard.gather_bytes -> GatheredResults, the host side of bench.py's sharded_candidate_list at N > 1"""
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_bench_gather_worker, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(3))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] is True and got[1] is None and got[2] is None
