"""CPU checks of the benchmark's host logic (bench.py / bench_configs.py): everything bench.py uses exists, the candidate list
of configs[2] is dealt evenly, the shard packing for the host gather round-trips, the workload generators produce the shapes
BASELINE.json names. (The timed paths need a GPU.)"""
import ctypes as C
import re

import numpy as np


def test_bench_uses_only_helpers_that_exist():
    import bench
    import bench_configs as bc
    names = set(re.findall(r"\bbc\.([a-zA-Z_0-9]+)", open(bench.__file__).read()))
    assert names and all(hasattr(bc, n) for n in names), sorted(n for n in names if not hasattr(bc, n))


def test_workload_shapes():
    import bench
    import bench_configs as bc
    v = bench.make_batch(2, seed=1000)[0]
    assert v["kind"] == "DEL" and 50 <= v["svlen"] <= 1000 and 1700 <= len(v["consensus"]) <= 1900 and all(len(r) == 150 for r in v["reads1"])
    kinds = [bc.c3_record(i)[0] for i in range(10)]
    assert kinds.count("DEL") == 5 and kinds.count("INS") == 5
    v3 = bc.c3_variant(1)
    assert v3["kind"] == "INS" and v3["svlen"] == bc.c3_record(1)[1]
    v4 = bc.c4_variant(0)
    assert v4["kind"] == "INS" and 1000 <= v4["svlen"] <= 10000 and v4["k"] == 1000 and len(v4["consensus"]) >= 3900
    shards = bc.shard_records(5000, 8)
    assert sorted(i for s in shards for i in s) == list(range(5000)) and {len(s) for s in shards} == {625}


def test_shard_packing_round_trip():
    import bench_configs as bc

    class R:
        pass
    rng = np.random.default_rng(1)
    res, parts = [], []
    for i in range(4):
        r = R()
        r.polished_len, r.n_cigar, r.rounds, r.substituted_bases, r.inserted_bases, r.deleted_bases, r.status = 5 + i, 1 + i, 4, i, 2 * i, 3, 0
        res.append(r)
        parts.append((None, rng.integers(0, 4, 20).astype(np.uint8), None, C.create_string_buffer(b"MDMIM" + b"\0" * 10, 16),
                      rng.integers(1, 99, 16).astype(np.uint32)))
    out = bc.unpack_shard(bc.pack_shard(dict(n=4, res=res, parts=parts)))
    for i, (p, o, c, k) in enumerate(out):
        assert np.array_equal(p, parts[i][1][:5 + i]) and bytes(o) == b"MDMIM"[:1 + i] and np.array_equal(c, parts[i][4][:1 + i])
        assert list(k) == [4, i, 2 * i, 3]
    assert bc.unpack_shard(bc.pack_shard(dict(n=0, res=[], parts=[]))) == []


def test_shard_packing_reads_the_c_structs_and_gathers_in_input_order():
    import bench_configs as bc
    from sviper_b200 import api
    rng = np.random.default_rng(2)
    owners = [[0, 3, 4], [1, 2]]
    bufs, want = [], {}
    for own in owners:
        n = len(own)
        res = (api.Result * n)()
        parts = []
        for i, j in enumerate(own):
            res[i].status = 0; res[i].polished_len = 7 + j; res[i].n_cigar = 1 + j % 3; res[i].rounds = 3 + j
            res[i].substituted_bases = j; res[i].inserted_bases = 2 * j; res[i].deleted_bases = 5
            parts.append((None, rng.integers(0, 4, 32).astype(np.uint8), None, C.create_string_buffer(b"MIDMI" + b"\0" * 10, 16),
                          rng.integers(1, 99, 16).astype(np.uint32)))
            want[j] = (parts[-1][1][:7 + j].copy(), b"MIDMI"[:1 + j % 3], parts[-1][4][:1 + j % 3].copy(), [3 + j, j, 2 * j, 5])
        bufs.append(bc.pack_shard(dict(n=n, res=res, parts=parts)))
    g = bc.GatheredResults(5, owners, bufs)
    assert len(g) == 5
    for j in range(5):
        pol, ops, cnts, k4 = g.record(j)
        assert np.array_equal(pol, want[j][0]) and bytes(ops) == want[j][1] and np.array_equal(cnts, want[j][2]) and list(k4) == want[j][3]
    import pytest
    with pytest.raises(ValueError):
        bc.GatheredResults(6, owners, bufs)


def test_whole_program_comparison_logic(tmp_path, monkeypatch):
    """bench.py's whole_program key with the reference binary standing in for the GPU program (no GPU here): identical VCFs
    are reported as such, multi-threaded and in the strict single-threaded comparison; a perturbed record is reported by both."""
    import os
    import pytest
    import bench_configs as bc
    ref = os.path.join(bc.ROOT, "oracle", "_ref", "sviper")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/sviper not built")
    root = tmp_path / "root"
    (root / "oracle" / "_ref").mkdir(parents=True)
    (root / "sviper_b200" / "host").mkdir(parents=True)
    os.symlink(ref, str(root / "oracle" / "_ref" / "sviper"))
    stand_in = root / "sviper_b200" / "host" / "sviper_b200"
    stand_in.write_text('#!/bin/bash\n%s "$@" || exit 1\nout=""; while [ $# -gt 0 ]; do if [ "$1" = "-o" ]; then out=$2; fi; shift; done\n'
                        'if [ -n "$PERTURB" ]; then sed -i \'0,/\\t9[0-9.]*\\tPASS/s/\\t9[0-9.]*\\tPASS/\\t12.5\\tPASS/\' $out.vcf; fi\n' % ref)
    stand_in.chmod(0o755)
    monkeypatch.setattr(bc, "ROOT", str(root))
    r = bc.run_whole_program(2, 4, 0)
    assert r["vcf_identical"] and r["determinism"]["records_differing_multi_threaded"] == [] and r["pass_records"] == 2 and r["records"] == 7
    assert r["strict_single_thread"]["vcf_identical"] and r["strict_single_thread"]["records"] == 7
    monkeypatch.setenv("PERTURB", "1")
    r = bc.run_whole_program(2, 4, 0)
    assert not r["vcf_identical"] and len(r["determinism"]["records_differing_multi_threaded"]) == 1
    assert not r["strict_single_thread"]["vcf_identical"]
