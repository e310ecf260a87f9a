"""bench_configs.py -- the other named configs of BASELINE.json for bench.py (configs[2..4] = SURVEY 8d C3 / C4 / C5) and the
whole-program comparison. Split from bench.py for size only; bench.py prints their results as extra keys of its JSON line.

Everything here that executes oracle/ (tests.refcases -> oracle/_ref/ref_harness) does so as the MEASURED CPU baseline.
"""
import math
import os
import subprocess
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
READ_LEN, K_FLANK, COV = 150, 400, 30


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def c3_record(i):
    """what the candidate VCF says about record i of configs[2]: (type, SVLEN)"""
    rng = np.random.default_rng(3 * 1000003 + i)
    return ("DEL" if i % 2 == 0 else "INS"), int(round(math.exp(rng.uniform(math.log(50), math.log(1000)))))


def c3_variant(i):
    """configs[2]: 5,000 mixed DEL / INS of 50..1000 bp (log-uniform), -k 400; record i of the candidate list"""
    from sviper_b200 import synth
    kind, svlen = c3_record(i)
    return synth.make_variant(3 * 100003 + i, kind, svlen=svlen, k=K_FLANK, cov=COV, cons_err=0.05)


def shard_records(n_total, world):
    """which records of configs[2] each rank polishes: sviper_b200.shard.deal on the work the candidate VCF lets one estimate"""
    from sviper_b200 import shard
    return shard.deal([shard.window_work(*c3_record(i), flanking_region=K_FLANK) for i in range(n_total)], world)


def c4_variant(i):
    """configs[3]: insertions of 1..10 kb (log-uniform), -k 1000: windows of 4..13 kb"""
    from sviper_b200 import synth
    rng = np.random.default_rng(4 * 1000003 + i)
    svlen = int(round(math.exp(rng.uniform(math.log(1000), math.log(10000)))))
    return synth.make_variant(4 * 100003 + i, "INS", svlen=svlen, k=1000, cov=COV, cons_err=0.05)


def generate(fn, indices, procs):
    """synthetic variants from forked worker processes (call this BEFORE the process touches CUDA / NCCL)"""
    indices = list(indices)
    if procs <= 1 or len(indices) < 64:
        return [fn(i) for i in indices]
    import multiprocessing as mp
    with mp.get_context("fork").Pool(procs) as pool:
        return pool.map(fn, indices, chunksize=max(1, len(indices) // (procs * 8)))


def algorithmic_cells(variants, res):
    """2 strands x |read| x |window| per read per alignment round (mapping.h:111-119)"""
    return float(sum(2 * READ_LEN * len(v["consensus"]) * (sum(1 for x in v["reads1"] if len(x)) + sum(1 for x in v["reads2"] if len(x)))
                     * r["alignment_rounds"] for v, r in zip(variants, res)))


def cpu_align_gcups(reads, wins, task_read, task_win, nproc, max_tasks_per_proc):
    """The UNMODIFIED reference's map_single_read (mapping.h:84-137: both strands, full matrix, traceback) on a sample of
    (read, window) tasks, one ref_harness process per host core. None when oracle/_ref is absent."""
    import tempfile
    from tests import refcases as rc
    if not os.path.exists(rc.HARNESS):
        return None
    nproc = max(1, nproc)
    n = min(len(task_read), nproc * max_tasks_per_proc)
    pick = np.linspace(0, len(task_read) - 1, n).astype(int)
    tmp = tempfile.mkdtemp(prefix="svp_align_")
    files, cells = [], 0
    for p in range(nproc):
        path = os.path.join(tmp, "align%d.txt" % p)
        with open(path, "w") as f:
            for t in pick[p::nproc]:
                r, w = reads[task_read[t]], wins[task_win[t]]
                f.write("ALIGN %s %s %s\n" % (rc.s(w), rc.s(r), "I" * len(r)))
                cells += 2 * len(r) * len(w)
        files.append(path)
    t0 = time.perf_counter()
    procs = [subprocess.Popen([rc.HARNESS, f], stdout=subprocess.DEVNULL) for f in files]
    for p in procs:
        p.wait()
    dt = time.perf_counter() - t0
    for f in files:
        os.remove(f)
    os.rmdir(tmp)
    return dict(gcups=cells / dt / 1e9, cells=cells, seconds=dt, processes=nproc, tasks=int(n))


def timed_reruns(ctx, flush, n):
    """n HBM-resident passes (svp_polish_rerun), CUDA events on the launching stream, L2 flushed before each"""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(n):
        flush.zero_()
        e0.record(); ctx.polish_rerun(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return ms


class GpuBusySampler(threading.Thread):
    """nvidia-smi utilization.gpu of one GPU, four times a second"""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.vals, self.stop_flag = gpu_index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.check_output(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=utilization.gpu",
                                             "--format=csv,noheader,nounits"], text=True, timeout=5)
                self.vals.append(float(o.strip()))
            except Exception:
                pass
            time.sleep(0.25)

    def mean(self):
        return float(np.mean(self.vals)) if self.vals else None

    def peak(self):
        return float(np.max(self.vals)) if self.vals else None


def _result_fields(res, n, names):
    """int32 fields of a ctypes array of svp_result structs as numpy columns (one strided view per field, no per-record
    Python objects); falls back to attribute access for plain objects (tests)."""
    import ctypes as C
    if n and isinstance(res, C.Array):
        raw = np.frombuffer(res, np.uint8, n * C.sizeof(res._type_)).reshape(n, C.sizeof(res._type_))
        return [raw[:, getattr(res._type_, f).offset:getattr(res._type_, f).offset + 4].copy().view(np.int32).ravel() for f in names]
    return [np.fromiter((getattr(res[i], f) for i in range(n)), np.int32, n) for f in names]


def pack_shard(pb):
    """What a rank hands to the gather, straight from the caller-side result buffers of a prepared batch: one uint8 array =
    header | polished lengths | CIGAR lengths | counters | polished bases | CIGAR ops | CIGAR counts."""
    n = pb["n"]
    status, plen, ncig, rounds, sub, ins, dele = _result_fields(pb["res"], n, ("status", "polished_len", "n_cigar", "rounds",
                                                                           "substituted_bases", "inserted_bases", "deleted_bases"))
    if np.any(status != 0):
        raise RuntimeError("a variant of the shard failed")
    cnt4 = np.stack([rounds, sub, ins, dele], axis=1).astype(np.int32).ravel() if n else np.zeros(0, np.int32)
    parts = pb["parts"]
    pol = np.concatenate([parts[i][1][:plen[i]] for i in range(n)]) if n else np.zeros(0, np.uint8)
    ops = np.concatenate([np.frombuffer(parts[i][3], np.uint8, int(ncig[i])) for i in range(n)]) if n else np.zeros(0, np.uint8)
    cnts = np.concatenate([parts[i][4][:ncig[i]] for i in range(n)]) if n else np.zeros(0, np.uint32)
    head = np.array([n, len(pol), len(ops)], np.int64)
    return np.concatenate([head.view(np.uint8), plen.view(np.uint8), ncig.view(np.uint8), cnt4.view(np.uint8), pol, ops,
                           cnts.astype(np.uint32).view(np.uint8)])


class Shard:
    """a packed shard, unpacked as views (structure of arrays): record i = (polished bases, CIGAR ops, CIGAR counts, counters)"""

    def __init__(self, buf):
        n, npol, nops = (int(x) for x in buf[:24].view(np.int64))
        o = 24
        self.n = n
        self.plen = buf[o:o + 4 * n].view(np.int32); o += 4 * n
        self.ncig = buf[o:o + 4 * n].view(np.int32); o += 4 * n
        self.cnt4 = buf[o:o + 16 * n].view(np.int32).reshape(n, 4); o += 16 * n
        self.pol = buf[o:o + npol]; o += npol
        self.ops = buf[o:o + nops]; o += nops
        self.cnts = buf[o:o + 4 * nops].view(np.uint32)
        self.pe, self.ce = np.cumsum(self.plen, dtype=np.int64), np.cumsum(self.ncig, dtype=np.int64)

    def record(self, i):
        pe, ce, pl, nc = self.pe[i], self.ce[i], self.plen[i], self.ncig[i]
        return self.pol[pe - pl:pe], self.ops[ce - nc:ce], self.cnts[ce - nc:ce], self.cnt4[i]


def unpack_shard(buf):
    """-> list of (polished bases, CIGAR ops, CIGAR counts, (rounds, subst, ins, del)) views, in shard order"""
    sh = Shard(buf)
    return [sh.record(i) for i in range(sh.n)]


class LocalResults:
    """world size 1: nothing to gather, the result buffers of the call are already in input order"""

    def __init__(self, pb):
        self.pb = pb
        status, self.plen, self.ncig = _result_fields(pb["res"], pb["n"], ("status", "polished_len", "n_cigar"))
        if np.any(status != 0):
            raise RuntimeError("a variant of the shard failed")

    def __len__(self):
        return self.pb["n"]

    def record(self, j):
        r, part = self.pb["res"][j], self.pb["parts"][j]
        return (part[1][:self.plen[j]], np.frombuffer(part[3], np.uint8, int(self.ncig[j])), part[4][:self.ncig[j]],
                (r.rounds, r.substituted_bases, r.inserted_bases, r.deleted_bases))


class GatheredResults:
    """The shards of all ranks on the host of rank 0 with an index in INPUT order (the reference writes its VCF from one vector
    in input order, io.h:89-90): record j of the candidate list = shard rank_of[j], position local_of[j]. Built with array
    operations only -- 5,000 per-record Python objects cost more than a rank's whole device time at 8 GPUs."""

    def __init__(self, n_total, owners, bufs):
        self.shards = [Shard(b) for b in bufs]
        self.rank_of = np.full(n_total, -1, np.int32)
        self.local_of = np.zeros(n_total, np.int32)
        for r, (own, sh) in enumerate(zip(owners, self.shards)):
            own = np.asarray(own, np.int64)
            if len(own) != sh.n:
                raise ValueError("rank %d returned %d results for %d records" % (r, sh.n, len(own)))
            self.rank_of[own] = r
            self.local_of[own] = np.arange(len(own), dtype=np.int32)
        if np.any(self.rank_of < 0):
            raise ValueError("records without a result")

    def __len__(self):
        return len(self.rank_of)

    def record(self, j):
        return self.shards[self.rank_of[j]].record(self.local_of[j])


def run_sharded_c3(ctx, rank, world, flush, gloo, n_total, shard, t_gen, barrier, max_over_ranks, sum_over_ranks):
    """BASELINE.json configs[2] as north_star states it: ONE candidate list of n_total mixed DEL / INS records, dealt to the
    ranks by shard_records() (balanced by the work the VCF lets one estimate), no data-path collective; rank 0 gathers
    every rank's results on the HOST and puts them into input order (the reference writes its VCF in input order from one
    vector, io.h:89-90). Strong scaling: total work is fixed. Timed end to end -- host buffers in, results gathered on
    rank 0 -- and, separately, HBM-resident (device events, max over ranks)."""
    import torch
    import torch.distributed as dist
    from sviper_b200 import shard as shardlib
    owners = shard_records(n_total, world)
    assert len(owners[rank]) == len(shard)
    ctx.set_lanes(4)
    pb = ctx.prepare(shard, with_final=True)
    ctx.run_prepared(pb)                                   # warm-up: sizes every arena for this shard
    barrier(world)
    times, gathered = [], None
    for _ in range(2):
        flush.zero_(); torch.cuda.synchronize()
        barrier(world)
        t0 = time.perf_counter()
        # host buffers -> device -> host buffers, 4 lanes, one blocking call for the shard (queueing the shard variant by
        # variant through the asynchronous scheduler measured the same: 306 vs 313 ms for 5,000 records, 40.0 vs 39.9 ms for 625)
        ctx.run_prepared(pb)
        t_dev = time.perf_counter() - t0
        if world > 1:
            # host gather to rank 0: payload through shared memory, sizes / completion through gloo (shard.gather_bytes)
            parts = shardlib.gather_bytes(pack_shard(pb), rank, world, group=gloo)
            if rank == 0:
                gathered = GatheredResults(n_total, owners, parts)      # input order
        else:
            gathered = LocalResults(pb)       # one rank: the caller's result buffers ARE the input-ordered vector
        times.append((time.perf_counter() - t0, t_dev))
    e2e_s = max_over_ranks(min(t[0] for t in times), world)
    # HBM-resident passes (every round + final alignment from resident inputs), device events
    ctx.polish_rerun()
    mine_ms = float(np.mean(timed_reruns(ctx, flush, 3)))
    all_ms = [None] * world
    if world > 1:
        dist.all_gather_object(all_ms, mine_ms, group=gloo)
    else:
        all_ms = [mine_ms]
    res = ctx.collect(pb)
    tot_cells = sum_over_ranks(algorithmic_cells(shard, res), world)
    if rank != 0:
        return None
    # (outside the timed region) what was gathered is what this rank's library call returned, record by record
    for local, j in enumerate(owners[0]):
        pol, ops, cnts, k4 = gathered.record(j)
        r = res[local]
        assert np.array_equal(pol, r["polished"]) and list(k4) == [r["rounds"], r["subst"], r["ins"], r["dele"]]
        assert [(chr(o), int(c)) for o, c in zip(ops, cnts)] == r.get("cigar", [])
    return {"workload": "BASELINE.json configs[2]: one candidate list of %d mixed DEL/INS (50-1000 bp, 30x 2x150 bp, -k 400), "
                        "dealt to the ranks in order of decreasing estimated work (snake), results gathered on the host of rank 0 into "
                        "input order" % n_total,
            "scaling": "strong", "variants": n_total, "n_gpus": world,
            "e2e_variants_per_s": n_total / e2e_s, "e2e_seconds": e2e_s,
            "e2e_includes": "staging + H2D + all kernels + D2H per rank; N > 1: packing the shard's results into one byte array, "
                            "host gather to rank 0 (payload through /dev/shm, sizes and completion through gloo), input-order "
                            "index over the gathered arrays (N = 1: the call's result buffers are already the input-ordered vector)",
            "gather_seconds_rank0": times[-1][0] - times[-1][1],
            "resident_variants_per_s": n_total / (max(all_ms) * 1e-3), "resident_ms_per_rank": all_ms,
            "rank_imbalance_max_over_mean": max(all_ms) / (sum(all_ms) / len(all_ms)),
            "alignment_gcups_resident": tot_cells / (max(all_ms) * 1e-3) / 1e9,
            "mean_alignment_rounds_rank0": float(np.mean([r["alignment_rounds"] for r in res])),
            "generation_seconds_rank0": t_gen, "gathered_records": len(gathered)}


def run_c4(ctx, flush, vs, t_gen, cores):
    """configs[3]: insertions of 1-10 kb with -k 1000 (windows of 4-13 kb, final alignment ~3 k x 13 k), one GPU"""
    import torch
    n = len(vs)
    ctx.set_lanes(4)
    pb = ctx.prepare(vs, with_final=True)
    ctx.run_prepared(pb)
    flush.zero_(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    ctx.run_prepared(pb)
    e2e_s = time.perf_counter() - t0
    res = ctx.collect(pb)
    ctx.polish_rerun()
    ms = float(np.mean(timed_reruns(ctx, flush, 2)))
    cells = algorithmic_cells(vs, res)
    out = {"workload": "BASELINE.json configs[3]: %d INS of 1-10 kb (log-uniform), -k 1000, 30x 2x150 bp; windows %d-%d bp, "
                       "%d read pairs" % (n, min(len(v["consensus"]) for v in vs), max(len(v["consensus"]) for v in vs),
                                          sum(len(v["reads1"]) for v in vs)),
           "variants_per_s": n / (ms * 1e-3), "ms_per_pass": ms, "e2e_variants_per_s": n / e2e_s,
           "alignment_gcups": cells / (ms * 1e-3) / 1e9,
           "mean_alignment_rounds": float(np.mean([r["alignment_rounds"] for r in res])), "generation_seconds": t_gen}
    # CPU baseline: one polish_to_perfection of such a variant is ~10 minutes of one core (521 s), so the bounded
    # sample is the reference's map_single_read on reads x windows of this shape; variants/s is derived from its GCUPS
    reads, wins, tr, tw = [], [], [], []
    for w, v in enumerate(vs[:cores]):
        wins.append(v["consensus"])
        for r in v["reads1"][:600]:
            tr.append(len(reads)); tw.append(w); reads.append(r)
    cpu = cpu_align_gcups(reads, wins, tr, tw, cores, 600)     # ~10-20 s on all cores
    if cpu:
        out["cpu_baseline"] = {"value": cpu["gcups"] * 1e9 / (cells / n), "unit": "variants/s", "cores": cpu["processes"],
                               "kind": "reference",
                               "sample": "map_single_read of the unmodified reference on %d (read, window) tasks of this config, "
                                         "%.1f s, %.3f GCUPS on all cores; variants/s = GCUPS / algorithmic cells per variant "
                                         "(alignment is >97%% of the reference's time)" % (cpu["tasks"], cpu["seconds"], cpu["gcups"])}
    return out


def run_c5_full(ctx, flush, make_c5_tasks, cores, peak_gcups):
    """configs[4] at FULL size: 1,048,576 reads of 150 bp vs 4,096 windows of 800-2,000 bp, score pass (svp_align_batch)"""
    t0 = time.perf_counter()
    reads, wins, tr, tw = make_c5_tasks(4096, 256)
    t_gen = time.perf_counter() - t0
    cells = float(2 * READ_LEN * sum(len(wins[w]) for w in tw))
    ctx.align_batch(reads, wins, tr, tw, traceback=False)
    ctx.align_rerun(False)
    ms = []
    for _ in range(3):
        flush.zero_()
        ms.append(ctx.align_rerun(False))
    ms = float(np.mean(ms))
    out = {"workload": "BASELINE.json configs[4] at full size: %d reads of 150 bp vs %d windows of 800-2,000 bp, score pass of "
                       "svp_align_batch, HBM-resident" % (len(tr), len(wins)),
           "tasks": len(tr), "cells": cells, "kernel_ms": ms, "gcups": cells / (ms * 1e-3) / 1e9,
           "frac_of_dpx_peak": cells / (ms * 1e-3) / 1e9 / peak_gcups if peak_gcups else None, "generation_seconds": t_gen}
    cpu = cpu_align_gcups(reads, wins, tr, tw, cores, 4000)    # ~10-20 s on all cores
    if cpu:
        out["cpu_baseline"] = {"value": cpu["gcups"], "unit": "GCUPS", "cores": cpu["processes"], "kind": "reference",
                               "sample": "map_single_read of the unmodified reference on %d of these tasks, one process per "
                                         "host core, %.1f s" % (cpu["tasks"], cpu["seconds"])}
    return out


def run_whole_program(n_records, cores, local_gpu):
    """The programs a user runs, on files: the UNMODIFIED reference binary (oracle/_ref/sviper, all host cores) against the
    pipeline program (sviper_b200/host/sviper_b200: same reference sources + shims, OpenMP loop replaced by the prepare /
    device / finish pipeline) on a synthetic contig of configs[1]'s shape: n_records DELs 7 kb apart, 30x 2x150 bp short
    reads, 20x long reads at 10% error, written as FASTA / BAM / VCF. Wall seconds include process start, CUDA context
    creation and all file I/O."""
    import shutil
    import tempfile
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "sviper")
    gpu_bin = os.path.join(ROOT, "sviper_b200", "host", "sviper_b200")
    if not (os.path.exists(ref_bin) and os.path.exists(gpu_bin)):
        return {"unavailable": "oracle/_ref/sviper or sviper_b200/host/sviper_b200 not built"}
    from tests import e2e_dataset
    tmp = tempfile.mkdtemp(prefix="svp_wp_")
    try:
        t0 = time.perf_counter()
        d = e2e_dataset.make_c2(tmp, n_records=n_records, seed=2)
        t_gen = time.perf_counter() - t0
        base = ["-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", "400", "-t", str(cores)]
        env = dict(os.environ, SVIPER_B200_STATS="1", SVIPER_B200_DEVICES=str(local_gpu))
        busy = GpuBusySampler(local_gpu); busy.start()
        t0 = time.perf_counter()
        p = subprocess.run([gpu_bin] + base + ["-o", os.path.join(tmp, "gpu")], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                           text=True, env=env)
        t_gpu = time.perf_counter() - t0
        busy.stop_flag = True; busy.join(timeout=3)
        if p.returncode != 0:
            return {"unavailable": "sviper_b200 exited with %d: %s" % (p.returncode, p.stderr[-300:])}
        stats = {}
        for line in p.stderr.split("\n"):
            if line.startswith("[sviper_b200] variants="):
                stats = dict(kv.split("=") for kv in line.split()[1:])
        t0 = time.perf_counter()
        subprocess.check_call([ref_bin] + base + ["-o", os.path.join(tmp, "ref")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        t_ref = time.perf_counter() - t0
        recs = lambda path: [x for x in open(path) if not x.startswith("#")]
        gpu_rows = recs(os.path.join(tmp, "gpu.vcf"))
        ref_rows = recs(os.path.join(tmp, "ref.vcf"))
        rows = [x.split("\t") for x in ref_rows]
        same = gpu_rows == ref_rows
        differing = [rows[i][2] for i, (a, b) in enumerate(zip(gpu_rows, ref_rows)) if a != b]
        # What the reference writes for a record depends on which records the SAME thread's BAM handles read before it
        # (profiles/r2_reference_determinism.txt, single-threaded: sv7 alone QUAL 95, after sv6 96.67, after sv50 98.33), so
        # with several threads -- records dealt to threads by timing -- neither program repeats its own output exactly. The
        # strict comparison therefore runs both programs with ONE thread on the first records of the same list.
        n_strict = min(12, len(ref_rows))
        sub = os.path.join(tmp, "strict.vcf")
        with open(sub, "w") as f:
            k = 0
            for line in open(d["vcf"]):
                if line.startswith("#") or k < n_strict:
                    f.write(line)
                    k += 0 if line.startswith("#") else 1
        sbase = ["-c", sub] + base[2:-1] + ["1"]
        t0 = time.perf_counter()
        ps = subprocess.run([gpu_bin] + sbase + ["-o", os.path.join(tmp, "sgpu")], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                            text=True, env=env)
        ts_gpu = time.perf_counter() - t0
        if ps.returncode != 0:
            return {"unavailable": "sviper_b200 -t 1 exited with %d: %s" % (ps.returncode, ps.stderr[-300:])}
        t0 = time.perf_counter()
        subprocess.check_call([ref_bin] + sbase + ["-o", os.path.join(tmp, "sref")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        ts_ref = time.perf_counter() - t0
        strict = {"what": "both programs with -t 1 on the first %d records of the list (fixed order of records per BAM handle)" % n_strict,
                  "records": n_strict, "vcf_identical": recs(os.path.join(tmp, "sgpu.vcf")) == recs(os.path.join(tmp, "sref.vcf")),
                  "reference_seconds": ts_ref, "gpu_program_seconds": ts_gpu}
        determinism = {"records_differing_multi_threaded": differing,
                       "note": "the reference's line for a record depends on the records its thread read before; compare "
                               "strict_single_thread, and profiles/r2_reference_determinism.txt"}
        return {"workload": "configs[1] shape as files: %d DELs of 50-1000 bp on a %d bp contig, 30x 2x150 bp + 20x long reads at "
                            "10%% error, -k 400, -t %d; FASTA/BAM/VCF written by tests/e2e_dataset.py" % (n_records, d["contig_len"], cores),
                "reference_seconds": t_ref, "gpu_program_seconds": t_gpu, "speedup": t_ref / t_gpu,
                "reference_variants_per_s": n_records / t_ref, "gpu_program_variants_per_s": n_records / t_gpu,
                "vcf_identical": same, "determinism": determinism, "strict_single_thread": strict, "records": len(rows), "pass_records": sum(r[6] == "PASS" for r in rows), "host_threads": cores,
                "mean_device_batch": float(stats.get("mean_device_batch", 0)), "device_batches": int(stats.get("device_batches", 0)),
                "peak_parked_variants_per_thread": int(stats.get("peak_parked_per_thread", 0)),
                "gpu_busy_percent_mean": busy.mean(), "gpu_busy_percent_max": busy.peak(), "dataset_seconds": t_gen,
                "note": "what is left on the host is the reference's own long-read MSA (build_consensus) and BAM I/O"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def pin_rank_to_core_slice(local, world):
    """Gives this rank (its Python thread and the library's lane / worker threads, which inherit the mask) its OWN slice of
    the host cores: the cores NVML reports as local to the rank's GPU, divided among the ranks whose GPUs share them.
    (Round 1 'pinned' all ranks to all cores.) Returns the number of cores in the slice."""
    try:
        import pynvml
        pynvml.nvmlInit()
        allowed = sorted(os.sched_getaffinity(0))
        words_n = (max(allowed) + 64) // 64

        def near(g):
            w = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(g), words_n)
            c = [64 * i + b for i, x in enumerate(w) for b in range(64) if (x >> b) & 1]
            c = [x for x in c if x in allowed]
            return tuple(c) if c else tuple(allowed)

        mine = near(local)
        sharing = [g for g in range(world) if near(g) == mine]
        k, n = sharing.index(local), len(sharing)
        per = max(1, len(mine) // n)
        cpus = list(mine[k * per:(k + 1) * per]) or list(mine)
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0
