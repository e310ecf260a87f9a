"""bench.py -- variants polished/sec (and alignment GCUPS) on N B200s, next to the reference CPU path.

A "step" is one pass of the hot path (polish_to_perfection for every variant of the batch + the final
convex-gap alignment, include/sviper/sviper.h:426-466) over one batch of synthetic variants of
BASELINE.json configs[1]: DELs of 50..1000 bp (log-uniform), 30x 2x150 short reads, -k 400.
Variants are independent, so for N > 1 every rank gets its own shard of the same shape (weak scaling),
there is no data-path collective; torch.distributed is only used for the barrier and the max-over-ranks.

  python bench.py --gpus N --steps K --warmup W          the CUDA path (this repo)
  python bench.py --impl reference ...                   the reference's own CPU code on the host cores
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

READ_LEN, K_FLANK, COV = 150, 400, 30
K1_DRAM_BYTES_PER_LAUNCH = 1.7e6   # dram__bytes_read.sum + dram__bytes_write.sum, profiles/r1_k1_score_ncu_raw.csv


def make_batch(n_variants, seed):
    """configs[1] shape: DELs 50..1000 bp (log-uniform), 30x 2x150 reads, 5% noisy long-read consensus"""
    from sviper_b200 import synth
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n_variants):
        svlen = int(round(math.exp(rng.uniform(math.log(50), math.log(1000)))))
        v = synth.make_variant(seed * 100003 + i, "DEL", svlen=svlen, k=K_FLANK, cov=COV, cons_err=0.05)
        out.append(v)
    return out


def batch_cells(variants, rounds):
    """algorithmic K1 cells: 2 strands x |read| x |window| per read per round (mapping.h:111-119)"""
    tot = 0
    for v, r in zip(variants, rounds):
        nreads = sum(1 for x in v["reads1"] if len(x)) + sum(1 for x in v["reads2"] if len(x))
        tot += 2 * READ_LEN * len(v["consensus"]) * nreads * r   # window length changes by a few bases per round
    return tot


def make_c5_tasks(n_windows, reads_per_window, seed=5, read_len=READ_LEN):
    """BASELINE.json configs[4] at reduced task count: window lengths from {800..2000}, every read a
    150 bp sample of its window with 1% substitutions and 0.1% single-base indels, half reverse-complemented."""
    rng = np.random.default_rng(seed)
    comp = np.array([3, 2, 1, 0, 4], np.uint8)
    lens = rng.choice([800, 1000, 1200, 1400, 1600, 1800, 2000], n_windows)
    wins = [rng.integers(0, 4, int(L)).astype(np.uint8) for L in lens]
    reads, tw = [], []
    for w, win in enumerate(wins):
        st = rng.integers(0, len(win) - read_len - 8, reads_per_window)
        blk = win[st[:, None] + np.arange(read_len + 8)[None, :]]
        sub = rng.random(blk.shape) < 0.01
        blk = np.where(sub, (blk + 1 + rng.integers(0, 3, blk.shape)) % 4, blk).astype(np.uint8)
        indel = np.nonzero(rng.random(reads_per_window) < 0.14)[0]          # ~0.1% per base over 150 bases
        for i in indel:
            p = int(rng.integers(1, read_len - 1))
            r = blk[i]
            blk[i] = np.concatenate([r[:p], r[p + 1:], r[:1]]) if rng.random() < 0.5 else \
                np.concatenate([r[:p], rng.integers(0, 4, 1).astype(np.uint8), r[p:-1]])
        rd = blk[:, :read_len]
        flip = rng.random(reads_per_window) < 0.5
        rd = np.where(flip[:, None], comp[rd[:, ::-1]], rd)
        reads.extend(np.ascontiguousarray(x) for x in rd)
        tw.extend([w] * reads_per_window)
    return reads, wins, np.arange(len(reads), dtype=np.int32), np.array(tw, np.int32)


class ClockSampler(threading.Thread):
    """samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)"""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows, self.stop_flag = [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                o = subprocess.check_output(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q,
                                             "--format=csv,noheader,nounits"], text=True, timeout=5)
                self.rows.append([x.strip() for x in o.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]),
                "samples": len(self.rows), "reasons": reasons}


# ---------------------------------------------------------------------------------------------
# CPU arms. The oracle / reference harness is executed here ONLY as the measured CPU baseline.
# ---------------------------------------------------------------------------------------------
def cpu_port_run(variants):
    """oracle/liboracle.so (C restatement), OpenMP over variants. Returns (seconds, threads)."""
    from oracle import oracle as orc
    from sviper_b200 import api
    L = orc.lib()
    nv = len(variants)
    keep = []
    cfgs = (orc.OrcConfig * nv)()
    arrs = {k: (C.c_void_p * nv)() for k in ("b1", "q1", "o1", "b2", "q2", "o2", "cons", "ref")}
    npairs = (C.c_int * nv)(); Ls = (C.c_int * nv)(); refL = (C.c_int * nv)()
    for i, v in enumerate(variants):
        r1, o1 = api.pack(v["reads1"]); q1, _ = api.pack(v["quals1"])
        r2, o2 = api.pack(v["reads2"]); q2, _ = api.pack(v["quals2"])
        cons = np.ascontiguousarray(v["consensus"], np.uint8); ref = np.ascontiguousarray(v["ref_part"], np.uint8)
        keep.append((r1, o1, q1, r2, o2, q2, cons, ref))
        c = orc.make_config(v["flank"], v["ins_mean"], v["ins_sd"], api.baseq_mean(v["quals1"], v["quals2"]))
        cfgs[i] = c
        for k, a in zip(("b1", "q1", "o1", "b2", "q2", "o2", "cons", "ref"), (r1, q1, o1, r2, q2, o2, cons, ref)):
            arrs[k][i] = a.ctypes.data
        npairs[i] = len(v["reads1"]); Ls[i] = len(cons); refL[i] = len(ref)
    ncig = (C.c_int * nv)(); rounds = (C.c_int * nv)()
    t0 = time.perf_counter()
    L.orc_polish_many(nv, cfgs, npairs, arrs["b1"], arrs["q1"], arrs["o1"], arrs["b2"], arrs["q2"], arrs["o2"],
                      arrs["cons"], Ls, arrs["ref"], refL, ncig, rounds)
    return time.perf_counter() - t0, L.orc_num_threads(), list(rounds)


def cpu_reference_run(variants, nproc):
    """the UNMODIFIED reference (oracle/_ref/ref_harness over /root/reference's own headers): one process per
    host core, each polishing its share of the variants (the reference parallelises over variants the same
    way, src/sviper.cpp:43-47). Returns seconds."""
    import tempfile
    from tests import refcases as rc
    nproc = max(1, min(nproc, len(variants)))
    tmp = tempfile.mkdtemp(prefix="svp_ref_")
    files = []
    for p in range(nproc):
        path = os.path.join(tmp, "cases%d.txt" % p)
        with open(path, "w") as f:
            for v in variants[p::nproc]:
                rc.write_polish(f, v)
                f.write("FINAL %s @\n" % rc.s(v["ref_part"]))          # '@': the sequence the POLISH above returned
        files.append(path)
    t0 = time.perf_counter()
    procs = [subprocess.Popen([rc.HARNESS, f], stdout=subprocess.DEVNULL) for f in files]
    for p in procs:
        p.wait()
    dt = time.perf_counter() - t0
    for f in files:
        os.remove(f)
    os.rmdir(tmp)
    return dt


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def dist_setup(n_gpus):
    import torch
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world):
    import torch
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world):
    import torch
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def workload_text(n):
    return ("BASELINE.json configs[1]: %d DELs 50-1000bp per GPU, 30x 2x150bp short reads, -k 400 "
            "(window = 500 + ~800 + 500), 5%% noisy long-read consensus, synthetic, fixed seeds; "
            "a step = all polishing rounds + final alignment + CIGAR for the batch" % n)


LANES_NOTE = ("the batch runs as 4 concurrent parts on the GPU (library default, svp_set_lanes); per-kernel times and "
              "the roofline are taken in a separate single-part pass of the same steps")


def run_gpu(args):
    import torch
    import bench_configs as bc
    from sviper_b200 import api
    # synthetic inputs of the other configs come from forked worker processes: before this process touches CUDA / NCCL
    env_rank, env_world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    single_extras = env_world == 1 and not args.quick
    cores_per_rank = max(1, host_cores() // max(1, env_world))
    t0 = time.perf_counter()
    n_c3 = 0 if args.quick else args.c3_variants
    c3_shard = bc.generate(bc.c3_variant, bc.shard_records(n_c3, env_world)[env_rank], min(8, cores_per_rank)) if n_c3 else []
    c3_gen = time.perf_counter() - t0
    t0 = time.perf_counter()
    c4_list = bc.generate(bc.c4_variant, range(args.c4_variants), min(8, cores_per_rank)) if (single_extras and args.c4_variants) else []
    c4_gen = time.perf_counter() - t0
    rank, world, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    gloo = None
    if world > 1:
        import torch.distributed as dist
        gloo = dist.new_group(backend="gloo")
    pinned_cpus = bc.pin_rank_to_core_slice(local, world) if world > 1 else 0
    stream = torch.cuda.current_stream().cuda_stream          # the library launches on torch's current stream
    ctx = api.Context(local, stream)
    # weak scaling with EXACTLY fixed per-GPU work: every rank polishes its own copy of the same batch (nothing is shared
    # between GPUs, so nothing is gained by the repetition). --distinct-shards gives rank r the batch of seed + r instead;
    # those batches differ by +-3.5 % in work and the max-over-ranks then measures the heaviest one.
    variants = make_batch(args.variants, seed=args.seed + (rank if args.distinct_shards else 0))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass

    # ---- warm-up (also makes the batch HBM-resident for the rerun path) ------------------------
    res = None
    for _ in range(max(1, args.warmup)):
        res = ctx.polish_batch(variants, with_final=True)
    rounds = [r["rounds"] for r in res]
    # K1 work: only the rounds that re-align (a no-change round is followed by a vote-only indel round)
    cells = batch_cells(variants, [r["alignment_rounds"] for r in res])
    for _ in range(2):
        ctx.polish_rerun()

    # ---- value: HBM-resident inputs, CUDA events on the launching stream, max over ranks ------
    # L2 policy: a 256 MiB buffer (2x the 126 MB L2) is overwritten between timed steps, outside the event pairs
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sampler = ClockSampler(local); sampler.start()
    ctx.stats(reset=1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(world)
    step_ms = []
    for _ in range(args.steps):
        flush.zero_()
        e0.record()
        ctx.polish_rerun()
        e1.record()
        torch.cuda.synchronize()
        step_ms.append(e0.elapsed_time(e1))
    barrier(world)
    dev_ms = max_over_ranks(sum(step_ms), world)
    n_launch, _ = ctx.stats(reset=1)

    # ---- e2e: host buffers in, results out, through the public C ABI ----------------------------
    # host buffers (the caller's reads / windows / result arrays) are laid out once, as the reference holds them in
    # memory at sviper.h:431. (a) the synchronous call, one svp_polish_batch per step: pack + H2D + all kernels + D2H,
    # nothing overlaps between steps. (b) the asynchronous scheduler a maintainer binds (svp_submit_polish / svp_flush /
    # svp_wait_any): step s + 1 is submitted before the results of step s are waited for, so the lanes run
    # out of phase and the staging / copies of one device batch overlap the kernels of the others. Every step's inputs
    # are copied host -> device and its results device -> host inside the timed region in both; (b) is `e2e.value`.
    pbs = [ctx.prepare(variants, with_final=True) for _ in range(3)]
    pb = pbs[0]
    ctx.run_prepared(pb)
    barrier(world)
    e2e_sync_local = 0.0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.run_prepared(pb)                    # returns after the results are in the caller's host buffers
        e2e_sync_local += time.perf_counter() - t0
    e2e_sync_s = max_over_ranks(e2e_sync_local, world)
    res_e2e = ctx.collect(pb)
    assert all(np.array_equal(a["polished"], b["polished"]) for a, b in zip(res, res_e2e))
    nvar = len(variants)

    def pipelined(steps):
        done = [0] * steps
        for s in range(steps):
            ctx.submit_prepared(pbs[s % 3], tag_base=s * 100000)
            # before the buffers of step s - 2 are reused by step s + 1, all of its results must be back
            while s >= 2 and done[s - 2] < nvar:
                done[ctx.wait_any() // 100000] += 1
        while sum(done) < steps * nvar:
            done[ctx.wait_any() // 100000] += 1
        assert ctx.wait_any() is None

    pipelined(3)                                 # warm-up of the lanes' own staging areas and arenas
    barrier(world)
    flush.zero_(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    pipelined(args.steps)
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    for pbx in pbs[:min(3, args.steps)]:
        assert all(np.array_equal(a["polished"], b["polished"]) and a["cigar"] == b["cigar"] for a, b in zip(res, ctx.collect(pbx)))
    eng = ctx.engine_stats()

    # ---- per-kernel times: the same steps with the batch as ONE part (svp_set_lanes(1)), so that every kernel is
    # timed without kernels of other parts running beside it. These feed `roofline` and `kernel_ms_per_step` (and
    # are what the serialised ncu launch list under profiles/ is compared with); `value` and `e2e` above come from
    # the default configuration, where the parts of the batch overlap.
    ctx.set_lanes(1)
    res1 = ctx.polish_batch(variants, with_final=True)
    assert all(np.array_equal(a["polished"], b["polished"]) for a, b in zip(res, res1))
    for _ in range(2):
        ctx.polish_rerun()
    ctx.stats(reset=1); ctx.stats(reset=2)
    ctx.k1_cells(reset=1); ctx.k1_failed_certificates(reset=1)
    single_ms = []
    for _ in range(args.steps):
        flush.zero_()
        e0.record()
        ctx.polish_rerun()
        e1.record()
        torch.cuda.synchronize()
        single_ms.append(e0.elapsed_time(e1))
    _, kms = ctx.stats(reset=1)
    k1_reg_cells, k1_mix_cells = ctx.k1_cells(reset=1, split=True)   # cells the score kernel computed in the per-kernel pass
    k1_computed = k1_reg_cells + k1_mix_cells
    k1_failed = ctx.k1_failed_certificates(reset=1)
    ctx.stats(reset=3)
    sampler.stop_flag = True; sampler.join(timeout=3)
    clocks = sampler.summary()

    h2d = sum(len(v["consensus"]) + len(v["ref_part"]) + 2 * sum(len(r) for r in v["reads1"]) +
              2 * sum(len(r) for r in v["reads2"]) for v in variants)
    d2h = sum(len(r["polished"]) + 4 * len(r["cov"]) + 5 * r["n_cigar"] + 64 for r in res)
    k5_cells = float(sum(len(v["ref_part"]) * len(r["polished"]) for v, r in zip(variants, res)))
    tot_variants = sum_over_ranks(len(variants), world)
    tot_cells = sum_over_ranks(cells, world)
    k3_bytes = float(sum((6 * (sum(len(x) for x in v["reads1"]) + sum(len(x) for x in v["reads2"])) +
                          57 * len(v["consensus"])) * r["alignment_rounds"] for v, r in zip(variants, res)))

    # ---- configs[4] kernel microbenchmark (score pass only, svp_align_batch / svp_align_rerun) at 1/16 scale ----
    c5 = None
    if rank == 0:
        c5_reads, c5_wins, c5_tr, c5_tw = make_c5_tasks(256, 256)
        ctx.align_batch(c5_reads, c5_wins, c5_tr, c5_tw, traceback=False)
        c5_cells = float(sum(2 * len(c5_reads[t]) * len(c5_wins[w]) for t, w in zip(c5_tr, c5_tw)))
        for _ in range(2):
            ctx.align_rerun(False)
        c5_ms = []
        for _ in range(max(5, args.steps // 2)):
            flush.zero_()
            c5_ms.append(ctx.align_rerun(False))      # CUDA events around the score kernel on the library's stream
        c5 = {"tasks": len(c5_tr), "kernel_ms": float(np.mean(c5_ms)), "gcups": c5_cells / (np.mean(c5_ms) * 1e-3) / 1e9}

    out = None
    if rank == 0:
        probe = ctx.probe_dpx()
        # peak cells/s of the 5-instruction packed recurrence (IADD, 2x VIADDMNMX, VIMNMX3, IADD; 2 cells each)
        peak_gcups = probe["mix"] / 5.0 * 2.0 / 1e9
        # roofline = the dominant kernel, k1_score_kernel<8,19,regular>: the full-batch rounds that align both strands of
        # every read (cells COMPUTED over their own time). Rounds >= 2 run a pair's two known strands as one packed task
        # (half the cells, less than one wave of blocks at 100 variants: latency-bound) and are reported beside it;
        # config.alignment_gcups is the algorithmic figure (2 x |read| x |window| per read and round) over the whole step.
        step_gcups = cells / (sum(single_ms) / args.steps * 1e-3) / 1e9
        k1_gcups_full = k1_reg_cells / (kms["k1_score"] * 1e-3) / 1e9
        # the honest figure: EVERY k1_score launch of the step -- regular rounds + the score pass of the mixed rounds
        # (list building, mixed launch, regular launch beside it, certificate failures) -- cells computed over their time
        k1_ms_all = kms["k1_score"] + kms["k1_score_mixed"]
        k1_gcups = k1_computed / (k1_ms_all * 1e-3) / 1e9
        # pileup/vote kernel: HBM-bound; algorithmic bytes = records read + profile/cov/decision written
        out = {
            "metric": "variants_polished_per_sec", "value": tot_variants * args.steps / (dev_ms * 1e-3),
            "unit": "variants/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "s16x2 (alignment) + f64 (pileup, final alignment)", "data": "synthetic",
            "config": {"workload": workload_text(args.variants),
                       "variants_per_gpu": args.variants, "mean_rounds": float(np.mean(rounds)),
                       "mean_alignment_rounds": float(np.mean([r["alignment_rounds"] for r in res])),
                       "alignment_gcups": tot_cells * args.steps / (dev_ms * 1e-3) / 1e9,
                       "lanes": LANES_NOTE, "host_cpus_pinned_per_rank": pinned_cpus,
                       "shards": ("rank r polishes the batch of seed + r" if args.distinct_shards else
                                  "every rank polishes its own copy of the same batch (fixed per-GPU work)"),
                       "l2_policy": "L2 flushed between timed steps (256 MiB buffer overwritten outside the event pairs)"},
            "e2e": {"value": tot_variants * args.steps / e2e_s, "unit": "variants/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "api": "svp_submit_polish / svp_flush / svp_wait_any, step s+1 submitted before step s is waited for; "
                           "host buffers in, results copied back into host buffers, all inside the timed region",
                    "sync_value": tot_variants * args.steps / e2e_sync_s,
                    "sync_api": "one blocking svp_polish_batch per step",
                    "mean_device_batch": eng["variants"] / max(1, eng["batches"])},
            "gpu_launches": int(n_launch),
            "clocks": clocks,
            "roofline": {"bound": "dpx_int16x2", "achieved": k1_gcups, "peak": peak_gcups, "unit": "GCUPS",
                         "frac": k1_gcups / peak_gcups if peak_gcups else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one k1_score launch (ncu --set full capture
                         # under profiles/): the kernel is compute-bound, its DRAM traffic is negligible
                         "traffic": K1_DRAM_BYTES_PER_LAUNCH,
                         "kernel": "k1_score_kernel (every launch of the step: regular + mixed rounds)",
                         "kernel_ms_per_step": k1_ms_all / args.steps,
                         "cells_computed_per_step": k1_computed / args.steps,
                         "full_wave": {"what": "regular rounds only (both strands of every read, >= 1.7 waves of blocks)",
                                       "gcups": k1_gcups_full, "frac": k1_gcups_full / peak_gcups if peak_gcups else None,
                                       "ms_per_step": kms["k1_score"] / args.steps,
                                       "cells_computed_per_step": k1_reg_cells / args.steps},
                         "whole_step_algorithmic": {"what": "algorithmic cells (2 x |read| x |window| per read per alignment "
                                                            "round) over the whole single-part step, every kernel included",
                                                    "gcups": step_gcups, "frac": step_gcups / peak_gcups if peak_gcups else None},
                         "mixed_rounds": {"what": "score pass of rounds >= 2: pairs with certified strands as one packed task "
                                                  "(k1_score_kernel<.,.,mixed>) + the remaining regular tasks",
                                          "ms_per_step": kms["k1_score_mixed"] / args.steps,
                                          "cells_computed_per_step": k1_mix_cells / args.steps,
                                          "gcups": (k1_mix_cells / (kms["k1_score_mixed"] * 1e-3) / 1e9) if kms["k1_score_mixed"] else None,
                                          "certificates_failed_inside_a_mixed_launch_per_step": k1_failed / args.steps},
                         "score_pass_cells_computed_over_algorithmic": k1_computed / float(cells * args.steps) if cells else None,
                         "timed": "single-part pass of the same steps (svp_set_lanes(1)): %.3f ms per step" %
                                  (sum(single_ms) / args.steps),
                         "peak_source": "svp_probe_dpx measured in this run: %.3e packed lane-ops/s on the 5-op "
                                        "recurrence mix (viaddmax alone %.3e, vimax3 alone %.3e)" %
                                        (probe["mix"], probe["viaddmax"], probe["vimax3"]),
                         "hbm_peak_gbs_measured": peaks.get("hbm_gbs")},
            "kernel_ms_per_step": {k: v / args.steps for k, v in kms.items()},
            # K3 (pileup + vote): algorithmic bytes per alignment round = 6 B of records per aligned read base +
            # 57 B per consensus column written (profile 48, coverage 4, gap flag, decision, new base, old base read);
            # one block per variant, so the kernel is bound by per-variant latency, not by HBM
            "pileup_kernel": {"kernel": "k3_pileup_vote_kernel", "bound": "hbm",
                              "achieved_gbs": k3_bytes * args.steps / (kms["k3_pileup_vote"] * 1e-3) / 1e9,
                              "peak_gbs": peaks.get("hbm_gbs"), "algorithmic_bytes_per_step": k3_bytes,
                              "kernel_ms_per_step": kms["k3_pileup_vote"] / args.steps},
            # K5 (final convex-gap alignment, fp64): cells = |ref_part| x |polished| per variant; the kernel is bound by
            # instruction issue (4 DADD + 4 DSETP + selects per cell), so its yardstick is the measured fp64 add rate
            "final_alignment_kernel": {"kernel": "k5_final_kernel", "bound": "fp64 issue",
                                       "cells_per_step": k5_cells, "kernel_ms_per_step": kms["k5_final"] / args.steps,
                                       "gcells_per_s": k5_cells / (kms["k5_final"] / args.steps * 1e-3) / 1e9,
                                       "dadd_per_cell": 4, "dadd_rate_achieved": 4 * k5_cells / (kms["k5_final"] / args.steps * 1e-3),
                                       "dadd_rate_peak_measured": probe.get("dadd"),
                                       "frac_of_dadd_peak": (4 * k5_cells / (kms["k5_final"] / args.steps * 1e-3)) / probe["dadd"] if probe.get("dadd") else None,
                                       "share_of_step": kms["k5_final"] / sum(kms.values()) if sum(kms.values()) else None},
            "alignment_microbench": dict(c5, frac_of_dpx_peak=c5["gcups"] / peak_gcups if peak_gcups else None,
                                         workload="BASELINE.json configs[4] at 1/16 scale: 65,536 reads of 150 bp vs 256 "
                                                  "windows of 800-2,000 bp, score pass of svp_align_batch, HBM-resident"),
        }
        # ---- cpu_baseline: the reference CPU path on this box's host cores, bounded sample -----------
        cores = host_cores()
        sample = variants[:max(1, min(len(variants), cores))]
        from tests import refcases as rc
        if os.path.exists(rc.HARNESS):
            dt = cpu_reference_run(sample, cores); kind = "reference"; used = min(cores, len(sample))
        else:
            dt, used, _ = cpu_port_run(sample); kind = "port"
        out["cpu_baseline"] = {"value": len(sample) / dt, "unit": "variants/s", "cores": used, "kind": kind,
                               "sample": "%d variants of the same batch, one per host thread, %.1f s" % (len(sample), dt)}
    # ---- the other named configs (extra keys; each bounded so that the default run stays within a few minutes) ----
    if n_c3:
        c3 = bc.run_sharded_c3(ctx, rank, world, flush, gloo, n_c3, c3_shard, c3_gen, barrier, max_over_ranks, sum_over_ranks)
        if out is not None:
            out["sharded_candidate_list"] = c3
        del c3_shard
    if single_extras and out is not None:
        peak_gcups = out["roofline"]["peak"]
        if c4_list:
            out["large_insertions"] = bc.run_c4(ctx, flush, c4_list, c4_gen, host_cores())
            del c4_list
        if args.c5_full:
            out["alignment_microbench_full"] = bc.run_c5_full(ctx, flush, make_c5_tasks, host_cores(), peak_gcups)
    ctx.close()
    if single_extras and out is not None and args.whole_program_records:
        out["whole_program"] = bc.run_whole_program(args.whole_program_records, host_cores(), local)
    if world > 1:
        import torch.distributed as dist
        dist.barrier(); dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from tests import refcases as rc
    cores = host_cores()
    n = max(1, min(args.variants, cores))
    variants = make_batch(args.variants, seed=1000)[:n]
    use_ref = os.path.exists(rc.HARNESS)
    times = []
    for it in range(args.warmup + args.steps):
        if use_ref:
            dt = cpu_reference_run(variants, cores); used = min(cores, n)
        else:
            dt, used, _ = cpu_port_run(variants)
        if it >= args.warmup:
            times.append(dt)
    tot = sum(times)
    val = n * len(times) / tot
    print(json.dumps({
        "impl": "reference", "metric": "variants_polished_per_sec", "value": val, "unit": "variants/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot / len(times) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32 + f64 (CPU)",
        "data": "synthetic",
        "config": {"workload": workload_text(args.variants), "variants_per_gpu": args.variants,
                   "reference_sample": "each step = the first %d variants of that batch, one per host core "
                                       "(polish_to_perfection + final alignment of the unmodified reference)" % n},
        "cpu_baseline": {"value": val, "unit": "variants/s", "cores": used, "kind": "reference" if use_ref else "port",
                         "sample": "%d variants per step" % n},
        "e2e": {"value": val, "unit": "variants/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--seed", type=int, default=1000, help="seed of the synthetic batch")
    ap.add_argument("--distinct-shards", action="store_true", help="rank r polishes the batch of seed + r")
    ap.add_argument("--variants", type=int, default=100, help="variants per GPU per step (configs[1]: 100)")
    ap.add_argument("--quick", action="store_true", help="configs[1] only: skip the extra keys for the other configs")
    ap.add_argument("--c3-variants", type=int, default=5000, help="configs[2]: records of the sharded candidate list (0: skip)")
    ap.add_argument("--c4-variants", type=int, default=200, help="configs[3]: large insertions, N = 1 only (0: skip)")
    ap.add_argument("--no-c5-full", dest="c5_full", action="store_false", help="skip configs[4] at full size (N = 1 only)")
    ap.add_argument("--whole-program-records", type=int, default=100,
                    help="reference binary vs pipeline program on files, N = 1 only (0: skip)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
