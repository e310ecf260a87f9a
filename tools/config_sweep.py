"""tools/config_sweep.py -- the BASELINE.json configs that bench.py's headline line does not carry, on ONE B200:

  configs[4]  FULL size: 1,048,576 reads of 150 bp vs 4,096 windows of 800..2,000 bp through
                             svp_align_batch (score pass), GCUPS per window length and over the whole set, next to
                             the unmodified reference's map_single_read on the host cores (bounded sample)
  configs[2] (C3)            the shard one GPU gets at 8 GPUs (625 mixed DEL/INS 50..1000 bp, -k 400), and --c3-full
                             for all 5,000 variants on one GPU
  configs[3] (C4)            INS 1..10 kb with -k 1000 (windows of 4..13 kb), 25 variants = 1/8 of the 200

Same timing rules as bench.py: HBM-resident reruns, CUDA events, L2 flushed between timed steps, >= 3 warm-ups.
Prints one JSON object; the committed copy is profiles/r1_configs.json.

  python tools/config_sweep.py [--c5-windows 4096] [--c3 625] [--c4 25] [--c3-full]
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def timed_polish(ctx, torch, flush, variants, steps, warmup=3):
    res = ctx.polish_batch(variants, with_final=True)
    for _ in range(warmup):
        ctx.polish_rerun()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(steps):
        flush.zero_()
        e0.record(); ctx.polish_rerun(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    # end to end through the C ABI with host buffers
    pb = ctx.prepare(variants, with_final=True)
    ctx.run_prepared(pb)
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.run_prepared(pb)
    e2e = (time.perf_counter() - t0) / steps
    cells = bench.batch_cells(variants, [r["alignment_rounds"] for r in res])
    return dict(variants=len(variants), ms_per_step=float(np.mean(ms)), variants_per_s=len(variants) / (np.mean(ms) * 1e-3),
                e2e_variants_per_s=len(variants) / e2e, mean_rounds=float(np.mean([r["rounds"] for r in res])),
                mean_alignment_rounds=float(np.mean([r["alignment_rounds"] for r in res])),
                window_len_min=int(min(len(v["consensus"]) for v in variants)),
                window_len_max=int(max(len(v["consensus"]) for v in variants)),
                read_pairs=int(sum(len(v["reads1"]) for v in variants)),
                alignment_gcups=cells / (np.mean(ms) * 1e-3) / 1e9), res


def mixed_batch(n, seed, k=400):
    rng = np.random.default_rng(seed)
    from sviper_b200 import synth
    out = []
    for i in range(n):
        svlen = int(round(math.exp(rng.uniform(math.log(50), math.log(1000)))))
        out.append(synth.make_variant(seed * 100003 + i, "DEL" if i % 2 == 0 else "INS", svlen=svlen, k=k, cov=30,
                                      cons_err=0.05))
    return out


def big_ins_batch(n, seed):
    rng = np.random.default_rng(seed)
    from sviper_b200 import synth
    return [synth.make_variant(seed * 100003 + i, "INS",
                               svlen=int(round(math.exp(rng.uniform(math.log(1000), math.log(10000))))),
                               k=1000, cov=30, cons_err=0.05) for i in range(n)]


def cpu_map_single_read(reads, wins, tw, per_proc, nproc):
    """the unmodified reference's map_single_read (oracle/_ref/ref_harness ALIGN cases), one process per host core"""
    from tests import refcases as rc
    if not os.path.exists(rc.HARNESS):
        return None
    rng = np.random.default_rng(7)
    pick = rng.choice(len(reads), per_proc * nproc, replace=False)
    tmp = tempfile.mkdtemp(prefix="svp_c5_")
    files, cells = [], 0
    q = "?" * 150
    for p in range(nproc):
        path = os.path.join(tmp, "c%d.txt" % p)
        with open(path, "w") as f:
            for t in pick[p::nproc]:
                f.write("ALIGN %s %s %s\n" % (rc.s(wins[tw[t]]), rc.s(reads[t]), q[:len(reads[t])]))
                cells += 2 * len(reads[t]) * len(wins[tw[t]])
        files.append(path)
    t0 = time.perf_counter()
    procs = [subprocess.Popen([rc.HARNESS, f], stdout=subprocess.DEVNULL) for f in files]
    for p in procs:
        p.wait()
    dt = time.perf_counter() - t0
    for f in files:
        os.remove(f)
    os.rmdir(tmp)
    return dict(gcups=cells / dt / 1e9, cores=nproc, kind="reference",
                sample="%d tasks of the same set, %.1f s (includes the harness's case parsing and row dump)" % (len(pick), dt))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c5-windows", type=int, default=4096)
    ap.add_argument("--c3", type=int, default=625)
    ap.add_argument("--c4", type=int, default=25)
    ap.add_argument("--c3-full", action="store_true")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from sviper_b200 import api
    torch.cuda.set_device(0)
    ctx = api.Context(0, torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    probe = ctx.probe_dpx()
    peak = probe["mix"] / 5.0 * 2.0 / 1e9
    out = {"dpx_peak_gcups": peak, "clock_note": "same rules as bench.py (events, L2 flush, resident inputs)"}

    # ---- configs[4] full size ------------------------------------------------------------------
    if args.c5_windows:
        t0 = time.perf_counter()
        reads, wins, tr, tw = bench.make_c5_tasks(args.c5_windows, 256)
        gen_s = time.perf_counter() - t0
        wl = np.array([len(w) for w in wins])

        def run(sel_w):
            ws = np.nonzero(sel_w)[0]
            remap = -np.ones(len(wins), np.int64); remap[ws] = np.arange(len(ws))
            keep = np.nonzero(sel_w[tw])[0]
            r = [reads[i] for i in keep]; w = [wins[i] for i in ws]
            ttw = remap[tw[keep]].astype(np.int32); ttr = np.arange(len(keep), dtype=np.int32)
            t1 = time.perf_counter()
            o = ctx.align_batch(r, w, ttr, ttw, traceback=False)
            first_call_s = time.perf_counter() - t1
            cells = float(sum(2 * 150 * len(w[x]) for x in ttw))
            for _ in range(3):
                ctx.align_rerun(False)
            ms = []
            for _ in range(args.steps):
                flush.zero_()
                ms.append(ctx.align_rerun(False))
            g = cells / (np.mean(ms) * 1e-3) / 1e9
            return dict(tasks=len(keep), cells=cells, kernel_ms=float(np.mean(ms)), gcups=g, frac_of_dpx_peak=g / peak,
                        e2e_first_call_s=first_call_s, forward_fraction=float(np.mean(o["is_rc"] == 0)))
        c5 = {"all": run(np.ones(len(wins), bool)), "generate_s": gen_s}
        for L in (800, 1000, 1200, 1400, 1600, 1800, 2000):
            c5["L=%d" % L] = run(wl == L)
        c5["cpu_baseline"] = cpu_map_single_read(reads, wins, tw, 400, bench.host_cores())
        c5["workload"] = ("BASELINE.json configs[4]: %d reads of 150 bp (1%% substitutions, 0.1%% indels, half reverse-"
                          "complemented) vs %d windows of 800-2,000 bp, score pass of svp_align_batch, HBM-resident; "
                          "cells = 2 x |read| x |window| per task" % (len(reads), len(wins)))
        out["configs4_alignment_microbench"] = c5
        del reads, wins

    # ---- configs[2]: one GPU's shard at 8 GPUs -------------------------------------------------
    if args.c3:
        r, _ = timed_polish(ctx, torch, flush, mixed_batch(args.c3, 3), args.steps)
        r["workload"] = "BASELINE.json configs[2]: the %d-variant shard of the 5,000 mixed DEL/INS (50-1000 bp, -k 400) one GPU polishes at 8 GPUs" % args.c3
        out["configs2_shard"] = r
    if args.c3_full:
        r, _ = timed_polish(ctx, torch, flush, mixed_batch(5000, 3), max(2, args.steps // 2))
        r["workload"] = "BASELINE.json configs[2]: all 5,000 mixed DEL/INS on one GPU"
        out["configs2_full_one_gpu"] = r
    # ---- configs[3]: kb-scale insertions, -k 1000 ----------------------------------------------
    if args.c4:
        r, res = timed_polish(ctx, torch, flush, big_ins_batch(args.c4, 4), args.steps)
        r["workload"] = "BASELINE.json configs[3]: %d INS of 1-10 kb, -k 1000, 30x 2x150 reads (1/8 of the 200)" % args.c4
        out["configs3_large_insertions"] = r
    ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
