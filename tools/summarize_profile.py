#!/usr/bin/env python
"""Builds profiles/<tag>_summary.md from
     <tag>_launches.csv   ncu --metrics gpu__time_duration.sum --clock-control none --csv  (bench.py under ncu, SVP_LANES=1)
     <tag>_bench.json     the bench line of the same code without a profiler
     <tag>_<kernel>.ncu-rep (optional)  ncu --set full captures; their raw pages are exported next to the summary
   usage: tools/summarize_profile.py gpurun_out profiles r1
"""
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

BENCH_KEY = {"k1_score_kernel": "k1_score", "k1_band_kernel": "k1_trace", "k1_trace_kernel": "k1_trace",
             "k2_traceback_kernel": "k2_traceback", "k3_pileup_vote_kernel": "k3_pileup_vote",
             "k5_final_kernel": "k5_final", "k5_traceback_kernel": "k5_traceback", "pack_kernel": "pack",
             "window_refresh_kernel": "pack", "k1_score_kernel (mixed rounds)": "k1_score_mixed",
             "mix_plan_kernel": "k1_score_mixed", "plan_kernel": "other", "scan2_fit_kernel": "other", "fit_kernel": "other",
             "DeviceScanKernel": "other", "DeviceScanInitKernel": "other", "count_active_kernel": "k3_pileup_vote",
             "gather_kernel": "other"}
RAW_KEEP = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "sm__warps_active.avg.per_cycle_active", "smsp__issue_active.avg.per_cycle_active",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def short(name):
    m = re.search(r"(?:svp::|cub::|at::native::|at::)?([A-Za-z0-9_]+)(?:<|\()", name)
    return m.group(1) if m else name[:40]


def read_launches(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, seq = None, []
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") == "gpu__time_duration.sum":
                v = float(d["Metric Value"].replace(",", ""))
                u = d["Metric Unit"]
                v = v / 1e3 if u.startswith("n") else (v * 1e3 if u.startswith("m") else v)   # -> us
                seq.append((d["Kernel Name"], v, d.get("Grid Size", "")))
    return seq


def last_step(seq):
    """the launches of the last HBM-resident step: between the two last k5_traceback_kernel launches (a rerun has no gather)"""
    idx = [i for i, (k, _, _) in enumerate(seq) if "k5_traceback_kernel" in k]
    if len(idx) < 2:
        return seq
    return [x for x in seq[idx[-2] + 1: idx[-1] + 1] if "gather_kernel" not in x[0]]


def raw_page(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        return {}, out
    names, units, vals = rows[0], rows[1], rows[2]
    return {n: (v, u) for n, u, v in zip(names, units, vals)}, out


def main():
    src, dst, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    bench = json.loads(open(os.path.join(src, tag + "_bench.json")).read().strip().splitlines()[-1])
    seq = read_launches(os.path.join(src, tag + "_launches.csv"))
    step = last_step(seq)
    agg = collections.OrderedDict()
    mixed_round = False          # the score-pass launches behind the step's first mix_plan_kernel belong to mixed rounds
    for k, v, _ in step:
        name = short(k)
        if name == "mix_plan_kernel":
            mixed_round = True
        if name == "k1_score_kernel" and mixed_round:
            name = "k1_score_kernel (mixed rounds)"
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    kms = bench["kernel_ms_per_step"]
    ksum = sum(kms.values())
    by_bench = collections.defaultdict(float)
    for k, (n, us) in agg.items():
        if k in BENCH_KEY:
            by_bench[BENCH_KEY[k]] += us
    L = []
    L.append("# ncu launch list vs in-bench CUDA-event timing (%s)\n" % tag)
    L.append("`%s_launches.csv`: `SVP_LANES=1 ncu --metrics gpu__time_duration.sum --clock-control none` over "
             "`python bench.py --steps 2 --warmup 1` (serialised, cold-cache per-launch times; the batch as one part, which is "
             "also how the bench's per-kernel pass runs). The table is ONE HBM-resident step (`svp_polish_rerun`: the launches "
             "between the last two `k5_traceback_kernel`s). `%s_bench.json`: the bench line of the same code without a "
             "profiler. Shares must agree, absolutes need not.\n" % (tag, tag))
    L.append("| kernel | launches/step | ncu us/step | ncu share | bench key | bench ms/step (events) | bench share |")
    L.append("|---|---|---|---|---|---|---|")
    seen = set()
    for k, (n, us) in agg.items():
        bk = BENCH_KEY.get(k, "")
        bcol = ""
        bshare = ""
        if bk and bk not in seen:
            seen.add(bk)
            bcol = "%.3f" % kms[bk]
            bshare = "%.1f%% (ncu for this key: %.1f%%)" % (100 * kms[bk] / ksum, 100 * by_bench[bk] / tot)
        L.append("| %s | %d | %.1f | %.1f%% | %s | %s | %s |" % (k, n, us, 100 * us / tot, bk, bcol, bshare))
    L.append("| total | %d | %.1f | 100%% | | %.3f | |" % (sum(a[0] for a in agg.values()), tot, ksum))
    L.append("")
    L.append("Bench line (`%s_bench.json`): value %.0f %s, %.3f ms/step, e2e %.0f, roofline.frac %.3f (%s), gpu_launches %d.\n"
             % (tag, bench["value"], bench["unit"], bench["ms_per_step"], bench["e2e"]["value"], bench["roofline"]["frac"],
                bench["roofline"]["kernel"], bench["gpu_launches"]))
    reps = [f for f in sorted(os.listdir(src)) if f.startswith(tag + "_") and f.endswith(".ncu-rep")]
    if not reps:
        # no new full captures: keep the sections of the existing summary (their raw pages are already under dst)
        old = os.path.join(dst, tag + "_summary.md")
        if os.path.exists(old):
            txt = open(old).read()
            at = txt.find("## `")
            if at >= 0:
                L.append("(full captures below: taken before the mixed rounds were added; the regular kernel's inner loop and "
                         "register count are unchanged)\n")
                L.append(txt[at:].rstrip("\n"))
                L.append("")
    for f in reps:
        if True:
            met, raw = raw_page(os.path.join(src, f))
            name = f[:-len(".ncu-rep")]
            open(os.path.join(dst, name + "_ncu_raw.csv"), "w").write(raw)
            L.append("## `%s_ncu_raw.csv` (`ncu --set full --clock-control none --import-source on`, one launch)\n" % name)
            L.append("| metric | value |")
            L.append("|---|---|")
            for m in RAW_KEEP:
                if m in met:
                    L.append("| %s | %s %s |" % (m, met[m][0], met[m][1]))
            L.append("")
    for f in (tag + "_launches.csv", tag + "_bench.json"):
        open(os.path.join(dst, f), "w").write(open(os.path.join(src, f), errors="replace").read())
    open(os.path.join(dst, tag + "_summary.md"), "w").write("\n".join(L) + "\n")
    print("\n".join(L))


if __name__ == "__main__":
    main()
