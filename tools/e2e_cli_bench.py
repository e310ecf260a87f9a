"""tools/e2e_cli_bench.py -- whole-program wall time of the UNMODIFIED reference (oracle/_ref/sviper) and of the GPU
build of the same source (sviper_b200/host/sviper_b200_cli) on the same synthetic FASTA / BAM / VCF inputs
(tests/e2e_dataset.py, SURVEY 8d "End to end": variants with FILTER != SKIP / wall seconds), same thread count.
Everything outside the hot path (BAM region fetches, the long-read MSA of build_consensus, VCF writing) is the
reference's own host code in BOTH programs, so this is the Amdahl view of the drop-in, not a kernel number.
Prints one JSON object; the committed copy is profiles/r1_e2e_cli.json.

  python tools/e2e_cli_bench.py [--contig 200000] [--threads N] [--repeat 2]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import e2e_dataset  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "sviper")
GPU_BIN = os.path.join(ROOT, "sviper_b200", "host", "sviper_b200_cli")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--contig", type=int, default=200000)
    ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--repeat", type=int, default=2)
    a = ap.parse_args()
    tmp = tempfile.mkdtemp(prefix="svp_cli_")
    d = e2e_dataset.make(tmp, seed=11, contig_len=a.contig)
    common = ["-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", str(d["k"]), "-t", str(a.threads)]
    out = {"workload": "tests/e2e_dataset.py contig %d bp: %d DEL/INS of 100-800 bp + 5 FILTER-path records, 30x short / "
                       "20x long reads, -k 400" % (a.contig, d["n_sv"]), "threads": a.threads, "records": d["n_records"]}
    for name, binary in (("reference_cpu", REF_BIN), ("gpu_build", GPU_BIN)):
        best = None
        for _ in range(a.repeat):
            t0 = time.perf_counter()
            subprocess.check_call([binary] + common + ["-o", os.path.join(tmp, name)], stdout=subprocess.DEVNULL)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        out[name] = {"wall_s": best, "variants_per_s": (d["n_records"] - 1) / best}
    # parity is checked single-threaded: the unmodified reference is not run-to-run deterministic with several threads
    # (build_consensus, host code in front of the hot path in both programs; see tests/test_e2e_vcf.py)
    vcf = {}
    for name, binary in (("reference_cpu", REF_BIN), ("gpu_build", GPU_BIN)):
        subprocess.check_call([binary] + common[:-1] + ["1", "-o", os.path.join(tmp, name + "_t1")], stdout=subprocess.DEVNULL)
        vcf[name] = open(os.path.join(tmp, name + "_t1.vcf")).read()
    out["vcf_identical_single_threaded"] = vcf["reference_cpu"] == vcf["gpu_build"]
    out["reference_same_vcf_at_%d_threads_as_at_1" % a.threads] = open(os.path.join(tmp, "reference_cpu.vcf")).read() == vcf["reference_cpu"]
    out["gpu_build_same_vcf_at_%d_threads_as_at_1" % a.threads] = open(os.path.join(tmp, "gpu_build.vcf")).read() == vcf["gpu_build"]
    out["speedup"] = out["reference_cpu"]["wall_s"] / out["gpu_build"]["wall_s"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
