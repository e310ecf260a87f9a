"""End-to-end VCF parity (SURVEY.md section 8c level 2, config C1 of 8d): the UNMODIFIED reference program
(oracle/_ref/sviper = /root/reference/src/sviper.cpp compiled as is) and the GPU build of the SAME source
(sviper_b200/host/sviper_b200_cli: identical translation unit, with sviper_b200/host/shim in front of the include path so
that polish_to_perfection and the final convex alignment run on the device) on the same synthetic FASTA / BAM / VCF
inputs. The output VCFs must be identical byte for byte (POS, END, SVLEN, SEQ, QUAL, FILTER incl. SKIP / FAIL0 / FAIL1 / FAIL2 / FAIL4), and
so must every per-variant log line that reports the hot path's side effects (substituted / deleted / inserted bases,
rounds, the coverage profile)."""
import os
import subprocess

import pytest

from tests import e2e_dataset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "sviper")
GPU_BIN = os.path.join(ROOT, "sviper_b200", "host", "sviper_b200_cli")


def test_bam_writer_is_read_by_the_reference(tmp_path):
    """CPU: the Python-written BAM / BAI / FASTA index are accepted by the reference (FILTER paths that need no polishing)"""
    if not os.path.exists(REF_BIN):
        pytest.skip("reference binary not built")
    d = e2e_dataset.make(str(tmp_path), seed=3, contig_len=60000, only_filter_paths=True)
    subprocess.check_call([REF_BIN, "-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-o",
                           str(tmp_path / "ref"), "-t", "2", "-k", "400"], stdout=subprocess.DEVNULL)
    rows = [x.split("\t") for x in open(str(tmp_path / "ref.vcf")) if not x.startswith("#")]
    assert [r[6] for r in rows] == ["FAIL4", "SKIP", "FAIL2", "FAIL1", "FAIL0"]


def hot_path_lines(log):
    keep = []
    for line in open(log):
        if line.startswith(("DONE POLISHING", "SUCCESS", "ERROR", "--- Built a consensus", "--- Extracted",
                            "--- Short read base qualities")) or (line[:1].isdigit() and " " in line):
            keep.append(line)
    return sorted(keep)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,threads", [(1, 4), (2, 1)])
def test_vcf_identical_to_reference(tmp_path, seed, threads):
    assert os.path.exists(REF_BIN) and os.path.exists(GPU_BIN), "build() must have produced both binaries"
    d = e2e_dataset.make(str(tmp_path), seed=seed)
    common = ["-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", str(d["k"])]
    subprocess.check_call([REF_BIN] + common + ["-o", str(tmp_path / "ref"), "-t", "8"], stdout=subprocess.DEVNULL)
    subprocess.check_call([GPU_BIN] + common + ["-o", str(tmp_path / "gpu"), "-t", str(threads)], stdout=subprocess.DEVNULL)
    ref_vcf = open(str(tmp_path / "ref.vcf")).read()
    gpu_vcf = open(str(tmp_path / "gpu.vcf")).read()
    assert gpu_vcf == ref_vcf
    rows = [x.split("\t") for x in ref_vcf.split("\n") if x and not x.startswith("#")]
    assert len(rows) == d["n_records"]
    assert sum(r[6] == "PASS" for r in rows) >= 5 and {r[6] for r in rows} >= {"PASS", "SKIP", "FAIL0", "FAIL1", "FAIL2", "FAIL4"}
    a, b = hot_path_lines(str(tmp_path / "ref.log")), hot_path_lines(str(tmp_path / "gpu.log"))
    assert len(a) >= 30 and a == b
