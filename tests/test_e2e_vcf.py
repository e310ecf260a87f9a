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
    with open(log) as f:
        lines = f.readlines()
    for line in lines:
        if line.startswith(("DONE POLISHING", "SUCCESS", "ERROR", "--- Built a consensus", "--- Extracted",
                            "--- Short read base qualities")) or (line[:1].isdigit() and " " in line):
            keep.append(line)
    return sorted(keep)


def variant_blocks(log):
    """per-variant blocks of a sviper log, keyed by the variant id"""
    blocks, cur = {}, None
    with open(log) as f:
        lines = f.readlines()
    for line in lines:
        if line.startswith(" PROCESS Variant "):
            cur = line.split()[2]
            blocks[cur] = []
        elif cur is not None:
            blocks[cur].append(line)
    return blocks


def run(binary, d, prefix, threads):
    subprocess.check_call([binary, "-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", str(d["k"]),
                           "-o", prefix, "-t", str(threads)], stdout=subprocess.DEVNULL)
    return open(prefix + ".vcf").read()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2])
def test_vcf_identical_to_reference(tmp_path, seed):
    """Both programs with -t 1: the reference itself is only run-to-run deterministic single-threaded (see below)."""
    assert os.path.exists(REF_BIN) and os.path.exists(GPU_BIN), "build() must have produced both binaries"
    d = e2e_dataset.make(str(tmp_path), seed=seed)
    ref_vcf = run(REF_BIN, d, str(tmp_path / "ref"), 1)
    gpu_vcf = run(GPU_BIN, d, str(tmp_path / "gpu"), 1)
    assert gpu_vcf == ref_vcf
    rows = [x.split("\t") for x in ref_vcf.split("\n") if x and not x.startswith("#")]
    assert len(rows) == d["n_records"]
    assert sum(r[6] == "PASS" for r in rows) >= 5
    assert {r[6] for r in rows} >= {"PASS", "SKIP", "FAIL0", "FAIL1", "FAIL2", "FAIL4"}
    a, b = hot_path_lines(str(tmp_path / "ref.log")), hot_path_lines(str(tmp_path / "gpu.log"))
    assert len(a) >= 30 and a == b


@pytest.mark.gpu
def test_multithreaded_gpu_build(tmp_path):
    """-t 4: the OpenMP threads' submissions are polished as shared device batches (BatchScheduler). The UNMODIFIED
    reference is not run-to-run deterministic with more than one thread -- build_consensus (the long-read MSA, host code
    that runs before the hot path in both programs) now and then returns a consensus of a different length for the same
    variant (observed: reference -t 16 vs -t 1 on a 200 kb contig differ in one record) -- so records are compared where
    the hot path saw the same input (same MSA length in the log); a difference there is reported as a warning, the hard
    gate is the single-threaded test above."""
    import warnings
    d = e2e_dataset.make(str(tmp_path), seed=5)
    ref_vcf = run(REF_BIN, d, str(tmp_path / "ref"), 1)
    gpu_vcf = run(GPU_BIN, d, str(tmp_path / "gpu"), 4)
    rr = {x.split("\t")[2]: x for x in ref_vcf.split("\n") if x and not x.startswith("#")}
    gg = {x.split("\t")[2]: x for x in gpu_vcf.split("\n") if x and not x.startswith("#")}
    assert list(rr) == list(gg)                                              # same records, input order (io.h:89-90)
    assert [x.split("\t")[6] for x in rr.values()] == [x.split("\t")[6] for x in gg.values()]   # FILTER column
    rb, gb = variant_blocks(str(tmp_path / "ref.log")), variant_blocks(str(tmp_path / "gpu.log"))
    msa = lambda blk: [x for x in blk if x.startswith("--- Built a consensus")]
    comparable = [k for k in rr if k in rb and k in gb and msa(rb[k]) == msa(gb[k])]
    assert len(comparable) >= len(rr) - 2
    bad = [k for k in comparable if rr[k] != gg[k]]
    if bad:
        warnings.warn("records differ although the MSA length is the same (MSA content may still differ): %s" % bad)
