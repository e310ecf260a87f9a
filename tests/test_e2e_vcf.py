"""End-to-end VCF parity (SURVEY.md section 8c level 2, config C1 of 8d): the UNMODIFIED reference program
(oracle/_ref/sviper = /root/reference/src/sviper.cpp compiled as is) and the two GPU programs -- sviper_b200_cli (the SAME
translation unit with sviper_b200/host/shim in front of the include path, so that polish_to_perfection and the final convex
alignment run on the device) and sviper_b200 (the reference's main() sequence with the OpenMP loop replaced by the prepare /
device / finish pipeline, host/sviper_b200_main.cpp) -- on the same synthetic FASTA / BAM / VCF inputs. The output VCFs must
be identical byte for byte (POS, END, SVLEN, SEQ, QUAL, FILTER incl. SKIP / FAIL0 / FAIL1 / FAIL2 / FAIL4 / FAIL5 / FAIL6 / FAIL7), and
so must every per-variant log line that reports the hot path's side effects (substituted / deleted / inserted bases,
rounds, the coverage profile)."""
import os
import subprocess

import pytest

from tests import e2e_dataset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "sviper")
GPU_BIN = os.path.join(ROOT, "sviper_b200", "host", "sviper_b200_cli")
PIPE_BIN = os.path.join(ROOT, "sviper_b200", "host", "sviper_b200")


def test_bam_writer_is_read_by_the_reference(tmp_path):
    """CPU: the Python-written BAM / BAI / FASTA index are accepted by the reference (FILTER paths that need no polishing)"""
    if not os.path.exists(REF_BIN):
        pytest.skip("reference binary not built")
    d = e2e_dataset.make(str(tmp_path), seed=3, contig_len=60000, only_filter_paths=True)
    subprocess.check_call([REF_BIN, "-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-o",
                           str(tmp_path / "ref"), "-t", "2", "-k", "400"], stdout=subprocess.DEVNULL)
    rows = [x.split("\t") for x in open(str(tmp_path / "ref.vcf")) if not x.startswith("#")]
    assert [r[6] for r in rows] == ["FAIL4", "SKIP", "FAIL2", "FAIL1", "FAIL0", "FAIL6", "FAIL7"]


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


def rows_of(vcf_text):
    return [x.split("\t") for x in vcf_text.split("\n") if x and not x.startswith("#")]


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2])
def test_vcf_identical_to_reference(tmp_path, seed):
    """The unmodified reference program against both GPU programs -- the unmodified src/sviper.cpp over the shims
    (sviper_b200_cli) and the prepare / device / finish pipeline (sviper_b200): byte-identical VCFs and identical hot-path
    log lines. Single-threaded runs are compared with the reference's single-threaded run, strictly. What the REFERENCE writes
    for a record depends on which records the same thread's BAM handles read before it (profiles/r2_reference_determinism.txt:
    single-threaded, one record alone QUAL 95, after another record 96.67), so with several threads -- records dealt to threads
    by timing -- the reference does not repeat its own output on larger lists (three different VCFs in seven runs), and the
    GPU programs run that same host code. A multi-threaded GPU run therefore has to agree, record by record, with a line the
    reference has written for that record: in its single-threaded run or in one of up to three fresh multi-threaded ones.
    (On these small lists every run made so far was byte-identical.)"""
    assert os.path.exists(REF_BIN) and os.path.exists(GPU_BIN) and os.path.exists(PIPE_BIN), "build() must have produced the binaries"
    d = e2e_dataset.make(str(tmp_path), seed=seed)
    ref_vcf = run(REF_BIN, d, str(tmp_path / "ref"), 1)
    rows = rows_of(ref_vcf)
    assert len(rows) == d["n_records"]
    assert sum(r[6] == "PASS" for r in rows) >= 5
    assert {r[6] for r in rows} >= {"PASS", "SKIP", "FAIL0", "FAIL1", "FAIL2", "FAIL4", "FAIL5", "FAIL6", "FAIL7"}
    # the call whose best long read is too short to crop comes back untouched (sviper.h:327): PASS without a QUAL
    assert any(r[6] == "PASS" and r[5] == "0" for r in rows)
    # the 1.2 kb insertion is polished and kept
    svlen = lambda r: abs(int([x for x in r[7].split(";") if x.startswith("SVLEN=")][0][6:]))
    assert any(r[6] == "PASS" and r[4] == "<INS>" and svlen(r) >= 1000 and float(r[5]) > 50 for r in rows)
    a = hot_path_lines(str(tmp_path / "ref.log"))
    assert len(a) >= 30
    for name, binary in (("cli1", GPU_BIN), ("pipe1", PIPE_BIN)):
        vcf = run(binary, d, str(tmp_path / name), 1)
        assert vcf == ref_vcf, name
        assert hot_path_lines(str(tmp_path / (name + ".log"))) == a, name
    seen = [{tuple(r)} for r in rows]           # per record: every line the reference has written for it so far
    for name, binary in (("cli4", GPU_BIN), ("pipe4", PIPE_BIN)):
        for attempt in range(3):
            vcf = run(binary, d, str(tmp_path / name), 4)
            if vcf == ref_vcf:
                assert hot_path_lines(str(tmp_path / (name + ".log"))) == a, name
                break
            got = rows_of(vcf)
            assert len(got) == len(rows), name
            if all(tuple(g) in s for g, s in zip(got, seen)):
                break
            for r, s in zip(rows_of(run(REF_BIN, d, str(tmp_path / ("ref4_%s_%d" % (name, attempt))), 4)), seen):
                s.add(tuple(r))
            if all(tuple(g) in s for g, s in zip(got, seen)):
                break
        else:
            raise AssertionError("%s: a record differs from every line the reference wrote for it in %d runs" % (name, 1 + 3))


@pytest.mark.gpu
def test_pipeline_parks_variants_and_batches_them(tmp_path):
    """The pipeline program with one host thread still has many variants in flight: the thread parks each at the device seam
    and prepares the next (the reference holds one variant per thread). By default a free lane starts on whatever is queued
    (the device is far faster than one host thread prepares); SVP_BATCH_TARGET + SVP_LINGER_MS keep the lanes back until
    several variants wait, so here they are polished as shared device batches. The VCF does not change."""
    d = e2e_dataset.make(str(tmp_path), seed=7, contig_len=100000)
    ref_vcf = run(REF_BIN, d, str(tmp_path / "ref"), 1)
    env = dict(os.environ, SVIPER_B200_STATS="1", SVP_BATCH_TARGET="6", SVP_LINGER_MS="20000")
    p = subprocess.run([PIPE_BIN, "-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", str(d["k"]),
                        "-o", str(tmp_path / "pipe"), "-t", "1"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=env)
    assert p.returncode == 0, p.stderr
    assert open(str(tmp_path / "pipe.vcf")).read() == ref_vcf
    stats = dict(kv.split("=") for kv in [x for x in p.stderr.split("\n") if x.startswith("[sviper_b200] variants=")][0].split()[1:])
    assert int(stats["peak_parked_per_thread"]) >= 6 and float(stats["mean_device_batch"]) >= 3.0
