"""The CIGAR walks of polish_variant's prepare side: record_supports_variant
(variant.h:239-275), get_read_region_boundaries (basics.h:189-241), cut_down_high_coverage (basics.h:403-442), merge_alignments (merge_split_alignments.h:378-561).
tests/golden/prep_expected.txt holds what the UNMODIFIED reference printed for prep_cases.txt (make_golden_prep.py).
Checked against it: the product's C ABI through ctypes, and the C++ host mirror (sviper_b200/host/prepare.hpp) through
harness_b200 --prep. Host arithmetic, no GPU needed."""
import os
import re
import subprocess

import numpy as np
import pytest

from sviper_b200 import api
from tests import refcases as rc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")
BIN = os.path.join(ROOT, "sviper_b200", "host", "harness_b200")


def cigar(s):
    return [(o, int(c)) for c, o in re.findall(r"(\d+)([A-Z=])", s)]


def read_cases(path):
    cases = []
    with open(path) as f:
        lines = iter(f.read().split("\n"))
    for line in lines:
        p = line.split()
        if not p:
            continue
        if p[0] == "SUPPORT":
            cases.append(dict(cmd="SUPPORT", sv_type=api.SV_DEL if p[1] == "DEL" else api.SV_INS, pos=int(p[2]),
                              svlen=int(p[3]), begin=int(p[4]), cigar=cigar(p[5])))
        elif p[0] == "REGION":
            cases.append(dict(cmd="REGION", begin=int(p[1]), cigar=cigar(p[2]), rb=int(p[3]), re=int(p[4])))
        else:
            assert p[0] == "CUTCOV"
            recs = []
            for _ in range(int(p[2])):
                q = next(lines).split()
                recs.append(dict(name=q[0], begin_pos=int(q[1]), mapq=int(q[2]), cigar=cigar(q[3])))
            cases.append(dict(cmd="CUTCOV", cov=int(p[1]), records=recs))
    return cases


def run_case(c):
    """one case through the C ABI, formatted like the reference harness's output line"""
    if c["cmd"] == "SUPPORT":
        return "SUPPORT supports=%d" % api.record_supports_variant(c["begin"], c["cigar"], c["sv_type"], c["pos"], c["svlen"])
    if c["cmd"] == "REGION":
        return "REGION begin=%d end=%d" % api.get_read_region_boundaries(c["begin"], c["cigar"], c["rb"], c["re"])
    kept = api.cut_down_high_coverage(c["records"], c["cov"])
    return "CUTCOV n=%d keep=%s" % (len(kept), ",".join(r["name"] for r in kept))


CASES = read_cases(os.path.join(G, "prep_cases.txt"))
with open(os.path.join(G, "prep_expected.txt")) as _f:
    EXPECTED = _f.read().strip("\n").split("\n")


def test_fixture_shape():
    assert len(CASES) == len(EXPECTED) == 480
    assert 60 < sum(e == "SUPPORT supports=1" for e in EXPECTED) < 180
    assert sum(e == "REGION begin=0 end=0" for e in EXPECTED) == 60        # the begin >= end error path
    thinned = [(c, e) for c, e in zip(CASES, EXPECTED) if c["cmd"] == "CUTCOV"]
    assert sum(int(e.split()[1][2:]) < len(c["records"]) for c, e in thinned) > 20


def test_c_abi_matches_reference():
    for c, e in zip(CASES, EXPECTED):
        assert run_case(c) == e


def test_cpp_host_matches_reference():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "sviper_b200", "csrc"), "../host/harness_b200"])
    out = subprocess.check_output([BIN, "--prep", os.path.join(G, "prep_cases.txt")], text=True,
                                  stderr=subprocess.DEVNULL)
    assert out == open(os.path.join(G, "prep_expected.txt")).read()


def test_c_abi_errors():
    L = api.lib()
    cnt = np.array([10, 5, 10], np.uint32)
    vp = api.VariantPos(api.SV_DEL, 5, 10, 5)
    assert L.svp_record_supports_variant(0, b"MND", api._p(cnt, api.u32p), 3, vp) == -2    # unknown op: the reference throws
    vp = api.VariantPos(3, 5, 10, 5)                                                        # DUP: never supported
    assert L.svp_record_supports_variant(0, b"MDM", api._p(cnt, api.u32p), 3, vp) == 0
    with pytest.raises(ValueError):                                                         # region ends before the record
        api.get_read_region_boundaries(1000, [("M", 100)], 10, 20)
    far = [dict(begin_pos=0, mapq=60, cigar=[("M", 150)]), dict(begin_pos=40000, mapq=60, cigar=[("M", 150)])]
    with pytest.raises(ValueError):                                                         # int16 offset of the reference wraps
        api.cut_down_high_coverage(far, 10)
    assert api.cut_down_high_coverage([], 10) == []
    assert L.svp_alignment_length_in_ref(b"SMIDMH", api._p(np.array([5, 10, 2, 3, 7, 4], np.uint32), api.u32p), 6) == 20


# ---- merge_alignments (merge_split_alignments.h:378-561): tests/golden/merge_*.txt ---------------------------------
def read_merge_cases(path):
    cases = []
    with open(path) as f:
        lines = iter(f.read().split("\n"))
    for line in lines:
        p = line.split()
        if not p:
            continue
        assert p[0] == "MERGE"
        recs = []
        for _ in range(int(p[1])):
            q = next(lines).split()
            recs.append(dict(name=q[0], flag=int(q[1]), rid=int(q[2]), begin_pos=int(q[3]), mapq=int(q[4]), cigar=cigar(q[5]),
                             seq=q[6]))
        cases.append(recs)
    return cases


def merge_line(recs):
    out = api.merge_alignments(recs)
    return "MERGE n=%d " % len(out) + " ".join("%s:%d:%d:%d:%s:%s" % (
        r["name"], r["flag"], r["begin_pos"], r["mapq"], "".join("%d%s" % (c, o) for o, c in r["cigar"]), r["seq"]) for r in out)


def test_merge_alignments_c_abi_matches_reference():
    cases = read_merge_cases(os.path.join(G, "merge_cases.txt"))
    exp = open(os.path.join(G, "merge_expected.txt")).read().strip("\n").split("\n")
    assert len(cases) == len(exp) >= 100
    for recs, e in zip(cases, exp):
        assert merge_line(recs) == e
    # most groups really are merged: the expected line holds a D / I element that no input record has
    inputs_have = lambda recs, el: any(el in [(o, c) for o, c in r["cigar"]] for r in recs)
    n_new = 0
    for recs, e in zip(cases, exp):
        for tok in e.split()[2:]:
            big = [(o, c) for o, c in cigar(tok.split(":")[4]) if o in "DI" and c >= 20]
            n_new += any(not inputs_have(recs, el) for el in big)
    assert n_new >= 60


def test_merge_alignments_cpp_host_matches_reference():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "sviper_b200", "csrc"), "../host/harness_b200"])
    out = subprocess.check_output([BIN, "--prep", os.path.join(G, "merge_cases.txt")], text=True)
    assert out == open(os.path.join(G, "merge_expected.txt")).read()


def test_merge_alignments_refuses_what_is_undefined_in_the_reference():
    """a primary whose 500 hard-clipped bases would have to come from a supplementary record that only holds 100: the
    reference copies prefix(seq, 500) beyond the sequence (heap garbage in its output, seen while generating the
    fixture); the C ABI answers SVP_ERR_ARG"""
    recs = [dict(name="r", flag=0, rid=0, begin_pos=5000, mapq=60, cigar=[("H", 500), ("M", 300)], seq="A" * 300),
            dict(name="r", flag=2048, rid=0, begin_pos=4000, mapq=60, cigar=[("M", 100), ("H", 700)], seq="C" * 100)]
    with pytest.raises(ValueError):
        api.merge_alignments(recs)
    recs[1]["seq"] = "C" * 100
    recs[1]["cigar"] = [("N", 100)]                                      # unknown operation: the reference throws
    with pytest.raises(ValueError):
        api.merge_alignments(recs)


@pytest.mark.skipif(not os.path.exists(rc.HARNESS), reason="reference harness not built")
def test_fresh_cases_against_live_reference(tmp_path):
    import importlib.util
    spec = importlib.util.spec_from_file_location("mgp", os.path.join(G, "make_golden_prep.py"))
    mgp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mgp)
    rng = np.random.default_rng(4242)
    p = str(tmp_path / "prep.txt")
    with open(p, "w") as f:
        for i in range(120):
            f.write(mgp.support_case(rng, i)); f.write(mgp.region_case(rng, i))
            if i % 6 == 0:
                f.write(mgp.cutcov_case(rng, i + 1))
    exp = subprocess.check_output([rc.HARNESS, p], text=True, stderr=subprocess.DEVNULL).strip("\n").split("\n")
    cases = read_cases(p)
    assert len(cases) == len(exp) == 260
    for c, e in zip(cases, exp):
        assert run_case(c) == e
