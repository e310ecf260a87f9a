"""Pins the CPU oracle (oracle/*.c) against the UNMODIFIED reference:
 - SeqAn's own golden vectors for the semi-global Gotoh alignment
   (extern/seqan/tests/align/test_alignment_algorithms_global.h:2084-2151, first three cases of cases.txt)
 - committed outputs of oracle/_ref/ref_harness (tests/golden/expected.txt, made by make_golden.py)
 - when oracle/_ref/ref_harness is present, fresh seeded cases run through it on the spot."""
import os

import numpy as np
import pytest

from tests import refcases as rc

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = rc.read_cases(os.path.join(G, "cases.txt"))
with open(os.path.join(G, "expected.txt")) as _f:
    EXPECTED = [rc.parse_line(x) for x in _f.read().strip().split("\n")]


def test_fixture_shape():
    assert len(CASES) == len(EXPECTED) == 225


def test_seqan_golden_vectors():
    want = [(2, "AT-G-T", "ATAGAT"), (7, "AAAAAGGGGTTTT", "--AAA---GTT--"),
            (6, "AAAAAATTTTTGGG-----", "---TTTTTTTTGGGGGGGG")]
    for c, w in zip(CASES[:3], want):
        sc, ops, _ = rc.orc.align_semiglobal(c["ref"], c["read"], *c["sc"])
        assert (sc,) + rc.orc.rows_from_ops(ops, c["ref"], c["read"]) == w


@pytest.mark.parametrize("i", range(len(CASES)))
def test_oracle_matches_reference_fixture(i):
    rc.check_case(CASES[i], EXPECTED[i])


def test_convex_quirks():
    # 10-base gap costs GO*(1 + 1/1 + 1/2 + ... + 1/9): SURVEY.md section 0 item 5 (1957103.1746031747)
    a = np.random.default_rng(5).integers(0, 4, 60).astype(np.uint8)
    b = np.concatenate([a[:25], a[35:]])
    sc, ops = rc.orc.align_convex(a, b)
    c = rc.orc.cigar(ops)
    assert len(c) == 3 and c[1] == ("D", 10) and c[0][1] + c[2][1] == 50
    want = 0.0
    for _ in range(c[0][1]):
        want += 90000.0
    want += -100000.0
    for n in range(1, 10):
        want += -100000.0 / n
    for _ in range(c[2][1]):
        want += 90000.0
    assert sc == want
    # non-positive final score: scout keeps its default cell -> nD mI and the denormal "infinity"
    sc, ops = rc.orc.align_convex(np.zeros(9, np.uint8), np.ones(4, np.uint8))
    assert sc == 1.1125369292536007e-308 and rc.orc.cigar_str(ops) == "9D4I"


@pytest.mark.skipif(not os.path.exists(rc.HARNESS), reason="reference harness not built")
def test_fresh_cases_against_live_reference(tmp_path):
    p = str(tmp_path / "cases.txt")
    rng = np.random.default_rng(99)
    with open(p, "w") as f:
        for sd in range(3):
            rc.write_round(f, rc.small_variant(500 + sd, L_core=200, n_pairs=50, err=0.1), sd % 2)
        rc.write_polish(f, rc.small_variant(600, L_core=200, n_pairs=60, err=0.1))
        for _ in range(20):
            a = rng.integers(0, 5, int(rng.integers(3, 200))).astype(np.uint8)
            b = rc.synth.mutate(rng, a, 0.2)
            if len(b):
                f.write("FINAL %s %s\n" % (rc.s(a), rc.s(b)))
    out = rc.run_harness(p).strip().split("\n")
    for c, line in zip(rc.read_cases(p), out):
        rc.check_case(c, rc.parse_line(line))
