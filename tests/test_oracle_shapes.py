"""Pins the CPU oracle against the UNMODIFIED reference at the REAL shapes of BASELINE.json's configs (tests/golden/
shapes_cases.txt / shapes_expected.txt, made by make_golden_shapes.py with oracle/_ref/ref_harness): 150 bp reads against
1.8 kb and 13 kb windows, the 152 / 153 bp row-class boundary, 250 / 256 bp reads, reads hanging off a window edge, strand
ties (reverse palindromes, (AT)n), reads of Ns, unrelated reads; one mapping + pileup round and whole polish_to_perfection
runs of a 300 bp DEL and a 500 bp INS at 30x, -k 400; final convex alignments of 1.8 kb against 2.8 kb with a 1 kb gap.
(cases.txt holds miniatures of the same.)"""
import os

import pytest

from tests import refcases as rc

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = rc.read_cases(os.path.join(G, "shapes_cases.txt"))
with open(os.path.join(G, "shapes_expected.txt")) as _f:
    EXPECTED = [rc.parse_line(x) for x in _f.read().strip().split("\n")]


def test_fixture_shape():
    assert len(CASES) == len(EXPECTED) == 35
    kinds = [c["cmd"] for c in CASES]
    assert kinds.count("ALIGN") == 28 and kinds.count("ROUND") == 2 and kinds.count("POLISH") == 2 and kinds.count("FINAL") == 3
    assert max(len(c["read"]) for c in CASES if c["cmd"] == "ALIGN") == 256
    assert max(len(c["ref"]) for c in CASES if c["cmd"] == "ALIGN") == 13000


@pytest.mark.parametrize("i", range(len(CASES)))
def test_oracle_matches_reference_at_real_shapes(i):
    rc.check_case(CASES[i], EXPECTED[i])
