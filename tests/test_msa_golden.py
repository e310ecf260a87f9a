"""SURVEY 8f row f2 (build_consensus, basics.h:484-520): the C restatement of the two pairwise DP fills of the T-Coffee
library generation (oracle/oracle_msa.c) against what the UNMODIFIED reference printed (tests/golden/msa_*.txt, made by
tests/golden/make_golden_msa.py with oracle/_ref/ref_harness): alignment scores, end cells and the COMPLETE trace matrices
(64-bit FNV-1a of every byte). The CUDA fills are compared with the same restatement in tests/test_gpu_msa.py."""
import os

import numpy as np
import pytest

from tests import refcases as rc

orc = rc.orc
G = os.path.join(os.path.dirname(__file__), "golden")


def enc(s):
    return np.zeros(0, np.uint8) if s == "*" else np.array(["ACGTN".index(c) for c in s], np.uint8)


def load():
    lines = open(os.path.join(G, "msa_cases.txt")).read().split("\n")
    exp = open(os.path.join(G, "msa_expected.txt")).read().strip().split("\n")
    cases, i = [], 0
    while i < len(lines):
        p = lines[i].split()
        i += 1
        if not p:
            continue
        if p[0] == "GFILL":
            cases.append(("GFILL", enc(p[1]), enc(p[2])))
        elif p[0] == "SWFILL":
            k = int(p[3]); xs = [int(x) for x in p[4:4 + 2 * k]]
            cases.append(("SWFILL", enc(p[1]), enc(p[2]), list(zip(xs[0::2], xs[1::2]))))
        elif p[0] == "MSA":
            n = int(p[1]); rows = [lines[i + j].split() for j in range(n)]
            i += n
            cases.append(("MSA", [enc(r[0]) for r in rows], [float(r[1]) for r in rows]))
    assert len(cases) == len(exp)
    return cases, [dict(kv.split("=", 1) for kv in e.split()[1:]) for e in exp]


def test_fixture_inventory():
    cases, exp = load()
    kinds = [c[0] for c in cases]
    assert kinds.count("GFILL") == 70 and kinds.count("SWFILL") == 70 and kinds.count("MSA") == 110
    assert sum(1 for c in cases if c[0] == "MSA" and max(len(s) for s in c[1]) >= 800) >= 6


def test_global_fill_matches_the_reference():
    cases, exp = load()
    n = 0
    for c, e in zip(cases, exp):
        if c[0] != "GFILL":
            continue
        trace, sc, (S, H, V) = orc.msa_global_fill(c[1], c[2])
        assert sc == int(e["score"]) == int(e["libscore"]) and (S, H, V) == (int(e["S"]), int(e["H"]), int(e["V"]))
        assert len(trace) == int(e["tracelen"]) and orc.fnv1a(trace) == int(e["hash"])
        n += 1
    assert n == 70


def test_smith_waterman_clump_fill_matches_the_reference():
    cases, exp = load()
    n = 0
    for c, e in zip(cases, exp):
        if c[0] != "SWFILL":
            continue
        trace, sc, row, col, d = orc.msa_sw_fill(c[1], c[2], c[3])
        assert (sc, row, col) == (int(e["score"]), int(e["row"]), int(e["col"]))
        if sc > 0:
            assert d == int(e["dir"])
        assert len(trace) == int(e["tracelen"]) and orc.fnv1a(trace) == int(e["hash"])
        n += 1
    assert n == 70


def test_live_reference_on_fresh_cases(tmp_path):
    """fresh random pairs through the live reference harness (when it is present: this container)"""
    if not os.path.exists(rc.HARNESS):
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(int.from_bytes(os.urandom(4), "little"))
    path = str(tmp_path / "c.txt")
    pairs = []
    with open(path, "w") as f:
        for _ in range(12):
            a = rng.integers(0, 4, int(rng.integers(1, 500))).astype(np.uint8)
            b = rc.synth.mutate(rng, a, 0.12) if rng.random() < 0.7 else rng.integers(0, 4, int(rng.integers(1, 300))).astype(np.uint8)
            if len(b) == 0:
                b = np.array([3], np.uint8)
            cells = [(int(rng.integers(1, len(a) + 1)), int(rng.integers(1, len(b) + 1))) for _ in range(int(rng.integers(0, 30)))]
            pairs.append((a, b, cells))
            f.write("GFILL %s %s\n" % (rc.s(a), rc.s(b)))
            f.write("SWFILL %s %s %d %s\n" % (rc.s(a), rc.s(b), len(cells), " ".join("%d %d" % c for c in cells)))
    out = [dict(kv.split("=", 1) for kv in e.split()[1:]) for e in rc.run_harness(path).strip().split("\n")]
    for k, (a, b, cells) in enumerate(pairs):
        g, s = out[2 * k], out[2 * k + 1]
        trace, sc, shv = orc.msa_global_fill(a, b)
        assert sc == int(g["score"]) and orc.fnv1a(trace) == int(g["hash"]) and shv == (int(g["S"]), int(g["H"]), int(g["V"]))
        trace, sc, row, col, d = orc.msa_sw_fill(a, b, cells)
        assert (sc, row, col) == (int(s["score"]), int(s["row"]), int(s["col"])) and orc.fnv1a(trace) == int(s["hash"])


def test_staged_msa_over_the_oracle_fills_matches_the_reference():
    """The host side of the GPU build_consensus (sviper_b200/host/shim/sviper_b200_msa.hpp: SeqAn's own traceback on trace
    matrices it did not fill, four local rounds with recorded forbidden cells, fragment order, distance matrix) with the
    oracle's C fills plugged in (oracle/_ref/msa_replay): all 110 consensus strings equal the unmodified reference's."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "msa_replay")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/msa_replay not built (needs the reference tree)")
    out = subprocess.check_output([exe, os.path.join(G, "msa_cases.txt")], text=True).strip().split("\n")
    exp = [x for x in open(os.path.join(G, "msa_expected.txt")).read().strip().split("\n") if x.startswith("MSA ")]
    assert out == exp and len(out) == 110
