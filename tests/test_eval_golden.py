"""POS / END / SVLEN / SEQ / QUAL / FILTER=FAIL5 as the reference derives them from the final CIGAR
(assign_quality + refine_variant, sviper.h:468-489).
tests/golden/eval_expected.txt holds what the UNMODIFIED reference printed for eval_cases.txt (make_golden_eval.py).
Checked against it: the oracle's C restatement (oracle/oracle_eval.c) and the product's C ABI
(svp_assign_quality / svp_refine_variant / svp_refine_info / svp_format_quality -- host arithmetic, no GPU needed)."""
import os
import re

import numpy as np
import pytest

from sviper_b200 import api
from tests import refcases as rc

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def parse_case(line):
    p = line.split()
    assert p[0] == "EVAL"
    cig = [(o, int(c)) for c, o in re.findall(r"(\d+)([A-Z=])", p[8])]
    return dict(sv_type=api.SV_DEL if p[1] == "DEL" else api.SV_INS, ref_pos=int(p[2]), ref_pos_end=int(p[3]),
                sv_length=int(p[4]), k=int(p[5]), begin=int(p[6]), info=p[7], cigar=cig,
                seq="" if p[9] == "*" else p[9], filter="PASS")


with open(os.path.join(G, "eval_cases.txt")) as _f:
    CASES = [parse_case(x) for x in _f.read().strip().split("\n")]
with open(os.path.join(G, "eval_expected.txt")) as _f:
    EXPECTED = [rc.parse_line(x) for x in _f.read().strip().split("\n")]


def test_fixture_shape():
    assert len(CASES) == len(EXPECTED) == 330
    assert sum(e["filter"] == "FAIL5" for e in EXPECTED) > 50 and sum(e["has"] == "1" for e in EXPECTED) > 150


def test_oracle_matches_reference():
    for c, e in zip(CASES, EXPECTED):
        o = rc.orc.evaluate_final_mapping(c["sv_type"], c["ref_pos"], c["ref_pos_end"], c["sv_length"], c["info"],
                                          c["cigar"], c["seq"], c["begin"], c["k"])
        assert (o["has"], o["pos"], o["end"], o["svlen"], o["filter"], o["info"]) == \
            (int(e["has"]), int(e["pos"]), int(e["end"]), int(e["svlen"]), e["filter"], e["info"])
        assert o["qual"] == float.fromhex(e["qual"])


def test_c_abi_matches_reference():
    for c, e in zip(CASES, EXPECTED):
        got = api.evaluate_final_mapping(c, c["cigar"], rc.orc.enc(c["seq"]), c["begin"], c["k"])
        assert (got["ref_pos"], got["ref_pos_end"], got["sv_length"], got["filter"], got["info"]) == \
            (int(e["pos"]), int(e["end"]), int(e["svlen"]), e["filter"], e["info"])
        assert got["quality"] == float.fromhex(e["qual"])      # the 1e-6 relative bar of BASELINE.json is met with ==
        assert got["quality_str"] == e["qualstr"]


def test_c_abi_errors():
    L = api.lib()
    cnt = np.array([10, 5], np.uint32)
    vp = api.VariantPos(api.SV_DEL, 100, 110, 10)
    assert L.svp_refine_variant(0, b"MN", api._p(cnt, api.u32p), 2, vp, None) == -2       # unknown op: the reference throws
    vp = api.VariantPos(3, 100, 110, 10)                                                   # DUP: not supported
    assert L.svp_refine_variant(0, b"MD", api._p(cnt, api.u32p), 2, vp, None) == -2
    vp = api.VariantPos(api.SV_INS, 5, 5, 4)
    pol = np.zeros(3, np.uint8)
    assert L.svp_refine_info(b"X", vp, api._p(pol), 3, 1, None, 0) == -2                   # insertion beyond the sequence


@pytest.mark.skipif(not os.path.exists(rc.HARNESS), reason="reference harness not built")
def test_fresh_cases_against_live_reference(tmp_path):
    import importlib.util
    spec = importlib.util.spec_from_file_location("mge", os.path.join(G, "make_golden_eval.py"))
    mge = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mge)
    rng = np.random.default_rng(777)
    p = str(tmp_path / "eval.txt")
    with open(p, "w") as f:
        for i in range(200):
            f.write(mge.one_case(rng, i))
    exp = [rc.parse_line(x) for x in rc.run_harness(p).strip().split("\n")]
    with open(p) as f:
        cases = [parse_case(x) for x in f.read().strip().split("\n")]
    for c, e in zip(cases, exp):
        got = api.evaluate_final_mapping(c, c["cigar"], rc.orc.enc(c["seq"]), c["begin"], c["k"])
        assert (got["ref_pos"], got["ref_pos_end"], got["sv_length"], got["filter"], got["info"], got["quality_str"]) == \
            (int(e["pos"]), int(e["end"]), int(e["svlen"]), e["filter"], e["info"], e["qualstr"])
