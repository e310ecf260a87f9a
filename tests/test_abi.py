"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every function that
include/sviper_b200.h declares, and refuses to work without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "sviper_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(svp_[a-z_0-9]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from sviper_b200 import api
    if not os.path.exists(api.LIB_PATH):
        api.build()
    return api.lib()


def test_header_functions_exported(lib):
    names = declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), "missing export " + n


def test_python_mirror_lists_every_export():
    from sviper_b200 import api
    assert sorted(api.EXPORTS) == declared_functions()


def test_struct_layouts_match_header(lib):
    from sviper_b200 import api
    # sizes computed by hand from the header (LP64): catches accidental field drift
    assert C.sizeof(api.AlignIn) == 72 and C.sizeof(api.AlignOut) == 64
    assert C.sizeof(api.WindowIn) == 120 and C.sizeof(api.Result) == 104 and C.sizeof(api.FinalIn) == 32


def test_assign_quality_host_arithmetic(lib):
    # evaluate_final_mapping.h:17-28: max((1 - ((n-1)/(2k))/0.15) * 100, 0)
    assert lib.svp_assign_quality(3, 400) == (1.0 - ((3 - 1.0) / 800.0) / 0.15) * 100.0
    assert lib.svp_assign_quality(500, 400) == 0.0


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from sviper_b200 import api
    with pytest.raises(RuntimeError):
        api.Context(0)


def test_product_does_not_touch_the_oracle():
    """the oracle is test infrastructure: nothing under sviper_b200/ may import, link or call it"""
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "sviper_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"liboracle|from oracle|import oracle|orc_[a-z_]+\s*\(", txt):
                    bad.append(f)
    assert not bad, bad
