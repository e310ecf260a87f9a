"""The C++ host side above the C ABI (sviper_b200/host/polisher.hpp): the mirror of the reference's interface
for the hot path (polish_to_perfection, final alignment, the per-variant loop -> BatchScheduler)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "sviper_b200", "host")
BIN = os.path.join(HOST, "harness_b200")
G = os.path.join(ROOT, "tests", "golden")


def test_header_compiles_as_cxx14(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('#include "sviper_b200/host/polisher.hpp"\nint main() { return 0; }\n')
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-pedantic", "-Werror", "-fsyntax-only", "-I", ROOT, str(src)])


def test_driver_builds():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "sviper_b200", "csrc"), "../host/harness_b200"])
    assert os.path.exists(BIN)


def test_cpp_evaluate_final_mapping_matches_reference_fixture():
    """POS / END / SVLEN / SEQ / QUAL / FAIL5 through polisher.hpp's evaluate_final_mapping (host arithmetic of the
    C ABI, no GPU): identical text to what the unmodified reference printed (tests/golden/eval_expected.txt)"""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "sviper_b200", "csrc"), "../host/harness_b200"])
    out = subprocess.check_output([BIN, "--eval", os.path.join(G, "eval_cases.txt")], text=True)
    assert out == open(os.path.join(G, "eval_expected.txt")).read()
    assert out.count("\n") == 330


@pytest.mark.gpu
def test_cpp_host_matches_reference_fixture():
    """the same case file through the reference harness (committed expected.txt) and through the C++ host
    side + BatchScheduler + CUDA library: POLISH and FINAL lines must be identical text (hex floats included)"""
    out = subprocess.check_output([BIN, os.path.join(G, "cases.txt")], text=True).strip().split("\n")
    exp = open(os.path.join(G, "expected.txt")).read().strip().split("\n")
    assert len(out) == len(exp)
    n = 0
    for o, e in zip(out, exp):
        if o == "SKIP":
            continue
        assert o == e
        n += 1
    assert n == 64


def test_gpu_programs_route_the_hot_path_to_the_library():
    """sviper_b200_cli = the reference's unmodified src/sviper.cpp, sviper_b200 = the pipeline program; both compiled from
    the reference's sources with host/shim in front of the include path and linked with --gc-sections. Their imports from
    the library are the asynchronous scheduler and the stand-alone final alignment; the reference's CPU polishing loop
    (renamed away by the shim) and SeqAn's convex-gap DP (sviper.h:457-460) are not part of the programs: nothing refers
    to them, i.e. the shim's overloads were chosen and there is no CPU hot path to fall back to."""
    ref = os.path.join(ROOT, "oracle", "_ref", "sviper")
    for name in ("sviper_b200_cli", "sviper_b200"):
        exe = os.path.join(HOST, name)
        if not os.path.exists(exe):
            pytest.skip("%s not built (needs the reference tree at build time)" % name)
        syms = subprocess.check_output(["nm", "-C", exe], text=True)
        imports = sorted(set(line.split()[-1] for line in syms.split("\n") if " U svp_" in line))
        assert {"svp_submit_polish", "svp_wait_any", "svp_final_align_batch"} <= set(imports), imports
        assert "polish_to_perfection_reference_cpu" not in syms
        assert "onvex" not in syms
    if os.path.exists(ref):
        rs = subprocess.check_output(["nm", "-C", ref], text=True)
        assert "onvex" in rs and "polish_to_perfection" in rs          # the reference does instantiate both


def test_pipeline_fibers_with_a_simulated_device(tmp_path):
    """host/pipeline.hpp + fiber.hpp on the CPU (tests/cpp/pipeline_test.cpp): a collector thread stands in for the device.
    A body (the reference's polish_variant in the real program) always continues on the thread that started it, a thread
    never runs two bodies at once (per-thread BAM handles, sviper.h:133-135), parked bodies per thread stay under the cap,
    bodies that never reach the device and a throwing body are handled, the idle hook fires."""
    exe = str(tmp_path / "pipeline_test")
    lib = os.path.join(ROOT, "sviper_b200", "csrc")
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-Wall", "-pedantic", "-Werror", "-fopenmp", "-I", ROOT,
                           os.path.join(ROOT, "tests", "cpp", "pipeline_test.cpp"), "-o", exe, "-L", lib, "-lsviper_b200",
                           "-Wl,-rpath," + lib, "-lpthread"])
    for args in (("4", "2000", "8"), ("1", "300", "64"), ("8", "10000", "16")):
        out = subprocess.check_output([exe] + list(args), text=True, timeout=300)
        assert out.strip().endswith("ok"), out
