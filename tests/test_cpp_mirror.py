"""The C++ host-side mirror of the reference's interface (include/rav1e_cuda.hpp) and the reference's own unit tests
restated against it (tests/cpp/reference_units.cpp): get_sad_same / get_satd_same (src/dist.rs:341-465), test_plane_pad
(src/frame/plane.rs:707-752), roundtrips (src/transform/mod.rs:589-725), test_divu_pair / test_tx_log_scale
(src/quantize.rs:160-203).  Compiled with g++ against librav1e_cuda.so:
the C ABI consumed from compiled code, no Python in the path.  Without a GPU the host-only part runs and the Context
constructor must throw (no CPU fallback); on a GPU box the whole program runs."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "reference_units")


@pytest.fixture(scope="module")
def exe():
    import __graft_entry__ as g
    from rav1e_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        g.build()
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "reference_units.cpp"), "-L", libdir, "-lrav1e_cuda",
                           "-Wl,-rpath," + libdir, "-o", EXE])
    return EXE


def _run(exe, *args):
    p = subprocess.run([exe, *args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    return p.returncode, p.stdout


def test_host_side_of_the_cpp_mirror(exe):
    rc, out = _run(exe)
    assert rc == 0 and "0 failed" in out, out


def test_cpp_mirror_fails_loudly_without_a_device(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    rc, out = _run(exe, "nodevice")
    assert rc == 0 and "0 failed" in out, out


@pytest.mark.gpu
def test_reference_unit_tests_through_the_cpp_mirror(exe):
    rc, out = _run(exe, "gpu")
    assert rc == 0 and "0 failed" in out, out
    assert int(out.strip().splitlines()[-1].split()[1]) > 48 + 4 * 22 + 2 * 37 * 4


@pytest.fixture(scope="module")
def example():
    """examples/lookahead_me.c: the headline call sequence from plain C99"""
    import __graft_entry__ as g
    from rav1e_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        g.build()
    libdir = os.path.dirname(capi.LIB_PATH)
    out = os.path.join(ROOT, "build", "lookahead_me")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "lookahead_me.c"), "-L", libdir, "-lrav1e_cuda",
                           "-Wl,-rpath," + libdir, "-o", out])
    return out


def test_c_example_builds_and_refuses_to_run_without_a_device(example):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    rc, out = _run(example)
    assert rc == 2 and "no CPU fallback" in out, out


@pytest.mark.gpu
def test_c_example_finds_the_true_motion(example):
    rc, out = _run(example, "6")
    assert rc == 0 and "5 frame pairs" in out, out
