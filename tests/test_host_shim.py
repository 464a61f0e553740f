"""The C++ host side (npge_b200/host): npge::gpu::AnchorFinder behind the
reference's Processor interface.  The binary mirrors src/test/anchor_finder.cpp
of the reference; the --no-gpu subset (options, rules, Meta registration,
workers=0, to_atgcn) runs everywhere."""
import subprocess

import pytest

from npge_b200 import build


@pytest.fixture(scope="module")
def binary():
    return build.build_shim(force=False)


def run(binary, *args):
    p = subprocess.run([binary] + list(args), capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert " 0 failures" in p.stdout
    return p.stdout


def test_host_side_without_a_device(binary):
    out = run(binary, "--no-gpu")
    assert "no-gpu subset" in out


def test_shim_library_exports_the_processor(binary):
    syms = subprocess.run(["nm", "-DC", build.SHIM_LIB], capture_output=True, text=True).stdout
    assert "npge::gpu::AnchorFinder::run_impl() const" in syms
    assert "npge::gpu::AnchorFinder::name_impl() const" in syms
    assert "npge_anchor_find" in syms  # bound to the C ABI, not to a copy of the kernels
    assert "npge_anchor_multi_find_device" in syms  # set_devices: the several-GPU host


def test_reference_spelling_of_the_processors_compiles_and_runs(binary):
    """The NPGE_B200_REAL_HEADERS branch (string option rules, 3-argument add_gopt,
    add_opt_check(boost::bind(pattern_length, this, _1))) is built against
    host/ref_api and passes the same device-free cases."""
    out = run(build.SHIM_TEST_REFAPI, "--no-gpu")
    assert "no-gpu subset" in out


@pytest.mark.gpu
def test_reference_unit_cases_on_the_gpu(binary):
    out = run(binary)
    assert "no-gpu" not in out
    out = run(build.SHIM_TEST_REFAPI)   # and in the reference's own spelling
    assert "no-gpu" not in out
