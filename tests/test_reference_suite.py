"""The reference's OWN unit tests (test/*_test.c, compiled unmodified by oracle/Makefile `reftests`) run against this
repo's library: include/*.h + librslqr.so stand in for the reference's headers and library (SURVEY 8b drop-in boundary).

* matrix / utils / binarytree / lqrdata / nddata / solver exercise the host layer only: they run on CPU;
* linalg / linalg_custom / nested_dissection compute (dense helper API and the stage functions ndlqr_SolveLeaf ...
  ndlqr_UpdateShurFactor against the golden vectors of sample_problem.json): they need the GPU.

The binaries are prebuilt in the container that has /root/reference (they travel to the GPU box); without them the
tests skip."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "tests")
CPU_TESTS = ["matrix", "utils", "binarytree", "lqrdata", "nddata", "solver"]
GPU_TESTS = ["linalg", "linalg_custom", "nested_dissection"]


def _run(name):
    exe = os.path.join(BIN, name + "_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/tests not built (needs /root/reference at build time)")
    if not os.path.isdir("/root/repo/tests/golden"):
        pytest.skip("the fixtures' compiled-in path /root/repo/tests/golden does not exist here")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "ALL TESTS PASSED" in out and "TEST FAILED" not in out, (name, r.returncode, out[-3000:])


@pytest.mark.parametrize("name", CPU_TESTS)
def test_reference_unit_test_host_layer(name):
    _run(name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", GPU_TESTS)
def test_reference_unit_test_on_gpu(name):
    _run(name)
