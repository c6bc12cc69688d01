"""C++20 drop-in facade (include/genogrove_b200/grove.hpp) — runs the compiled tests/cpp binaries.

* facade_tests.bin : the reference's hot-path test cases restated against the facade (GPU does all searching)
* ref_parity.bin   : the same scenarios on the facade AND on the unmodified reference compiled into the same
                     binary (traces must be identical), plus INTEGRATION.md's reference-side binding checked
                     key*-for-key* against the reference's own intersect.  Built only where /root/reference
                     exists; the GPU box runs the prebuilt file.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")


def _build():
    r = subprocess.run(["make", "-C", CPP], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def _run(name, timeout=900):
    path = os.path.join(CPP, name)
    if not os.path.exists(path):
        _build()
    return subprocess.run([path], capture_output=True, text=True, timeout=timeout)


def test_facade_header_compiles_and_scenarios_hold_on_the_reference():
    """CPU: the facade + tests compile as C++20; without a GPU the parity binary still checks that every
    expectation written into the scenarios holds on the unmodified reference (exit code 77 = no device)."""
    _build()
    assert os.path.exists(os.path.join(CPP, "facade_tests.bin"))
    if not os.path.exists(os.path.join(CPP, "ref_parity.bin")):
        pytest.skip("reference checkout not available: ref_parity.bin not built")
    from genogrove_b200 import _lib
    if _lib.lib().gg_device_count() > 0:
        pytest.skip("a CUDA device is present: covered by the gpu tests")
    r = _run("ref_parity.bin")
    assert r.returncode == 77, r.stdout + r.stderr
    assert " 0 failures" in r.stdout


@pytest.mark.gpu
def test_facade_scenarios_on_gpu():
    r = _run("facade_tests.bin")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert " 0 failures" in r.stdout


@pytest.mark.gpu
def test_facade_matches_unmodified_reference_on_gpu():
    if not os.path.exists(os.path.join(CPP, "ref_parity.bin")) and not os.path.isdir("/root/reference"):
        pytest.skip("ref_parity.bin was not built (no reference checkout)")
    r = _run("ref_parity.bin")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "0 differ" in r.stdout and " 0 failures" in r.stdout
