"""CPU: the C-ABI shared library loads and exports every symbol include/gg_b200.h declares; without a GPU the
product fails loudly instead of falling back to any CPU path."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gg_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gg_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    from genogrove_b200 import _lib
    assert declared_symbols() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol():
    from genogrove_b200 import _lib
    L = _lib.lib()
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in gg_b200.h but not exported by libgg_b200.so"
    assert L.gg_version() == 100


def test_library_is_sm100a_native_and_has_no_torch_dependency():
    from genogrove_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    ldd = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "libtorch" not in ldd and "libpython" not in ldd and "libc10" not in ldd
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass, "TMA bulk copy (cp.async.bulk) missing from the count kernel"


def test_no_cpu_fallback_without_a_gpu():
    """On a machine without a CUDA device the product must refuse to work (no CPU path)."""
    from genogrove_b200 import _lib, DeviceIndex, GGError, Grove
    L = _lib.lib()
    if L.gg_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = ctypes.c_void_p()
    assert L.gg_index_create(ctypes.byref(h), 0) == 2  # GG_ERR_CUDA
    with pytest.raises(GGError):
        DeviceIndex(0)
    with pytest.raises(GGError):
        Grove(3)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "genogrove_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), f
                assert "liboracle" not in text and "libgg_ref" not in text, f
    for f in os.listdir(os.path.join(ROOT, "include")):
        p = os.path.join(ROOT, "include", f)
        if os.path.isfile(p):
            assert "oracle" not in open(p).read().lower()
