"""The C-ABI library loads and exports every symbol include/mrz_gpu.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mrz_gpu.h")
LIB = os.path.join(ROOT, "modern-rzip_b200", "libmrz_gpu.so")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mrz_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "modern-rzip_b200"), "libmrz_gpu.so"], stdout=subprocess.DEVNULL)
    return C.CDLL(LIB)


def test_every_declared_symbol_is_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), s


def test_python_binding_lists_the_same_symbols():
    import mrz_py
    assert sorted(mrz_py.DECLARED_SYMBOLS) == declared_symbols()


def test_no_cpu_fallback(lib):
    """Without a CUDA device every compute entry point refuses to run."""
    lib.mrz_device_count.restype = C.c_int
    n = lib.mrz_device_count()
    if n > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    rc = lib.mrz_ctx_create(0, C.c_int64(1 << 20), C.byref(h))
    assert rc == -1 and not h.value      # MRZ_E_NODEVICE
    lib.mrz_last_error.restype = C.c_char_p
    assert lib.mrz_last_error(None)


def test_product_does_not_reference_the_oracle():
    """Nothing under modern-rzip_b200/ or include/ may link, load or include anything from oracle/."""
    bad = []
    for base in ("modern-rzip_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f.endswith((".c", ".h", ".cu", ".cuh", ".py", "Makefile")):
                    if re.search(r"oracle/|rzip_oracle|librzip_oracle", open(os.path.join(dp, f), errors="ignore").read().replace("oracle/spec_model.c", "")):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
