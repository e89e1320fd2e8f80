"""The C-ABI shared libraries load on a CPU-only box and export every symbol their headers declare;
without a GPU the product refuses to run (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re
import subprocess

import pytest

import m17_cxx_demod_b200 as m

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header, prefix):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(" + prefix + r"\w+)\s*\(", src)))


def test_libm17b_exports_every_declared_symbol():
    lib = m.load_library()
    names = declared("m17b.h", "m17b_")
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"libm17b.so does not export {n}"
    assert lib.m17b_version() == 0x00010000


def test_libm17gen_exports_every_declared_symbol():
    from m17_cxx_demod_b200 import gen
    lib, cuda = gen.lib(), gen.cuda_lib()          # (the device-side modulator has its own library; loading needs no GPU)
    for n in declared("m17gen.h", "m17gen_"):
        assert hasattr(cuda if n == "m17gen_batch_device" else lib, n), f"libm17gen(_cuda).so does not export {n}"


def test_signatures_use_plain_c_types_only():
    src = open(os.path.join(ROOT, "include", "m17b.h")).read()
    assert "torch" not in src and "std::" not in src and "at::" not in src
    assert 'extern "C"' in src


def test_frame_struct_layout_matches_python_dtype():
    assert m.FRAME_DTYPE.itemsize == 80
    assert m.FRAME_DTYPE.fields["payload"][1] == 24 and m.FRAME_DTYPE.fields["lsf"][1] == 50


def test_product_is_built_for_sm_100a():
    out = subprocess.run(["cuobjdump", "-lelf", m.library_path()], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_no_cpu_fallback_without_device():
    if m.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(m.M17BError, match="no usable CUDA device"):
        m.Demodulator(4, 1024)


def test_product_never_references_the_oracle():
    """Nothing under the package may import, link or execute oracle/ (tests and bench legs only)."""
    pkg = os.path.join(ROOT, "m17_cxx_demod_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "libm17port" not in text and "libm17ref" not in text, f
    out = subprocess.run(["ldd", m.library_path()], capture_output=True, text=True).stdout
    assert "m17port" not in out and "m17ref" not in out
