"""The C-ABI library loads and exports every symbol include/mbb200.h declares; without a
GPU the compute entries fail loudly (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from multibody_jl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mbb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mbb200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_all_exported():
    names = _declared()
    assert len(names) >= 20
    L = _lib.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mbb200.h but not exported"
    assert sorted(_lib.SYMBOLS) == names, "the ctypes binding must bind exactly the header's entry points"


def test_abi_version_and_struct_layout():
    assert _lib.lib().mbb200_abi_version() == 2
    assert C.sizeof(_lib.Opts) == 96            # 2 x int32 + 7 x double + int64 + 2 x double + 2 x int32
    assert C.sizeof(_lib.ImageHeader) == 72     # 8 + 48 + 2 x int32 + int64


def test_no_visible_internal_symbols():
    import subprocess
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert all(s.startswith("mbb200_") for s in exported), exported


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.Mbb200Error, match="no CPU fallback"):
        _lib.lib()


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_without_a_device_every_compute_entry_refuses():
    L = _lib.lib()
    assert L.mbb200_init(1) == -1                      # MBB200_ERR_NO_DEVICE
    assert b"no CUDA device" in L.mbb200_last_error()
    assert L.mbb200_device_count() == 0
    h = C.c_void_p()
    img = _lib.make_image("pendulum", 2, 2, __import__("numpy").array([0.0, 1.0, 0.0]))
    buf = C.create_string_buffer(img, len(img))
    assert L.mbb200_model_load(buf, len(img), C.byref(h)) == -5     # MBB200_ERR_NOT_INIT
    v = C.c_double()
    assert L.mbb200_measure_fp64_peak(0, C.byref(v)) == -2
