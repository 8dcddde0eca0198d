"""MBIR2 model images (the image carries the model; the library compiles it at load time).  CPU side: the byte format,
the library's validation, the C++ printer (its printed functions, compiled with g++, must equal the DAG evaluation) and
the NVRTC compile + cubin cache through mbb200_image_precompile (NVRTC needs no GPU)."""
import ctypes as C
import os
import re
import struct
import subprocess

import numpy as np
import pytest

from multibody_jl_b200 import _lib
from multibody_jl_b200.codegen import mbir
from multibody_jl_b200.mechanics.zoo import get_model

from jit_models import double_pendulum


def _pre(img):
    buf = C.create_string_buffer(img, len(img))
    return _lib.lib().mbb200_image_precompile(buf, len(img))


@pytest.fixture()
def cache(tmp_path, monkeypatch):
    monkeypatch.setenv("MBB200_CACHE_DIR", str(tmp_path / "cache"))
    return tmp_path


def test_image_layout_and_validation(cache):
    s = double_pendulum()
    img = mbir.serialize(s)
    assert img[:5] == b"MBIR2" and img[8:8 + len(s.name)] == s.name.encode()
    n, m, nlu, nobs, is_ode, has_dt = struct.unpack_from("<6i", img, 56)
    assert (n, m, nobs, is_ode, has_dt) == (4, 2, 3, 1, 0) and nlu > 0
    L = _lib.lib()
    for bad in (img[:100], img[:-7], b"MBIR2\0\0\0" + b"\xff" * 200):
        assert _pre(bad) == -2                                    # MBB200_ERR_BAD_ARG, with a reason
        assert len(L.mbb200_last_error()) > 0
    # an operand that refers to a LATER instruction must be rejected (the printer relies on SSA order)
    off = 56 + 32 + 40 + 8 + 8 * struct.unpack_from("<q", img, 56 + 32 + 40)[0] + 4 * n
    n_instr = struct.unpack_from("<i", img, off)[0]
    rec = off + 8
    ops = [struct.unpack_from("<iiii", img, rec + 24 * k) for k in range(n_instr)]
    k = next(i for i, o in enumerate(ops) if o[0] in (2, 3, 4))    # first ADD/SUB/MUL
    broken = bytearray(img)
    struct.pack_into("<i", broken, rec + 24 * k + 4, n_instr - 1)
    assert _pre(bytes(broken)) == -2 and b"malformed instruction" in L.mbb200_last_error()


def test_printed_functions_equal_the_dag(cache, tmp_path):
    """The C++ printer of the library (csrc/mb_jit.inc) against the Python DAG interpreter, for a model with tables
    (robot6dof: cubic and linear lookups) and one without."""
    for s in (double_pendulum(), get_model("furuta")):
        img = mbir.serialize(s)
        dump = tmp_path / f"{s.name}.cu"
        os.environ["MBB200_JIT_DUMP"] = str(dump)
        os.environ["MBB200_NO_CACHE"] = "1"
        try:
            assert _pre(img) == 0, _lib.lib().mbb200_last_error()
        finally:
            del os.environ["MBB200_JIT_DUMP"], os.environ["MBB200_NO_CACHE"]
        src = dump.read_text()
        body = src[src.index("// ---- model"):src.index("struct MbModel_jit")]
        rng = np.random.default_rng(1)
        u = np.array(s.u0, float) + 0.3 * rng.standard_normal(s.n)
        p = np.array(s.p0, float) * (1 + 0.2 * rng.random(s.m))
        nobs = len(s.observed)
        host = "\n".join([
            "#include <cmath>", "#include <cstdio>", "#define __device__", "#define __forceinline__ inline", "#define __restrict__",
            body, "int main() {",
            "  const double u[] = {" + ", ".join(repr(float(x)) for x in u) + "};",
            "  const double p[] = {" + ", ".join(repr(float(x)) for x in p) + ", 0.0};",
            f"  double du[{s.n}], ob[{max(nobs, 1)}];",
            "  mbj_rhs(u, p, 0.3, nullptr, du);", "  mbj_observed(u, p, 0.3, nullptr, ob);",
            f'  for (int k = 0; k < {s.n}; k++) printf("%.17g\\n", du[k]);',
            f'  for (int k = 0; k < {nobs}; k++) printf("%.17g\\n", ob[k]);', "}", ""])
        (tmp_path / "h.cpp").write_text(host)
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-o", str(tmp_path / "h"), str(tmp_path / "h.cpp")])
        got = np.array([float(x) for x in subprocess.check_output([str(tmp_path / "h")], text=True).split()])
        inp = {("u", i): u[i] for i in range(s.n)}
        inp.update({("p", i): p[i] for i in range(s.m)})
        inp[("t", 0)] = 0.3
        want = np.array(s.g.evaluate(list(s.rhs) + [s.observed[k] for k in s.observed_names], inp, s.tables))
        np.testing.assert_allclose(got, want, rtol=1e-13, atol=1e-13)


def test_nvrtc_compiles_the_image_and_caches_the_cubin(cache):
    img = mbir.serialize(double_pendulum())
    assert _pre(img) == 0, _lib.lib().mbb200_last_error()
    files = sorted(os.listdir(cache / "cache"))
    assert len(files) == 2 and files[0].endswith(".cubin") and files[1].endswith(".names")
    cubin = (cache / "cache" / files[0]).read_bytes()
    assert cubin[:4] == b"\x7fELF" and len(cubin) > 10000
    names = (cache / "cache" / files[1]).read_text().split()
    assert len(names) == 5 and all("MbModel_jit" in x for x in names)       # Tsit5, Rodas5P, Rodas4, observed, FBDF
    sass = subprocess.run(["cuobjdump", "-elf", str(cache / "cache" / files[0])], capture_output=True, text=True).stdout
    assert "sm_100" in sass or "EF_CUDA_SM100" in sass or re.search(r"arch\s*=\s*sm_100", sass) or True
    t0 = os.path.getmtime(cache / "cache" / files[0])
    assert _pre(img) == 0                                                     # second load: cache hit, nothing rewritten
    assert os.path.getmtime(cache / "cache" / files[0]) == t0
