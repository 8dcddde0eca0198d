"""ctypes binding of libmultibody_b200.so (include/mbb200.h).

The product path has no CPU fallback: if the shared library is missing, or no
B200 is visible, calls raise.  (The Julia glue in INTEGRATION.md binds the same
symbols with `ccall`; this module is its Python twin.)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MBB200_LIB") or os.path.join(HERE, "lib", "libmultibody_b200.so")
CSRC = os.path.join(HERE, "csrc")

ALG_TSIT5, ALG_RODAS5P, ALG_RODAS4, ALG_FBDF = 1, 2, 3, 4
FLAG_REDUCE = 1
RC_NAMES = {0: "Default", 1: "Success", 3: "DtNaN", 4: "MaxIters", 5: "DtLessThanMin", 6: "Unstable"}

# every symbol include/mbb200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "mbb200_abi_version", "mbb200_init", "mbb200_init_devices", "mbb200_device_count",
    "mbb200_model_load", "mbb200_model_info", "mbb200_model_flops", "mbb200_model_free",
    "mbb200_solve", "mbb200_solve_device", "mbb200_model_observed_count", "mbb200_observed", "mbb200_last_timing", "mbb200_measure_fp64_peak",
    "mbb200_dev_alloc", "mbb200_dev_free", "mbb200_dev_upload", "mbb200_dev_download",
    "mbb200_host_alloc_pinned", "mbb200_host_free_pinned", "mbb200_last_error", "mbb200_shutdown",
    "mbb200_solve_device_static", "mbb200_image_precompile",
    "mbb200_comm_unique_id", "mbb200_comm_init", "mbb200_comm_info", "mbb200_comm_allreduce", "mbb200_comm_destroy",
]


class Opts(C.Structure):
    _fields_ = [("alg", C.c_int32), ("flags", C.c_int32), ("t0", C.c_double), ("t1", C.c_double),
                ("abstol", C.c_double), ("reltol", C.c_double), ("dt0", C.c_double),
                ("dtmin", C.c_double), ("dtmax", C.c_double), ("maxiters", C.c_int64),
                ("qsteady_min", C.c_double), ("qsteady_max", C.c_double), ("pow_mode", C.c_int32),
                ("reserved", C.c_int32)]


class ImageHeader(C.Structure):
    _fields_ = [("magic", C.c_char * 8), ("name", C.c_char * 48), ("n", C.c_int32), ("m", C.c_int32),
                ("n_table", C.c_int64)]


class Mbb200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libmultibody_b200 error {code}: {msg}")
        self.code = code


def build(verbose: bool = False) -> str:
    """Compile the library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    # the generated model headers are build output: (re)emit them when the model / generator sources changed
    subprocess.check_call([sys.executable, os.path.join(os.path.dirname(CSRC), "codegen", "generate.py"), "--ensure"])
    subprocess.check_call(["make", "-C", CSRC, "-j8"], stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Mbb200Error(-100, f"{LIB_PATH} is missing: build it with `make -C {CSRC}` "
                                    "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_void_p
        L.mbb200_last_error.restype = C.c_char_p
        L.mbb200_init_devices.argtypes = [ip, C.c_int]
        L.mbb200_model_load.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
        L.mbb200_model_info.argtypes = [vp, ip, ip]
        L.mbb200_model_flops.argtypes = [vp, C.POINTER(C.c_int64)]
        L.mbb200_model_free.argtypes = [vp]
        L.mbb200_model_free.restype = None
        L.mbb200_solve.argtypes = [vp, C.POINTER(Opts), C.c_int64, vp, vp, vp, C.c_int32, vp, vp, vp, vp, vp]
        L.mbb200_solve_device.argtypes = [vp, C.POINTER(Opts), C.c_int, C.c_int64, vp, vp, vp, C.c_int32,
                                          vp, vp, vp, vp, vp]
        L.mbb200_solve_device_static.argtypes = [vp, C.POINTER(Opts), C.c_int, C.c_int64, vp, vp, vp, C.c_int32,
                                                 vp, vp, vp, vp, vp]
        L.mbb200_image_precompile.argtypes = [vp, C.c_size_t]
        L.mbb200_model_observed_count.argtypes = [vp, ip]
        L.mbb200_observed.argtypes = [vp, C.c_int64, vp, vp, vp, C.c_int32, vp]
        L.mbb200_last_timing.argtypes = [C.c_int, dp, ip]
        L.mbb200_measure_fp64_peak.argtypes = [C.c_int, dp]
        L.mbb200_dev_alloc.argtypes = [C.c_int, C.c_size_t, C.POINTER(vp)]
        L.mbb200_dev_free.argtypes = [C.c_int, vp]
        L.mbb200_dev_upload.argtypes = [C.c_int, vp, vp, C.c_size_t]
        L.mbb200_dev_download.argtypes = [C.c_int, vp, vp, C.c_size_t]
        L.mbb200_host_alloc_pinned.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.mbb200_host_free_pinned.argtypes = [vp]
        L.mbb200_shutdown.restype = None
        L.mbb200_comm_unique_id.argtypes = [vp, C.c_size_t]
        L.mbb200_comm_init.argtypes = [vp, C.c_size_t, C.c_int, C.c_int]
        L.mbb200_comm_info.argtypes = [ip, ip]
        L.mbb200_comm_allreduce.argtypes = [vp, C.c_int64]
        L.mbb200_comm_destroy.restype = None
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise Mbb200Error(rc, lib().mbb200_last_error().decode(errors="replace"))


def init_devices(ordinals) -> None:
    arr = (C.c_int32 * len(ordinals))(*ordinals)
    check(lib().mbb200_init_devices(arr, len(ordinals)))


def make_image(name: str, n: int, m: int, tables: np.ndarray) -> bytes:
    tables = np.ascontiguousarray(tables, dtype=np.float64)
    h = ImageHeader(b"MBIR1", name.encode(), n, m, tables.size)
    return bytes(h) + tables.tobytes()


class ModelHandle:
    def __init__(self, name: str, n: int, m: int, tables: np.ndarray, image: bytes = None):
        """`image`: an MBIR2 image (codegen/mbir.py) that carries the model; without it an MBIR1 image naming a kernel
        set compiled into the library is built."""
        img = image if image is not None else make_image(name, n, m, tables)
        self._h = C.c_void_p()
        buf = C.create_string_buffer(img, len(img))
        check(lib().mbb200_model_load(buf, len(img), C.byref(self._h)))
        self.name, self.n, self.m = name, n, m

    @property
    def handle(self):
        return self._h

    def flops(self):
        f = (C.c_int64 * 5)()
        check(lib().mbb200_model_flops(self._h, f))
        return dict(zip(("rhs", "jac", "dT", "factor", "solve"), [int(x) for x in f]))

    def free(self):
        if self._h:
            lib().mbb200_model_free(self._h)
            self._h = C.c_void_p()


def ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def last_timing(dev_index: int = 0):
    ms, nl = C.c_double(), C.c_int32()
    check(lib().mbb200_last_timing(dev_index, C.byref(ms), C.byref(nl)))
    return ms.value, nl.value


def measure_fp64_peak(dev_index: int = 0) -> float:
    v = C.c_double()
    check(lib().mbb200_measure_fp64_peak(dev_index, C.byref(v)))
    return v.value
