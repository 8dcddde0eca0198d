"""ORACLE -- test infrastructure only (tests/, __graft_entry__.smoke(), bench.py's
cpu_baseline / --impl reference leg).  ctypes binding of oracle/_build/libmbo.so."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libmbo.so")
ALG = {"tsit5": 1, "rodas5p": 2, "rodas4": 3, "fbdf": 4}
_lib = None


class Opts(C.Structure):
    _fields_ = [("alg", C.c_int32), ("jac_mode", C.c_int32), ("abstol", C.c_double),
                ("reltol", C.c_double), ("dt0", C.c_double), ("dtmin", C.c_double),
                ("dtmax", C.c_double), ("maxiters", C.c_int64),
                ("qsteady_min", C.c_double), ("qsteady_max", C.c_double), ("pow_mode", C.c_int32),
                ("reserved", C.c_int32)]


def build(force=False):
    if force and os.path.exists(LIB):
        os.remove(LIB)
    # the oracle compiles the same generated model headers as the product: make sure they exist and are current
    subprocess.check_call([sys.executable, os.path.join(os.path.dirname(HERE), "multibody.jl_b200", "codegen", "generate.py"), "--ensure"])
    subprocess.check_call(["make", "-C", HERE, "-j8"], stdout=subprocess.DEVNULL)   # no-op when up to date
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def dims(model):
    n, m, nnz = C.c_int(), C.c_int(), C.c_int()
    if lib().mbo_model_dims(model.encode(), C.byref(n), C.byref(m), C.byref(nnz)) != 0:
        raise KeyError(model)
    return n.value, m.value, nnz.value


def eval_fn(model, what, u, p, t, tb):
    n, m, nnz = dims(model)
    out = np.zeros({0: n, 1: nnz, 2: n, 3: max(lib().mbo_nobs(model.encode()), 1)}[what])
    u = np.ascontiguousarray(u, dtype=np.float64); p = np.ascontiguousarray(p, dtype=np.float64)
    tb = np.ascontiguousarray(tb, dtype=np.float64)
    lib().mbo_eval(model.encode(), what, _dp(u), _dp(p), C.c_double(t), _dp(tb), _dp(out))
    return out


def factor_solve(model, u, p, t, tb, c, b):
    """x = (c*M - J)^-1 b through the product's generated factor/solve code (CPU compile)."""
    x = np.array(b, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64); p = np.ascontiguousarray(p, dtype=np.float64)
    tb = np.ascontiguousarray(tb, dtype=np.float64)
    lib().mbo_factor_solve(model.encode(), _dp(u), _dp(p), C.c_double(t), _dp(tb), C.c_double(c), _dp(x))
    return x


def pivmax(model, u, p, t, tb, c):
    """Largest |1/pivot| stored by the generated static-pivot factor() at this state (what the kernels' guard sees)."""
    f = lib().mbo_pivmax
    f.restype = C.c_double
    u = np.ascontiguousarray(u, dtype=np.float64); p = np.ascontiguousarray(p, dtype=np.float64)
    tb = np.ascontiguousarray(tb, dtype=np.float64)
    return f(model.encode(), _dp(u), _dp(p), C.c_double(t), _dp(tb), C.c_double(c))


def mass_diag(model):
    n, m, nnz = dims(model)
    jr = np.zeros(nnz, np.int32); jc = np.zeros(nnz, np.int32); ms = np.zeros(n, np.int32)
    lib().mbo_pattern(model.encode(), _ip(jr), _ip(jc), _ip(ms))
    return ms


def jac_dense(model, u, p, t, tb):
    n, m, nnz = dims(model)
    jr = np.zeros(nnz, np.int32); jc = np.zeros(nnz, np.int32); ms = np.zeros(n, np.int32)
    lib().mbo_pattern(model.encode(), _ip(jr), _ip(jc), _ip(ms))
    J = np.zeros((n, n))
    J[jr, jc] = eval_fn(model, 1, u, p, t, tb)
    return J


def solve(model, alg, u0, p, tb, tspan, saveat, abstol=1e-6, reltol=1e-3, jac_mode=0,
          dt0=0.0, dtmin=0.0, dtmax=0.0, maxiters=0, nthreads=1, qsteady=(0.0, 0.0), pow_mode=0):
    """u0: [n][N], p: [m][N] (SoA).  Returns dict(u=[nsave][n][N], t, retcode, stats=[5][N])."""
    n, m, _ = dims(model)
    u0 = np.ascontiguousarray(u0, dtype=np.float64).reshape(n, -1)
    N = u0.shape[1]
    p = np.ascontiguousarray(p, dtype=np.float64).reshape(max(m, 0), -1) if m else np.zeros((1, N))
    tb = np.ascontiguousarray(tb, dtype=np.float64)
    saveat = np.ascontiguousarray(saveat, dtype=np.float64)
    ns = len(saveat)
    out = np.full((ns, n, N), np.nan)
    tf = np.zeros(N); rc = np.zeros(N, np.int32); st = np.zeros((5, N), np.int32)
    o = Opts(ALG[alg], jac_mode, abstol, reltol, dt0, dtmin, dtmax, maxiters, qsteady[0], qsteady[1], pow_mode, 0)
    f = lib().mbo_solve_ensemble
    f.argtypes = [C.c_char_p, C.POINTER(Opts), C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double),
                  C.POINTER(C.c_double), C.c_double, C.c_double, C.POINTER(C.c_double), C.c_int,
                  C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int]
    r = f(model.encode(), C.byref(o), N, _dp(u0), _dp(p), _dp(tb), tspan[0], tspan[1], _dp(saveat), ns,
          _dp(out), _dp(tf), _ip(rc), _ip(st), nthreads)
    if r != 0:
        raise RuntimeError(f"oracle: unknown model {model}")
    return dict(u=out, t=tf, retcode=rc, stats=st)
