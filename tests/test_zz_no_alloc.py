"""mbb200_solve_device_static: the allocation-free variant of mbb200_solve_device (scratch in a static
device array; no counterpart in the reference).

Collected last on purpose (file name).  Called through the C ABI
and compared with the existing entry point on two calls in a row -- the second with a different
ensemble, save grid and algorithm state, so anything the shared scratch array kept from the first
call (work counter, save times) would show."""
import ctypes as C

import numpy as np
import pytest

from multibody_jl_b200 import _lib, api, ensembles

pytestmark = pytest.mark.gpu
RTOL = 1e-6      # the tolerance of tests/test_gpu_parity.py (same kernel on both sides: expected bit-equal)


class _DevBuf:
    """Device buffer owned by the CALLER (mbb200_dev_alloc), as the entry point's contract says."""

    def __init__(self, nbytes):
        self.p = C.c_void_p()
        self.nbytes = max(int(nbytes), 8)
        _lib.check(_lib.lib().mbb200_dev_alloc(0, self.nbytes, C.byref(self.p)))

    def upload(self, a):
        a = np.ascontiguousarray(a)
        _lib.check(_lib.lib().mbb200_dev_upload(0, self.p, _lib.ptr(a), a.nbytes))
        return self

    def download(self, shape, dtype):
        out = np.empty(shape, dtype)
        if out.nbytes:
            _lib.check(_lib.lib().mbb200_dev_download(0, _lib.ptr(out), self.p, out.nbytes))
        return out

    def free(self):
        _lib.lib().mbb200_dev_free(0, self.p)


def _solve_both(config, N, n_keep, offset, algname):
    spec = ensembles.make_ensemble(config, N, offset=offset)
    sys_ = api.multibody(spec.model)
    h = api._ensure(api.EnsembleB200(), sys_)
    alg = {"tsit5": api.Tsit5(), "rodas5p": api.Rodas5P()}[algname]
    opts = api.make_opts(alg, spec.tspan, spec.abstol, spec.reltol, reduce=True)
    sv = np.ascontiguousarray(spec.saveat[:n_keep])
    n, m, ns = sys_.n, sys_.m, len(sv)
    L = _lib.lib()
    bufs = []

    def buf(nbytes):
        b = _DevBuf(nbytes)
        bufs.append(b)
        return b
    d_u0 = buf(spec.u0.nbytes).upload(spec.u0)
    d_p = buf(spec.p.nbytes).upload(spec.p) if m else None
    res = {}
    try:
        for name, fn in (("existing", L.mbb200_solve_device), ("static", L.mbb200_solve_device_static)):
            d_out, d_tf = buf(8 * ns * n * N), buf(8 * N)
            d_rc, d_st, d_red = buf(4 * N), buf(4 * 5 * N), buf(8 * (2 * ns * n + 1))
            d_out.upload(np.full((ns, n, N), -7.0))            # poison: every value must be written by the call
            _lib.check(fn(h.handle, C.byref(opts), 0, N, d_u0.p, d_p.p if m else None, _lib.ptr(sv), ns,
                          d_out.p, d_tf.p, d_rc.p, d_st.p, d_red.p))
            res[name] = dict(u=d_out.download((ns, n, N), np.float64), t=d_tf.download((N,), np.float64),
                             rc=d_rc.download((N,), np.int32), st=d_st.download((5, N), np.int32),
                             red=d_red.download((2 * ns * n + 1,), np.float64))
    finally:
        for b in bufs:
            b.free()
    return res


def _check(res):
    a, b = res["existing"], res["static"]
    assert np.all(a["rc"] == 1) and np.array_equal(a["rc"], b["rc"])
    scale = np.maximum(np.abs(a["u"]).max(axis=(0, 2), keepdims=True), 1e-3)
    err = np.abs(a["u"] - b["u"]) / scale
    assert np.isfinite(b["u"]).all() and err.max() < RTOL, err.max()
    np.testing.assert_allclose(b["t"], a["t"], rtol=0, atol=1e-12)
    assert np.array_equal(a["st"], b["st"])                     # same step counts: the same trajectories were integrated
    np.testing.assert_allclose(b["red"], a["red"], rtol=1e-9, atol=1e-9)
    assert b["red"][0] == a["rc"].size                          # every trajectory dispensed exactly once


def test_static_scratch_entry_matches_the_existing_one_on_two_calls_in_a_row():
    # call 1: 2000 spring-damper trajectories, 37 save points (Tsit5)
    _check(_solve_both("C2", 2000, 37, 0, "tsit5"))
    # call 2 right after it: other model, other size, SHORTER save grid, Rosenbrock kernel -- the scratch array still
    # holds call 1's work counter (2000 + one per thread) and its 37 save times
    _check(_solve_both("C4", 300, 21, 4096, "rodas5p"))
    # and once more with the first shape (counter left far beyond N by call 2)
    _check(_solve_both("C2", 777, 100, 13, "tsit5"))


def test_static_scratch_entry_rejects_sizes_beyond_its_limit():
    spec = ensembles.make_ensemble("C1", 4)
    sys_ = api.multibody(spec.model)
    h = api._ensure(api.EnsembleB200(), sys_)
    opts = api.make_opts(api.Tsit5(), spec.tspan, 1e-8, 1e-8)
    ns = 4097                                                   # MBB200_STATIC_MAX_SAVE + 1
    sv = np.linspace(spec.tspan[0], spec.tspan[1], ns)
    dummy = _DevBuf(64)
    try:
        rc = _lib.lib().mbb200_solve_device_static(h.handle, C.byref(opts), 0, 4, dummy.p, dummy.p, _lib.ptr(sv), ns,
                                                   dummy.p, dummy.p, dummy.p, dummy.p, None)
        assert rc == -2                                         # MBB200_ERR_BAD_ARG, nothing launched
        assert b"MBB200_STATIC_MAX_SAVE" in _lib.lib().mbb200_last_error()
    finally:
        dummy.free()
