"""MBIR2 on the GPU: a model that is NOT compiled into the library (the reference's double pendulum,
test/basic_tests.jl:274-325) is carried by its image, compiled at load time and solved -- through the Python host and
through the C++ driver of the ABI (tests/abi_driver.cpp)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from multibody_jl_b200 import _lib, api
from multibody_jl_b200.codegen import mbir

from jit_models import double_pendulum

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ensemble(s, N):
    rng = np.random.default_rng(2)
    U = np.repeat(np.array(s.u0, float)[:, None], N, axis=1)
    U[0] += 0.3 * rng.standard_normal(N)
    P = np.repeat(np.array(s.p0, float)[:, None], N, axis=1)
    P[0] *= 1.0 + 0.5 * rng.random(N)
    U[:, 0], P[:, 0] = s.u0, s.p0                       # trajectory 0 = the reference's own test case
    return np.ascontiguousarray(U), np.ascontiguousarray(P)


def test_load_time_compiled_model_matches_scipy_and_the_reference_assertions(tmp_path, monkeypatch):
    monkeypatch.setenv("MBB200_CACHE_DIR", str(tmp_path / "cache"))
    from scipy.integrate import solve_ivp
    s = double_pendulum()
    with pytest.raises(_lib.Mbb200Error):               # really not in the library's registry
        _lib.ModelHandle(s.name, s.n, s.m, s.tables.pack())
    N = 96
    U, P = _ensemble(s, N)
    sv = np.array([0.5, 1.0, 2.0])
    ep = api.EnsembleProblem(api.ODEProblem(s, tspan=(0.0, 2.0)), u0_matrix=U, p_matrix=P)
    for alg in (api.Tsit5(), api.Rodas5P(), api.Rodas4(), api.FBDF()):
        sol = api.solve(ep, alg, api.EnsembleB200(), trajectories=N, saveat=sv, abstol=1e-9, reltol=1e-9)
        assert np.all(sol.retcode == 1)
        bound = 2e-5 if isinstance(alg, api.FBDF) else 2e-7            # a BDF's global error sits further above its tolerance
        for j in (0, 17, 95):
            ref = solve_ivp(lambda t, u: s.eval_rhs(u, P[:, j], t), (0.0, 2.0), U[:, j], method="DOP853", rtol=1e-11, atol=1e-12, t_eval=sv)
            assert np.abs(sol.u[:, :, j] - ref.y.T).max() < bound, (type(alg).__name__, j)
    # the reference's own run: Rodas4, default tolerances, (0, 35); rev1.phi -> -pi/2, body2.r_0[2] -> -2 (rtol 1e-2)
    ep35 = api.EnsembleProblem(api.ODEProblem(s, tspan=(0.0, 35.0)))
    sol = api.solve(ep35, api.Rodas4(), api.EnsembleB200(), trajectories=3, saveat=np.array([35.0]))
    assert api.successful_retcode(sol)
    assert abs(sol.state("rev1.phi")[0, 0] + np.pi / 2) < 0.01 * np.pi / 2
    assert abs(sol.observed("body2.r_0[2]")["body2.r_0[2]"][0, 0] + 2.0) < 0.02
    assert len(os.listdir(tmp_path / "cache")) == 2     # one cubin + its kernel names


def test_cpp_driver_of_the_abi_loads_the_image_and_agrees_with_the_python_host(tmp_path, monkeypatch):
    monkeypatch.setenv("MBB200_CACHE_DIR", str(tmp_path / "cache"))
    exe = tmp_path / "abi_driver"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-o", str(exe), os.path.join(ROOT, "tests", "abi_driver.cpp"), "-ldl"])
    s = double_pendulum()
    N = 500
    U, P = _ensemble(s, N)
    sv = np.linspace(0.2, 2.0, 10)
    (tmp_path / "model.mbir").write_bytes(mbir.serialize(s))
    job = struct.pack("<4iq4d", 2, s.n, s.m, len(sv), N, 0.0, 2.0, 1e-8, 1e-8) + U.tobytes() + P.tobytes() + sv.tobytes()
    (tmp_path / "job.bin").write_bytes(job)
    out = subprocess.run([str(exe), _lib.LIB_PATH, str(tmp_path / "model.mbir"), str(tmp_path / "job.bin"), str(tmp_path / "res.bin")],
                         capture_output=True, text=True, env=dict(os.environ))
    assert out.returncode == 0, out.stderr + out.stdout
    assert f"{N} successful" in out.stdout
    raw = np.fromfile(tmp_path / "res.bin", dtype=np.uint8)
    u = raw[:8 * len(sv) * s.n * N].view(np.float64).reshape(len(sv), s.n, N)
    ep = api.EnsembleProblem(api.ODEProblem(s, tspan=(0.0, 2.0)), u0_matrix=U, p_matrix=P)
    sol = api.solve(ep, api.Rodas5P(), api.EnsembleB200(), trajectories=N, saveat=sv, abstol=1e-8, reltol=1e-8)
    assert np.array_equal(sol.u, u)                      # same image, same library: bit-identical


def test_image_models_point_gravity_and_spring_with_mass(tmp_path, monkeypatch):
    """The component variants of tests/test_components_extra.py on the device (carried by their images): point gravity
    (test/basic_tests.jl:124-168) and `Spring(c, m)` / `LineForceWithMass` (src/forces.jl:279-326)."""
    monkeypatch.setenv("MBB200_CACHE_DIR", str(tmp_path / "cache"))
    from scipy.integrate import solve_ivp
    for name, t1 in (("point_gravity", 5.0), ("spring_with_mass", 3.0)):
        s = api.multibody(name)
        N = 40
        rng = np.random.default_rng(6)
        U = np.repeat(np.array(s.u0, float)[:, None], N, axis=1)
        P = np.repeat(np.array(s.p0, float)[:, None], N, axis=1) * (1.0 + 0.2 * rng.random((s.m, N)))
        P[:, 0] = s.p0
        sv = np.array([0.5 * t1, t1])
        ep = api.EnsembleProblem(api.ODEProblem(s, tspan=(0.0, t1)), u0_matrix=U, p_matrix=P)
        sol = api.solve(ep, api.Tsit5(), api.EnsembleB200(), trajectories=N, saveat=sv, abstol=1e-10, reltol=1e-10)
        assert np.all(sol.retcode == 1)
        for j in (0, 13):
            ref = solve_ivp(lambda t, u: s.eval_rhs(u, P[:, j], t), (0.0, t1), U[:, j], method="DOP853", rtol=1e-11, atol=1e-13, t_eval=sv)
            assert np.abs(sol.u[:, :, j] - ref.y.T).max() < 1e-7, (name, j)
    s = api.multibody("point_gravity")
    assert abs(sol.u.shape[1] - 2) == 0                      # (last loop item was the 2-state spring)
