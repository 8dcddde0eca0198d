#!/usr/bin/env python
"""bench.py -- ensemble trajectories/sec on BASELINE.json's headline config (C4: Robot6DOF
parameter sweep, Rodas5P, abstol = reltol = 1e-8, saveat 0:0.1:2, 1 Mi trajectories per GPU).

    python bench.py --gpus N --steps K --warmup W            # our arm (one rank per GPU)
    python bench.py --impl reference --gpus N ...            # CPU arm: the oracle on host cores

A step = one pass of the hot path over the whole ensemble shard of this rank.
`value` = trajectories/s with inputs resident in HBM (device time, CUDA events on the
launching stream, max over ranks); `e2e` = the same through `mbb200_solve` with pinned HOST
buffers (H2D + kernels + D2H inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

METRIC = "ensemble trajectories/sec (Robot6DOF Rodas5P)"      # BASELINE.json metric (config C4)


def metric_name(spec):
    return METRIC if spec.config == "C4" else f"ensemble trajectories/sec ({spec.model} {spec.alg})"
UNIT = "trajectories/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C4")
    ap.add_argument("--alg", default="", choices=["", "tsit5", "rodas5p", "rodas4", "fbdf"],
                    help="solver override (default: the config's own; the headline metric is quoted on the default)")
    ap.add_argument("--trajectories", type=int, default=0, help="per-GPU ensemble size (0 = the config's full size)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="trajectories in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="tuning runs only: skip the host-buffer leg (the line then has no e2e)")
    ap.add_argument("--links", type=int, default=0, help="C5 / C5S only: rope length 20 or 50 (BASELINE config 5 sizes)")
    ap.add_argument("--scaling", default="auto", choices=["auto", "strong", "weak"],
                    help="N > 1: strong = the config's ensemble (1 Mi) sharded over the ranks (BASELINE config 4, default); "
                         "weak = that many trajectories per rank")
    ap.add_argument("--dump", default="", help="(reference arm) write the sample's results to this .npz for the parity check")
    return ap.parse_args()


def workload_name(spec, N):
    return (f"{spec.config} {spec.model} {spec.alg} abstol=reltol={spec.reltol:g} tspan={spec.tspan} "
            f"saveat={len(spec.saveat)} trajectories_per_gpu={N}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(len(r) > 3 + j and r[3 + j].lower().startswith("active") for r in self.rows)]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


def flops_per_trajectory(flops, n, n_save, stats, alg="rodas5p"):
    """Algorithmic FLOPs (BASELINE.md section 4 / SURVEY.md 8d), from the measured step counts.
    Rodas5P: F_step = F_factor(jac + structured LU) + F_dT + 8 F_rhs + 8 F_solve + 86 n + 4 n.
    Tsit5:   F_step = 6 F_rhs + 56 n (the two are equal: no linear algebra)."""
    steps = stats[0].astype(np.float64) + stats[1].astype(np.float64)
    if alg == "tsit5":
        f_step = dense_step = 6 * flops["rhs"] + 56 * n
        f_interp = (7 * 2 + 7 * 8 + 2) * n
        return steps * f_step + n_save * f_interp + 3 * flops["rhs"], f_step, dense_step, steps
    if alg == "fbdf":
        # per step attempt: one factor (jac + LU); per Newton iteration (= linear solve, stats[4]): rhs + solve + 8 n;
        # predictor / differences / error norms ~ 40 n
        iters = stats[4].astype(np.float64)
        f_step = flops["factor"] + 40 * n
        dense_step = flops["jac"] + (2 * n ** 3) // 3 + 40 * n
        per_it = flops["rhs"] + flops["solve"] + 8 * n
        return steps * f_step + iters * per_it + n_save * 12 * n + 3 * flops["rhs"], f_step, dense_step, steps
    nst = 6 if alg == "rodas4" else 8
    f_step = flops["factor"] + flops["dT"] + nst * flops["rhs"] + nst * flops["solve"] + 86 * n + 4 * n
    f_interp = (8 * 3 * 2 + 10) * n
    dense_step = flops["jac"] + flops["dT"] + (2 * n ** 3) // 3 + nst * flops["rhs"] + 2 * nst * n * n + 86 * n + 4 * n
    return steps * f_step + n_save * f_interp + 2 * flops["rhs"], f_step, dense_step, steps


def run_reference(args):
    """CPU arm: the oracle (CPU restatement of the reference algorithm; the Julia reference
    itself cannot run here) on all host cores, a bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import mbpkg
    mbpkg.load()
    import mbo
    from multibody_jl_b200 import ensembles
    from multibody_jl_b200.mechanics.zoo import get_model
    mbo.build()
    cores = os.cpu_count() or 1
    sample = args.cpu_sample or (cores * (2 if args.links <= 20 else 1) if args.links else max(256, 128 * cores))
    spec = ensembles.make_ensemble(args.config, sample, links=args.links or None)
    spec.alg = args.alg or spec.alg
    tb = get_model(spec.model).tables.pack()

    # the ropes' generated oracle units are compiled at -O0 (198k lines: -O1 takes 5.5 min of gcc); the loop-based code of
    # the same equations (csrc/mb_chain.h, -O2, finite-difference Jacobian as the reference's rope examples use) is 7x
    # faster on the CPU (profiles/r02_rope_long.md) and is therefore the CPU arm of C5 / C5E
    cpu_model = {"rope6": "rope6c", "rope3e": "rope3ec"}.get(spec.model, spec.model)

    def step():
        t = time.perf_counter()
        r = mbo.solve(cpu_model, spec.alg, spec.u0, spec.p, tb, spec.tspan, spec.saveat, abstol=spec.abstol,
                      reltol=spec.reltol, nthreads=cores)
        dt_ = time.perf_counter() - t
        assert (r["retcode"] == 1).all()
        if args.dump:
            np.savez(args.dump, u=r["u"], retcode=r["retcode"], stats=r["stats"])
        return dt_
    for _ in range(args.warmup):
        step()
    dt = sum(step() for _ in range(args.steps))
    v = sample * args.steps / dt
    full = spec.n_total if not args.trajectories else args.trajectories
    print(json.dumps({
        "impl": "reference", "metric": metric_name(spec), "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(spec, full), "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "model": cpu_model,
                         "sample": f"first {sample} trajectories of the same workload per step (same RNG stream), all {cores} host threads, "
                                   f"{dt / args.steps:.1f} s per step"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import ctypes as C
    import mbpkg
    mbpkg.load()
    from multibody_jl_b200 import _lib, api, ensembles
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    from multibody_jl_b200 import distributed as mbdist
    spec0 = ensembles.make_ensemble(args.config, 1, links=args.links or None)
    spec0.alg = args.alg or spec0.alg
    scaling = args.scaling if args.scaling != "auto" else "strong"
    if world == 1:
        scaling = "weak" if args.scaling == "auto" else scaling       # one GPU: both mean the same run
    if scaling == "strong":     # the whole ensemble (BASELINE config 4: 1 Mi trajectories) sharded by trajectory index
        n_job = args.trajectories or spec0.n_total
        lo, hi = mbdist.shard_range(n_job, rank, world)
    else:                       # fixed work per GPU
        n_job = (args.trajectories or spec0.n_total) * world
        lo, hi = rank * (n_job // world), (rank + 1) * (n_job // world)
    N = hi - lo
    spec = ensembles.make_ensemble(args.config, N, offset=lo, links=args.links or None)
    spec.alg = args.alg or spec.alg
    sys_ = api.multibody(spec.model)
    n, m, ns = sys_.n, sys_.m, len(spec.saveat)
    ens = api.EnsembleB200(devices=[local])
    h = api._ensure(ens, sys_)
    if world > 1:
        # the path's only collective runs inside the library (ncclAllReduce behind mbb200_comm_init); torchrun's
        # rendezvous is used just to hand rank 0's communicator id to the other ranks
        def bcast(payload):
            box = [payload]
            dist.broadcast_object_list(box, src=0)
            return box[0]
        mbdist.comm_init(rank, world, bcast)
    flops = h.flops()
    alg = {"tsit5": api.Tsit5(), "rodas5p": api.Rodas5P(), "rodas4": api.Rodas4(), "fbdf": api.FBDF()}[spec.alg]
    opts = api.make_opts(alg, spec.tspan, spec.abstol, spec.reltol, reduce=True)
    L = _lib.lib()

    # ---- device-resident inputs/outputs (torch is only the allocator here) --------------------
    d_u0 = torch.from_numpy(spec.u0).to(dev)
    d_p = torch.from_numpy(spec.p).to(dev)
    d_out = torch.empty((ns, n, N), dtype=torch.float64, device=dev)
    d_tf = torch.empty(N, dtype=torch.float64, device=dev)
    d_rc = torch.empty(N, dtype=torch.int32, device=dev)
    d_st = torch.empty((5, N), dtype=torch.int32, device=dev)
    d_red = torch.zeros(2 * ns * n + 1, dtype=torch.float64, device=dev)
    sv = np.ascontiguousarray(spec.saveat)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        _lib.check(L.mbb200_solve_device(h.handle, C.byref(opts), 0, N, d_u0.data_ptr(), d_p.data_ptr(), _lib.ptr(sv), ns,
                                         d_out.data_ptr(), d_tf.data_ptr(), d_rc.data_ptr(), d_st.data_ptr(),
                                         d_red.data_ptr()))
        ms, nl = _lib.last_timing(0)      # with a communicator d_red now holds the statistic summed over all ranks
        return ms, nl

    for _ in range(args.warmup):
        device_step()
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    t_wall = time.perf_counter()
    dev_ms, launches = 0.0, 0
    for _ in range(args.steps):
        ms, nl = device_step()
        dev_ms += ms
        launches += nl
    barrier()
    wall = time.perf_counter() - t_wall
    clk = clocks.stop()
    tms = torch.tensor([dev_ms, wall * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    dev_ms_max, wall_ms_max = tms.tolist()
    value = n_job * args.steps / (dev_ms_max * 1e-3)
    red_count = float(d_red[0].item())          # successful trajectories of the whole job (all-reduced in the library)

    rc = d_rc.cpu().numpy()
    stats = d_st.cpu().numpy()
    ok = int((rc == 1).sum())
    ftraj, f_step, dense_step, steps = flops_per_trajectory(flops, n, ns, stats, spec.alg)
    peak = _lib.measure_fp64_peak(0)
    achieved = float(ftraj.sum()) * args.steps / (dev_ms * 1e-3)
    # DRAM bytes of the dominant kernel per launch: ncu-measured bytes per trajectory (one
    # `--set full` capture, profiles/traffic.json) x trajectories of this launch
    traffic, hbm = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath)).get(spec.config)
        # a measured-traffic entry is only used for the kernel it was captured on: the kernel string must name this
        # model (or this rope length) and this solver's tableau; anything else is stale and `traffic` stays null
        chain = sys_.meta.get("chain")
        want = (f"chain_warp_kernel<{chain[0]}, {int(chain[1])}," if chain else f"MbModel_{spec.model},")
        tab = {"rodas5p": ("TabRodas5P", ", 2,"), "rodas4": ("TabRodas4", ", 3,"), "tsit5": (", 1,",), "fbdf": ("bdf_kernel",)}[spec.alg]
        # ... and the entry names the mangled symbol of the kernel it was captured on (template arguments = CTA size
        # and row mode included): it must be present in the library this run loaded
        sym_ok = True
        if tj and tj.get("symbol"):
            with open(_lib.LIB_PATH, "rb") as fh:
                sym_ok = tj["symbol"].encode() in fh.read()
        if tj and sym_ok and want in tj.get("kernel", "") and any(x in tj["kernel"] for x in tab):
            traffic = tj["dram_bytes_per_trajectory"] * N
            peak_hbm = 6530.6
            ppath = os.path.join(ROOT, "MEASURED_PEAKS.json")
            if os.path.exists(ppath):
                peak_hbm = json.load(open(ppath)).get("hbm_gbs", peak_hbm)
            gbs = traffic * args.steps / (dev_ms * 1e-3) / 1e9
            hbm = {"achieved": gbs, "peak": peak_hbm, "unit": "GB/s", "frac": gbs / peak_hbm, "kernel": tj["kernel"],
                   "note": "ncu dram bytes per trajectory of this kernel (profiles/r02_ncu_summary.md) x trajectories / launch time: "
                           "thread-local / per-warp working state cycling through the L2, not algorithmic traffic"}
        elif tj:
            print(f"bench: profiles/traffic.json[{spec.config}] was captured on '{tj.get('kernel')}', not on this kernel; traffic = null",
                  file=sys.stderr)

    # ---- end to end through the host-buffer ABI call (pinned host memory) ----------------------
    h_u0 = torch.from_numpy(spec.u0).pin_memory()
    h_p = torch.from_numpy(spec.p).pin_memory()
    h_out = torch.empty((ns, n, N), dtype=torch.float64).pin_memory()
    h_tf = torch.empty(N, dtype=torch.float64).pin_memory()
    h_rc = torch.empty(N, dtype=torch.int32).pin_memory()
    h_st = torch.empty((5, N), dtype=torch.int32).pin_memory()
    h_red = np.zeros(2 * ns * n + 1)

    def e2e_step():
        _lib.check(L.mbb200_solve(h.handle, C.byref(opts), N, h_u0.data_ptr(), h_p.data_ptr(), _lib.ptr(sv), ns,
                                  h_out.data_ptr(), h_tf.data_ptr(), h_rc.data_ptr(), h_st.data_ptr(), _lib.ptr(h_red)))
    if not args.no_e2e:
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(0 if args.no_e2e else args.steps):
        e2e_step()
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = None if args.no_e2e else n_job * args.steps / e2e_t.item()
    same = None if args.no_e2e else bool(np.array_equal(h_rc.numpy(), rc))
    # the same call with ordinary pageable buffers (what a Julia Array is): page-locked inside mbb200_solve
    e2e_pageable = None
    if not args.no_e2e:
        g_u0, g_p = np.array(spec.u0), np.array(spec.p)
        g_out = np.empty((ns, n, N)); g_tf = np.empty(N); g_rc = np.empty(N, np.int32); g_st = np.empty((5, N), np.int32)
        barrier()
        t0 = time.perf_counter()
        _lib.check(L.mbb200_solve(h.handle, C.byref(opts), N, _lib.ptr(g_u0), _lib.ptr(g_p), _lib.ptr(sv), ns,
                                  _lib.ptr(g_out), _lib.ptr(g_tf), _lib.ptr(g_rc), _lib.ptr(g_st), _lib.ptr(h_red)))
        barrier()
        tp = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tp, op=dist.ReduceOp.MAX)
        e2e_pageable = n_job / tp.item()
        same = same and bool(np.array_equal(g_rc, rc))
        del g_out
    h2d = 8 * N * (n + m) + 8 * ns
    d2h = 8 * N * n * ns + N * (8 + 4 + 20) + 8 * (2 * ns * n + 1)

    out = {
        "metric": metric_name(spec), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(spec, N), "full_size": n_job >= spec.n_total, "trajectories_job": n_job,
                   "l2": "inputs+outputs per step (%.1f GB) exceed the 126 MB L2; no flush needed" % ((h2d + d2h) / 1e9),
                   "parallelism": f"trajectory shards x{world}, no data-path collective; ncclAllReduce of the ensemble statistic inside the library"},
        "successful_trajectories_job": red_count,
        "wall_ms_per_step": wall_ms_max / args.steps,
        "successful_trajectories": ok, "mean_steps_per_trajectory": float(steps.mean()),
        "mean_rejects_per_trajectory": float(stats[1].mean()),
        "gpu_launches": launches, "clocks": clk,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "matches_device_run": same, "host_buffers": "pinned", "pageable_value": e2e_pageable},
        "roofline": {"bound": "fp64", "achieved": achieved / 1e12, "peak": peak / 1e12, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "peak_source": "FP64 DFMA probe measured in this run (MEASURED_PEAKS.json holds HBM/BF16 only)",
                     "flops_per_step_structured": f_step, "flops_per_step_dense_equiv": dense_step,
                     "hbm_gbs_algorithmic": (h2d + d2h) * args.steps / (dev_ms * 1e-3) / 1e9,
                     "dram_measured": hbm},
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the oracle on the host cores, in a fresh interpreter (no torch/OpenMP worker threads
        # spinning next to the oracle's pthreads); bounded sample of the same workload
        cores = os.cpu_count() or 1
        sample = args.cpu_sample or max(256, 256 * cores)
        sample = min(sample, N)
        dump = os.path.join("/tmp", f"mbb200_parity_{os.getpid()}.npz")
        if args.links and not args.cpu_sample:      # long ropes cost seconds per trajectory on a core: one or two per thread
            sample = min(N, cores * (2 if args.links <= 20 else 1))
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--config", args.config,
                            "--links", str(args.links), "--steps", "1", "--warmup", "0", "--cpu-sample", str(sample), "--dump", dump],
                           capture_output=True, text=True, env={k: v for k, v in os.environ.items()
                                                                if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")})
        try:
            out["cpu_baseline"] = json.loads(r.stdout.strip().splitlines()[-1])["cpu_baseline"]
        except Exception as e:   # keep the GPU line even if the CPU leg failed
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"failed: {e}: {r.stderr[-200:]}"}
        # parity of THIS run: the oracle's sample (first `sample` trajectories of the stream) against the device result
        try:
            ref = np.load(dump)
            gpu_u = d_out[:, :, :sample].cpu().numpy()
            scale = np.maximum(np.abs(ref["u"]).max(axis=(0, 2), keepdims=True), 1e-3)
            err = np.abs(gpu_u - ref["u"]) / scale
            out["parity"] = {"n": int(sample), "max_scaled_err": float(err.max()), "tolerance": 1e-6,
                             "retcodes_equal": bool(np.array_equal(ref["retcode"], rc[:sample])),
                             "naccept_gpu_mean": float(stats[0, :sample].mean()), "naccept_oracle_mean": float(ref["stats"][0].mean()),
                             "against": "oracle (CPU restatement; Rosenbrock path unpinned below 3.5e-4 vs reference-run outputs, profiles/r02_pin_table.md)"}
            os.remove(dump)
        except Exception as e:
            out["parity"] = {"n": 0, "error": str(e)}
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
