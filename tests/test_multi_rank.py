"""N > 1 path on CPU: two gloo ranks shard an ensemble by trajectory index, solve their
shards independently (here with the oracle standing in for the device) and all-reduce the
ensemble statistic; the result must equal the single-rank answer."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, N, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import mbpkg
    mbpkg.load()
    import torch.distributed as dist
    import mbo
    from multibody_jl_b200 import ensembles
    import torch
    from multibody_jl_b200.distributed import host_statistics, shard_range
    from multibody_jl_b200.mechanics.zoo import get_model
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(N, rank, world)
    spec = ensembles.make_ensemble("C2", hi - lo, offset=lo)
    r = mbo.solve(spec.model, spec.alg, spec.u0, spec.p, get_model(spec.model).tables.pack(), spec.tspan,
                  spec.saveat[:5], abstol=1e-8, reltol=1e-8)
    red = host_statistics(r["u"], r["retcode"])
    # the product sums the statistic inside the library (ncclAllReduce behind mbb200_comm_init, GPU test
    # test_library_communicator_*); here gloo plays that part so the sharding logic runs on CPU
    t = torch.from_numpy(red)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put((lo, hi, red))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_reduce_to_the_single_rank_statistic(oracle):
    from multibody_jl_b200 import ensembles
    from multibody_jl_b200.distributed import finalize_statistics, host_statistics, shard_range
    from multibody_jl_b200.mechanics.zoo import get_model
    N, world = 37, 2                       # ragged split: 18 + 19
    assert [shard_range(N, r, world) for r in range(world)] == [(0, 18), (18, 37)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, N, q)) for r in range(world)]
    for p in procs:
        p.start()
    lo, hi, red = q.get(timeout=240)
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    spec = ensembles.make_ensemble("C2", N)
    whole = oracle.solve(spec.model, spec.alg, spec.u0, spec.p, get_model(spec.model).tables.pack(), spec.tspan,
                         spec.saveat[:5], abstol=1e-8, reltol=1e-8)
    want = host_statistics(whole["u"], whole["retcode"])
    np.testing.assert_allclose(red, want, rtol=1e-12, atol=1e-12)
    st = finalize_statistics(red, 5, 14)
    assert st["count"] == N and st["mean"].shape == (5, 14) and (st["var"] >= 0).all()


def test_shard_ranges_cover_everything():
    from multibody_jl_b200.distributed import shard_range
    for N in (0, 1, 7, 1 << 20):
        for world in (1, 2, 3, 8):
            r = [shard_range(N, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == N and all(r[k][1] == r[k + 1][0] for k in range(world - 1))
