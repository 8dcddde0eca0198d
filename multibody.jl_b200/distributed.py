"""One process per GPU: trajectory sharding and the path's only collective.

Trajectory i depends only on (u0_i, p_i) and the shared read-only tables, so ranks never
exchange trajectory data.  The single collective is the all-reduce of
the ensemble statistic `[count, sum(x), sum(x^2)]` (SciML `reduction` / `u_init` of an
EnsembleProblem): 2 * n_save * n + 1 doubles, NCCL over NVLink on GPUs, gloo in CPU tests.
"""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of rank `rank`: [lo, hi), same split as mbb200_solve uses across the
    devices of one process (csrc/mb_runtime.cu)."""
    return n_total * rank // world, n_total * (rank + 1) // world


def allreduce_statistics(red, group=None):
    """In-place sum over ranks of the raw statistic vector (torch tensor on the rank's device,
    or a numpy array that is staged through a CPU tensor)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return red
    if isinstance(red, np.ndarray):
        t = torch.from_numpy(red)
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        red[:] = t.cpu().numpy()
        return red
    dist.all_reduce(red, op=dist.ReduceOp.SUM, group=group)
    return red


def finalize_statistics(red: np.ndarray, n_save: int, n: int) -> Dict[str, np.ndarray]:
    """count / mean / variance per (save point, state) from the raw reduced vector."""
    red = np.asarray(red, dtype=np.float64)
    cnt = float(red[0])
    s1 = red[1:1 + n_save * n].reshape(n_save, n)
    s2 = red[1 + n_save * n:1 + 2 * n_save * n].reshape(n_save, n)
    c = max(cnt, 1.0)
    return dict(count=cnt, sum=s1, sumsq=s2, mean=s1 / c, var=np.maximum(s2 / c - (s1 / c) ** 2, 0.0), raw=red)


def host_statistics(u: np.ndarray, retcode: np.ndarray) -> np.ndarray:
    """The same raw vector computed on the host from saved states u [n_save][n][N] (used to
    check the device reduction kernel)."""
    ok = np.asarray(retcode) == 1
    x = u[:, :, ok]
    return np.concatenate([[float(ok.sum())], x.sum(axis=2).ravel(), (x * x).sum(axis=2).ravel()])
