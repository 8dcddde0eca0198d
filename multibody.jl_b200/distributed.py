"""One process per GPU: trajectory sharding and the path's only collective.

Trajectory i depends only on (u0_i, p_i) and the shared read-only tables, so ranks never
exchange trajectory data.  The single collective is the all-reduce of
the ensemble statistic `[count, sum(x), sum(x^2)]` (SciML `reduction` / `u_init` of an
EnsembleProblem): 2 * n_save * n + 1 doubles.  It runs INSIDE the library
(`mbb200_comm_init` + `ncclAllReduce` on the solve stream, csrc/mb_runtime.cu); this module only
distributes the communicator id.  No PyTorch here: the launcher hands in a `broadcast` callable
(bench.py passes one built on its torchrun rendezvous, a Julia host would use MPI.jl).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Dict, Optional, Tuple

import numpy as np

NCCL_ID_BYTES = 128


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of rank `rank`: [lo, hi), same split as mbb200_solve uses across the
    devices of one process (csrc/mb_runtime.cu)."""
    return n_total * rank // world, n_total * (rank + 1) // world


def comm_init(rank: int, world: int, broadcast: Callable[[Optional[bytes]], bytes]) -> None:
    """Create the library's NCCL communicator over `world` processes (call after the devices are
    initialised).  `broadcast(payload_or_None) -> bytes` must return rank 0's payload on every
    rank (rank 0 passes the id, the others None)."""
    from . import _lib
    L = _lib.lib()
    buf = C.create_string_buffer(NCCL_ID_BYTES)
    if rank == 0:
        _lib.check(L.mbb200_comm_unique_id(buf, NCCL_ID_BYTES))
        ident = broadcast(bytes(buf.raw))
    else:
        ident = broadcast(None)
    if len(ident) != NCCL_ID_BYTES:
        raise ValueError("communicator id must be 128 bytes")
    _lib.check(L.mbb200_comm_init(C.create_string_buffer(ident, NCCL_ID_BYTES), NCCL_ID_BYTES, rank, world))


def comm_active() -> bool:
    from . import _lib
    return bool(_lib.lib().mbb200_comm_info(None, None))


def allreduce_statistics(red: np.ndarray) -> np.ndarray:
    """In-place sum over the ranks of the library communicator (no-op in a single process)."""
    from . import _lib
    red = np.ascontiguousarray(red, dtype=np.float64)
    _lib.check(_lib.lib().mbb200_comm_allreduce(_lib.ptr(red), red.size))
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
