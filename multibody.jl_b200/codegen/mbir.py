"""MBIR2: the model image that carries a model.

What the front end (this Python package here, the Julia glue of INTEGRATION.md on the reference's side) hands to
`mbb200_model_load` for a model that is NOT compiled into the library: the scheduled expression DAGs of the reduced
system's functions plus its tables.  The library prints them as straight-line CUDA C, appends its own integrator
templates and compiles the lot with NVRTC for sm_100a at load time (csrc/mb_jit.inc), caching the cubin by image hash.
This is the "IR (.mbir) -> CUDA C -> NVRTC -> cubin" half of SURVEY.md section 7.1 / 8f-1 that does not need Julia.

Byte layout (little endian, no padding beyond what is listed):

    char    magic[8]   = "MBIR2\\0\\0\\0"
    char    name[48]   NUL padded
    int32   n, m, nlu, nobs, is_ode, has_dt, reserved0, reserved1
    int64   flops[5]   algorithmic flop counts: rhs, jac, dT, factor (jac + LU), solve  
    int64   n_table    doubles of table data that follow (MBIR1 layout: t0, h, nk, then per table nk (value, z) pairs)
    float64 table[n_table]
    int32   mass[n]    1 = differential state, 0 = algebraic
    5 functions, in this order:  rhs(u,p,t) -> du[n];  dT(u,p,t) -> dT[n];  observed(u,p,t) -> obs[nobs];
                                 factor(u,p,t,c) -> LU[nlu];  solve(LU,c,p,b) -> b[n]
      int32 n_instr, n_out
      n_instr x { int32 op, a, b, c;  float64 val }        operands are indices of EARLIER instructions
      n_out   x int32                                      instruction that holds output k

    op codes = codegen/dag.py: CONST 0 (val), INPUT 1 (a = kind: u 0, p 1, t 2, c 3, lu 4, b 5; b = index), ADD 2, SUB 3,
    MUL 4, DIV 5, NEG 6, SIN 7, COS 8, SQRT 9, ATAN2 10, ABS 11, MIN 12, MAX 13, SELGE 14 (a >= 0 ? b : c),
    INTERP 15 (a = time operand, b = table, c = derivative order | (linear ? 256 : 0)).

The pivot sequence of the structured LU is not a separate field: `factor` and `solve` are already the eliminated
straight-line programs (codegen/model_emit.py), which is what keeps the device code free of pivot searches.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Sequence

import numpy as np

from .dag import Graph, CONST, INPUT, INTERP, ADD, SUB, MUL, DIV, ATAN2, MIN, MAX, SELGE
from .model_emit import ModelCode

KIND = {"u": 0, "p": 1, "t": 2, "c": 3, "lu": 4, "b": 5}
MAGIC = b"MBIR2\0\0\0"


def _function(g: Graph, outs: Sequence[int]) -> bytes:
    order = g.schedule(list(outs)) if outs else []
    pos: Dict[int, int] = {}
    rec: List[bytes] = []
    for n in order:
        o = g.op[n]
        a = b = c = -1
        val = 0.0
        if o == CONST:
            val = g.val[n]
        elif o == INPUT:
            kind, idx = g.payload[n]
            a, b = KIND[kind], int(idx)
        elif o == INTERP:
            kind, table, deriv = g.payload[n]
            a, b, c = pos[g.a[n]], int(table), int(deriv) | (256 if kind == "linear" else 0)
        else:
            a = pos[g.a[n]]
            if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                b = pos[g.b[n]]
            elif o == SELGE:
                b, c = pos[g.b[n]], pos[g.c[n]]
        pos[n] = len(rec)
        rec.append(struct.pack("<iiiid", o, a, b, c, val))
    head = struct.pack("<ii", len(rec), len(outs))
    return head + b"".join(rec) + b"".join(struct.pack("<i", pos[o]) for o in outs)


def serialize(sys, code: ModelCode = None) -> bytes:
    """Model image of a ReducedSystem (mechanics/builder.py)."""
    code = code or ModelCode(sys, pivots=sys.meta.get("pivots"))
    g = sys.g
    n, m = sys.n, sys.m
    obs = [sys.observed[k] for k in sys.observed_names]
    has_dt = int(any(d != g.zero for d in code.dT_nodes))
    is_ode = int(all(sys.mass_diag))
    tables = np.ascontiguousarray(sys.tables.pack(), dtype=np.float64)
    out = [MAGIC, sys.name.encode()[:47].ljust(48, b"\0"),
           struct.pack("<8i", n, m, code.nlu, len(obs), is_ode, has_dt, 0, 0),
           struct.pack("<5q", code.F_rhs, code.F_jac, code.F_dT, code.F_factor, code.F_solve),
           struct.pack("<q", tables.size), tables.tobytes(),
           np.asarray(sys.mass_diag, dtype="<i4").tobytes(),
           _function(g, list(sys.rhs)), _function(g, list(code.dT_nodes)), _function(g, obs),
           _function(g, list(code.lu_nodes)), _function(code.solve_graph, list(code.solve_out))]
    return b"".join(out)
