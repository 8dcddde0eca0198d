"""Frame / rotation algebra on DAG expressions.

Conventions follow the reference exactly (src/orientation.jl): `R` rotates
world -> local (`resolve2`, :120; `resolve1` is R^T, :133), `w` is the angular
velocity resolved in the local frame, composition is `R2 = Rrel * R1`,
`w2 = Rrel*w1 + wrel` (:72-74, :169-176).
"""
from __future__ import annotations

from typing import List, Sequence

from ..codegen.dag import Ex, sin, cos, sqrt

Vec = List  # 3-list of Ex or float
Mat = List  # 3x3 nested list


def vadd(a, b): return [a[0] + b[0], a[1] + b[1], a[2] + b[2]]
def vsub(a, b): return [a[0] - b[0], a[1] - b[1], a[2] - b[2]]
def vscale(s, a): return [s * a[0], s * a[1], s * a[2]]
def vneg(a): return [-a[0], -a[1], -a[2]]
def dot(a, b): return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def cross(a, b):
    return [a[1] * b[2] - a[2] * b[1],
            a[2] * b[0] - a[0] * b[2],
            a[0] * b[1] - a[1] * b[0]]


def matvec(R, v):
    """resolve2: R*v (world -> local when R is a frame orientation), orientation.jl:120"""
    return [R[i][0] * v[0] + R[i][1] * v[1] + R[i][2] * v[2] for i in range(3)]


def matTvec(R, v):
    """resolve1: R'*v (local -> world), orientation.jl:133"""
    return [R[0][i] * v[0] + R[1][i] * v[1] + R[2][i] * v[2] for i in range(3)]


def matmul(A, B):
    return [[A[i][0] * B[0][j] + A[i][1] * B[1][j] + A[i][2] * B[2][j]
             for j in range(3)] for i in range(3)]


def transpose(A):
    return [[A[j][i] for j in range(3)] for i in range(3)]


def eye():
    return [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]


def planar_rotation(n: Sequence[float], phi: Ex):
    """R = n n' + (I - n n') cos(phi) - skew(n) sin(phi)   (orientation.jl:153-160)."""
    c, s = cos(phi), sin(phi)
    sk = [[0.0, -n[2], n[1]], [n[2], 0.0, -n[0]], [-n[1], n[0], 0.0]]
    R = []
    for i in range(3):
        row = []
        for j in range(3):
            ee = n[i] * n[j]
            d = (1.0 if i == j else 0.0) - ee
            row.append(ee + d * c - sk[i][j] * s)
        R.append(row)
    return R


def _axis_rot(k: int, a: Ex):
    """A_k(alpha) = Rot_k(alpha)^T: elementary world->local rotation about axis k
    (orientation.jl:313-323 `axis_rotation`)."""
    c, s = cos(a), sin(a)
    if k == 1:
        return [[1.0, 0.0, 0.0], [0.0, c, s], [0.0, -s, c]]
    if k == 2:
        return [[c, 0.0, -s], [0.0, 1.0, 0.0], [s, 0.0, c]]
    return [[c, s, 0.0], [-s, c, 0.0], [0.0, 0.0, 1.0]]


def _unit(k: int):
    e = [0.0, 0.0, 0.0]
    e[k - 1] = 1.0
    return e


def axes_rotations(seq: Sequence[int], ang: Sequence[Ex], der: Sequence[Ex]):
    """Euler/Cardan rotation and its local angular velocity (orientation.jl:288-299):
    R = A_s3(a3) A_s2(a2) A_s1(a1);
    w = e_s3 d3 + A_s3 e_s2 d2 + A_s3 A_s2 e_s1 d1."""
    A1 = _axis_rot(seq[0], ang[0])
    A2 = _axis_rot(seq[1], ang[1])
    A3 = _axis_rot(seq[2], ang[2])
    R = matmul(A3, matmul(A2, A1))
    w3 = vscale(der[2], _unit(seq[2]))
    w2 = vscale(der[1], matvec(A3, _unit(seq[1])))
    w1 = vscale(der[0], matvec(A3, matvec(A2, _unit(seq[0]))))
    return R, vadd(w3, vadd(w2, w1))


def from_Q(Q: Sequence[Ex]):
    """Rotation matrix (world->local) of a unit quaternion [s,i,j,k]
    (orientation.jl:258-266: `RotMatrix(q)'`)."""
    s, i, j, k = Q
    R = [[1.0 - 2.0 * (j * j + k * k), 2.0 * (i * j - k * s), 2.0 * (i * k + j * s)],
         [2.0 * (i * j + k * s), 1.0 - 2.0 * (i * i + k * k), 2.0 * (j * k - i * s)],
         [2.0 * (i * k - j * s), 2.0 * (j * k + i * s), 1.0 - 2.0 * (i * i + j * j)]]
    return transpose(R)


def norm(v):
    return sqrt(dot(v, v))
