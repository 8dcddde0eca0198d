/* Loop-based right-hand side of `Rope` chains of any length (the reference: src/components.jl:574-626;
 * docs/src/examples/ropes_and_cables.md:24-36 stiff, :48-60 flexible).
 *
 * The straight-line code generator is O(n^3) in code size on a chain (mbm_rope6.h: 198k lines), so ropes
 * with 20-50 links (BASELINE.json config 5) get this hand-written O(N) formulation instead: Featherstone's
 * articulated-body algorithm in body coordinates over a list of joints
 *     spherical joint with Euler angles, sequence [1,2,3]  (`Spherical(state=true)`, src/joints.jl:213-259,
 *         R_rel = A3 A2 A1 and w_rel = S(phi) phid of src/orientation.jl:288-299)   3 degrees of freedom
 *     prismatic joint along `dir` (`Prismatic`, src/joints.jl:139-151; flexible ropes only)   1 degree of freedom
 * with the bodies of `Rope`: `BodyShape(m = m/n, r = l/n dir)` links (stiff) or `Body(m/n, r_cm = l/(2n) dir)` +
 * `Prismatic` with `Translational.Spring(c n)`/`Damper(d n)` (flexible, :594-611), air resistance as in
 * `Body` (src/components.jl:445-458), joint damping resolved as the reference does (src/joints.jl:218-222),
 * and a `Body(m = body.m)` at the end.  States, parameters and their order are those of the generated models
 * `rope6` / `rope3e` (mechanics/zoo.py), against which this file is checked (tests/test_chain.py).
 *
 * One source for both sides: the CUDA kernels (csrc/mb_chain_kernel.cuh) and the CPU oracle include it.
 * MBC_FN is the function qualifier; per-joint scratch is addressed as ws[k * stride] (stride 1 on the CPU,
 * 32 on the GPU where the 32 lanes of a warp evaluate 32 Jacobian columns side by side).
 */
#ifndef MB_CHAIN_H
#define MB_CHAIN_H
#include <math.h>
#ifndef MBC_FN
#define MBC_FN static inline
#endif

typedef struct {
  int nlinks;      /* N */
  int elastic;     /* 0: stiff links, 1: Prismatic + spring/damper links */
  double dir[3];   /* unit direction of the rope in its own frame */
  double l, m;     /* total length and mass */
} MbcRope;

#define MBC_G 9.80665          /* World(): g along -y (src/components.jl:63) */
#define MBC_IC 0.001           /* default inertia of Body (src/components.jl:345) */
#define MBC_LINK_DOUBLES 64    /* scratch per joint: E 9 | v 6 | c 6 | Y = L^-1 U' 18 | L 9 | z = L^-1 uu 3 | S 9 | r 3 */

MBC_FN int mbc_njoints(const MbcRope* R) { return R->elastic ? 2 * R->nlinks + 1 : R->nlinks + 1; }
MBC_FN int mbc_nstates(const MbcRope* R) { return R->elastic ? 8 * R->nlinks + 6 : 6 * (R->nlinks + 1); }
MBC_FN int mbc_nparams(const MbcRope* R) { return R->elastic ? 5 : 3; }
/* joint j: type (0 spherical, 1 prismatic) and offset of its [q, qd] block in the state vector */
MBC_FN int mbc_jtype(const MbcRope* R, int j) { return R->elastic ? (j & 1) : 0; }
MBC_FN int mbc_joff(const MbcRope* R, int j) { return R->elastic ? 8 * (j >> 1) + ((j & 1) ? 6 : 0) : 6 * j; }

MBC_FN void mbc_cross(const double* a, const double* b, double* o) {
  const double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
MBC_FN void mbc_mv(const double* M, const double* v, double* o) {       /* o = M v, M row-major 3x3 */
  const double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2], y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2],
               z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
MBC_FN void mbc_mtv(const double* M, const double* v, double* o) {      /* o = M' v */
  const double x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2], y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2],
               z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
/* body carried by joint j: mass and centre of mass (Ic = MBC_IC * 1 about the centre of mass) */
MBC_FN double mbc_body(const MbcRope* R, int j, const double* p, double* cm) {
  const int N = R->nlinks;
  const double li = R->l / N;
  cm[0] = cm[1] = cm[2] = 0.0;
  if (R->elastic && (j & 1)) return 0.0;                 /* the prismatic joint carries nothing itself */
  const int i = R->elastic ? (j >> 1) : j;
  if (i == N) return p[1];                                /* Body(m = body.m) at the rope's end */
  cm[0] = 0.5 * li * R->dir[0]; cm[1] = 0.5 * li * R->dir[1]; cm[2] = 0.5 * li * R->dir[2];
  return R->m / N;
}

/* o = skew(r) M  (every column m of M becomes r x m), M row-major 3x3 */
MBC_FN void mbc_skewmul(const double* r, const double* M, double* o) {
  double t[9];
  for (int k = 0; k < 3; k++) {
    const double c[3] = {M[k], M[3 + k], M[6 + k]};
    double x[3];
    mbc_cross(r, c, x);
    t[k] = x[0]; t[3 + k] = x[1]; t[6 + k] = x[2];
  }
  for (int k = 0; k < 9; k++) o[k] = t[k];
}
/* o = E' M E */
MBC_FN void mbc_rot_block(const double* E, const double* M, double* o) {
  double t[9], r[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) t[3 * i + j] = E[i] * M[j] + E[3 + i] * M[3 + j] + E[6 + i] * M[6 + j];          /* E' M */
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r[3 * i + j] = t[3 * i] * E[j] + t[3 * i + 1] * E[3 + j] + t[3 * i + 2] * E[6 + j]; /* (E'M) E */
  for (int k = 0; k < 9; k++) o[k] = r[k];
}
/* Cholesky factor of the joint-space inertia D (3x3, symmetric positive definite): L = [l00, l10, l11, l20, l21, l22].
 * D is ill-conditioned on a rope (bending inertia ~ m r^2 against the 0.001 twist inertia), so its inverse is never
 * formed: an explicit inverse by cofactors lost 6-8 digits here (profiles/r02_rope_long.md); triangular solves are
 * backward stable. */
/* The diagonal of L is stored as its reciprocal (slots 0, 2, 5): the sweep applies L^-1 to nine vectors per joint,
 * and a multiplication is worth ~10 divisions on the FP64 pipe. */
MBC_FN void mbc_chol3(const double* D, double* L) {
  L[0] = 1.0 / sqrt(D[0]);
  L[1] = D[3] * L[0];
  L[3] = D[6] * L[0];
  L[2] = 1.0 / sqrt(D[4] - L[1] * L[1]);
  L[4] = (D[7] - L[3] * L[1]) * L[2];
  L[5] = 1.0 / sqrt(D[8] - L[3] * L[3] - L[4] * L[4]);
}
MBC_FN void mbc_lsolve3(const double* L, double* x) {          /* x <- L^-1 x */
  x[0] = x[0] * L[0];
  x[1] = (x[1] - L[1] * x[0]) * L[2];
  x[2] = (x[2] - L[3] * x[0] - L[4] * x[1]) * L[5];
}
MBC_FN void mbc_ltsolve3(const double* L, double* x) {         /* x <- L^-T x */
  x[2] = x[2] * L[5];
  x[1] = (x[1] - L[4] * x[2]) * L[2];
  x[0] = (x[0] - L[1] * x[1] - L[3] * x[2]) * L[0];
}

#define MBC_W(j, k) ws[(size_t)((j) * MBC_LINK_DOUBLES + (k)) * stride]
enum { MBC_E = 0, MBC_V = 9, MBC_C = 15, MBC_U = 21, MBC_DI = 39, MBC_UU = 48, MBC_S = 51, MBC_R = 60 };

/* du = f(u + eps e_jp, p) of the rope described by R (jp < 0: no perturbation; the perturbed form is what a
 * forward-difference Jacobian column needs, without a copy of the state per column).
 * p = [d_joint, body.m, air_resistance (, c, d)]. */
#define MBC_UK(k) (u[k] + (((k) == jp) ? eps : 0.0))
MBC_FN void mbc_rope_rhs(const MbcRope* R, const double* u, int jp, double eps, const double* p, double* du, double* ws,
                         int stride) {
  const int NJ = mbc_njoints(R), N = R->nlinks;
  const double li = R->l / N, d_joint = p[0], air = p[2];
  const double* dir = R->dir;
  /* ---- pass 1: velocities and velocity-product accelerations, base to tip ---------------------------- */
  double vp[6] = {0, 0, 0, 0, 0, 0};
  for (int j = 0; j < NJ; j++) {
    const int o = mbc_joff(R, j);
    double E[9], r[3], v[6], c[6];
    if (mbc_jtype(R, j) == 0) {
      const double s1 = sin(MBC_UK(o)), c1 = cos(MBC_UK(o)), s2 = sin(MBC_UK(o + 1)), c2 = cos(MBC_UK(o + 1)), s3 = sin(MBC_UK(o + 2)), c3 = cos(MBC_UK(o + 2));
      const double q1 = MBC_UK(o + 3), q2 = MBC_UK(o + 4), q3 = MBC_UK(o + 5);
      /* E = A3 A2 A1 (src/orientation.jl:288-299, 313-323) */
      const double A21[9] = {c2, s2 * s1, -s2 * c1, 0.0, c1, s1, s2, -c2 * s1, c2 * c1};
      for (int k = 0; k < 3; k++) {
        E[k] = c3 * A21[k] + s3 * A21[3 + k];
        E[3 + k] = -s3 * A21[k] + c3 * A21[3 + k];
        E[6 + k] = A21[6 + k];
      }
      /* w_rel = S1 phid1 + S2 phid2 + S3 phid3 with S1 = A3 A2 e1, S2 = A3 e2, S3 = e3 */
      const double S[9] = {c2 * c3, -c2 * s3, s2, s3, c3, 0.0, 0.0, 0.0, 1.0};
      const double wrel[3] = {S[0] * q1 + S[3] * q2, S[1] * q1 + S[4] * q2, S[2] * q1 + q3};
      /* (dS/dt) phid */
      const double sd[3] = {q1 * (-s2 * c3 * q2 - c2 * s3 * q3) + q2 * (c3 * q3), q1 * (s2 * s3 * q2 - c2 * c3 * q3) - q2 * (s3 * q3),
                            q1 * (c2 * q2)};
      const int stiff_shift = (!R->elastic && j > 0);
      r[0] = stiff_shift ? li * dir[0] : 0.0; r[1] = stiff_shift ? li * dir[1] : 0.0; r[2] = stiff_shift ? li * dir[2] : 0.0;
      double t[3], wl[3];
      mbc_cross(vp, r, t);                                   /* w_p x r */
      t[0] += vp[3]; t[1] += vp[4]; t[2] += vp[5];
      mbc_mv(E, vp, wl);
      mbc_mv(E, t, v + 3);
      v[0] = wl[0] + wrel[0]; v[1] = wl[1] + wrel[1]; v[2] = wl[2] + wrel[2];
      mbc_cross(v, wrel, c);                                  /* w x w_rel */
      c[0] += sd[0]; c[1] += sd[1]; c[2] += sd[2];
      mbc_cross(v + 3, wrel, c + 3);                          /* v x w_rel */
      for (int k = 0; k < 9; k++) MBC_W(j, MBC_S + k) = S[k];
    } else {
      const double s = MBC_UK(o), sdot = MBC_UK(o + 1);
      for (int k = 0; k < 9; k++) E[k] = (k % 4 == 0) ? 1.0 : 0.0;
      r[0] = dir[0] * s; r[1] = dir[1] * s; r[2] = dir[2] * s;
      double t[3], vj[3] = {dir[0] * sdot, dir[1] * sdot, dir[2] * sdot};
      mbc_cross(vp, r, t);
      v[0] = vp[0]; v[1] = vp[1]; v[2] = vp[2];
      v[3] = vp[3] + t[0] + vj[0]; v[4] = vp[4] + t[1] + vj[1]; v[5] = vp[5] + t[2] + vj[2];
      c[0] = c[1] = c[2] = 0.0;
      mbc_cross(v, vj, c + 3);                                /* w x (dir sdot) */
    }
    for (int k = 0; k < 9; k++) MBC_W(j, MBC_E + k) = E[k];
    for (int k = 0; k < 6; k++) { MBC_W(j, MBC_V + k) = v[k]; MBC_W(j, MBC_C + k) = c[k]; vp[k] = v[k]; }
    for (int k = 0; k < 3; k++) MBC_W(j, MBC_R + k) = r[k];
  }
  /* ---- pass 2: articulated-body inertias and bias forces, tip to base -------------------------------- */
  double Ac[9] = {0}, Bc[9] = {0}, Cc[9] = {0}, pc[6] = {0};     /* contribution of the child, in this joint's frame */
  for (int j = NJ - 1; j >= 0; j--) {
    const int o = mbc_joff(R, j);
    double cm[3], v[6], c[6], E[9], r[3];
    const double m = mbc_body(R, j, p, cm);
    for (int k = 0; k < 6; k++) { v[k] = MBC_W(j, MBC_V + k); c[k] = MBC_W(j, MBC_C + k); }
    for (int k = 0; k < 9; k++) E[k] = MBC_W(j, MBC_E + k);
    for (int k = 0; k < 3; k++) r[k] = MBC_W(j, MBC_R + k);
    double A[9], B[9], C[9], pA[6];
    for (int k = 0; k < 9; k++) { A[k] = Ac[k]; B[k] = Bc[k]; C[k] = Cc[k]; }
    for (int k = 0; k < 6; k++) pA[k] = pc[k];
    if (m != 0.0) {
      const double cc = cm[0] * cm[0] + cm[1] * cm[1] + cm[2] * cm[2];
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) A[3 * a + b] += ((a == b) ? (MBC_IC + m * cc) : 0.0) - m * cm[a] * cm[b];
      B[1] += -m * cm[2]; B[2] += m * cm[1]; B[3] += m * cm[2]; B[5] += -m * cm[0]; B[6] += -m * cm[1]; B[7] += m * cm[0];
      C[0] += m; C[4] += m; C[8] += m;
      /* p = v x* (I v) - f_ext */
      double wxc[3], h[3], n[3], t[3];
      mbc_cross(v, cm, wxc);
      h[0] = m * (v[3] + wxc[0]); h[1] = m * (v[4] + wxc[1]); h[2] = m * (v[5] + wxc[2]);
      mbc_cross(cm, h, n);
      n[0] += MBC_IC * v[0]; n[1] += MBC_IC * v[1]; n[2] += MBC_IC * v[2];
      mbc_cross(v, n, t); pA[0] += t[0]; pA[1] += t[1]; pA[2] += t[2];
      mbc_cross(v + 3, h, t); pA[0] += t[0]; pA[1] += t[1]; pA[2] += t[2];
      mbc_cross(v, h, t); pA[3] += t[0]; pA[4] += t[1]; pA[5] += t[2];
      /* air resistance of Body (src/components.jl:445-458): force -m c |v| v on the body, moment arm r_cm */
      /* (links only: the Body at the rope's end is the user's own component and has none) */
      const double air_j = (j == NJ - 1) ? 0.0 : air;
      const double nv = sqrt(v[3] * v[3] + v[4] * v[4] + v[5] * v[5] + 1e-300), k_ = m * air_j * nv;
      const double F[3] = {-k_ * v[3], -k_ * v[4], -k_ * v[5]};
      mbc_cross(cm, F, t);
      pA[0] -= t[0]; pA[1] -= t[1]; pA[2] -= t[2]; pA[3] -= F[0]; pA[4] -= F[1]; pA[5] -= F[2];
    }
    /* joint-space inertia D = S' IA S = L L';  Y = L^-1 (IA S)' (nd x 6);  z = L^-1 (tau - S' pA) */
    double Y[18], L[9] = {0}, z[3] = {0, 0, 0};
    int nd;
    if (mbc_jtype(R, j) == 0) {
      nd = 3;
      double S[9], D[9], U[18], t[3];
      for (int k = 0; k < 9; k++) S[k] = MBC_W(j, MBC_S + k);
      for (int a = 0; a < 3; a++) {                           /* U(:,a) = IA [S_a; 0] */
        mbc_mv(A, S + 3 * a, t);
        U[0 + a] = t[0]; U[3 + a] = t[1]; U[6 + a] = t[2];
        mbc_mtv(B, S + 3 * a, t);
        U[9 + a] = t[0]; U[12 + a] = t[1]; U[15 + a] = t[2];
      }
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) D[3 * a + b] = S[3 * a] * U[b] + S[3 * a + 1] * U[3 + b] + S[3 * a + 2] * U[6 + b];
      mbc_chol3(D, L);
      for (int rr = 0; rr < 6; rr++) {
        t[0] = U[3 * rr]; t[1] = U[3 * rr + 1]; t[2] = U[3 * rr + 2];
        mbc_lsolve3(L, t);
        Y[rr] = t[0]; Y[6 + rr] = t[1]; Y[12 + rr] = t[2];
      }
      /* joint damping as the reference resolves it: fric = Rrel (d w_rel), tau_a = -S_a . fric */
      const double q1 = MBC_UK(o + 3), q2 = MBC_UK(o + 4), q3 = MBC_UK(o + 5);
      const double wrel[3] = {S[0] * q1 + S[3] * q2 + S[6] * q3, S[1] * q1 + S[4] * q2 + S[7] * q3, S[2] * q1 + S[5] * q2 + S[8] * q3};
      double fr[3];
      mbc_mv(E, wrel, fr);
      for (int a = 0; a < 3; a++)
        z[a] = -d_joint * (S[3 * a] * fr[0] + S[3 * a + 1] * fr[1] + S[3 * a + 2] * fr[2])
               - (S[3 * a] * pA[0] + S[3 * a + 1] * pA[1] + S[3 * a + 2] * pA[2]);
      mbc_lsolve3(L, z);
    } else {
      nd = 1;
      double t[3], t2[3];
      for (int k = 0; k < 18; k++) Y[k] = 0.0;
      mbc_mv(B, dir, t);
      mbc_mv(C, dir, t2);
      L[0] = 1.0 / sqrt(dir[0] * t2[0] + dir[1] * t2[1] + dir[2] * t2[2]);       /* reciprocal, as in mbc_chol3 */
      Y[0] = t[0] * L[0]; Y[1] = t[1] * L[0]; Y[2] = t[2] * L[0]; Y[3] = t2[0] * L[0]; Y[4] = t2[1] * L[0]; Y[5] = t2[2] * L[0];
      const double fs = -((N * p[3]) * (MBC_UK(o) - li) + (N * p[4]) * MBC_UK(o + 1));   /* Spring(c n, s_rel0 = l/n) + Damper(d n) */
      z[0] = (fs - (dir[0] * pA[3] + dir[1] * pA[4] + dir[2] * pA[5])) * L[0];
    }
    for (int k = 0; k < 18; k++) MBC_W(j, MBC_U + k) = Y[k];
    for (int k = 0; k < 9; k++) MBC_W(j, MBC_DI + k) = L[k];
    for (int k = 0; k < 3; k++) MBC_W(j, MBC_UU + k) = z[k];
    if (j == 0) break;
    /* Ia = IA - Y'Y,  pa = pA + Ia c + Y' z */
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) {
        double sa = 0, sb = 0, sc = 0;
        for (int k = 0; k < nd; k++) {
          sa += Y[6 * k + a] * Y[6 * k + b];
          sb += Y[6 * k + a] * Y[6 * k + 3 + b];
          sc += Y[6 * k + 3 + a] * Y[6 * k + 3 + b];
        }
        A[3 * a + b] -= sa; B[3 * a + b] -= sb; C[3 * a + b] -= sc;
      }
    double t1[3], t2[3];
    mbc_mv(A, c, t1); mbc_mv(B, c + 3, t2);
    pA[0] += t1[0] + t2[0]; pA[1] += t1[1] + t2[1]; pA[2] += t1[2] + t2[2];
    mbc_mtv(B, c, t1); mbc_mv(C, c + 3, t2);
    pA[3] += t1[0] + t2[0]; pA[4] += t1[1] + t2[1]; pA[5] += t1[2] + t2[2];
    for (int rr = 0; rr < 6; rr++)
      for (int k = 0; k < nd; k++) pA[rr] += Y[6 * k + rr] * z[k];
    /* to the parent frame: rotate by E', then shift the reference point by r */
    mbc_rot_block(E, A, A); mbc_rot_block(E, B, B); mbc_rot_block(E, C, C);
    mbc_mtv(E, pA, pc); mbc_mtv(E, pA + 3, pc + 3);
    double Q[9], T1[9], Bt[9], Qt[9], T2[9];
    mbc_skewmul(r, C, Q);                                     /* rx C */
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) { Bt[3 * a + b] = B[3 * b + a]; Qt[3 * a + b] = Q[3 * b + a]; }
    mbc_skewmul(r, Bt, T1);                                   /* rx B' */
    mbc_skewmul(r, Qt, T2);                                   /* rx (rx C)' ; its transpose is rx C rx' */
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) {
        Ac[3 * a + b] = A[3 * a + b] + T1[3 * a + b] + T1[3 * b + a] + T2[3 * b + a];
        Bc[3 * a + b] = B[3 * a + b] + Q[3 * a + b];
        Cc[3 * a + b] = C[3 * a + b];
      }
    mbc_cross(r, pc + 3, t1);
    pc[0] += t1[0]; pc[1] += t1[1]; pc[2] += t1[2];
  }
  /* ---- pass 3: accelerations, base to tip; the base accelerates with -g (gravity) --------------------- */
  double ap[6] = {0.0, 0.0, 0.0, 0.0, MBC_G, 0.0};
  for (int j = 0; j < NJ; j++) {
    const int o = mbc_joff(R, j);
    double E[9], r[3], a[6], t[3];
    for (int k = 0; k < 9; k++) E[k] = MBC_W(j, MBC_E + k);
    for (int k = 0; k < 3; k++) r[k] = MBC_W(j, MBC_R + k);
    mbc_cross(ap, r, t);
    t[0] += ap[3]; t[1] += ap[4]; t[2] += ap[5];
    mbc_mv(E, ap, a); mbc_mv(E, t, a + 3);
    for (int k = 0; k < 6; k++) a[k] += MBC_W(j, MBC_C + k);
    const int nd = mbc_jtype(R, j) == 0 ? 3 : 1;
    double qdd[3] = {0, 0, 0};
    for (int b = 0; b < nd; b++) {
      double s_ = MBC_W(j, MBC_UU + b);
      for (int rr = 0; rr < 6; rr++) s_ -= MBC_W(j, MBC_U + 6 * b + rr) * a[rr];
      qdd[b] = s_;
    }
    if (nd == 3) {
      double L[6];
      for (int k = 0; k < 6; k++) L[k] = MBC_W(j, MBC_DI + k);
      mbc_ltsolve3(L, qdd);
    } else {
      qdd[0] *= MBC_W(j, MBC_DI);
    }
    if (nd == 3) {
      for (int k = 0; k < 3; k++) {
        a[k] += MBC_W(j, MBC_S + k) * qdd[0] + MBC_W(j, MBC_S + 3 + k) * qdd[1] + MBC_W(j, MBC_S + 6 + k) * qdd[2];
        du[o + k] = MBC_UK(o + 3 + k);
        du[o + 3 + k] = qdd[k];
      }
    } else {
      a[3] += R->dir[0] * qdd[0]; a[4] += R->dir[1] * qdd[0]; a[5] += R->dir[2] * qdd[0];
      du[o] = MBC_UK(o + 1);
      du[o + 1] = qdd[0];
    }
    for (int k = 0; k < 6; k++) ap[k] = a[k];
  }
}
#endif /* MB_CHAIN_H */
