/*
 * CPU ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference's dense bundle adjustment, included twice
 * by ba_oracle.c (REAL = float and REAL = double).  Only tests/, the smoke check
 * and bench.py's cpu_baseline / `--impl reference` legs may call it; the product
 * path (glorie-slam_b200/) never does.
 *
 * Every function cites the reference lines it follows (paths relative to
 * /root/reference).  Per-pixel arithmetic is done in REAL exactly in the
 * reference's order; reductions over pixels are accumulated in double so that the
 * oracle is a stable "truth" for both the reference (fp32 tree reductions on the
 * GPU) and the B200 kernels (warp-shuffle reductions) -- documented in DESIGN.md.
 *
 * Two semantics modes:
 *   ORACLE_MODE_DROID_CUDA (0): src/lib/droid_kernels.cu::ba_cuda
 *   ORACLE_MODE_GEOM_PY    (1): src/geom/ba.py::BA + src/geom/chol.py::schur_solve
 */

#ifndef REAL
#error "include from ba_oracle.c"
#endif

#define FN2(a, b) a##_##b
#define FN1(a, b) FN2(a, b)
#define FN(name) FN1(name, SUFFIX)

/* src/lib/droid_kernels.cu:58-68 (actSO3) / lietorch include/so3.h:55-60 */
static void FN(act_so3)(const REAL *q, const REAL *X, REAL *Y) {
  REAL uv[3];
  uv[0] = (REAL)2.0 * (q[1] * X[2] - q[2] * X[1]);
  uv[1] = (REAL)2.0 * (q[2] * X[0] - q[0] * X[2]);
  uv[2] = (REAL)2.0 * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  Y[1] = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  Y[2] = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}

/* src/lib/droid_kernels.cu:79-94 (adjSE3: transposed adjoint action);
 * lietorch include/se3.h:58-67,84-86 (AdjT) */
static void FN(adjT_se3)(const REAL *t, const REAL *q, const REAL *X, REAL *Y) {
  REAL qinv[4] = {-q[0], -q[1], -q[2], q[3]};
  REAL u[3], v[3];
  FN(act_so3)(qinv, &X[0], &Y[0]);
  FN(act_so3)(qinv, &X[3], &Y[3]);
  u[0] = t[2] * X[1] - t[1] * X[2];
  u[1] = t[0] * X[2] - t[2] * X[0];
  u[2] = t[1] * X[0] - t[0] * X[1];
  FN(act_so3)(qinv, u, v);
  Y[3] += v[0];
  Y[4] += v[1];
  Y[5] += v[2];
}

/* src/lib/droid_kernels.cu:96-107 (relSE3): G_ij = T_j * T_i^-1 */
static void FN(rel_se3)(const REAL *ti, const REAL *qi, const REAL *tj, const REAL *qj, REAL *tij, REAL *qij) {
  qij[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
  qij[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
  qij[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
  qij[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
  FN(act_so3)(qij, ti, tij);
  tij[0] = tj[0] - tij[0];
  tij[1] = tj[1] - tij[1];
  tij[2] = tj[2] - tij[2];
}

static void FN(cross3)(const REAL *a, const REAL *b, REAL *c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

/* src/lib/droid_kernels.cu:110-175,877-895 (expSO3, expSE3, retrSE3):
 * thresholds theta^2 < 1e-8 and theta > 1e-4, no quaternion renormalisation. */
static void FN(retr_droid)(const REAL *xi, const REAL *t, const REAL *q, REAL *t1, REAL *q1) {
  REAL dq[4], dt[3];
  const REAL *phi = xi + 3;
  REAL theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  REAL theta_p4 = theta_sq * theta_sq;
  REAL theta = (REAL)sqrt((double)theta_sq);
  REAL imag, real;
  if (theta_sq < (REAL)1e-8) {
    imag = (REAL)0.5 - (REAL)(1.0 / 48.0) * theta_sq + (REAL)(1.0 / 3840.0) * theta_p4;
    real = (REAL)1.0 - (REAL)(1.0 / 8.0) * theta_sq + (REAL)(1.0 / 384.0) * theta_p4;
  } else {
    imag = (REAL)sin(0.5 * (double)theta) / theta;
    real = (REAL)cos(0.5 * (double)theta);
  }
  dq[0] = imag * phi[0];
  dq[1] = imag * phi[1];
  dq[2] = imag * phi[2];
  dq[3] = real;
  dt[0] = xi[0];
  dt[1] = xi[1];
  dt[2] = xi[2];
  if (theta > (REAL)1e-4) {
    REAL tau[3] = {xi[0], xi[1], xi[2]}, c[3];
    REAL a = ((REAL)1 - (REAL)cos((double)theta)) / theta_sq;
    REAL b;
    FN(cross3)(phi, tau, c);
    dt[0] += a * c[0];
    dt[1] += a * c[1];
    dt[2] += a * c[2];
    b = (theta - (REAL)sin((double)theta)) / (theta * theta_sq);
    FN(cross3)(phi, c, tau);
    dt[0] += b * tau[0];
    dt[1] += b * tau[1];
    dt[2] += b * tau[2];
  }
  q1[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  q1[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
  q1[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
  q1[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  FN(act_so3)(dq, t, t1);
  t1[0] += dt[0];
  t1[1] += dt[1];
  t1[2] += dt[2];
}

static void FN(qnormalize)(REAL *q) {
  REAL n = (REAL)sqrt((double)(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]));
  q[0] /= n;
  q[1] /= n;
  q[2] /= n;
  q[3] /= n;
}

/* lietorch retr = Exp(a) * X: groups.py:153-156; include/se3.h:134-142 (Exp),
 * include/so3.h:153-170 (SO3::Exp, EPS = 1e-6, common.h:8), so3.h:172-190
 * (left_jacobian), so3.h:31-37 (constructor renormalises), se3.h:45-47 (mul). */
static void FN(retr_lietorch)(const REAL *xi, const REAL *t, const REAL *q_in, REAL *t1, REAL *q1) {
  const REAL *tau = xi, *phi = xi + 3;
  REAL q[4] = {q_in[0], q_in[1], q_in[2], q_in[3]};
  REAL theta2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  REAL theta = (REAL)sqrt((double)theta2);
  REAL imag, real, c1, c2, dq[4], dt[3], a[3], b[3];
  if (theta < (REAL)1e-6) {
    REAL theta4 = theta2 * theta2;
    imag = (REAL)0.5 - (REAL)(1.0 / 48.0) * theta2 + (REAL)(1.0 / 3840.0) * theta4;
    real = (REAL)1 - (REAL)(1.0 / 8.0) * theta2 + (REAL)(1.0 / 384.0) * theta4;
  } else {
    imag = (REAL)sin(0.5 * (double)theta) / theta;
    real = (REAL)cos(0.5 * (double)theta);
  }
  dq[0] = imag * phi[0];
  dq[1] = imag * phi[1];
  dq[2] = imag * phi[2];
  dq[3] = real;
  FN(qnormalize)(dq);
  c1 = (theta < (REAL)1e-6) ? (REAL)0.5 - (REAL)(1.0 / 24.0) * theta2 : ((REAL)1.0 - (REAL)cos((double)theta)) / theta2;
  c2 = (theta < (REAL)1e-6) ? (REAL)(1.0 / 6.0) - (REAL)(1.0 / 120.0) * theta2
                            : (theta - (REAL)sin((double)theta)) / (theta2 * theta);
  FN(cross3)(phi, tau, a);
  FN(cross3)(phi, a, b);
  dt[0] = tau[0] + c1 * a[0] + c2 * b[0];
  dt[1] = tau[1] + c1 * a[1] + c2 * b[1];
  dt[2] = tau[2] + c1 * a[2] + c2 * b[2];
  FN(qnormalize)(q);
  /* Eigen quaternion product dq * q, then SO3 ctor normalises */
  q1[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  q1[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
  q1[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
  q1[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  FN(qnormalize)(q1);
  FN(act_so3)(dq, t, t1);
  t1[0] += dt[0];
  t1[1] += dt[1];
  t1[2] += dt[2];
}

/* Per-edge linearisation.
 * droid_cuda: src/lib/droid_kernels.cu:176-424 (projective_transform_kernel).
 * geom_py   : src/geom/projective_ops.py:96-125 + src/geom/ba.py:40-71; differs only in the
 *             validity rule (MIN_DEPTH 0.2, Z < 0.1 -> d = 1; projective_ops.py:6,44,112).
 * Outputs per edge: Hs[4][36] (ii,ij,ji,jj), vs[2][6], Eii/Eij[6][HW], Cii/wi[HW]. */
static void FN(linearize_edge)(const REAL *poses, const REAL *disps, const REAL *intr, const float *target,
                               const float *weight, int ix, int jx, int HW, int wd, int mode, double *Hs4,
                               double *vs2, REAL *Eii, REAL *Eij, REAL *Cii, REAL *wi) {
  const REAL fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  REAL tij[3], qij[4];
  double hij[78], vi[6], vj[6];
  int k, n, m, l;
  if (ix == jx) { /* stereo pair: droid_kernels.cu:219-229, projective_ops.py:105 */
    tij[0] = (REAL)-0.1; tij[1] = 0; tij[2] = 0;
    qij[0] = 0; qij[1] = 0; qij[2] = 0; qij[3] = 1;
  } else {
    FN(rel_se3)(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
  }
  for (l = 0; l < 78; l++) hij[l] = 0;
  for (n = 0; n < 6; n++) vi[n] = vj[n] = 0;

  for (k = 0; k < HW; k++) {
    const int i = k / wd, j = k % wd;
    const REAL u = (REAL)j, v = (REAL)i;
    REAL Xi[4], Xj[4], Jx[12], *Ji = Jx, *Jj = Jx + 6, Jz;
    REAL x, y, h, d, d2, wu, wv, ru, rv, wu_pose, wv_pose;
    int invalid;
    Xi[0] = (u - cx) / fx;
    Xi[1] = (v - cy) / fy;
    Xi[2] = 1;
    Xi[3] = disps[(size_t)ix * HW + k];
    FN(act_so3)(qij, Xi, Xj); /* actSE3: droid_kernels.cu:70-77 */
    Xj[3] = Xi[3];
    Xj[0] += Xi[3] * tij[0];
    Xj[1] += Xi[3] * tij[1];
    Xj[2] += Xi[3] * tij[2];
    x = Xj[0]; y = Xj[1]; h = Xj[3];
    if (mode == 0) { /* droid_kernels.cu:26,302-306 */
      invalid = Xj[2] < (REAL)0.25;
      d = invalid ? (REAL)0.0 : (REAL)1.0 / Xj[2];
    } else { /* projective_ops.py:44,112 */
      invalid = !(Xj[2] > (REAL)0.2);
      d = (Xj[2] < (REAL)0.1) ? (REAL)1.0 : (REAL)1.0 / Xj[2];
    }
    d2 = d * d;
    wu = invalid ? (REAL)0.0 : (REAL).001 * (REAL)weight[k];
    wv = invalid ? (REAL)0.0 : (REAL).001 * (REAL)weight[HW + k];
    ru = (REAL)target[k] - (fx * d * x + cx);
    rv = (REAL)target[HW + k] - (fy * d * y + cy);

    /* u row: droid_kernels.cu:312-342 */
    Jj[0] = fx * (h * d);
    Jj[1] = fx * 0;
    Jj[2] = fx * (-x * h * d2);
    Jj[3] = fx * (-x * y * d2);
    Jj[4] = fx * (1 + x * x * d2);
    Jj[5] = fx * (-y * d);
    Jz = fx * (tij[0] * d - tij[2] * (x * d2));
    Cii[k] = wu * Jz * Jz;
    wi[k] = wu * ru * Jz;
    /* stereo self-edges drop the pose terms only in the CUDA path (droid_kernels.cu:323) */
    wu_pose = (mode == 0 && ix == jx) ? (REAL)0 : wu;
    FN(adjT_se3)(tij, qij, Jj, Ji);
    for (n = 0; n < 6; n++) Ji[n] *= -1;
    l = 0;
    for (n = 0; n < 12; n++)
      for (m = 0; m <= n; m++) hij[l++] += (double)(wu_pose * Jx[n] * Jx[m]);
    for (n = 0; n < 6; n++) {
      vi[n] += (double)(wu_pose * ru * Ji[n]);
      vj[n] += (double)(wu_pose * ru * Jj[n]);
      Eii[(size_t)n * HW + k] = wu_pose * Jz * Ji[n];
      Eij[(size_t)n * HW + k] = wu_pose * Jz * Jj[n];
    }
    /* v row: droid_kernels.cu:345-375 */
    Jj[0] = fy * 0;
    Jj[1] = fy * (h * d);
    Jj[2] = fy * (-y * h * d2);
    Jj[3] = fy * (-1 - y * y * d2);
    Jj[4] = fy * (x * y * d2);
    Jj[5] = fy * (x * d);
    Jz = fy * (tij[1] * d - tij[2] * (y * d2));
    Cii[k] += wv * Jz * Jz;
    wi[k] += wv * rv * Jz;
    wv_pose = (mode == 0 && ix == jx) ? (REAL)0 : wv;
    FN(adjT_se3)(tij, qij, Jj, Ji);
    for (n = 0; n < 6; n++) Ji[n] *= -1;
    l = 0;
    for (n = 0; n < 12; n++)
      for (m = 0; m <= n; m++) hij[l++] += (double)(wv_pose * Jx[n] * Jx[m]);
    for (n = 0; n < 6; n++) {
      vi[n] += (double)(wv_pose * rv * Ji[n]);
      vj[n] += (double)(wv_pose * rv * Jj[n]);
      Eii[(size_t)n * HW + k] += wv_pose * Jz * Ji[n];
      Eij[(size_t)n * HW + k] += wv_pose * Jz * Jj[n];
    }
  }
  /* unpack the 12x12 lower triangle into the four 6x6 blocks: droid_kernels.cu:400-423 */
  l = 0;
  for (n = 0; n < 12; n++)
    for (m = 0; m <= n; m++) {
      double s = hij[l++];
      if (n < 6 && m < 6) {
        Hs4[0 * 36 + n * 6 + m] = s;
        Hs4[0 * 36 + m * 6 + n] = s;
      } else if (n >= 6 && m < 6) {
        Hs4[1 * 36 + m * 6 + (n - 6)] = s;
        Hs4[2 * 36 + (n - 6) * 6 + m] = s;
      } else {
        Hs4[3 * 36 + (n - 6) * 6 + (m - 6)] = s;
        Hs4[3 * 36 + (m - 6) * 6 + (n - 6)] = s;
      }
    }
  for (n = 0; n < 6; n++) {
    vs2[n] = vi[n];
    vs2[6 + n] = vj[n];
  }
}

/*
 * Full BA.  droid_cuda: src/lib/droid_kernels.cu:1314-1437 (ba_cuda) with accum_cuda :948-998,
 * SparseBlock :1117-1219, schur_block :1222-1311, EvT6x1 :1095-1115, retraction :898-946.
 * geom_py: src/geom/ba.py:33-121 + src/geom/chol.py:58-85.
 *
 * State (poses [N,7], disps [N,HW]) is REAL and updated in place.  dbg_* (may be NULL)
 * receive the FIRST iteration's intermediate quantities for field-by-field parity tests.
 */
int FN(oracle_ba)(REAL *poses, REAL *disps, const float *intr_f, const float *disps_sens, const float *targets,
                  const float *weights, const float *eta, const int64_t *ii, const int64_t *jj, int N, int E, int ht,
                  int wd, int t0, int t1, int iters, float lm, float ep, int motion_only, int depth_only, int mode,
                  double *dx_out, double *dz_out, int64_t *kx_out, int *K_out, int *fail_out, double *dbg_Hs,
                  double *dbg_vs, double *dbg_C, double *dbg_w, double *dbg_M, double *dbg_rhs) {
  const int HW = ht * wd, P = t1 - t0, n6 = 6 * P;
  const REAL alpha = (REAL)0.05; /* droid_kernels.cu:1397, ba.py:35 */
  REAL intr[4] = {(REAL)intr_f[0], (REAL)intr_f[1], (REAL)intr_f[2], (REAL)intr_f[3]};
  int *fmap = (int *)malloc(sizeof(int) * (size_t)(N > 0 ? N : 1));
  int64_t *kx = (int64_t *)malloc(sizeof(int64_t) * (size_t)(N > 0 ? N : 1));
  int K = 0, f, n, itr, nfail = 0, e, *eptr, *eidx;
  double *Hs, *vs, *A, *S, *b, *vS, *dx;
  REAL *Eii = NULL, *Eij = NULL, *Cii = NULL, *wi = NULL, *C = NULL, *w = NULL, *Q = NULL, *Ei = NULL;
  double *dz = NULL;
  if (P <= 0 || N <= 0) { free(fmap); free(kx); return -1; }

  /* index prep: droid_kernels.cu:1337-1345 (_unique sorted+inverse over cat(ts, ii));
   * geom: ba.py:74 (unique(ii) only) */
  for (f = 0; f < N; f++) fmap[f] = 0;
  if (mode == 0)
    for (f = t0; f < t1 && f < N; f++) if (f >= 0) fmap[f] = 1;
  for (n = 0; n < E; n++) if (ii[n] >= 0 && ii[n] < N) fmap[ii[n]] = 1;
  for (f = 0; f < N; f++) {
    if (fmap[f]) { kx[K] = f; fmap[f] = K; K++; } else fmap[f] = -1;
  }
  if (K_out) *K_out = K;
  if (kx_out) for (f = 0; f < K; f++) kx_out[f] = kx[f];
  /* edges grouped by source frame, in edge order (what accum_cuda's argsort+CSR yields,
   * droid_kernels.cu:948-984) */
  eptr = (int *)calloc((size_t)K + 2, sizeof(int));
  eidx = (int *)malloc(sizeof(int) * (size_t)(E > 0 ? E : 1));
  for (n = 0; n < E; n++) eptr[fmap[ii[n]] + 2]++;
  for (f = 0; f < K; f++) eptr[f + 2] += eptr[f + 1];
  for (n = 0; n < E; n++) eidx[eptr[fmap[ii[n]] + 1]++] = n;
  /* now edges of frame f are eidx[eptr[f] .. eptr[f+1]) */

  Hs = (double *)calloc((size_t)(E > 0 ? E : 1) * 4 * 36, sizeof(double));
  vs = (double *)calloc((size_t)(E > 0 ? E : 1) * 12, sizeof(double));
  A = (double *)malloc(sizeof(double) * (size_t)n6 * n6);
  S = (double *)malloc(sizeof(double) * (size_t)n6 * n6);
  b = (double *)malloc(sizeof(double) * (size_t)n6);
  vS = (double *)malloc(sizeof(double) * (size_t)n6);
  dx = (double *)calloc((size_t)n6, sizeof(double));
  Eii = (REAL *)malloc(sizeof(REAL) * (size_t)(E > 0 ? E : 1) * 6 * HW);
  Eij = (REAL *)malloc(sizeof(REAL) * (size_t)(E > 0 ? E : 1) * 6 * HW);
  Cii = (REAL *)malloc(sizeof(REAL) * (size_t)(E > 0 ? E : 1) * HW);
  wi = (REAL *)malloc(sizeof(REAL) * (size_t)(E > 0 ? E : 1) * HW);
  if (!motion_only) {
    C = (REAL *)malloc(sizeof(REAL) * (size_t)K * HW);
    w = (REAL *)malloc(sizeof(REAL) * (size_t)K * HW);
    Q = (REAL *)malloc(sizeof(REAL) * (size_t)K * HW);
    Ei = (REAL *)malloc(sizeof(REAL) * (size_t)P * 6 * HW);
    dz = (double *)calloc((size_t)K * HW, sizeof(double));
  }

  for (itr = 0; itr < iters; itr++) {
    int i, j;
    /* K1: per-edge linearisation (droid_kernels.cu:1360-1373) */
#pragma omp parallel for schedule(dynamic, 1)
    for (e = 0; e < E; e++)
      FN(linearize_edge)(poses, disps, intr, targets + (size_t)e * 2 * HW, weights + (size_t)e * 2 * HW, (int)ii[e],
                         (int)jj[e], HW, wd, mode, Hs + (size_t)e * 144, vs + (size_t)e * 12,
                         Eii + (size_t)e * 6 * HW, Eij + (size_t)e * 6 * HW, Cii + (size_t)e * HW,
                         wi + (size_t)e * HW);
    if (itr == 0 && dbg_Hs) memcpy(dbg_Hs, Hs, sizeof(double) * (size_t)E * 144);
    if (itr == 0 && dbg_vs) memcpy(dbg_vs, vs, sizeof(double) * (size_t)E * 12);

    /* H1: pose-pose block system (droid_kernels.cu:1377-1384, update_lhs :1131-1156 drops
     * negative indices, setFromTriplets sums duplicates; ba.py:82-90 also drops >= P) */
    memset(A, 0, sizeof(double) * (size_t)n6 * n6);
    memset(b, 0, sizeof(double) * (size_t)n6);
    for (e = 0; e < E; e++) {
      const int pi = (int)ii[e] - t0, pj = (int)jj[e] - t0;
      const int rows[4] = {pi, pi, pj, pj}, cols[4] = {pi, pj, pi, pj};
      int blk, r, c;
      for (blk = 0; blk < 4; blk++) {
        if (rows[blk] < 0 || cols[blk] < 0 || rows[blk] >= P || cols[blk] >= P) continue;
        for (r = 0; r < 6; r++)
          for (c = 0; c < 6; c++)
            A[(size_t)(6 * rows[blk] + r) * n6 + 6 * cols[blk] + c] += Hs[(size_t)e * 144 + blk * 36 + r * 6 + c];
      }
      if (pi >= 0 && pi < P) for (r = 0; r < 6; r++) b[6 * pi + r] += vs[(size_t)e * 12 + r];
      if (pj >= 0 && pj < P) for (r = 0; r < 6; r++) b[6 * pj + r] += vs[(size_t)e * 12 + 6 + r];
    }

    memset(S, 0, sizeof(double) * (size_t)n6 * n6);
    memset(vS, 0, sizeof(double) * (size_t)n6);

    if (!motion_only) {
      /* depth diagonal: droid_kernels.cu:1397-1401 (accum_cuda = sum over edges with ii == kx[j]);
       * ba.py:92-103 */
#pragma omp parallel for schedule(static)
      for (f = 0; f < K; f++) {
        const size_t fr = (size_t)kx[f];
        int k2, q2;
        for (k2 = 0; k2 < HW; k2++) {
          double cs = 0, ws = 0;
          REAL m, c, ww;
          for (q2 = eptr[f]; q2 < eptr[f + 1]; q2++) {
            const int e2 = eidx[q2];
            cs += (double)Cii[(size_t)e2 * HW + k2];
            ws += (double)wi[(size_t)e2 * HW + k2];
          }
          m = (disps_sens[fr * HW + k2] > 0) ? (REAL)1 : (REAL)0;
          c = (REAL)cs + m * alpha + ((REAL)1 - m) * (REAL)eta[(size_t)f * HW + k2];
          ww = (REAL)ws - m * alpha * (disps[fr * HW + k2] - (REAL)disps_sens[fr * HW + k2]);
          C[(size_t)f * HW + k2] = c;
          w[(size_t)f * HW + k2] = ww;
          Q[(size_t)f * HW + k2] = (REAL)1.0 / c;
        }
      }
      if (itr == 0 && dbg_C) for (size_t z = 0; z < (size_t)K * HW; z++) dbg_C[z] = (double)C[z];
      if (itr == 0 && dbg_w) for (size_t z = 0; z < (size_t)K * HW; z++) dbg_w[z] = (double)w[z];

      /* Ei = accum(Eii, ii, ts): droid_kernels.cu:1403; ba.py:88 (scatter of Ei by ii) */
#pragma omp parallel for schedule(static)
      for (i = 0; i < P; i++) {
        size_t z;
        int q2;
        const int fi = (t0 + i >= 0 && t0 + i < N) ? fmap[t0 + i] : -1;
        for (z = 0; z < (size_t)6 * HW; z++) {
          double s = 0;
          if (fi >= 0)
            for (q2 = eptr[fi]; q2 < eptr[fi + 1]; q2++) s += (double)Eii[(size_t)eidx[q2] * 6 * HW + z];
          Ei[(size_t)i * 6 * HW + z] = (REAL)s;
        }
      }

      /* Schur complement, grouped by depth frame: droid_kernels.cu:1222-1311 (rows r = pose rows
       * ts ++ edge rows; pairs sharing kk; S += E_r Q E_s^T, v += E_r Q w); chol.py:73-75. */
#pragma omp parallel for schedule(dynamic, 1)
      for (f = 0; f < K; f++) {
        /* rows of this depth frame: (pose index, pointer to 6xHW row) */
        int cnt = 0, a2, b2, q2;
        const int deg = eptr[f + 1] - eptr[f];
        int *pidx = (int *)malloc(sizeof(int) * (size_t)(deg + 1));
        const REAL **rp = (const REAL **)malloc(sizeof(REAL *) * (size_t)(deg + 1));
        const REAL *q = Q + (size_t)f * HW, *wf = w + (size_t)f * HW;
        if (kx[f] >= t0 && kx[f] < t1) { pidx[cnt] = (int)kx[f] - t0; rp[cnt] = Ei + (size_t)(kx[f] - t0) * 6 * HW; cnt++; }
        for (q2 = eptr[f]; q2 < eptr[f + 1]; q2++) {
          const int e2 = eidx[q2];
          if (jj[e2] >= t0 && jj[e2] < t1) { pidx[cnt] = (int)jj[e2] - t0; rp[cnt] = Eij + (size_t)e2 * 6 * HW; cnt++; }
        }
        for (a2 = 0; a2 < cnt; a2++) {
          double vacc[6] = {0, 0, 0, 0, 0, 0};
          int r, c, k2;
          for (b2 = 0; b2 < cnt; b2++) {
            double blk[36];
            for (r = 0; r < 36; r++) blk[r] = 0;
            for (k2 = 0; k2 < HW; k2++) {
              REAL ea[6], eb[6];
              for (r = 0; r < 6; r++) { ea[r] = rp[a2][(size_t)r * HW + k2] * q[k2]; eb[r] = rp[b2][(size_t)r * HW + k2]; }
              for (r = 0; r < 6; r++) for (c = 0; c < 6; c++) blk[r * 6 + c] += (double)(ea[r] * eb[c]);
            }
            for (r = 0; r < 6; r++)
              for (c = 0; c < 6; c++) {
                double *dst = &S[(size_t)(6 * pidx[a2] + r) * n6 + 6 * pidx[b2] + c];
#pragma omp atomic
                *dst += blk[r * 6 + c];
              }
          }
          for (k2 = 0; k2 < HW; k2++) {
            const REAL qw = q[k2] * wf[k2]; /* Ev6x1: droid_kernels.cu:1074-1080 */
            for (r = 0; r < 6; r++) vacc[r] += (double)(qw * rp[a2][(size_t)r * HW + k2]);
          }
          for (r = 0; r < 6; r++) {
#pragma omp atomic
            vS[6 * pidx[a2] + r] += vacc[r];
          }
        }
        free(pidx);
        free((void *)rp);
      }
    }

    /* damping + solve: droid_kernels.cu:1192-1213,1407 (after Schur) vs chol.py:66-68 (before) */
    for (i = 0; i < n6; i++) {
      if (mode == 1) A[(size_t)i * n6 + i] += (double)ep + (double)lm * A[(size_t)i * n6 + i];
      for (j = 0; j < n6; j++) A[(size_t)i * n6 + j] -= S[(size_t)i * n6 + j];
      b[i] -= vS[i];
    }
    if (itr == 0 && dbg_M) memcpy(dbg_M, A, sizeof(double) * (size_t)n6 * n6); /* pre-damping in droid mode */
    if (itr == 0 && dbg_rhs) memcpy(dbg_rhs, b, sizeof(double) * (size_t)n6);
    if (mode == 0)
      for (i = 0; i < n6; i++) A[(size_t)i * n6 + i] += (double)ep + (double)lm * A[(size_t)i * n6 + i];
    {
      int ok = oracle_chol_solve(A, b, dx, n6); /* fp64 LLT; zero update on failure (:1207-1210) */
      if (!ok) { nfail++; for (i = 0; i < n6; i++) dx[i] = 0; }
    }
    /* the reference hands dx back as fp32 (droid_kernels.cu:1204-1205) */
    if (sizeof(REAL) == sizeof(float)) for (i = 0; i < n6; i++) dx[i] = (double)(float)dx[i];

    if (!motion_only) {
      /* back-substitution: EvT6x1 (droid_kernels.cu:1095-1115, skips index <= 0) + :1418;
       * chol.py:81 (no skip) */
#pragma omp parallel for schedule(static)
      for (f = 0; f < K; f++) {
        int k2, q2, r;
        for (k2 = 0; k2 < HW; k2++) {
          double acc = 0;
          if (kx[f] >= t0 && kx[f] < t1) {
            const int p = (int)kx[f] - t0;
            if (mode == 1 || p > 0) {
              REAL dw = 0;
              for (r = 0; r < 6; r++) dw += Ei[(size_t)p * 6 * HW + (size_t)r * HW + k2] * (REAL)dx[6 * p + r];
              acc += (double)dw;
            }
          }
          for (q2 = eptr[f]; q2 < eptr[f + 1]; q2++) {
            const int e2 = eidx[q2];
            {
              const int p = (int)jj[e2] - t0;
              if (p >= P || p < 0 || (mode == 0 && p == 0)) continue;
              {
                REAL dw = 0;
                for (r = 0; r < 6; r++) dw += Eij[(size_t)e2 * 6 * HW + (size_t)r * HW + k2] * (REAL)dx[6 * p + r];
                acc += (double)dw;
              }
            }
          }
          dz[(size_t)f * HW + k2] = (double)(Q[(size_t)f * HW + k2] * (w[(size_t)f * HW + k2] - (REAL)acc));
        }
      }
    }

    /* retraction: droid_kernels.cu:1386-1393,1420-1431 ; ba.py:111-119 */
    if (motion_only || !depth_only) {
      for (i = 0; i < P; i++) {
        REAL xi[6], t1v[3], q1v[4], *pp = poses + (size_t)7 * (t0 + i);
        for (j = 0; j < 6; j++) xi[j] = (REAL)dx[6 * i + j];
        if (mode == 0) FN(retr_droid)(xi, pp, pp + 3, t1v, q1v);
        else FN(retr_lietorch)(xi, pp, pp + 3, t1v, q1v);
        pp[0] = t1v[0]; pp[1] = t1v[1]; pp[2] = t1v[2];
        pp[3] = q1v[0]; pp[4] = q1v[1]; pp[5] = q1v[2]; pp[6] = q1v[3];
      }
    }
    if (!motion_only) {
      for (f = 0; f < K; f++) {
        int k2;
        for (k2 = 0; k2 < HW; k2++) {
          REAL dnew = disps[(size_t)kx[f] * HW + k2] + (REAL)dz[(size_t)f * HW + k2];
          disps[(size_t)kx[f] * HW + k2] = dnew;
        }
      }
      if (mode == 1) { /* ba.py:119 disps.clamp(min=0.0): the whole tensor, not only the rows in kx */
        size_t q;
        for (q = 0; q < (size_t)N * HW; q++)
          if (disps[q] < 0) disps[q] = 0;
      }
    }
  }

  if (dx_out) memcpy(dx_out, dx, sizeof(double) * (size_t)n6);
  if (dz_out && dz) memcpy(dz_out, dz, sizeof(double) * (size_t)K * HW);
  if (fail_out) *fail_out = nfail;
  free(eptr); free(eidx); free(fmap); free(kx); free(Hs); free(vs); free(A); free(S); free(b); free(vS); free(dx);
  free(Eii); free(Eij); free(Cii); free(wi); free(C); free(w); free(Q); free(Ei); free(dz);
  return 0;
}

/*
 * DSPO stage 2 (disparity + per-frame scale/shift, poses fixed):
 * src/geom/ba.py:127-216 (BA_with_scale_shift) + src/geom/chol.py:58-85.
 * One call = one iteration, functional on disps / wqs like the reference.
 * The dense [2M x 2M] system of the reference is block diagonal (ll,ll scatter, ba.py:191-204);
 * the oracle solves the per-frame 2x2 blocks and reproduces the reference's all-or-nothing
 * failure rule (one non-SPD block zeroes the whole update, chol.py:10-15).
 * valid_mask is one byte per frame pixel.  eta has M = |unique(ii)| rows.
 */
int FN(oracle_dspo)(const REAL *poses, REAL *disps, const float *intr_f, const float *targets, const float *weights,
                    const float *eta, const int64_t *ii, const int64_t *jj, const float *mono, REAL *scales,
                    REAL *shifts, const unsigned char *valid_mask, int N, int E, int ht, int wd, int ignore_frames,
                    float lm, float ep, float alpha_f, int interleaved, double *dz_out, double *dwq_out,
                    int *fail_out) {
  const int HW = ht * wd;
  REAL intr[4] = {(REAL)intr_f[0], (REAL)intr_f[1], (REAL)intr_f[2], (REAL)intr_f[3]};
  const REAL fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const REAL sa = (REAL)sqrtf(alpha_f); /* torch.tensor(alpha).sqrt() is float32 even in the fp64 run (ba.py:141) */
  int *fmap = (int *)malloc(sizeof(int) * (size_t)N);
  int64_t *kx = (int64_t *)malloc(sizeof(int64_t) * (size_t)N);
  int M = 0, f, n, fail = 0;
  double *Cp, *wp, *dzv, *dwq;
  for (f = 0; f < N; f++) fmap[f] = 0;
  for (n = 0; n < E; n++) fmap[ii[n]] = 1;
  for (f = 0; f < N; f++) if (fmap[f]) { kx[M] = f; fmap[f] = M; M++; } else fmap[f] = -1;
  Cp = (double *)calloc((size_t)M * HW, sizeof(double));
  wp = (double *)calloc((size_t)M * HW, sizeof(double));
  dzv = (double *)calloc((size_t)M * HW, sizeof(double));
  dwq = (double *)calloc((size_t)M * 2, sizeof(double));

  /* reprojection part: only J_z is used (ba.py:151-152,181-186) */
  for (n = 0; n < E; n++) {
    const int ix = (int)ii[n], jx = (int)jj[n], kf = fmap[ix];
    REAL tij[3], qij[4];
    int k;
    if (ix == jx) { tij[0] = (REAL)-0.1; tij[1] = 0; tij[2] = 0; qij[0] = qij[1] = qij[2] = 0; qij[3] = 1; }
    else FN(rel_se3)(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
    for (k = 0; k < HW; k++) {
      const int i = k / wd, j = k % wd;
      REAL Xi[4], Xj[3], x, y, d, d2, wu, wv, ru, rv, Jzu, Jzv, tu, tv, wgu, wgv;
      int invalid;
      Xi[0] = ((REAL)j - cx) / fx; Xi[1] = ((REAL)i - cy) / fy; Xi[2] = 1; Xi[3] = disps[(size_t)ix * HW + k];
      FN(act_so3)(qij, Xi, Xj);
      Xj[0] += Xi[3] * tij[0]; Xj[1] += Xi[3] * tij[1]; Xj[2] += Xi[3] * tij[2];
      x = Xj[0]; y = Xj[1];
      invalid = !(Xj[2] > (REAL)0.2);
      d = (Xj[2] < (REAL)0.1) ? (REAL)1.0 : (REAL)1.0 / Xj[2];
      d2 = d * d;
      if (interleaved) { /* [E,ht,wd,2] as BA_with_scale_shift receives it */
        tu = (REAL)targets[((size_t)n * HW + k) * 2]; tv = (REAL)targets[((size_t)n * HW + k) * 2 + 1];
        wgu = (REAL)weights[((size_t)n * HW + k) * 2]; wgv = (REAL)weights[((size_t)n * HW + k) * 2 + 1];
      } else {
        tu = (REAL)targets[(size_t)n * 2 * HW + k]; tv = (REAL)targets[(size_t)n * 2 * HW + HW + k];
        wgu = (REAL)weights[(size_t)n * 2 * HW + k]; wgv = (REAL)weights[(size_t)n * 2 * HW + HW + k];
      }
      wu = invalid ? (REAL)0 : (REAL).001 * wgu;
      wv = invalid ? (REAL)0 : (REAL).001 * wgv;
      ru = tu - (fx * d * x + cx);
      rv = tv - (fy * d * y + cy);
      Jzu = fx * (tij[0] * d - tij[2] * (x * d2));
      Jzv = fy * (tij[1] * d - tij[2] * (y * d2));
      Cp[(size_t)kf * HW + k] += (double)(wu * Jzu * Jzu + wv * Jzv * Jzv);   /* ba.py:186 */
      wp[(size_t)kf * HW + k] += (double)(-(wu * ru * Jzu) - (wv * rv * Jzv)); /* ba.py:185 */
    }
  }
  /* prior part + per-frame 2x2 Schur system (ba.py:154-179,197-206; chol.py:58-85) */
  for (f = 0; f < M; f++) {
    const size_t fr = (size_t)kx[f];
    const REAL s = scales[fr], q = shifts[fr];
    const int ignore = kx[f] < ignore_frames;
    double H00 = 0, H01 = 0, H11 = 0, u0 = 0, u1 = 0, S00 = 0, S01 = 0, S11 = 0, g0 = 0, g1 = 0;
    double a00, a01, a11, r0, r1, l00, l10, l11, y0, y1, dws, dwq1;
    int k, bad = 0;
    for (k = 0; k < HW; k++) {
      const REAL mo = (REAL)mono[fr * HW + k];
      const int inval = (mo < (REAL)1e-6) || ignore;
      const int vdm = valid_mask[fr * HW + k] != 0;
      const REAL rdep = sa * (disps[fr * HW + k] - (s * mo + q)); /* un-boosted sqrt(alpha): ba.py:155 */
      const REAL sap = vdm ? sa * (REAL)10 : sa;                   /* ba.py:159-160 */
      REAL Jd = sap, Js = -mo * sap, Jq = -sap, Cc, ww, Qc;
      if (inval && vdm) Jd = 0; /* ba.py:166 */
      if (inval) { Js = 0; Jq = 0; }
      H00 += (double)(Js * Js); H01 += (double)(Js * Jq); H11 += (double)(Jq * Jq);
      u0 -= (double)(Js * rdep); u1 -= (double)(Jq * rdep);
      Cc = (REAL)Cp[(size_t)f * HW + k] + Jd * Jd + (REAL)eta[(size_t)f * HW + k]; /* ba.py:199 */
      ww = -(REAL)wp[(size_t)f * HW + k] - Jd * rdep;                                /* ba.py:202 */
      Qc = (REAL)1.0 / Cc;
      S00 += (double)((Js * Jd) * Qc * (Js * Jd));
      S01 += (double)((Js * Jd) * Qc * (Jq * Jd));
      S11 += (double)((Jq * Jd) * Qc * (Jq * Jd));
      g0 += (double)((Js * Jd) * Qc * ww);
      g1 += (double)((Jq * Jd) * Qc * ww);
    }
    a00 = H00 + ((double)ep + (double)lm * H00) - S00; /* damping on H before Schur: chol.py:66-68 */
    a11 = H11 + ((double)ep + (double)lm * H11) - S11;
    a01 = H01 - S01;
    r0 = u0 - g0;
    r1 = u1 - g1;
    if (!(a00 > 0)) bad = 1;
    l00 = bad ? 1 : sqrt(a00);
    l10 = a01 / l00;
    if (!(a11 - l10 * l10 > 0)) bad = 1;
    l11 = bad ? 1 : sqrt(a11 - l10 * l10);
    if (bad) { fail = 1; continue; }
    y0 = r0 / l00; y1 = (r1 - l10 * y0) / l11;
    dwq1 = y1 / l11; dws = (y0 - l10 * dwq1) / l00;
    if (sizeof(REAL) == sizeof(float)) { dws = (double)(float)dws; dwq1 = (double)(float)dwq1; }
    dwq[2 * f] = dws; dwq[2 * f + 1] = dwq1;
    for (k = 0; k < HW; k++) {
      const REAL mo = (REAL)mono[fr * HW + k];
      const int inval = (mo < (REAL)1e-6) || ignore;
      const int vdm = valid_mask[fr * HW + k] != 0;
      const REAL rdep = sa * (disps[fr * HW + k] - (s * mo + q));
      const REAL sap = vdm ? sa * (REAL)10 : sa;
      REAL Jd = sap, Js = -mo * sap, Jq = -sap, Cc, ww;
      if (inval && vdm) Jd = 0;
      if (inval) { Js = 0; Jq = 0; }
      Cc = (REAL)Cp[(size_t)f * HW + k] + Jd * Jd + (REAL)eta[(size_t)f * HW + k];
      ww = -(REAL)wp[(size_t)f * HW + k] - Jd * rdep;
      dzv[(size_t)f * HW + k] = (double)(((REAL)1.0 / Cc) * (ww - ((Js * Jd) * (REAL)dws + (Jq * Jd) * (REAL)dwq1)));
    }
  }
  if (fail) { /* CholeskySolver exception -> zeros for dwq; dz = Q*(w - 0): chol.py:13-15,81 */
    for (f = 0; f < M; f++) {
      const size_t fr = (size_t)kx[f];
      const REAL s = scales[fr], q = shifts[fr];
      const int ignore = kx[f] < ignore_frames;
      int k;
      dwq[2 * f] = dwq[2 * f + 1] = 0;
      for (k = 0; k < HW; k++) {
        const REAL mo = (REAL)mono[fr * HW + k];
        const int inval = (mo < (REAL)1e-6) || ignore;
        const int vdm = valid_mask[fr * HW + k] != 0;
        const REAL rdep = sa * (disps[fr * HW + k] - (s * mo + q));
        const REAL sap = vdm ? sa * (REAL)10 : sa;
        REAL Jd = sap, Cc, ww;
        if (inval && vdm) Jd = 0;
        Cc = (REAL)Cp[(size_t)f * HW + k] + Jd * Jd + (REAL)eta[(size_t)f * HW + k];
        ww = -(REAL)wp[(size_t)f * HW + k] - Jd * rdep;
        dzv[(size_t)f * HW + k] = (double)(((REAL)1.0 / Cc) * ww);
      }
    }
  }
  /* retraction: ba.py:211-214 */
  for (f = 0; f < M; f++) {
    const size_t fr = (size_t)kx[f];
    int k;
    for (k = 0; k < HW; k++) {
      disps[fr * HW + k] = disps[fr * HW + k] + (REAL)dzv[(size_t)f * HW + k];
    }
    scales[fr] += (REAL)dwq[2 * f];
    shifts[fr] += (REAL)dwq[2 * f + 1];
  }
  { /* ba.py:214 disps.clamp(min=0.0): the whole tensor, not only the rows in kx */
    size_t q;
    for (q = 0; q < (size_t)N * HW; q++)
      if (disps[q] < 0) disps[q] = 0;
  }
  if (dz_out) memcpy(dz_out, dzv, sizeof(double) * (size_t)M * HW);
  if (dwq_out) memcpy(dwq_out, dwq, sizeof(double) * (size_t)M * 2);
  if (fail_out) *fail_out = fail;
  free(fmap); free(kx); free(Cp); free(wp); free(dzv); free(dwq);
  return M;
}

/*
 * frame_distance: mean induced-flow magnitude between keyframe pairs, beta-mix of the full
 * relative motion and its translation-only part; 1000 when fewer than 75 % of the pixels land in
 * front of the target camera.  src/lib/droid_kernels.cu:518-657 (frame_distance_kernel),
 * called from src/depth_video.py:166-196 and src/factor_graph.py:323-462.
 */
int FN(oracle_frame_distance)(const REAL *poses, const REAL *disps, const float *intr_f, const int64_t *ii,
                              const int64_t *jj, int n_pairs, int ht, int wd, float beta_f, REAL *dist) {
  const int HW = ht * wd;
  const REAL fx = (REAL)intr_f[0], fy = (REAL)intr_f[1], cx = (REAL)intr_f[2], cy = (REAL)intr_f[3];
  const REAL beta = (REAL)beta_f;
  int n;
#pragma omp parallel for schedule(dynamic, 16)
  for (n = 0; n < n_pairs; n++) {
    const int ix = (int)ii[n], jx = (int)jj[n];
    REAL tij[3], qij[4];
    double accum = 0, valid = 0, total = 0;
    int k;
    FN(rel_se3)(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
    for (k = 0; k < HW; k++) {
      const REAL u = (REAL)(k % wd), v = (REAL)(k / wd);
      REAL Xi[4], Xj[3], du, dv, d;
      Xi[0] = (u - cx) / fx; Xi[1] = (v - cy) / fy; Xi[2] = 1; Xi[3] = disps[(size_t)ix * HW + k];
      FN(act_so3)(qij, Xi, Xj);  /* actSE3 */
      Xj[0] += Xi[3] * tij[0]; Xj[1] += Xi[3] * tij[1]; Xj[2] += Xi[3] * tij[2];
      du = fx * (Xj[0] / Xj[2]) + cx - u;
      dv = fy * (Xj[1] / Xj[2]) + cy - v;
      d = (REAL)sqrt((double)(du * du + dv * dv));
      total += (double)beta;
      if (Xj[2] > (REAL)0.25) { accum += (double)(beta * d); valid += (double)beta; }
      /* translation only (droid_kernels.cu:611-628) */
      Xj[0] = Xi[0] + Xi[3] * tij[0]; Xj[1] = Xi[1] + Xi[3] * tij[1]; Xj[2] = Xi[2] + Xi[3] * tij[2];
      du = fx * (Xj[0] / Xj[2]) + cx - u;
      dv = fy * (Xj[1] / Xj[2]) + cy - v;
      d = (REAL)sqrt((double)(du * du + dv * dv));
      total += (double)((REAL)1 - beta);
      if (Xj[2] > (REAL)0.25) { accum += (double)(((REAL)1 - beta) * d); valid += (double)((REAL)1 - beta); }
    }
    dist[n] = (valid / (total + 1e-8) < 0.75) ? (REAL)1000.0 : (REAL)(accum / valid);
  }
  return 0;
}

/*
 * depth_filter: for every listed keyframe and pixel, the number of the 6 neighbouring keyframes
 * (ix-1, ix-2, ix-3, ix+3, ix+4, ix+5) whose disparity map agrees with the reprojected depth at one
 * of the 4 pixels around the reprojection.  src/lib/droid_kernels.cu:661-775,1494-1518; caller
 * src/depth_video.py:326-361 (update_valid_depth_mask).  The comparisons are done in double like
 * the reference (its `1.0/dj` literals promote to double).
 */
int FN(oracle_depth_filter)(const REAL *poses, const REAL *disps, const float *intr_f, const int64_t *inds,
                            const float *thresh, int n_frames, int n_inds, int ht, int wd, float *counter) {
  const int HW = ht * wd;
  const REAL fx = (REAL)intr_f[0], fy = (REAL)intr_f[1], cx = (REAL)intr_f[2], cy = (REAL)intr_f[3];
  int b;
#pragma omp parallel for schedule(dynamic, 1)
  for (b = 0; b < n_inds; b++) {
    const int ix = (int)inds[b];
    const double t = (double)thresh[b];
    int nb, k;
    for (k = 0; k < HW; k++) counter[(size_t)b * HW + k] = 0.f;
    for (nb = 0; nb < 6; nb++) {
      const int jx = (nb < 3) ? ix - nb - 1 : ix + nb; /* droid_kernels.cu:696 */
      REAL tij[3], qij[4];
      if (jx < 0 || jx >= n_frames) continue;
      FN(rel_se3)(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
      for (k = 0; k < HW; k++) {
        const int i = k / wd, j = k % wd;
        REAL Xi[4], Xj[3], uj, vj, dj;
        int u0, v0;
        Xi[0] = ((REAL)j - cx) / fx; Xi[1] = ((REAL)i - cy) / fy; Xi[2] = 1; Xi[3] = disps[(size_t)ix * HW + k];
        FN(act_so3)(qij, Xi, Xj);
        Xj[0] += Xi[3] * tij[0]; Xj[1] += Xi[3] * tij[1]; Xj[2] += Xi[3] * tij[2];
        uj = fx * (Xj[0] / Xj[2]) + cx;
        vj = fy * (Xj[1] / Xj[2]) + cy;
        dj = Xi[3] / Xj[2];
        if (!(uj > (REAL)-1e8 && uj < (REAL)1e8 && vj > (REAL)-1e8 && vj < (REAL)1e8)) continue; /* NaN / overflow: never inside */
        u0 = (int)floor((double)uj);
        v0 = (int)floor((double)vj);
        if (u0 >= 0 && v0 >= 0 && u0 < wd - 1 && v0 < ht - 1) {
          const REAL *dp = disps + (size_t)jx * HW;
          const double inv = 1.0 / (double)dj;
          const double d00 = dp[v0 * wd + u0], d01 = dp[v0 * wd + u0 + 1], d10 = dp[(v0 + 1) * wd + u0],
                       d11 = dp[(v0 + 1) * wd + u0 + 1];
          if (fabs(inv - 1.0 / d00) < t || fabs(inv - 1.0 / d01) < t || fabs(inv - 1.0 / d10) < t ||
              fabs(inv - 1.0 / d11) < t)
            counter[(size_t)b * HW + k] += 1.0f;
        }
      }
    }
  }
  return 0;
}

/*
 * align_scale_and_shift: per keyframe, weighted least squares  min sum w (s p + q - t)^2  and the
 * weighted mean absolute error of the fit.  src/utils/common.py:402-438 (float32 torch sums there).
 */
int FN(oracle_align_scale_shift)(const REAL *pred, const REAL *target, const REAL *weights, int n, int HW, REAL *scale,
                                 REAL *shift, REAL *err) {
  int b;
  for (b = 0; b < n; b++) {
    const REAL *p = pred + (size_t)b * HW, *t = target + (size_t)b * HW, *w = weights + (size_t)b * HW;
    double a00 = 0, a01 = 0, a11 = 0, b0 = 0, b1 = 0, es = 0, det, s, q;
    int k;
    for (k = 0; k < HW; k++) {
      a00 += (double)(w[k] * p[k] * p[k]); a01 += (double)(w[k] * p[k]); a11 += (double)w[k];
      b0 += (double)(w[k] * p[k] * t[k]);  b1 += (double)(w[k] * t[k]);
    }
    det = a00 * a11 - a01 * a01;
    s = (a11 * b0 - a01 * b1) / det;
    q = (-a01 * b0 + a00 * b1) / det;
    for (k = 0; k < HW; k++) es += fabs((double)((REAL)s * p[k] + (REAL)q - t[k])) * (double)w[k];
    scale[b] = (REAL)s; shift[b] = (REAL)q; err[b] = (REAL)(es / a11);
  }
  return 0;
}

/*
 * iproj: back-projection of every pixel of every keyframe, transformed by the given pose:
 * points = (R (x, y, 1) + t d) / d.  src/lib/droid_kernels.cu:779-850,1520-1545; caller
 * src/neural_point.py:156 (passes the INVERSE poses to get world points).
 */
int FN(oracle_iproj)(const REAL *poses, const REAL *disps, const float *intr_f, int n_frames, int ht, int wd, REAL *points) {
  const int HW = ht * wd;
  const REAL fx = (REAL)intr_f[0], fy = (REAL)intr_f[1], cx = (REAL)intr_f[2], cy = (REAL)intr_f[3];
  int b;
#pragma omp parallel for schedule(static)
  for (b = 0; b < n_frames; b++) {
    int k;
    for (k = 0; k < HW; k++) {
      REAL Xi[4], Xj[3];
      Xi[0] = ((REAL)(k % wd) - cx) / fx; Xi[1] = ((REAL)(k / wd) - cy) / fy; Xi[2] = 1; Xi[3] = disps[(size_t)b * HW + k];
      FN(act_so3)(poses + 7 * b + 3, Xi, Xj);
      points[((size_t)b * HW + k) * 3 + 0] = (Xj[0] + Xi[3] * poses[7 * b + 0]) / Xi[3];
      points[((size_t)b * HW + k) * 3 + 1] = (Xj[1] + Xi[3] * poses[7 * b + 1]) / Xi[3];
      points[((size_t)b * HW + k) * 3 + 2] = (Xj[2] + Xi[3] * poses[7 * b + 2]) / Xi[3];
    }
  }
  return 0;
}

/*
 * projmap: where the pixels of keyframe ii[n] land in keyframe jj[n].  coords [n,ht,wd,3]: (u', v', 0),
 * the pixel's own (u, v) when Z <= 0.01; valid [n,ht,wd,1] = Z > 0.25.
 * src/lib/droid_kernels.cu:427-516,1466-1492.
 */
int FN(oracle_projmap)(const REAL *poses, const REAL *disps, const float *intr_f, const int64_t *ii, const int64_t *jj,
                       int n_pairs, int ht, int wd, REAL *coords, REAL *valid) {
  const int HW = ht * wd;
  const REAL fx = (REAL)intr_f[0], fy = (REAL)intr_f[1], cx = (REAL)intr_f[2], cy = (REAL)intr_f[3];
  int n;
#pragma omp parallel for schedule(static)
  for (n = 0; n < n_pairs; n++) {
    const int ix = (int)ii[n], jx = (int)jj[n];
    REAL tij[3], qij[4];
    int k;
    FN(rel_se3)(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
    for (k = 0; k < HW; k++) {
      REAL Xi[4], Xj[3], u = (REAL)(k % wd), v = (REAL)(k / wd), Z;
      REAL *c = coords + ((size_t)n * HW + k) * 3;
      Xi[0] = (u - cx) / fx; Xi[1] = (v - cy) / fy; Xi[2] = 1; Xi[3] = disps[(size_t)ix * HW + k];
      FN(act_so3)(qij, Xi, Xj);
      Xj[0] += Xi[3] * tij[0]; Xj[1] += Xi[3] * tij[1]; Xj[2] += Xi[3] * tij[2];
      Z = Xj[2];
      c[0] = u; c[1] = v; c[2] = 0;
      if (Z > (REAL)0.01) {
        c[0] = fx * (Xj[0] / Z) + cx;
        c[1] = fy * (Xj[1] / Z) + cy;
      }
      valid[(size_t)n * HW + k] = (Z > (REAL)0.25) ? 1 : 0;
    }
  }
  return 0;
}

#undef FN
#undef FN1
#undef FN2
