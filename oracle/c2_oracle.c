/*
 * TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, single-threaded restatement of the reference's semiseparable recursions.  It exists to check
 * the CUDA path (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline leg) and must never be linked
 * into, imported by or executed from the product path.
 *
 * Every function follows the loop structure and operation order of the reference function it cites
 * (paths relative to /root/reference).  Parity is PINNED: tests/test_oracle.py checks every function
 * here against the reference's own C++ core compiled into oracle/_ref (in the build container) and
 * against the committed fixtures under tests/golden/ that were generated from that build.
 *
 * Layout conventions (all FP64, C-contiguous):
 *   t (N)   c (J)   a,d (N)   U,V,W (N,J) row-major   Y,Z (N,K) row-major
 *   S (N,J*J): row n = column-major flattening of the J x J matrix (index j + J*i for element (j,i))
 *              c++/include/celerite2/forward.hpp:97-103,116-120
 *   F (N,J*K): row n = column-major flattening of the J x K matrix (index j + J*k)
 *              c++/include/celerite2/internal.hpp:127-129 ("Inner" is a default, i.e. column-major, Eigen matrix)
 *   general_matmul F (M,J*K): row m = ROW-major flattening (index j*K + k), forward.hpp:299 ("Inner" is RowMajor there)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define C2O_MAXJ 64

/* c++/include/celerite2/forward.hpp:69-135 (S != NULL) and interface.hpp:37-48 (S == NULL).
 * Returns 0 or the first n >= 1 with d[n] <= 0; rows after that are left untouched.
 * d may alias a, W may alias V. */
int64_t c2o_factor(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *d,
                   double *W, double *S) {
  double *Sn  = (double *)calloc((size_t)(J * J), sizeof(double)); /* column-major J x J */
  double *tmp = (double *)malloc(sizeof(double) * (size_t)J);
  double *p   = (double *)malloc(sizeof(double) * (size_t)J);
  int64_t flag = 0;
  if (S) memset(S, 0, sizeof(double) * (size_t)(J * J)); /* S.row(0).setZero(), forward.hpp:92 */

  d[0] = a[0];                                           /* forward.hpp:107 */
  for (int64_t j = 0; j < J; ++j) W[j] = V[j] / d[0];    /* forward.hpp:108 */

  for (int64_t n = 1; n < N; ++n) {
    const double dt = t[n - 1] - t[n];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt); /* forward.hpp:112 */
    /* Sn += d(n-1) * W(n-1)^T W(n-1), forward.hpp:115 */
    for (int64_t i = 0; i < J; ++i)
      for (int64_t j = 0; j < J; ++j) Sn[j + J * i] += (d[n - 1] * W[(n - 1) * J + j]) * W[(n - 1) * J + i];
    /* Sn = diag(p) * Sn, forward.hpp:116 */
    for (int64_t i = 0; i < J; ++i)
      for (int64_t j = 0; j < J; ++j) Sn[j + J * i] = p[j] * Sn[j + J * i];
    if (S) memcpy(S + n * J * J, Sn, sizeof(double) * (size_t)(J * J)); /* forward.hpp:120 */
    /* Sn *= diag(p), forward.hpp:123 */
    for (int64_t i = 0; i < J; ++i)
      for (int64_t j = 0; j < J; ++j) Sn[j + J * i] *= p[i];
    /* tmp = U(n) * Sn; d(n) = a(n) - tmp * U(n)^T, forward.hpp:126-127 */
    double acc = 0.0;
    for (int64_t i = 0; i < J; ++i) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) s += U[n * J + j] * Sn[j + J * i];
      tmp[i] = s;
      acc += s * U[n * J + i];
    }
    const double dn = a[n] - acc;
    d[n]            = dn;
    if (dn <= 0.0) { /* forward.hpp:128 */
      flag = n;
      break;
    }
    for (int64_t j = 0; j < J; ++j) W[n * J + j] = (V[n * J + j] - tmp[j]) / dn; /* forward.hpp:131 */
  }
  free(Sn);
  free(tmp);
  free(p);
  return flag;
}

/* internal::forward<is_solve, do_update>, c++/include/celerite2/internal.hpp:105-146, wrapped as
 * solve_lower (forward.hpp:156-170: Z = Y first) and matmul_lower (forward.hpp:228-239: Z accumulated).
 * Z may alias Y.  F may be NULL (no workspace). */
static void c2o_forward(int is_solve, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *V,
                        const double *Y, double *Z, double *F) {
  double *Fn  = (double *)calloc((size_t)(J * K), sizeof(double)); /* column-major J x K */
  double *tmp = (double *)malloc(sizeof(double) * (size_t)K);
  double *p   = (double *)malloc(sizeof(double) * (size_t)J);
  if (F) memset(F, 0, sizeof(double) * (size_t)(J * K)); /* internal.hpp:125 */
  memcpy(tmp, Y, sizeof(double) * (size_t)K);            /* tmp = Y.row(0), internal.hpp:134 */
  for (int64_t n = 1; n < N; ++n) {
    const double dt = t[n - 1] - t[n];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
    /* update_f: Fn += V(n-1)^T * (is_solve ? Z(n-1) : tmp), internal.hpp:45-85,140 */
    for (int64_t k = 0; k < K; ++k) {
      const double r = is_solve ? Z[(n - 1) * K + k] : tmp[k];
      for (int64_t j = 0; j < J; ++j) Fn[j + J * k] += V[(n - 1) * J + j] * r;
    }
    memcpy(tmp, Y + n * K, sizeof(double) * (size_t)K);                  /* internal.hpp:141 */
    if (F) memcpy(F + n * J * K, Fn, sizeof(double) * (size_t)(J * K)); /* internal.hpp:142 */
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) Fn[j + J * k] = p[j] * Fn[j + J * k]; /* internal.hpp:143 */
    /* update_z: Z(n) -/+= U(n) * Fn, internal.hpp:87-103,144 */
    for (int64_t k = 0; k < K; ++k) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) s += U[n * J + j] * Fn[j + J * k];
      if (is_solve)
        Z[n * K + k] -= s;
      else
        Z[n * K + k] += s;
    }
  }
  free(Fn);
  free(tmp);
  free(p);
}

/* internal::backward<is_solve, do_update>, internal.hpp:148-189 (solve_upper forward.hpp:193-207,
 * matmul_upper forward.hpp:260-271). */
static void c2o_backward(int is_solve, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *V,
                         const double *Y, double *Z, double *F) {
  double *Fn  = (double *)calloc((size_t)(J * K), sizeof(double));
  double *tmp = (double *)malloc(sizeof(double) * (size_t)K);
  double *p   = (double *)malloc(sizeof(double) * (size_t)J);
  if (F) memset(F + (N - 1) * J * K, 0, sizeof(double) * (size_t)(J * K)); /* internal.hpp:168 */
  memcpy(tmp, Y + (N - 1) * K, sizeof(double) * (size_t)K);                /* internal.hpp:177 */
  for (int64_t n = N - 2; n >= 0; --n) {
    const double dt = t[n] - t[n + 1];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
    for (int64_t k = 0; k < K; ++k) {
      const double r = is_solve ? Z[(n + 1) * K + k] : tmp[k];
      for (int64_t j = 0; j < J; ++j) Fn[j + J * k] += U[(n + 1) * J + j] * r; /* internal.hpp:183 */
    }
    memcpy(tmp, Y + n * K, sizeof(double) * (size_t)K);
    if (F) memcpy(F + n * J * K, Fn, sizeof(double) * (size_t)(J * K));
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) Fn[j + J * k] = p[j] * Fn[j + J * k];
    for (int64_t k = 0; k < K; ++k) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) s += V[n * J + j] * Fn[j + J * k]; /* internal.hpp:187 */
      if (is_solve)
        Z[n * K + k] -= s;
      else
        Z[n * K + k] += s;
    }
  }
  free(Fn);
  free(tmp);
  free(p);
}

/* op: 0 solve_lower, 1 solve_upper, 2 matmul_lower, 3 matmul_upper.  For the solves Z is overwritten
 * (Z = Y, forward.hpp:168,205); for the matmuls the caller's Z is accumulated into (forward.hpp:238,270). */
void c2o_sweep(int32_t op, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *W,
               const double *Y, double *Z, double *F) {
  const int is_solve = (op == 0 || op == 1);
  if (is_solve && Z != Y) memcpy(Z, Y, sizeof(double) * (size_t)(N * K));
  if (op == 0 || op == 2)
    c2o_forward(is_solve, N, J, K, t, c, U, W, Y, Z, F);
  else
    c2o_backward(is_solve, N, J, K, t, c, U, W, Y, Z, F);
}

/* internal::forward_rev<is_solve>, internal.hpp:191-246.  bZ is the running cotangent (aliases bY for the
 * solves, reverse.hpp:116-117); outputs must be zeroed by the caller (done in c2o_sweep_rev). */
static void c2o_forward_rev(int is_solve, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *V,
                            const double *Y, const double *Z, const double *F, double *bZ, double *bt, double *bc, double *bU, double *bV,
                            double *bY) {
  double *bF = (double *)calloc((size_t)(J * K), sizeof(double));
  double *p  = (double *)malloc(sizeof(double) * (size_t)J);
  double *bp = (double *)malloc(sizeof(double) * (size_t)J);
  for (int64_t n = N - 1; n >= 1; --n) {
    const double dt = t[n - 1] - t[n];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
    const double *Fn = F + n * J * K;
    /* bU(n) -/+= bZ(n) * (P Fn)^T ; bF -/+= U(n)^T bZ(n), internal.hpp:232-233 */
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += bZ[n * K + k] * (p[j] * Fn[j + J * k]);
      if (is_solve)
        bU[n * J + j] -= s;
      else
        bU[n * J + j] += s;
    }
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) {
        const double s = U[n * J + j] * bZ[n * K + k];
        if (is_solve)
          bF[j + J * k] -= s;
        else
          bF[j + J * k] += s;
      }
    /* bp = diag(Fn bF^T) * p, internal.hpp:236-241 */
    double factor = 0.0;
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += Fn[j + J * k] * bF[j + J * k];
      bp[j] = s * p[j];
      bc[j] += dt * bp[j];
      factor += c[j] * bp[j];
    }
    bt[n] -= factor;
    bt[n - 1] += factor;
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) bF[j + J * k] = p[j] * bF[j + J * k];
    /* update_f<is_solve>::reverse, internal.hpp:55-84,244 */
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += (is_solve ? Z[(n - 1) * K + k] : Y[(n - 1) * K + k]) * bF[j + J * k];
      bV[(n - 1) * J + j] += s;
    }
    for (int64_t k = 0; k < K; ++k) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) s += V[(n - 1) * J + j] * bF[j + J * k];
      if (is_solve)
        bZ[(n - 1) * K + k] += s;
      else
        bY[(n - 1) * K + k] += s;
    }
  }
  free(bF);
  free(p);
  free(bp);
}

/* internal::backward_rev<is_solve>, internal.hpp:248-303 */
static void c2o_backward_rev(int is_solve, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U,
                             const double *V, const double *Y, const double *Z, const double *F, double *bZ, double *bt, double *bc,
                             double *bU, double *bV, double *bY) {
  double *bF = (double *)calloc((size_t)(J * K), sizeof(double));
  double *p  = (double *)malloc(sizeof(double) * (size_t)J);
  double *bp = (double *)malloc(sizeof(double) * (size_t)J);
  for (int64_t n = 0; n <= N - 2; ++n) {
    const double dt = t[n] - t[n + 1];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
    const double *Fn = F + n * J * K;
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += bZ[n * K + k] * (p[j] * Fn[j + J * k]);
      if (is_solve)
        bV[n * J + j] -= s;
      else
        bV[n * J + j] += s;
    }
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) {
        const double s = V[n * J + j] * bZ[n * K + k];
        if (is_solve)
          bF[j + J * k] -= s;
        else
          bF[j + J * k] += s;
      }
    double factor = 0.0;
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += Fn[j + J * k] * bF[j + J * k];
      bp[j] = s * p[j];
      bc[j] += dt * bp[j];
      factor += c[j] * bp[j];
    }
    bt[n + 1] -= factor;
    bt[n] += factor;
    for (int64_t k = 0; k < K; ++k)
      for (int64_t j = 0; j < J; ++j) bF[j + J * k] = p[j] * bF[j + J * k];
    for (int64_t j = 0; j < J; ++j) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += (is_solve ? Z[(n + 1) * K + k] : Y[(n + 1) * K + k]) * bF[j + J * k];
      bU[(n + 1) * J + j] += s;
    }
    for (int64_t k = 0; k < K; ++k) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) s += U[(n + 1) * J + j] * bF[j + J * k];
      if (is_solve)
        bZ[(n + 1) * K + k] += s;
      else
        bY[(n + 1) * K + k] += s;
    }
  }
  free(bF);
  free(p);
  free(bp);
}

/* solve_lower_rev / solve_upper_rev / matmul_lower_rev / matmul_upper_rev, reverse.hpp:87-217.
 * op numbering as in c2o_sweep.  All outputs are overwritten. */
void c2o_sweep_rev(int32_t op, int64_t N, int64_t J, int64_t K, const double *t, const double *c, const double *U, const double *W,
                   const double *Y, const double *Z, const double *F, const double *bZ, double *bt, double *bc, double *bU, double *bW,
                   double *bY) {
  const int is_solve = (op == 0 || op == 1);
  memset(bt, 0, sizeof(double) * (size_t)N);
  memset(bc, 0, sizeof(double) * (size_t)J);
  memset(bU, 0, sizeof(double) * (size_t)(N * J));
  memset(bW, 0, sizeof(double) * (size_t)(N * J));
  if (is_solve) {
    memcpy(bY, bZ, sizeof(double) * (size_t)(N * K)); /* bY = bZ, then bY is the running bZ: reverse.hpp:116-117 */
    if (op == 0)
      c2o_forward_rev(1, N, J, K, t, c, U, W, Y, Z, F, bY, bt, bc, bU, bW, bY);
    else
      c2o_backward_rev(1, N, J, K, t, c, U, W, Y, Z, F, bY, bt, bc, bU, bW, bY);
  } else {
    memset(bY, 0, sizeof(double) * (size_t)(N * K));
    /* bZ is a read-only input here (reverse.hpp:183,216); the cast only satisfies the shared signature */
    if (op == 2)
      c2o_forward_rev(0, N, J, K, t, c, U, W, Y, Z, F, (double *)bZ, bt, bc, bU, bW, bY);
    else
      c2o_backward_rev(0, N, J, K, t, c, U, W, Y, Z, F, (double *)bZ, bt, bc, bU, bW, bY);
  }
}

/* core::factor_rev, reverse.hpp:10-85.  a and V are unused by the reference (reverse.hpp:29-30). */
void c2o_factor_rev(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V,
                    const double *d, const double *W, const double *S, const double *bd, const double *bW, double *bt, double *bc,
                    double *ba, double *bU, double *bV) {
  (void)a;
  (void)V;
  double *bS   = (double *)calloc((size_t)(J * J), sizeof(double)); /* column-major */
  double *p    = (double *)malloc(sizeof(double) * (size_t)J);
  double *bp   = (double *)malloc(sizeof(double) * (size_t)J);
  double *x    = (double *)malloc(sizeof(double) * (size_t)J);
  double *bSWT = (double *)malloc(sizeof(double) * (size_t)J);
  memset(bt, 0, sizeof(double) * (size_t)N);
  memset(bc, 0, sizeof(double) * (size_t)J);
  for (int64_t n = 0; n < N; ++n) {
    ba[n] = bd[n];                                                            /* reverse.hpp:55 */
    for (int64_t j = 0; j < J; ++j) bV[n * J + j] = bW[n * J + j] / d[n];     /* reverse.hpp:56-57 */
  }
  for (int64_t n = N - 1; n > 0; --n) {
    const double dt = t[n - 1] - t[n];
    for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
    const double *Sn = S + n * J * J; /* column-major: Sn(j,i) = Sn[j + J*i] */
    /* Step 6, reverse.hpp:65-67 */
    double s = 0.0;
    for (int64_t j = 0; j < J; ++j) s += W[n * J + j] * bV[n * J + j];
    ba[n] -= s;
    for (int64_t j = 0; j < J; ++j) x[j] = bV[n * J + j] + 2.0 * ba[n] * U[n * J + j];
    for (int64_t i = 0; i < J; ++i) {
      double r = 0.0;
      for (int64_t j = 0; j < J; ++j) r += x[j] * Sn[j + J * i];
      bU[n * J + i] = -(r * p[i]);
    }
    for (int64_t i = 0; i < J; ++i)
      for (int64_t j = 0; j < J; ++j) bS[j + J * i] -= U[n * J + j] * (bV[n * J + i] + ba[n] * U[n * J + i]);
    /* Step 4, reverse.hpp:70-74: bp = diag(bS Sn + Sn^T bS) * p */
    double factor = 0.0;
    for (int64_t i = 0; i < J; ++i) {
      double r = 0.0;
      for (int64_t j = 0; j < J; ++j) r += bS[i + J * j] * Sn[j + J * i] + Sn[j + J * i] * bS[j + J * i];
      bp[i] = r * p[i];
      bc[i] += dt * bp[i];
      factor += c[i] * bp[i];
    }
    bt[n] -= factor;
    bt[n - 1] += factor;
    /* Step 3, reverse.hpp:77-80 */
    for (int64_t i = 0; i < J; ++i)
      for (int64_t j = 0; j < J; ++j) bS[j + J * i] = p[j] * bS[j + J * i] * p[i];
    double q = 0.0;
    for (int64_t j = 0; j < J; ++j) {
      double r = 0.0;
      for (int64_t i = 0; i < J; ++i) r += bS[j + J * i] * W[(n - 1) * J + i];
      bSWT[j] = r;
      q += W[(n - 1) * J + j] * r;
    }
    ba[n - 1] += q;
    for (int64_t i = 0; i < J; ++i) {
      double r = 0.0;
      for (int64_t j = 0; j < J; ++j) r += W[(n - 1) * J + j] * (bS[j + J * i] + bS[i + J * j]);
      bV[(n - 1) * J + i] += r;
    }
  }
  for (int64_t j = 0; j < J; ++j) bU[j] = 0.0; /* reverse.hpp:83 */
  {
    double s = 0.0;
    for (int64_t j = 0; j < J; ++j) s += bV[j] * W[j];
    ba[0] -= s; /* reverse.hpp:84 */
  }
  free(bS);
  free(p);
  free(bp);
  free(x);
  free(bSWT);
}

/* core::general_matmul_lower, forward.hpp:285-332 (upper == 0) and general_matmul_upper, forward.hpp:346-392.
 * Z (N,K) is accumulated into; rows of t1 outside the range covered by t2 are untouched.  F (M,J*K) may be NULL;
 * its rows are ROW-major (j*K + k). */
void c2o_general_matmul(int32_t upper, int64_t N, int64_t M, int64_t J, int64_t K, const double *t1, const double *t2, const double *c,
                        const double *U, const double *V, const double *Y, double *Z, double *F) {
  double *Fm = (double *)malloc(sizeof(double) * (size_t)(J * K));
  double *p  = (double *)malloc(sizeof(double) * (size_t)J);
  if (F) memset(F, 0, sizeof(double) * (size_t)(J * K)); /* F.row(0).setZero() in BOTH variants: forward.hpp:307,369 */
  if (!upper) {
    for (int64_t j = 0; j < J; ++j)
      for (int64_t k = 0; k < K; ++k) Fm[j * K + k] = V[j] * Y[k]; /* forward.hpp:311 */
    if (F) memcpy(F, Fm, sizeof(double) * (size_t)(J * K));       /* forward.hpp:313 */
    double tn = t2[0];
    int64_t n, m = 1;
    for (n = 0; n < N; ++n)
      if (t1[n] >= tn) break;
    for (; n < N; ++n) {
      tn = t1[n];
      while (m < M && t2[m] <= tn) {
        const double dt = t2[m - 1] - t2[m];
        for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
        for (int64_t j = 0; j < J; ++j)
          for (int64_t k = 0; k < K; ++k) Fm[j * K + k] = p[j] * Fm[j * K + k];
        for (int64_t j = 0; j < J; ++j)
          for (int64_t k = 0; k < K; ++k) Fm[j * K + k] += V[m * J + j] * Y[m * K + k];
        if (F) memcpy(F + m * J * K, Fm, sizeof(double) * (size_t)(J * K));
        m++;
      }
      const double dt = t2[m - 1] - tn;
      for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
      for (int64_t k = 0; k < K; ++k) {
        double s = 0.0;
        for (int64_t j = 0; j < J; ++j) s += (U[n * J + j] * p[j]) * Fm[j * K + k]; /* forward.hpp:330 */
        Z[n * K + k] += s;
      }
    }
  } else {
    for (int64_t j = 0; j < J; ++j)
      for (int64_t k = 0; k < K; ++k) Fm[j * K + k] = V[(M - 1) * J + j] * Y[(M - 1) * K + k]; /* forward.hpp:373 */
    double tn = t2[M - 1];
    int64_t n, m = M - 2;
    for (n = N - 1; n >= 0; --n)
      if (t1[n] < tn) break;
    for (; n >= 0; --n) {
      tn = t1[n];
      while (m >= 0 && t2[m] > tn) {
        const double dt = t2[m] - t2[m + 1];
        for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
        for (int64_t j = 0; j < J; ++j)
          for (int64_t k = 0; k < K; ++k) Fm[j * K + k] = p[j] * Fm[j * K + k];
        for (int64_t j = 0; j < J; ++j)
          for (int64_t k = 0; k < K; ++k) Fm[j * K + k] += V[m * J + j] * Y[m * K + k];
        if (F) memcpy(F + m * J * K, Fm, sizeof(double) * (size_t)(J * K));
        m--;
      }
      const double dt = tn - t2[m + 1];
      for (int64_t j = 0; j < J; ++j) p[j] = exp(c[j] * dt);
      for (int64_t k = 0; k < K; ++k) {
        double s = 0.0;
        for (int64_t j = 0; j < J; ++j) s += (U[n * J + j] * p[j]) * Fm[j * K + k]; /* forward.hpp:390 */
        Z[n * K + k] += s;
      }
    }
  }
  free(Fm);
  free(p);
}

/* core::to_dense, forward.hpp:19-42.  K_out (N,N) row-major. */
void c2o_to_dense(int64_t N, int64_t J, const double *t, const double *c, const double *a, const double *U, const double *V, double *K_out) {
  double p[C2O_MAXJ];
  for (int64_t m = 0; m < N; ++m) {
    for (int64_t j = 0; j < J; ++j) p[j] = 1.0;
    K_out[m * N + m] = a[m];
    for (int64_t n = m + 1; n < N; ++n) {
      double s = 0.0;
      for (int64_t j = 0; j < J; ++j) {
        p[j] *= exp(-c[j] * (t[n] - t[n - 1]));
        s += U[n * J + j] * V[m * J + j] * p[j];
      }
      K_out[n * N + m] = s;
      K_out[m * N + n] = s;
    }
  }
}

/* a/U/V assembly.
 *   layout 0 "blocked"     [real.. | cos.. | sin..]        Term::get_celerite_matrices, c++/include/celerite2/terms.hpp:112-139
 *   layout 1 "interleaved" [real.. | cos0 sin0 cos1 sin1]  driver get_celerite_matrices, python/celerite2/driver.cpp:422-477,
 *                                                          with the c fill of python/celerite2/terms.py:171-173
 * c_out (J) may be NULL (the numpy front-end fills c itself). */
void c2o_get_celerite_matrices(int32_t layout, int64_t N, int64_t Jr, int64_t Jc, const double *ar, const double *cr, const double *ac,
                               const double *bc, const double *cc, const double *dc, const double *x, const double *diag, double *c_out,
                               double *a_out, double *U_out, double *V_out) {
  const int64_t J = Jr + 2 * Jc;
  double sum      = 0.0;
  for (int64_t j = 0; j < Jr; ++j) sum += ar[j];
  for (int64_t j = 0; j < Jc; ++j) sum += ac[j];
  if (c_out) {
    for (int64_t j = 0; j < Jr; ++j) c_out[j] = cr[j];
    for (int64_t j = 0; j < Jc; ++j) {
      const int64_t ic = layout ? Jr + 2 * j : Jr + j;
      const int64_t is = layout ? Jr + 2 * j + 1 : Jr + Jc + j;
      c_out[ic]        = cc[j];
      c_out[is]        = cc[j];
    }
  }
  for (int64_t n = 0; n < N; ++n) {
    a_out[n] = diag[n] + sum;
    for (int64_t j = 0; j < Jr; ++j) {
      V_out[n * J + j] = 1.0;
      U_out[n * J + j] = ar[j];
    }
    for (int64_t j = 0; j < Jc; ++j) {
      const int64_t ic = layout ? Jr + 2 * j : Jr + j;
      const int64_t is = layout ? Jr + 2 * j + 1 : Jr + Jc + j;
      const double arg = dc[j] * x[n];
      const double cs = cos(arg), sn = sin(arg);
      V_out[n * J + ic] = cs;
      V_out[n * J + is] = sn;
      U_out[n * J + ic] = ac[j] * cs + bc[j] * sn;
      U_out[n * J + is] = ac[j] * sn - bc[j] * cs;
    }
  }
}
