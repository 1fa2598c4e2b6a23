// P3P (Lambda Twist, Persson & Nordberg ECCV 2018, fixed solution order) and the reference-exact
// single-point reprojection error, for device and host. Shared by csrc/solver_kernels.cu (one
// sample per thread), csrc/localize_kernels.cu (whole LO-MSAC runs on the device) and the host
// side of CalibratedAbsolutePoseEstimator::MinimalSolver (one sample: no launch). The kernels are
// compiled with -fmad=false, the host library with -ffp-contract=off, and this code uses only
// + - * / sqrt frexp ldexp, so the three give bit-identical poses.
//
// Replaces (reference): examples/calibrated_absolute_pose_estimator.cc:111-137 (+ pose_lib::p3p,
// external to the reference) and :154-167.
#ifndef RLB_EXAMPLES_P3P_H_
#define RLB_EXAMPLES_P3P_H_

#include <cfloat>
#include <cmath>

#ifdef __CUDACC__
#define RLB_P3P_HD __host__ __device__
#define RLB_P3P_HD_INLINE __host__ __device__ __forceinline__
#else
#define RLB_P3P_HD inline
#define RLB_P3P_HD_INLINE inline
#endif

// This file's functions must round like a strict-IEEE evaluation: no fused multiply-add contraction.
// nvcc device code gets that from -fmad=false on the command line (csrc/Makefile); for host compilers the
// setting is pinned HERE, so that a user who includes the drop-in headers with the reference's own flags
// (-O3 -march=native: GCC contracts by default) still gets the trajectories of the device-resident drivers.
#if defined(__clang__)
#pragma STDC FP_CONTRACT OFF
#elif defined(__GNUC__) && !defined(__CUDA_ARCH__)
#pragma GCC push_options
#pragma GCC optimize("fp-contract=off")
#define RLB_FP_CONTRACT_PUSHED_P3P 1
#endif

namespace rlb {
namespace p3pdev {

static RLB_P3P_HD_INLINE void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}
static RLB_P3P_HD_INLINE double dot3(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
static RLB_P3P_HD_INLINE double det3(const double* m) {
  return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
         m[2] * (m[3] * m[7] - m[4] * m[6]);
}
// tr(adj(A) B): coefficient of g in det(A + g B).
static RLB_P3P_HD double tr_adj_mul(const double* A, const double* B) {
  const double c00 = A[4] * A[8] - A[5] * A[7];
  const double c01 = -(A[3] * A[8] - A[5] * A[6]);
  const double c02 = A[3] * A[7] - A[4] * A[6];
  const double c10 = -(A[1] * A[8] - A[2] * A[7]);
  const double c11 = A[0] * A[8] - A[2] * A[6];
  const double c12 = -(A[0] * A[7] - A[1] * A[6]);
  const double c20 = A[1] * A[5] - A[2] * A[4];
  const double c21 = -(A[0] * A[5] - A[2] * A[3]);
  const double c22 = A[0] * A[4] - A[1] * A[3];
  return c00 * B[0] + c01 * B[1] + c02 * B[2] + c10 * B[3] + c11 * B[4] + c12 * B[5] +
         c20 * B[6] + c21 * B[7] + c22 * B[8];
}
static RLB_P3P_HD double cbrt_upper_pow2(double q) {
  int e;
  (void)frexp(q, &e);
  const int ce = (e >= 0) ? (e + 2) / 3 : -((-e) / 3);
  return ldexp(1.0, ce);
}
// One real root of x^3 + a x^2 + b x + c: monotone Newton from an upper bound on
// the convex side of the inflection point.
static RLB_P3P_HD double cubic_one_root(double a, double b, double c) {
  const double xi = -a / 3.0;
  const double p = b - a * a / 3.0;
  const double q = ((xi + a) * xi + b) * xi + c;
  if (q == 0.0) return xi;
  const double aq = fabs(q);
  double h;
  if (p >= 0.0) {
    h = cbrt_upper_pow2(aq);
    if (p > 0.0) {
      const double lin = aq / p;
      if (lin < h) h = lin;
    }
  } else {
    h = sqrt(-p) + cbrt_upper_pow2(aq);
  }
  for (int it = 0; it < 60; ++it) {
    const double g = (h * h + p) * h - aq;
    const double dg = 3.0 * h * h + p;
    if (!(dg > 0.0)) break;
    const double hn = h - g / dg;
    if (!(hn < h)) break;
    if (h - hn <= 1e-16 * h) {
      h = hn;
      break;
    }
    h = hn;
  }
  return (q < 0.0) ? xi + h : xi - h;
}
static RLB_P3P_HD bool null_vector(const double* m, double* v) {
  double c0[3], c1[3], c2[3];
  cross3(m + 0, m + 3, c0);
  cross3(m + 0, m + 6, c1);
  cross3(m + 3, m + 6, c2);
  const double n0 = dot3(c0, c0), n1 = dot3(c1, c1), n2 = dot3(c2, c2);
  const double* best = c0;
  double nb = n0;
  if (n1 > nb) {
    best = c1;
    nb = n1;
  }
  if (n2 > nb) {
    best = c2;
    nb = n2;
  }
  if (!(nb > 0.0)) return false;
  const double inv = 1.0 / sqrt(nb);
  v[0] = best[0] * inv;
  v[1] = best[1] * inv;
  v[2] = best[2] * inv;
  return true;
}
static RLB_P3P_HD void refine_lambda(double* L, double a12, double a13, double a23, double b12, double b13,
                              double b23) {
  for (int it = 0; it < 5; ++it) {
    const double l1 = L[0], l2 = L[1], l3 = L[2];
    const double r1 = l1 * l1 + l2 * l2 + b12 * l1 * l2 - a12;
    const double r2 = l1 * l1 + l3 * l3 + b13 * l1 * l3 - a13;
    const double r3 = l2 * l2 + l3 * l3 + b23 * l2 * l3 - a23;
    if (fabs(r1) + fabs(r2) + fabs(r3) < 1e-10) break;
    const double j00 = 2.0 * l1 + b12 * l2, j01 = 2.0 * l2 + b12 * l1;
    const double j10 = 2.0 * l1 + b13 * l3, j12 = 2.0 * l3 + b13 * l1;
    const double j21 = 2.0 * l2 + b23 * l3, j22 = 2.0 * l3 + b23 * l2;
    const double det = -j00 * j12 * j21 - j01 * j10 * j22;
    if (det == 0.0 || det != det) break;
    const double idet = 1.0 / det;
    const double d1 = (-j12 * j21 * r1 - j01 * j22 * r2 + j01 * j12 * r3) * idet;
    const double d2 = (-j10 * j22 * r1 + j00 * j22 * r2 - j00 * j12 * r3) * idet;
    const double d3 = (j10 * j21 * r1 - j00 * j21 * r2 - j01 * j10 * r3) * idet;
    L[0] = l1 - d1;
    L[1] = l2 - d2;
    L[2] = l3 - d3;
  }
}

// Reference-exact residual on ONE point in the reference's own formulation
// (calibrated_absolute_pose_estimator.cc:154-167): used for the 4th-point test.
static RLB_P3P_HD_INLINE double reproj_sq_error_v(const double* P /*col-major 3x4 [R|c]*/,
                                                           double X0, double X1, double X2, double u0,
                                                           double v0, double fx, double fy) {
  const double d0 = X0 - P[9], d1 = X1 - P[10], d2 = X2 - P[11];
  const double pz = P[2] * d0 + P[5] * d1 + P[8] * d2;
  if (pz < 0.0) return DBL_MAX;
  const double px = P[0] * d0 + P[3] * d1 + P[6] * d2;
  const double py = P[1] * d0 + P[4] * d1 + P[7] * d2;
  const double u = (px / pz) * fx;
  const double v = (py / pz) * fy;
  const double du = u - u0, dv = v - v0;
  return du * du + dv * dv;
}
static RLB_P3P_HD double reproj_sq_error(const double* P /*col-major 3x4 [R|c]*/, const double* X,
                                         const double* uv, double fx, double fy) {
  return reproj_sq_error_v(P, X[0], X[1], X[2], uv[0], uv[1], fx, fy);
}

// Lambda Twist + "first solution consistent with the 4th point wins"
// (calibrated_absolute_pose_estimator.cc:124-133). Returns true and fills P.
static RLB_P3P_HD bool p3p_pick(const double* y, const double* x, const double* X4, const double* uv4,
                         double fx, double fy, double sq_thr, double* P) {
  const double* y1 = y;
  const double* y2 = y + 3;
  const double* y3 = y + 6;
  const double* x1 = x;
  const double* x2 = x + 3;
  const double* x3 = x + 6;
  const double b12 = -2.0 * dot3(y1, y2);
  const double b13 = -2.0 * dot3(y1, y3);
  const double b23 = -2.0 * dot3(y2, y3);
  double d12[3] = {x1[0] - x2[0], x1[1] - x2[1], x1[2] - x2[2]};
  double d13[3] = {x1[0] - x3[0], x1[1] - x3[1], x1[2] - x3[2]};
  double d23[3] = {x2[0] - x3[0], x2[1] - x3[1], x2[2] - x3[2]};
  double d12xd13[3];
  cross3(d12, d13, d12xd13);
  const double a12 = dot3(d12, d12);
  const double a13 = dot3(d13, d13);
  const double a23 = dot3(d23, d23);
  if (!(a12 > 0.0) || !(a13 > 0.0) || !(a23 > 0.0)) return false;

  const double D1[9] = {a23, 0.5 * a23 * b12, 0.0, 0.5 * a23 * b12, a23 - a12,
                        -0.5 * a12 * b23, 0.0, -0.5 * a12 * b23, -a12};
  const double D2[9] = {a23, 0.0, 0.5 * a23 * b13, 0.0, -a13,
                        -0.5 * a13 * b23, 0.5 * a23 * b13, -0.5 * a13 * b23, a23 - a13};
  const double c3 = det3(D2);
  const double c2 = tr_adj_mul(D2, D1);
  const double c1 = tr_adj_mul(D1, D2);
  const double c0 = det3(D1);
  if (c3 == 0.0 || c3 != c3) return false;
  const double ic3 = 1.0 / c3;
  const double g = cubic_one_root(c2 * ic3, c1 * ic3, c0 * ic3);

  double D0[9];
  for (int i = 0; i < 9; ++i) D0[i] = D1[i] + g * D2[i];
  const double tr = D0[0] + D0[4] + D0[8];
  const double m2 = (D0[0] * D0[4] - D0[1] * D0[3]) + (D0[0] * D0[8] - D0[2] * D0[6]) +
                    (D0[4] * D0[8] - D0[5] * D0[7]);
  double disc = tr * tr - 4.0 * m2;
  if (disc < 0.0) disc = 0.0;
  const double sq = sqrt(disc);
  double s_pos, s_neg;
  if (tr >= 0.0) {
    s_pos = 0.5 * (tr + sq);
    s_neg = (s_pos != 0.0) ? m2 / s_pos : 0.0;
  } else {
    s_neg = 0.5 * (tr - sq);
    s_pos = (s_neg != 0.0) ? m2 / s_neg : 0.0;
  }
  if (!(s_pos > 0.0) || !(s_neg <= 0.0)) return false;

  double A[9], e1[3], e2[3];
  for (int i = 0; i < 9; ++i) A[i] = D0[i];
  A[0] -= s_pos;
  A[4] -= s_pos;
  A[8] -= s_pos;
  if (!null_vector(A, e1)) return false;
  for (int i = 0; i < 9; ++i) A[i] = D0[i];
  A[0] -= s_neg;
  A[4] -= s_neg;
  A[8] -= s_neg;
  if (!null_vector(A, e2)) return false;
  const double s = sqrt(-s_neg / s_pos);

  const double nn = dot3(d12xd13, d12xd13);
  if (!(nn > 0.0)) return false;
  double r0[3], r1[3];
  cross3(d13, d12xd13, r0);
  cross3(d12xd13, d12, r1);
  const double inn = 1.0 / nn;
  const double Xinv[9] = {r0[0] * inn, r0[1] * inn, r0[2] * inn,
                          r1[0] * inn, r1[1] * inn, r1[2] * inn,
                          d12xd13[0] * inn, d12xd13[1] * inn, d12xd13[2] * inn};

  int nsol = 0;
  for (int branch = 0; branch < 2; ++branch) {
    const double sg = (branch == 0) ? s : -s;
    const double n0 = e1[0] + sg * e2[0];
    const double n1 = e1[1] + sg * e2[1];
    const double n2 = e1[2] + sg * e2[2];
    if (n0 == 0.0) continue;
    const double w0 = -n1 / n0;
    const double w1 = -n2 / n0;
    const double qa = a23 * w1 * w1 - a12;
    const double qb = a23 * (2.0 * w0 * w1 + b12 * w1) - a12 * b23;
    const double qc = a23 * (w0 * w0 + 1.0 + b12 * w0) - a12;
    double taus[2];
    int nt = 0;
    if (qa == 0.0) {
      if (qb != 0.0) taus[nt++] = -qc / qb;
    } else {
      const double dsc = qb * qb - 4.0 * qa * qc;
      if (dsc >= 0.0) {
        const double sd = sqrt(dsc);
        const double qq = (qb >= 0.0) ? -0.5 * (qb + sd) : -0.5 * (qb - sd);
        double ta = qq / qa;
        double tb = (qq != 0.0) ? qc / qq : ta;
        if (tb < ta) {
          const double tmp = ta;
          ta = tb;
          tb = tmp;
        }
        taus[nt++] = ta;
        if (dsc > 0.0) taus[nt++] = tb;
      }
    }
    for (int k = 0; k < nt; ++k) {
      const double tau = taus[k];
      if (!(tau > 0.0)) continue;
      const double den = 1.0 + tau * (tau + b23);
      if (!(den > 0.0)) continue;
      double L[3];
      L[1] = sqrt(a23 / den);
      L[2] = tau * L[1];
      L[0] = w0 * L[1] + w1 * L[2];
      if (!(L[0] > 0.0)) continue;
      refine_lambda(L, a12, a13, a23, b12, b13, b23);
      if (!(L[0] > 0.0) || !(L[1] > 0.0) || !(L[2] > 0.0)) continue;
      double z1[3], z2[3], z3[3];
      for (int i = 0; i < 3; ++i) {
        z1[i] = L[0] * y1[i] - L[1] * y2[i];
        z2[i] = L[0] * y1[i] - L[2] * y3[i];
      }
      cross3(z1, z2, z3);
      double R[9], t[3];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
          R[3 * r + c] = z1[r] * Xinv[c] + z2[r] * Xinv[3 + c] + z3[r] * Xinv[6 + c];
      for (int r = 0; r < 3; ++r)
        t[r] = L[0] * y1[r] - (R[3 * r] * x1[0] + R[3 * r + 1] * x1[1] + R[3 * r + 2] * x1[2]);
      bool finite = true;
      for (int i = 0; i < 9; ++i) finite = finite && (R[i] - R[i] == 0.0);
      for (int i = 0; i < 3; ++i) finite = finite && (t[i] - t[i] == 0.0);
      if (!finite) continue;
      // [R | -R^T t], column-major (calibrated_absolute_pose_estimator.cc:126-127)
      double Q[12];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Q[r + 3 * c] = R[3 * r + c];
      for (int c = 0; c < 3; ++c) Q[9 + c] = -(R[c] * t[0] + R[3 + c] * t[1] + R[6 + c] * t[2]);
      const double e = reproj_sq_error(Q, X4, uv4, fx, fy);
      if (e < sq_thr) {
        for (int i = 0; i < 12; ++i) P[i] = Q[i];
        return true;
      }
      if (++nsol == 4) return false;
    }
  }
  return false;
}

}  // namespace p3pdev
}  // namespace rlb

#if defined(RLB_FP_CONTRACT_PUSHED_P3P)
#pragma GCC pop_options
#undef RLB_FP_CONTRACT_PUSHED_P3P
#endif

#endif  // RLB_EXAMPLES_P3P_H_
