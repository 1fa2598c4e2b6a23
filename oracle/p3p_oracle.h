// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under ransaclib_b200/ or include/
// may include, link or execute this file. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
//
// P3P (calibrated absolute pose from 3 bearing/point pairs), CPU f64.
//
// PARITY UNPINNED: the reference does not contain this arithmetic. It calls
// pose_lib::p3p() from PoseLib (examples/calibrated_absolute_pose_estimator.cc:120-121,
// header include examples/calibrated_absolute_pose_estimator.h:46-47), which is
// neither vendored nor version-pinned (examples/CMakeLists.txt:13
// find_package(PoseLib REQUIRED); .gitmodules lists only pybind11). PoseLib's
// p3p of that era restates Persson & Nordberg, "Lambda Twist: An Accurate Fast
// Robust Perspective Three Point (P3P) Solver", ECCV 2018. This file restates
// that published algorithm (quadric pencil -> one cubic root -> degenerate conic
// -> two quadratics -> <=4 depth triples -> R,t), with our own fixed solution
// ORDER (sign '+' branch first, then ascending tau), because PoseLib's order
// cannot be known here. It is pinned independently by tests/golden/p3p_cv2.json
// (solution SET agreement with OpenCV solveP3P) and by invariants.
//
// Only +,-,*,/,sqrt,frexp,ldexp are used, so that the CUDA restatement
// (compiled with -fmad=false) is bit-identical to this file compiled with
// -ffp-contract=off.
#ifndef RLB_ORACLE_P3P_ORACLE_H_
#define RLB_ORACLE_P3P_ORACLE_H_

#include <cmath>

namespace rlb_oracle {

struct P3PSolutions {
  int n;
  double R[4][9];  // row-major 3x3, lambda * y = R * x + t
  double t[4][3];
};

namespace p3p_detail {

inline void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}
inline double dot3(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

// det of symmetric-or-not 3x3 row-major.
inline double det3(const double* m) {
  return m[0] * (m[4] * m[8] - m[5] * m[7]) -
         m[1] * (m[3] * m[8] - m[5] * m[6]) +
         m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// tr(adj(A) * B) for 3x3 row-major A,B: the coefficient of g in det(A + g B).
inline double tr_adj_mul(const double* A, const double* B) {
  // adj(A)^T = cofactor matrix C; tr(adj(A) B) = sum_ij C_ij B_ij
  double c00 = A[4] * A[8] - A[5] * A[7];
  double c01 = -(A[3] * A[8] - A[5] * A[6]);
  double c02 = A[3] * A[7] - A[4] * A[6];
  double c10 = -(A[1] * A[8] - A[2] * A[7]);
  double c11 = A[0] * A[8] - A[2] * A[6];
  double c12 = -(A[0] * A[7] - A[1] * A[6]);
  double c20 = A[1] * A[5] - A[2] * A[4];
  double c21 = -(A[0] * A[5] - A[2] * A[3]);
  double c22 = A[0] * A[4] - A[1] * A[3];
  return c00 * B[0] + c01 * B[1] + c02 * B[2] + c10 * B[3] + c11 * B[4] +
         c12 * B[5] + c20 * B[6] + c21 * B[7] + c22 * B[8];
}

// Power-of-two upper bound of cbrt(q), q > 0, exact on every IEEE platform.
inline double cbrt_upper_pow2(double q) {
  int e;
  (void)std::frexp(q, &e);  // q < 2^e
  int ce = (e >= 0) ? (e + 2) / 3 : -((-e) / 3);
  return std::ldexp(1.0, ce);
}

// One real root of x^3 + a x^2 + b x + c by monotone Newton on the convex
// (or concave) side of the inflection point. Deterministic: no libm beyond
// sqrt/frexp/ldexp.
inline double cubic_one_root(double a, double b, double c) {
  const double xi = -a / 3.0;                    // inflection
  const double p = b - a * a / 3.0;              // f'(xi)
  const double q = ((xi + a) * xi + b) * xi + c; // f(xi)
  if (q == 0.0) return xi;
  // depressed: g(h) = h^3 + p h + q; root h with sign opposite to q.
  const double aq = std::fabs(q);
  double h;  // upper bound of |root|
  if (p >= 0.0) {
    h = cbrt_upper_pow2(aq);
    if (p > 0.0) {
      double lin = aq / p;
      if (lin < h) h = lin;
    }
  } else {
    h = std::sqrt(-p) + cbrt_upper_pow2(aq);
  }
  // Work on |h| with g~(h) = h^3 + p h - |q| (mirror when q > 0): root > 0,
  // convex for h > 0, start above the root -> monotone decrease.
  for (int it = 0; it < 60; ++it) {
    const double g = (h * h + p) * h - aq;
    const double dg = 3.0 * h * h + p;
    if (!(dg > 0.0)) break;
    const double hn = h - g / dg;
    if (!(hn < h)) break;  // converged (or stalled at rounding level)
    if (h - hn <= 1e-16 * h) {
      h = hn;
      break;
    }
    h = hn;
  }
  return (q < 0.0) ? xi + h : xi - h;
}

// Unit null-ish vector of a (near) rank-2 3x3 row-major matrix: the largest of
// the three row cross products.
inline bool null_vector(const double* m, double* v) {
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
  const double inv = 1.0 / std::sqrt(nb);
  v[0] = best[0] * inv;
  v[1] = best[1] * inv;
  v[2] = best[2] * inv;
  return true;
}

// A few Gauss-Newton steps on the three depth constraints
//   l_i^2 + l_j^2 + b_ij l_i l_j = a_ij   (Lambda Twist sec. 3.8).
inline void refine_lambda(double* L, double a12, double a13, double a23,
                          double b12, double b13, double b23) {
  for (int it = 0; it < 5; ++it) {
    const double l1 = L[0], l2 = L[1], l3 = L[2];
    const double r1 = l1 * l1 + l2 * l2 + b12 * l1 * l2 - a12;
    const double r2 = l1 * l1 + l3 * l3 + b13 * l1 * l3 - a13;
    const double r3 = l2 * l2 + l3 * l3 + b23 * l2 * l3 - a23;
    if (std::fabs(r1) + std::fabs(r2) + std::fabs(r3) < 1e-10) break;
    // Jacobian rows
    const double j00 = 2.0 * l1 + b12 * l2, j01 = 2.0 * l2 + b12 * l1;
    const double j10 = 2.0 * l1 + b13 * l3, j12 = 2.0 * l3 + b13 * l1;
    const double j21 = 2.0 * l2 + b23 * l3, j22 = 2.0 * l3 + b23 * l2;
    // J = [j00 j01 0; j10 0 j12; 0 j21 j22]
    const double det = -j00 * j12 * j21 - j01 * j10 * j22;
    if (det == 0.0 || det != det) break;
    const double idet = 1.0 / det;
    // inverse via adjugate
    const double d1 = (-j12 * j21 * r1 - j01 * j22 * r2 + j01 * j12 * r3) * idet;
    const double d2 = (-j10 * j22 * r1 + j00 * j22 * r2 - j00 * j12 * r3) * idet;
    const double d3 = (j10 * j21 * r1 - j00 * j21 * r2 - j01 * j10 * r3) * idet;
    L[0] = l1 - d1;
    L[1] = l2 - d2;
    L[2] = l3 - d3;
  }
}

}  // namespace p3p_detail

// y: 3 unit bearing vectors (y[3*i..]); x: 3 world points. Returns #solutions.
inline int p3p_lambda_twist(const double* y, const double* x,
                            P3PSolutions* out) {
  using namespace p3p_detail;
  out->n = 0;
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
  if (!(a12 > 0.0) || !(a13 > 0.0) || !(a23 > 0.0)) return 0;

  // D1 = M12*a23 - M23*a12 ; D2 = M13*a23 - M23*a13 (symmetric, row-major).
  const double D1[9] = {a23,  0.5 * a23 * b12,  0.0,
                        0.5 * a23 * b12, a23 - a12, -0.5 * a12 * b23,
                        0.0, -0.5 * a12 * b23, -a12};
  const double D2[9] = {a23, 0.0, 0.5 * a23 * b13,
                        0.0, -a13, -0.5 * a13 * b23,
                        0.5 * a23 * b13, -0.5 * a13 * b23, a23 - a13};

  // det(D1 + g D2) = c0 + c1 g + c2 g^2 + c3 g^3
  const double c3 = det3(D2);
  const double c2 = tr_adj_mul(D2, D1);
  const double c1 = tr_adj_mul(D1, D2);
  const double c0 = det3(D1);
  if (c3 == 0.0 || c3 != c3) return 0;
  const double ic3 = 1.0 / c3;
  const double g = cubic_one_root(c2 * ic3, c1 * ic3, c0 * ic3);

  double D0[9];
  for (int i = 0; i < 9; ++i) D0[i] = D1[i] + g * D2[i];

  // Eigen-decomposition of the rank-2 symmetric D0 with a known zero eigenvalue.
  // Non-zero eigenvalues: roots of s^2 - tr s + m2 = 0 (m2 = sum of principal minors).
  const double tr = D0[0] + D0[4] + D0[8];
  const double m2 = (D0[0] * D0[4] - D0[1] * D0[3]) +
                    (D0[0] * D0[8] - D0[2] * D0[6]) +
                    (D0[4] * D0[8] - D0[5] * D0[7]);
  double disc = tr * tr - 4.0 * m2;
  if (disc < 0.0) disc = 0.0;
  const double sq = std::sqrt(disc);
  // s_pos > 0 > s_neg for an indefinite degenerate conic (m2 < 0).
  double s_pos, s_neg;
  if (tr >= 0.0) {
    s_pos = 0.5 * (tr + sq);
    s_neg = (s_pos != 0.0) ? m2 / s_pos : 0.0;
  } else {
    s_neg = 0.5 * (tr - sq);
    s_pos = (s_neg != 0.0) ? m2 / s_neg : 0.0;
  }
  if (!(s_pos > 0.0) || !(s_neg <= 0.0)) return 0;

  double A[9], e1[3], e2[3];
  for (int i = 0; i < 9; ++i) A[i] = D0[i];
  A[0] -= s_pos; A[4] -= s_pos; A[8] -= s_pos;
  if (!null_vector(A, e1)) return 0;
  for (int i = 0; i < 9; ++i) A[i] = D0[i];
  A[0] -= s_neg; A[4] -= s_neg; A[8] -= s_neg;
  if (!null_vector(A, e2)) return 0;

  const double s = std::sqrt(-s_neg / s_pos);

  // X = [d12 d13 d12xd13] (columns); its inverse is shared by all solutions.
  // Xinv rows = (d13 x n, n x d12, n) / |n|^2 with n = d12 x d13.
  const double nn = dot3(d12xd13, d12xd13);
  if (!(nn > 0.0)) return 0;
  double r0[3], r1[3];
  cross3(d13, d12xd13, r0);
  cross3(d12xd13, d12, r1);
  const double inn = 1.0 / nn;
  const double Xinv[9] = {r0[0] * inn, r0[1] * inn, r0[2] * inn,
                          r1[0] * inn, r1[1] * inn, r1[2] * inn,
                          d12xd13[0] * inn, d12xd13[1] * inn, d12xd13[2] * inn};

  for (int branch = 0; branch < 2; ++branch) {
    const double sg = (branch == 0) ? s : -s;
    // (e1 + sg e2) . lambda = 0  ->  l1 = w0 l2 + w1 l3
    const double n0 = e1[0] + sg * e2[0];
    const double n1 = e1[1] + sg * e2[1];
    const double n2 = e1[2] + sg * e2[2];
    if (n0 == 0.0) continue;
    const double w0 = -n1 / n0;
    const double w1 = -n2 / n0;
    // lambda^T D1 lambda = 0 with l1 = l2 (w0 + w1 tau), l3 = tau l2:
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
        const double sd = std::sqrt(dsc);
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
      L[1] = std::sqrt(a23 / den);
      L[2] = tau * L[1];
      L[0] = w0 * L[1] + w1 * L[2];
      if (!(L[0] > 0.0)) continue;
      refine_lambda(L, a12, a13, a23, b12, b13, b23);
      if (!(L[0] > 0.0) || !(L[1] > 0.0) || !(L[2] > 0.0)) continue;

      // Y = [z1 z2 z1xz2], z1 = l1 y1 - l2 y2, z2 = l1 y1 - l3 y3; R = Y X^-1.
      double z1[3], z2[3], z3[3];
      for (int i = 0; i < 3; ++i) {
        z1[i] = L[0] * y1[i] - L[1] * y2[i];
        z2[i] = L[0] * y1[i] - L[2] * y3[i];
      }
      cross3(z1, z2, z3);
      double* R = out->R[out->n];
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
          R[3 * r + c] =
              z1[r] * Xinv[c] + z2[r] * Xinv[3 + c] + z3[r] * Xinv[6 + c];
        }
      }
      double* t = out->t[out->n];
      for (int r = 0; r < 3; ++r) {
        t[r] = L[0] * y1[r] -
               (R[3 * r] * x1[0] + R[3 * r + 1] * x1[1] + R[3 * r + 2] * x1[2]);
      }
      bool finite = true;
      for (int i = 0; i < 9; ++i) finite = finite && (R[i] - R[i] == 0.0);
      for (int i = 0; i < 3; ++i) finite = finite && (t[i] - t[i] == 0.0);
      if (!finite) continue;
      out->n += 1;
      if (out->n == 4) return 4;
    }
  }
  return out->n;
}

}  // namespace rlb_oracle

#endif  // RLB_ORACLE_P3P_ORACLE_H_
