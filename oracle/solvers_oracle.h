// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md). Never used by the
// product path; only tests/, smoke() and bench.py's CPU-baseline legs call it.
//
// Eigen-free CPU restatements (IEEE double, operation order as in the cited
// lines) of the three example solvers of the reference. They satisfy the
// reference's Solver / HybridSolver concept (README.md:105-154, 164-208) so the
// SAME classes are driven (a) by the reference's own unmodified RansacLib/*.h
// templates in oracle/_ref (built from /root/reference) and (b) by the restated
// LO-MSAC of oracle/lomsac_oracle.h.
//
// Eigen, Ceres and PoseLib are absent from this image and cannot be installed,
// so: fixed-size Eigen expressions are written out; pose_lib::p3p is the
// Lambda-Twist restatement of oracle/p3p_oracle.h (parity unpinned, see there);
// the Ceres refinement is a Levenberg-Marquardt restatement with Ceres' default
// trust-region constants (parity unpinned: Ceres is not in /root/reference).
#ifndef RLB_ORACLE_SOLVERS_ORACLE_H_
#define RLB_ORACLE_SOLVERS_ORACLE_H_

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

#include "p3p_oracle.h"

// The reference defines EvaluateModelOnPoint out-of-line in the solver's own
// translation unit (e.g. calibrated_absolute_pose_estimator.cc:154), so its
// scoring loop pays a call per point. The timed reference build keeps that
// structure with -DRLO_EVAL_NOINLINE; parity builds leave it inlinable.
#ifdef RLO_EVAL_NOINLINE
#define RLO_EVAL_ATTR __attribute__((noinline))
#else
#define RLO_EVAL_ATTR
#endif

namespace rlb_oracle {

// ----------------------------------------------------------------------------
// Model PODs. Line: (a,b,c) with a*x+b*y+c=0. Pose: 3x4 column-major [R | c]
// exactly like Eigen::Matrix<double,3,4> (calibrated_absolute_pose_estimator.h:55,68-70).
struct Line {
  double v[3];
};
struct Pose34 {
  double m[12];  // m[r + 3*c]
};
typedef std::vector<Line> Lines;
typedef std::vector<Pose34> Poses;

// Symmetric 2x2 eigenvector for the LARGEST eigenvalue (what Eigen's
// SelfAdjointEigenSolver<Matrix2d>::eigenvectors().col(1) holds: eigenvalues are
// sorted increasingly). Sign is implementation-defined in Eigen; squared
// residuals do not depend on it. line_estimator.cc:97-100.
inline bool sym2_largest_eigvec(double a, double b, double d, double* ex,
                                double* ey) {
  const double h = 0.5 * (a - d);
  const double r = std::sqrt(h * h + b * b);
  const double lam = 0.5 * (a + d) + r;
  // (A - lam I) v = 0 -> v = (b, lam - a) or (lam - d, b); take the larger.
  double v0x = b, v0y = lam - a;
  double v1x = lam - d, v1y = b;
  const double n0 = v0x * v0x + v0y * v0y;
  const double n1 = v1x * v1x + v1y * v1y;
  double vx = v0x, vy = v0y, nn = n0;
  if (n1 > n0) {
    vx = v1x;
    vy = v1y;
    nn = n1;
  }
  if (!(nn > 0.0)) {
    // isotropic covariance: any direction is an eigenvector.
    *ex = 0.0;
    *ey = 1.0;
    return (a == a) && (d == d);
  }
  const double inv = 1.0 / std::sqrt(nn);
  *ex = vx * inv;
  *ey = vy * inv;
  return true;
}

// ----------------------------------------------------------------------------
// A solver that is NOT one of the reference's examples: a 2-D circle (cx, cy, r) written against the
// solver concept of the reference's README.md:105-154. It exists to drive the reference's unmodified
// ransac.h with a user-defined residual, as the CPU side of the user-defined-functor parity test
// (examples/circle_estimator.h is the device-backed counterpart; the arithmetic below is written out
// independently, in the same order of operations).
struct Circle {
  double v[3];
};
typedef std::vector<Circle> Circles;

class CircleSolver {
 public:
  CircleSolver(int n, const double* xy) : n_(n), xy_(xy) {}
  inline int min_sample_size() const { return 3; }
  inline int non_minimal_sample_size() const { return 6; }
  inline int num_data() const { return n_; }

  // circumscribed circle of three points (perpendicular-bisector intersection)
  int MinimalSolver(const std::vector<int>& sample, Circles* out) const {
    out->clear();
    if (sample.size() < 3u) return 0;
    const double x1 = X(sample[0]), y1 = Y(sample[0]);
    const double x2 = X(sample[1]), y2 = Y(sample[1]);
    const double x3 = X(sample[2]), y3 = Y(sample[2]);
    const double a = x2 - x1, b = y2 - y1, c = x3 - x1, d = y3 - y1;
    const double e = a * (x1 + x2) + b * (y1 + y2);
    const double f = c * (x1 + x3) + d * (y1 + y3);
    const double g = 2.0 * (a * (y3 - y2) - b * (x3 - x2));
    if (g == 0.0) return 0;
    Circle m;
    m.v[0] = (d * e - b * f) / g;
    m.v[1] = (a * f - c * e) / g;
    const double dx = x1 - m.v[0], dy = y1 - m.v[1];
    m.v[2] = std::sqrt(dx * dx + dy * dy);
    out->push_back(m);
    return 1;
  }

  // algebraic (Kasa) fit: minimise sum (x^2 + y^2 + A x + B y + C)^2, 3x3 normal equations by Cramer
  int NonMinimalSolver(const std::vector<int>& sample, Circle* out) const {
    if (sample.size() < 6u) return 0;
    double sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0, sxz = 0, syz = 0, sz = 0;
    const double cnt = static_cast<double>(sample.size());
    for (size_t i = 0; i < sample.size(); ++i) {
      const double x = X(sample[i]), y = Y(sample[i]);
      const double z = x * x + y * y;
      sx += x;
      sy += y;
      sxx += x * x;
      syy += y * y;
      sxy += x * y;
      sxz += x * z;
      syz += y * z;
      sz += z;
    }
    // [sxx sxy sx; sxy syy sy; sx sy n] [A B C]' = -[sxz syz sz]'
    const double det = sxx * (syy * cnt - sy * sy) - sxy * (sxy * cnt - sy * sx) + sx * (sxy * sy - syy * sx);
    if (det == 0.0 || !(det == det)) return 0;
    const double r0 = -sxz, r1 = -syz, r2 = -sz;
    const double A = (r0 * (syy * cnt - sy * sy) - sxy * (r1 * cnt - sy * r2) + sx * (r1 * sy - syy * r2)) / det;
    const double B = (sxx * (r1 * cnt - r2 * sy) - r0 * (sxy * cnt - sy * sx) + sx * (sxy * r2 - r1 * sx)) / det;
    const double Cc = (sxx * (syy * r2 - sy * r1) - sxy * (sxy * r2 - r1 * sx) + r0 * (sxy * sy - syy * sx)) / det;
    const double cx = -0.5 * A, cy = -0.5 * B;
    const double rr = cx * cx + cy * cy - Cc;
    if (!(rr > 0.0)) return 0;
    out->v[0] = cx;
    out->v[1] = cy;
    out->v[2] = std::sqrt(rr);
    return 1;
  }

  double EvaluateModelOnPoint(const Circle& m, int i) const {
    const double dx = X(i) - m.v[0], dy = Y(i) - m.v[1];
    const double d = std::sqrt(dx * dx + dy * dy) - m.v[2];
    return d * d;
  }

  inline void LeastSquares(const std::vector<int>& sample, Circle* m) const { NonMinimalSolver(sample, m); }

 private:
  double X(int i) const { return xy_[2 * static_cast<size_t>(i)]; }
  double Y(int i) const { return xy_[2 * static_cast<size_t>(i) + 1]; }
  int n_;
  const double* xy_;
};

// ----------------------------------------------------------------------------
// examples/line_estimator.{h,cc}
class LineSolver {
 public:
  // xy: interleaved x0,y0,x1,y1,... (= column-major Eigen::Matrix2Xd, line_estimator.h:77)
  LineSolver(int n, const double* xy) : n_(n), xy_(xy) {}

  inline int min_sample_size() const { return 2; }          // line_estimator.h:53
  inline int non_minimal_sample_size() const { return 6; }  // line_estimator.h:55
  inline int num_data() const { return n_; }

  // line_estimator.cc:53-72
  int MinimalSolver(const std::vector<int>& sample, Lines* lines) const {
    lines->clear();
    if (sample.size() < 2u) return 0;
    Line l;
    if (!TwoPoint(xy_, sample[0], sample[1], &l)) return 0;
    lines->push_back(l);
    return 1;
  }

  static bool TwoPoint(const double* xy, int i0, int i1, Line* l) {
    const double x1 = xy[2 * i0], y1 = xy[2 * i0 + 1];
    const double x2 = xy[2 * i1], y2 = xy[2 * i1 + 1];
    // (x1,y1,1) x (x2,y2,1)
    double a = y1 * 1.0 - 1.0 * y2;
    double b = 1.0 * x2 - x1 * 1.0;
    double c = x1 * y2 - y1 * x2;
    const double nrm = std::sqrt(a * a + b * b);
    if (nrm == 0.0) return false;
    l->v[0] = a / nrm;
    l->v[1] = b / nrm;
    l->v[2] = c / nrm;
    return true;
  }

  // line_estimator.cc:74-107. Keeps the reference's quirk: the eigenvector of
  // the LARGEST eigenvalue (the line direction) is written as the normal.
  int NonMinimalSolver(const std::vector<int>& sample, Line* line) const {
    if (sample.size() < 6u) return 0;
    const int k = static_cast<int>(sample.size());
    double mx = 0.0, my = 0.0;
    for (int i = 0; i < k; ++i) {
      mx += xy_[2 * sample[i]];
      my += xy_[2 * sample[i] + 1];
    }
    mx /= static_cast<double>(k);
    my /= static_cast<double>(k);
    double cxx = 0.0, cxy = 0.0, cyy = 0.0;
    for (int i = 0; i < k; ++i) {
      const double dx = xy_[2 * sample[i]] - mx;
      const double dy = xy_[2 * sample[i] + 1] - my;
      cxx += dx * dx;
      cxy += dx * dy;
      cyy += dy * dy;
    }
    const double s = static_cast<double>(k - 1);
    cxx /= s;
    cxy /= s;
    cyy /= s;
    double ex, ey;
    if (!sym2_largest_eigvec(cxx, cxy, cyy, &ex, &ey)) return 0;
    line->v[0] = ex;
    line->v[1] = ey;
    line->v[2] = -(ex * mx + ey * my);
    return 1;
  }

  // line_estimator.cc:110-114
  RLO_EVAL_ATTR double EvaluateModelOnPoint(const Line& l, int i) const {
    const double r = l.v[0] * xy_[2 * i] + l.v[1] * xy_[2 * i + 1] + l.v[2] * 1.0;
    return r * r;
  }

  // line_estimator.h:69-72
  inline void LeastSquares(const std::vector<int>& sample, Line* line) const {
    NonMinimalSolver(sample, line);
  }

 private:
  int n_;
  const double* xy_;
};

// ----------------------------------------------------------------------------
// examples/hybrid_line_estimator.{h,cc}
class HybridLineSolver {
 public:
  // xy0: 2 x n0 column-major; xynn1: 4 x n1 column-major (x,y,nx,ny per point).
  HybridLineSolver(int n0, const double* xy0, int n1, const double* xynn1,
                   const std::vector<double>& priors)
      : n0_(n0), xy0_(xy0), n1_(n1), p1_(xynn1), priors_(priors) {}

  inline int num_minimal_solvers() const { return 2; }  // .h:57
  inline void min_sample_sizes(std::vector<std::vector<int>>* s) const {  // .h:59-64
    s->resize(2);
    (*s)[0] = {2, 0};
    (*s)[1] = {0, 1};
  }
  inline int num_data_types() const { return 2; }  // .h:66
  inline void num_data(std::vector<int>* nd) const {  // .h:68-72
    nd->resize(2);
    (*nd)[0] = n0_;
    (*nd)[1] = n1_;
  }
  inline void solver_probabilities(std::vector<double>* p) const { *p = priors_; }

  // Per-point sampling weights, the extra concept member HybridBiasedSampling reads
  // (RansacLib/hybrid_sampling.h:149-152). No reference example provides one.
  void set_weights(const std::vector<double>& w0, const std::vector<double>& w1) {
    weights_.clear();
    weights_.push_back(w0);
    weights_.push_back(w1);
  }
  void get_weights(std::vector<std::vector<double>>& w) const { w = weights_; }

  // .h:79-87
  inline int MinimalSolver(const std::vector<std::vector<int>>& sample,
                           const int solver_idx, Lines* lines) const {
    if (solver_idx == 0) return TwoPointSolver(sample[0], lines);
    return PointNormalSolver(sample[1], lines);
  }

  // hybrid_line_estimator.cc:100-109
  RLO_EVAL_ATTR double EvaluateModelOnPoint(const Line& l, int t, int i) const {
    double r;
    if (t == 0) {
      r = l.v[0] * xy0_[2 * i] + l.v[1] * xy0_[2 * i + 1] + l.v[2] * 1.0;
    } else {
      r = l.v[0] * p1_[4 * i] + l.v[1] * p1_[4 * i + 1] + l.v[2] * 1.0;
    }
    return r * r;
  }

  // hybrid_line_estimator.cc:58-97
  void LeastSquares(const std::vector<std::vector<int>>& sample,
                    Line* line) const {
    const int k0 = static_cast<int>(sample[0].size());
    const int k1 = static_cast<int>(sample[1].size());
    const int k = k0 + k1;
    if (k < 6) return;
    double mx = 0.0, my = 0.0;
    for (int i = 0; i < k0; ++i) {
      mx += xy0_[2 * sample[0][i]];
      my += xy0_[2 * sample[0][i] + 1];
    }
    for (int i = 0; i < k1; ++i) {
      mx += p1_[4 * sample[1][i]];
      my += p1_[4 * sample[1][i] + 1];
    }
    mx /= static_cast<double>(k);
    my /= static_cast<double>(k);
    double cxx = 0.0, cxy = 0.0, cyy = 0.0;
    for (int i = 0; i < k0; ++i) {
      const double dx = xy0_[2 * sample[0][i]] - mx;
      const double dy = xy0_[2 * sample[0][i] + 1] - my;
      cxx += dx * dx;
      cxy += dx * dy;
      cyy += dy * dy;
    }
    for (int i = 0; i < k1; ++i) {
      const double dx = p1_[4 * sample[1][i]] - mx;
      const double dy = p1_[4 * sample[1][i] + 1] - my;
      cxx += dx * dx;
      cxy += dx * dy;
      cyy += dy * dy;
    }
    const double s = static_cast<double>(k - 1);
    cxx /= s;
    cxy /= s;
    cyy /= s;
    double ex, ey;
    if (!sym2_largest_eigvec(cxx, cxy, cyy, &ex, &ey)) return;
    line->v[0] = ex;
    line->v[1] = ey;
    line->v[2] = -(ex * mx + ey * my);
  }

  // hybrid_line_estimator.cc:111-130
  int TwoPointSolver(const std::vector<int>& sample, Lines* lines) const {
    lines->clear();
    if (sample.size() < 2u) return 0;
    Line l;
    if (!LineSolver::TwoPoint(xy0_, sample[0], sample[1], &l)) return 0;
    lines->push_back(l);
    return 1;
  }

  // hybrid_line_estimator.cc:132-152
  int PointNormalSolver(const std::vector<int>& sample, Lines* lines) const {
    lines->clear();
    if (sample.size() < 1u) return 0;
    Line l;
    if (!PointNormal(p1_, sample[0], &l)) return 0;
    lines->push_back(l);
    return 1;
  }

  static bool PointNormal(const double* p, int i, Line* l) {
    double nx = p[4 * i + 2], ny = p[4 * i + 3];
    // Eigen normalize(): divides by the norm only if it is > 0.
    const double n2 = nx * nx + ny * ny;
    if (n2 > 0.0) {
      const double n = std::sqrt(n2);
      nx /= n;
      ny /= n;
    }
    const double c = -(nx * p[4 * i] + ny * p[4 * i + 1]);
    const double nrm = std::sqrt(nx * nx + ny * ny);
    if (nrm == 0.0) return false;
    l->v[0] = nx / nrm;
    l->v[1] = ny / nrm;
    l->v[2] = c / nrm;
    return true;
  }

 private:
  int n0_;
  const double* xy0_;
  int n1_;
  const double* p1_;
  std::vector<double> priors_;
  std::vector<std::vector<double>> weights_;
};

// ----------------------------------------------------------------------------
// Pose refinement: stands in for the Ceres problem of
// calibrated_absolute_pose_estimator.cc:170-212 (AutoDiffCostFunction<.,2,6>,
// DENSE_QR, default Levenberg-Marquardt options).
namespace lsq {

// Rotation matrix (row-major 3x3) -> angle-axis, via quaternion like
// Eigen::AngleAxisd(Matrix3d) does (.cc:172-173).
inline void rot_to_angle_axis(const double* R, double* aa) {
  double w, x, y, z;
  const double tr = R[0] + R[4] + R[8];
  if (tr > 0.0) {
    double t = std::sqrt(tr + 1.0);
    w = 0.5 * t;
    t = 0.5 / t;
    x = (R[7] - R[5]) * t;
    y = (R[2] - R[6]) * t;
    z = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    double q[3];
    q[i] = 0.5 * t;
    t = 0.5 / t;
    w = (R[3 * k + j] - R[3 * j + k]) * t;
    q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
    x = q[0];
    y = q[1];
    z = q[2];
  }
  double n = std::sqrt(x * x + y * y + z * z);
  if (n < DBL_MIN) {
    aa[0] = aa[1] = aa[2] = 0.0;
    // Eigen sets axis (1,0,0), angle 0 -> product 0.
    return;
  }
  if (w < 0.0) n = -n;
  const double angle = 2.0 * std::atan2(n, std::fabs(w));
  aa[0] = x / n * angle;
  aa[1] = y / n * angle;
  aa[2] = z / n * angle;
}

// Angle-axis -> rotation matrix (row-major), Rodrigues (.cc:204-210).
inline void angle_axis_to_rot(const double* aa, double* R) {
  const double angle = std::sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
  double ax = aa[0], ay = aa[1], az = aa[2];
  if (angle > 0.0) {
    ax /= angle;
    ay /= angle;
    az /= angle;
  }
  const double s = std::sin(angle), c = std::cos(angle);
  const double tx = (1.0 - c) * ax, ty = (1.0 - c) * ay, tz = (1.0 - c) * az;
  const double sx = s * ax, sy = s * ay, sz = s * az;
  R[0] = tx * ax + c;
  R[1] = tx * ay - sz;
  R[2] = tx * az + sy;
  R[3] = tx * ay + sz;
  R[4] = ty * ay + c;
  R[5] = ty * az - sx;
  R[6] = tx * az - sy;
  R[7] = ty * az + sx;
  R[8] = tz * az + c;
}

// Tiny forward-mode dual number with 6 partials (what ceres::Jet<double,6> is).
struct Jet6 {
  double a;
  double v[6];
};
inline Jet6 J(double a) {
  Jet6 r;
  r.a = a;
  for (int i = 0; i < 6; ++i) r.v[i] = 0.0;
  return r;
}
inline Jet6 Jvar(double a, int k) {
  Jet6 r = J(a);
  r.v[k] = 1.0;
  return r;
}
inline Jet6 operator+(const Jet6& x, const Jet6& y) {
  Jet6 r;
  r.a = x.a + y.a;
  for (int i = 0; i < 6; ++i) r.v[i] = x.v[i] + y.v[i];
  return r;
}
inline Jet6 operator-(const Jet6& x, const Jet6& y) {
  Jet6 r;
  r.a = x.a - y.a;
  for (int i = 0; i < 6; ++i) r.v[i] = x.v[i] - y.v[i];
  return r;
}
inline Jet6 operator*(const Jet6& x, const Jet6& y) {
  Jet6 r;
  r.a = x.a * y.a;
  for (int i = 0; i < 6; ++i) r.v[i] = x.a * y.v[i] + x.v[i] * y.a;
  return r;
}
inline Jet6 operator/(const Jet6& x, const Jet6& y) {
  Jet6 r;
  const double inv = 1.0 / y.a;
  r.a = x.a * inv;
  for (int i = 0; i < 6; ++i) r.v[i] = (x.v[i] - r.a * y.v[i]) * inv;
  return r;
}
inline Jet6 jsqrt(const Jet6& x) {
  Jet6 r;
  r.a = std::sqrt(x.a);
  const double d = 0.5 / r.a;
  for (int i = 0; i < 6; ++i) r.v[i] = x.v[i] * d;
  return r;
}
inline Jet6 jsin(const Jet6& x) {
  Jet6 r;
  r.a = std::sin(x.a);
  const double d = std::cos(x.a);
  for (int i = 0; i < 6; ++i) r.v[i] = x.v[i] * d;
  return r;
}
inline Jet6 jcos(const Jet6& x) {
  Jet6 r;
  r.a = std::cos(x.a);
  const double d = -std::sin(x.a);
  for (int i = 0; i < 6; ++i) r.v[i] = x.v[i] * d;
  return r;
}

// Residual of .cc:55-74 (ReprojectionError::operator()) on jets, with
// ceres::AngleAxisRotatePoint's two branches (theta^2 > epsilon or first order).
inline void reproj_residual(const Jet6* cam, double u, double v, double X,
                            double Y, double Z, double fx, double fy,
                            Jet6* res) {
  Jet6 p[3] = {J(X) - cam[3], J(Y) - cam[4], J(Z) - cam[5]};
  Jet6 pr[3];
  const Jet6 theta2 = cam[0] * cam[0] + cam[1] * cam[1] + cam[2] * cam[2];
  if (theta2.a > std::numeric_limits<double>::epsilon()) {
    const Jet6 theta = jsqrt(theta2);
    const Jet6 ct = jcos(theta), st = jsin(theta);
    const Jet6 inv = J(1.0) / theta;
    const Jet6 w[3] = {cam[0] * inv, cam[1] * inv, cam[2] * inv};
    const Jet6 wxp[3] = {w[1] * p[2] - w[2] * p[1], w[2] * p[0] - w[0] * p[2],
                         w[0] * p[1] - w[1] * p[0]};
    const Jet6 tmp = (w[0] * p[0] + w[1] * p[1] + w[2] * p[2]) * (J(1.0) - ct);
    for (int i = 0; i < 3; ++i) pr[i] = p[i] * ct + wxp[i] * st + w[i] * tmp;
  } else {
    const Jet6 wxp[3] = {cam[1] * p[2] - cam[2] * p[1],
                         cam[2] * p[0] - cam[0] * p[2],
                         cam[0] * p[1] - cam[1] * p[0]};
    for (int i = 0; i < 3; ++i) pr[i] = p[i] + wxp[i];
  }
  const Jet6 xp = J(fx) * pr[0] / pr[2];
  const Jet6 yp = J(fy) * pr[1] / pr[2];
  res[0] = J(u) - xp;
  res[1] = J(v) - yp;
}

// Solve (A) x = b for symmetric positive definite 6x6 by Cholesky; false if not SPD.
inline bool chol6_solve(const double* A, const double* b, double* x) {
  double L[36];
  for (int i = 0; i < 6; ++i) {
    for (int j = 0; j <= i; ++j) {
      double s = A[6 * i + j];
      for (int k = 0; k < j; ++k) s -= L[6 * i + k] * L[6 * j + k];
      if (i == j) {
        if (!(s > 0.0)) return false;
        L[6 * i + i] = std::sqrt(s);
      } else {
        L[6 * i + j] = s / L[6 * j + j];
      }
    }
  }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double s = b[i];
    for (int k = 0; k < i; ++k) s -= L[6 * i + k] * y[k];
    y[i] = s / L[6 * i + i];
  }
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < 6; ++k) s -= L[6 * k + i] * x[k];
    x[i] = s / L[6 * i + i];
  }
  return true;
}

// Levenberg-Marquardt with Ceres' default trust-region schedule
// (initial radius 1e4, max 50 iterations, function tol 1e-6, gradient tol 1e-10,
// parameter tol 1e-8, min relative decrease 1e-3, diagonal clamp [1e-6,1e32]).
// Returns false if the result is not usable (non-finite).
inline bool refine_pose(int k, const int* idx, const double* p2d,
                        const double* p3d, double fx, double fy, double* cam) {
  auto eval = [&](const double* c, double* JtJ, double* Jtr) -> double {
    Jet6 cj[6];
    for (int i = 0; i < 6; ++i) cj[i] = Jvar(c[i], i);
    double cost = 0.0;
    if (JtJ) {
      for (int i = 0; i < 36; ++i) JtJ[i] = 0.0;
      for (int i = 0; i < 6; ++i) Jtr[i] = 0.0;
    }
    for (int s = 0; s < k; ++s) {
      const int i = idx[s];
      Jet6 r[2];
      reproj_residual(cj, p2d[2 * i], p2d[2 * i + 1], p3d[3 * i],
                      p3d[3 * i + 1], p3d[3 * i + 2], fx, fy, r);
      cost += 0.5 * (r[0].a * r[0].a + r[1].a * r[1].a);
      if (JtJ) {
        for (int e = 0; e < 2; ++e) {
          for (int a = 0; a < 6; ++a) {
            Jtr[a] += r[e].v[a] * r[e].a;
            for (int b = 0; b <= a; ++b) JtJ[6 * a + b] += r[e].v[a] * r[e].v[b];
          }
        }
      }
    }
    if (JtJ) {
      for (int a = 0; a < 6; ++a)
        for (int b = a + 1; b < 6; ++b) JtJ[6 * a + b] = JtJ[6 * b + a];
    }
    return cost;
  };

  double x[6];
  for (int i = 0; i < 6; ++i) x[i] = cam[i];
  double JtJ[36], g[6];
  double cost = eval(x, JtJ, g);
  if (!(cost == cost) || cost > DBL_MAX) return false;
  double radius = 1e4;
  double decrease_factor = 2.0;
  for (int iter = 0; iter < 50; ++iter) {
    double gmax = 0.0;
    for (int i = 0; i < 6; ++i) gmax = std::max(gmax, std::fabs(g[i]));
    if (gmax <= 1e-10) break;
    double A[36], rhs[6], step[6];
    for (int i = 0; i < 36; ++i) A[i] = JtJ[i];
    for (int i = 0; i < 6; ++i) {
      double d = JtJ[7 * i];
      d = std::min(std::max(d, 1e-6), 1e32);
      A[7 * i] += d / radius;
      rhs[i] = -g[i];
    }
    bool ok = chol6_solve(A, rhs, step);
    double model_decrease = 0.0;
    if (ok) {
      // model cost change: -(g.s + 0.5 s^T JtJ s)
      double gs = 0.0, sJs = 0.0;
      for (int i = 0; i < 6; ++i) {
        gs += g[i] * step[i];
        double t = 0.0;
        for (int j = 0; j < 6; ++j) t += JtJ[6 * i + j] * step[j];
        sJs += step[i] * t;
      }
      model_decrease = -(gs + 0.5 * sJs);
      ok = model_decrease > 0.0;
    }
    if (!ok) {
      radius /= decrease_factor;
      decrease_factor *= 2.0;
      if (radius < 1e-32) break;
      continue;
    }
    double xn[6], sn = 0.0, xnorm = 0.0;
    for (int i = 0; i < 6; ++i) {
      xn[i] = x[i] + step[i];
      sn += step[i] * step[i];
      xnorm += x[i] * x[i];
    }
    sn = std::sqrt(sn);
    xnorm = std::sqrt(xnorm);
    if (sn <= 1e-8 * (xnorm + 1e-8)) break;
    const double cost_new = eval(xn, nullptr, nullptr);
    const double rho = (cost - cost_new) / model_decrease;
    if (cost_new == cost_new && rho > 1e-3) {
      const double delta = cost - cost_new;
      for (int i = 0; i < 6; ++i) x[i] = xn[i];
      const double prev = cost;
      cost = eval(x, JtJ, g);
      const double t = 2.0 * rho - 1.0;
      radius = radius / std::max(1.0 / 3.0, 1.0 - t * t * t);
      radius = std::min(radius, 1e16);
      decrease_factor = 2.0;
      if (std::fabs(delta) <= 1e-6 * prev) break;
    } else {
      radius /= decrease_factor;
      decrease_factor *= 2.0;
      if (radius < 1e-32) break;
    }
  }
  for (int i = 0; i < 6; ++i)
    if (!(x[i] - x[i] == 0.0)) return false;
  for (int i = 0; i < 6; ++i) cam[i] = x[i];
  return true;
}

}  // namespace lsq

// ----------------------------------------------------------------------------
// examples/calibrated_absolute_pose_estimator.{h,cc}
class PoseSolver {
 public:
  // p2d: n x 2 (u,v centred on the principal point); rays: n x 3 unit bearings;
  // p3d: n x 3. Row-major == std::vector<Eigen::Vector*d> memory.
  PoseSolver(double fx, double fy, double sq_thr, int n, const double* p2d,
             const double* rays, const double* p3d)
      : fx_(fx), fy_(fy), thr_(sq_thr), n_(n), p2d_(p2d), rays_(rays), p3d_(p3d) {}

  inline int min_sample_size() const { return 4; }          // .h:88
  inline int non_minimal_sample_size() const { return 6; }  // .h:90
  inline int num_data() const { return n_; }

  // .cc:111-137
  int MinimalSolver(const std::vector<int>& sample, Poses* poses) const {
    poses->clear();
    Pose34 P;
    if (Minimal4(sample[0], sample[1], sample[2], sample[3], &P)) {
      poses->push_back(P);
    }
    return static_cast<int>(poses->size());
  }

  bool Minimal4(int i0, int i1, int i2, int i3, Pose34* P) const {
    double y[9], x[9];
    const int id[3] = {i0, i1, i2};
    for (int i = 0; i < 3; ++i) {
      for (int c = 0; c < 3; ++c) {
        y[3 * i + c] = rays_[3 * id[i] + c];
        x[3 * i + c] = p3d_[3 * id[i] + c];
      }
    }
    P3PSolutions sols;
    if (p3p_lambda_twist(y, x, &sols) == 0) return false;
    for (int s = 0; s < sols.n; ++s) {
      const double* R = sols.R[s];
      const double* t = sols.t[s];
      Pose34 Q;
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Q.m[r + 3 * c] = R[3 * r + c];
      // c = -R^T t  (.cc:127)
      for (int c = 0; c < 3; ++c) {
        Q.m[9 + c] = -(R[c] * t[0] + R[3 + c] * t[1] + R[6 + c] * t[2]);
      }
      const double e = EvaluateModelOnPoint(Q, i3);
      if (e < thr_) {
        *P = Q;
        return true;
      }
    }
    return false;
  }

  // .cc:141-151
  int NonMinimalSolver(const std::vector<int>& sample, Pose34* pose) const {
    Poses poses;
    if (MinimalSolver(sample, &poses) == 1) {
      *pose = poses[0];
      LeastSquares(sample, pose);
      return 1;
    }
    return 0;
  }

  // .cc:154-167
  RLO_EVAL_ATTR double EvaluateModelOnPoint(const Pose34& P, int i) const {
    const double d0 = p3d_[3 * i] - P.m[9];
    const double d1 = p3d_[3 * i + 1] - P.m[10];
    const double d2 = p3d_[3 * i + 2] - P.m[11];
    const double pz = P.m[2] * d0 + P.m[5] * d1 + P.m[8] * d2;
    if (pz < 0.0) return std::numeric_limits<double>::max();
    const double px = P.m[0] * d0 + P.m[3] * d1 + P.m[6] * d2;
    const double py = P.m[1] * d0 + P.m[4] * d1 + P.m[7] * d2;
    const double u = (px / pz) * fx_;
    const double v = (py / pz) * fy_;
    const double du = u - p2d_[2 * i];
    const double dv = v - p2d_[2 * i + 1];
    return du * du + dv * dv;
  }

  // .cc:170-212
  void LeastSquares(const std::vector<int>& sample, Pose34* pose) const {
    double R[9];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) R[3 * r + c] = pose->m[r + 3 * c];
    double cam[6];
    lsq::rot_to_angle_axis(R, cam);
    cam[3] = pose->m[9];
    cam[4] = pose->m[10];
    cam[5] = pose->m[11];
    if (!lsq::refine_pose(static_cast<int>(sample.size()), sample.data(), p2d_,
                          p3d_, fx_, fy_, cam))
      return;
    lsq::angle_axis_to_rot(cam, R);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) pose->m[r + 3 * c] = R[3 * r + c];
    pose->m[9] = cam[3];
    pose->m[10] = cam[4];
    pose->m[11] = cam[5];
  }

  // .cc:214-227
  static void PixelsToViewingRays(double fx, double fy, int n, const double* p2d,
                                  double* rays) {
    for (int i = 0; i < n; ++i) {
      const double x = p2d[2 * i] / fx;
      const double y = p2d[2 * i + 1] / fy;
      const double z = 1.0;
      const double nrm = std::sqrt(x * x + y * y + z * z);
      rays[3 * i] = x / nrm;
      rays[3 * i + 1] = y / nrm;
      rays[3 * i + 2] = z / nrm;
    }
  }

  double fx() const { return fx_; }
  double fy() const { return fy_; }

 private:
  double fx_, fy_, thr_;
  int n_;
  const double* p2d_;
  const double* rays_;
  const double* p3d_;
};

// ----------------------------------------------------------------------------
// A hybrid POSE solver: the use the reference's README names for HybridRANSAC ("2D-3D matches and ... for pose
// estimation", RansacLib/hybrid_ransac.h:126-137 is the concept) but ships no example of. Two data types,
//   type 0: 2D-3D matches (u, v | X), squared reprojection error as CalibratedAbsolutePoseEstimator (.cc:154-167),
//   type 1: 3D-3D matches (a point measured in the camera frame, e.g. by a depth sensor | X), squared distance,
// and two minimal solvers: P3P + 4th point on type 0 ({4, 0}), three 3D-3D matches ({0, 3}: the frames spanned by
// the two triangles give the rotation). LeastSquares refines on the sample's 2D-3D matches (PoseSolver's LM).
// The product's counterpart is examples/hybrid_pose_estimator.h (same operations in the same order); this class
// drives BOTH the restated hybrid LO-MSAC and, under RLO_USE_REFERENCE, the reference's own template.
class HybridPoseSolver {
 public:
  // p2d: n0 x 2, rays: n0 x 3, p3d: n0 x 3 as PoseSolver; cam1 / world1: n1 x 3 each.
  HybridPoseSolver(double fx, double fy, double sq_thr0, int n0, const double* p2d, const double* rays,
                   const double* p3d, int n1, const double* cam1, const double* world1,
                   const std::vector<double>& priors)
      : pose_(fx, fy, sq_thr0, n0, p2d, rays, p3d), n0_(n0), n1_(n1), cam1_(cam1), world1_(world1), priors_(priors) {}

  inline int num_minimal_solvers() const { return 2; }
  inline void min_sample_sizes(std::vector<std::vector<int>>* s) const {
    s->resize(2);
    (*s)[0] = {4, 0};
    (*s)[1] = {0, 3};
  }
  inline int num_data_types() const { return 2; }
  inline void num_data(std::vector<int>* nd) const {
    nd->resize(2);
    (*nd)[0] = n0_;
    (*nd)[1] = n1_;
  }
  inline void solver_probabilities(std::vector<double>* p) const { *p = priors_; }

  inline int MinimalSolver(const std::vector<std::vector<int>>& sample, const int solver_idx, Poses* poses) const {
    poses->clear();
    Pose34 P;
    if (solver_idx == 0) {
      const std::vector<int>& s = sample[0];
      if (s.size() < 4u) return 0;
      if (!pose_.Minimal4(s[0], s[1], s[2], s[3], &P)) return 0;
    } else {
      const std::vector<int>& s = sample[1];
      if (s.size() < 3u) return 0;
      if (!FromTriangles(&world1_[3 * s[0]], &world1_[3 * s[1]], &world1_[3 * s[2]], &cam1_[3 * s[0]],
                         &cam1_[3 * s[1]], &cam1_[3 * s[2]], P.m))
        return 0;
    }
    poses->push_back(P);
    return 1;
  }

  // Right-handed orthonormal frame of a triangle: e1 along P0 -> P1, e3 normal to the triangle, e2 = e3 x e1.
  static bool TriangleFrame(const double* P0, const double* P1, const double* P2, double* e1, double* e2, double* e3) {
    const double a[3] = {P1[0] - P0[0], P1[1] - P0[1], P1[2] - P0[2]};
    const double b[3] = {P2[0] - P0[0], P2[1] - P0[1], P2[2] - P0[2]};
    const double na = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
    if (!(na > 0.0)) return false;
    e1[0] = a[0] / na;
    e1[1] = a[1] / na;
    e1[2] = a[2] / na;
    const double c[3] = {e1[1] * b[2] - e1[2] * b[1], e1[2] * b[0] - e1[0] * b[2], e1[0] * b[1] - e1[1] * b[0]};
    const double nc = std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    if (!(nc > 0.0)) return false;
    e3[0] = c[0] / nc;
    e3[1] = c[1] / nc;
    e3[2] = c[2] / nc;
    e2[0] = e3[1] * e1[2] - e3[2] * e1[1];
    e2[1] = e3[2] * e1[0] - e3[0] * e1[2];
    e2[2] = e3[0] * e1[1] - e3[1] * e1[0];
    return true;
  }

  // cam_i = R (world_i - c): R = F E^T with the two triangle frames as columns, c = world_0 - R^T cam_0.
  // m: R column-major, then c (the layout of Pose34).
  static bool FromTriangles(const double* A0, const double* A1, const double* A2, const double* B0, const double* B1,
                            const double* B2, double* m) {
    double e1[3], e2[3], e3[3], f1[3], f2[3], f3[3];
    if (!TriangleFrame(A0, A1, A2, e1, e2, e3)) return false;
    if (!TriangleFrame(B0, B1, B2, f1, f2, f3)) return false;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) m[r + 3 * c] = f1[r] * e1[c] + f2[r] * e2[c] + f3[r] * e3[c];
    for (int j = 0; j < 3; ++j) m[9 + j] = A0[j] - (m[3 * j] * B0[0] + m[3 * j + 1] * B0[1] + m[3 * j + 2] * B0[2]);
    return true;
  }

  RLO_EVAL_ATTR double EvaluateModelOnPoint(const Pose34& P, int t, int i) const {
    if (t == 0) return pose_.EvaluateModelOnPoint(P, i);
    const double d0 = world1_[3 * i] - P.m[9];
    const double d1 = world1_[3 * i + 1] - P.m[10];
    const double d2 = world1_[3 * i + 2] - P.m[11];
    const double ex = (P.m[0] * d0 + P.m[3] * d1 + P.m[6] * d2) - cam1_[3 * i];
    const double ey = (P.m[1] * d0 + P.m[4] * d1 + P.m[7] * d2) - cam1_[3 * i + 1];
    const double ez = (P.m[2] * d0 + P.m[5] * d1 + P.m[8] * d2) - cam1_[3 * i + 2];
    return ex * ex + ey * ey + ez * ez;
  }

  // Non-linear refinement on the sample's 2D-3D matches (six or more; otherwise the model stays).
  void LeastSquares(const std::vector<std::vector<int>>& sample, Pose34* pose) const {
    if (sample[0].size() < 6u) return;
    pose_.LeastSquares(sample[0], pose);
  }

 private:
  PoseSolver pose_;
  int n0_, n1_;
  const double* cam1_;
  const double* world1_;
  std::vector<double> priors_;
};

}  // namespace rlb_oracle

#endif  // RLB_ORACLE_SOLVERS_ORACLE_H_
