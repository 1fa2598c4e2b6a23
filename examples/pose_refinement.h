// Host-side non-linear pose refinement for CalibratedAbsolutePoseEstimator::LeastSquares.
//
// The reference builds a Ceres problem (AutoDiffCostFunction<ReprojectionError,2,6>,
// DENSE_QR, default Levenberg-Marquardt; examples/calibrated_absolute_pose_estimator.cc:43-97,
// 170-212). Ceres and Eigen are not available in this image, so this file carries a small
// dense LM over the same 6 parameters (angle-axis, camera centre), the same residual and
// forward-mode derivatives, and Ceres' default trust-region constants. It stays on the
// host by design (north_star: "Local optimisation keeps its least-squares refinement on
// the host, but all of its rescoring goes through the GPU scorer").
#ifndef RANSACLIB_B200_EXAMPLES_POSE_REFINEMENT_H_
#define RANSACLIB_B200_EXAMPLES_POSE_REFINEMENT_H_

#include <cfloat>
#include <cmath>

// The same arithmetic also runs inside the device-resident localization kernel
// (ransaclib_b200/csrc/localize_kernels.cu, compiled with -fmad=false): every function here is
// host+device, and the LM loop is a template over the "lineariser" so that the device can form
// cost / J^T J / J^T r with one point per lane while keeping the host's summation order.
#if defined(__CUDACC__)
#define RLB_HD __host__ __device__
#else
#define RLB_HD
#endif
// On the device the LM state below is ONE copy per warp in shared memory that all 32 lanes
// read and (redundantly, with identical values) write; the barrier keeps a lane from
// overwriting state another lane of the warp has not consumed yet. No-op on the host.
#if defined(__CUDA_ARCH__)
#define RLB_WARP_SYNC() __syncwarp()
#else
#define RLB_WARP_SYNC() ((void)0)
#endif
// Small dense loops stay rolled on the device (instruction-cache footprint); plain loops on the host.
#if defined(__CUDACC__) && defined(RLB_ROLL_LM) && RLB_ROLL_LM
#define RLB_ROLLED _Pragma("unroll 1")
#else
#define RLB_ROLLED
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
#define RLB_FP_CONTRACT_PUSHED_POSE_REF 1
#endif

namespace ransac_lib {
namespace calibrated_absolute_pose {
namespace refine {

// std::min / std::max semantics (usable in device code).
RLB_HD inline double Max(double a, double b) { return (a < b) ? b : a; }
RLB_HD inline double Min(double a, double b) { return (b < a) ? b : a; }

// Row-major rotation -> angle-axis through the quaternion, as Eigen::AngleAxisd(Matrix3d)
// does (calibrated_absolute_pose_estimator.cc:172-173).
RLB_HD inline void RotationToAngleAxis(const double* R, double* aa) {
  double w, x, y, z;
  const double trace = R[0] + R[4] + R[8];
  if (trace > 0.0) {
    double t = std::sqrt(trace + 1.0);
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
    return;
  }
  if (w < 0.0) n = -n;
  const double angle = 2.0 * std::atan2(n, std::fabs(w));
  aa[0] = x / n * angle;
  aa[1] = y / n * angle;
  aa[2] = z / n * angle;
}

// Angle-axis -> row-major rotation (Rodrigues; .cc:204-210).
RLB_HD inline void AngleAxisToRotation(const double* aa, double* R) {
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

// Value + ND partial derivatives (the role ceres::Jet<double,6> plays in the reference).
// ND = 0 carries the value only: its arithmetic on `.a` is the same sequence of operations,
// so a cost-only evaluation returns bit-identical residuals without the derivative work.
template <int ND>
struct DualN {
  double a;
  double d[ND > 0 ? ND : 1];
  RLB_HD static DualN Constant(double v) {
    DualN r;
    r.a = v;
    for (int i = 0; i < ND; ++i) r.d[i] = 0.0;
    return r;
  }
  // d/dx_k = 1; a k outside [0, ND) gives a constant (used when the partials are formed in slices).
  RLB_HD static DualN Variable(double v, int k) {
    DualN r;
    r.a = v;
    for (int i = 0; i < ND; ++i) r.d[i] = (i == k) ? 1.0 : 0.0;
    return r;
  }
};
typedef DualN<6> Dual;
typedef DualN<0> Value;

template <int ND>
RLB_HD inline DualN<ND> operator+(const DualN<ND>& x, const DualN<ND>& y) {
  DualN<ND> r;
  r.a = x.a + y.a;
  for (int i = 0; i < ND; ++i) r.d[i] = x.d[i] + y.d[i];
  return r;
}
template <int ND>
RLB_HD inline DualN<ND> operator-(const DualN<ND>& x, const DualN<ND>& y) {
  DualN<ND> r;
  r.a = x.a - y.a;
  for (int i = 0; i < ND; ++i) r.d[i] = x.d[i] - y.d[i];
  return r;
}
template <int ND>
RLB_HD inline DualN<ND> operator*(const DualN<ND>& x, const DualN<ND>& y) {
  DualN<ND> r;
  r.a = x.a * y.a;
  for (int i = 0; i < ND; ++i) r.d[i] = x.a * y.d[i] + x.d[i] * y.a;
  return r;
}
template <int ND>
RLB_HD inline DualN<ND> operator/(const DualN<ND>& x, const DualN<ND>& y) {
  DualN<ND> r;
  const double inv = 1.0 / y.a;
  r.a = x.a * inv;
  for (int i = 0; i < ND; ++i) r.d[i] = (x.d[i] - r.a * y.d[i]) * inv;
  return r;
}
template <int ND>
RLB_HD inline DualN<ND> Sqrt(const DualN<ND>& x) {
  DualN<ND> r;
  r.a = std::sqrt(x.a);
  const double g = 0.5 / r.a;
  for (int i = 0; i < ND; ++i) r.d[i] = x.d[i] * g;
  return r;
}
// sin and cos of the same argument: each libm function is evaluated once and serves both the
// value of one result and the derivative of the other.
template <int ND>
RLB_HD inline void SinCos(const DualN<ND>& x, DualN<ND>* sn, DualN<ND>* cs) {
  const double s = std::sin(x.a);
  const double c = std::cos(x.a);
  sn->a = s;
  cs->a = c;
  const double ms = -s;
  for (int i = 0; i < ND; ++i) sn->d[i] = x.d[i] * c;
  for (int i = 0; i < ND; ++i) cs->d[i] = x.d[i] * ms;
}

// ReprojectionError::operator() (.cc:55-74) with ceres::AngleAxisRotatePoint's two branches.
template <class D>
RLB_HD inline void Residual(const D* cam, double u, double v, double X, double Y, double Z,
                            double fx, double fy, D* res) {
  const D one = D::Constant(1.0);
  D p[3] = {D::Constant(X) - cam[3], D::Constant(Y) - cam[4], D::Constant(Z) - cam[5]};
  D q[3];
  const D theta2 = cam[0] * cam[0] + cam[1] * cam[1] + cam[2] * cam[2];
  if (theta2.a > DBL_EPSILON) {
    const D theta = Sqrt(theta2);
    D ct, st;
    SinCos(theta, &st, &ct);
    const D inv = one / theta;
    const D w[3] = {cam[0] * inv, cam[1] * inv, cam[2] * inv};
    const D wxp[3] = {w[1] * p[2] - w[2] * p[1], w[2] * p[0] - w[0] * p[2],
                      w[0] * p[1] - w[1] * p[0]};
    const D tmp = (w[0] * p[0] + w[1] * p[1] + w[2] * p[2]) * (one - ct);
    for (int i = 0; i < 3; ++i) q[i] = p[i] * ct + wxp[i] * st + w[i] * tmp;
  } else {
    const D wxp[3] = {cam[1] * p[2] - cam[2] * p[1], cam[2] * p[0] - cam[0] * p[2],
                      cam[0] * p[1] - cam[1] * p[0]};
    for (int i = 0; i < 3; ++i) q[i] = p[i] + wxp[i];
  }
  const D xp = D::Constant(fx) * q[0] / q[2];
  const D yp = D::Constant(fy) * q[1] / q[2];
  res[0] = D::Constant(u) - xp;
  res[1] = D::Constant(v) - yp;
}

// A x = b for a symmetric positive definite 6 x 6 A through A = L L^T. Every division by a diagonal entry of L is a
// multiplication by its reciprocal, formed once per column (six divisions instead of 27): on the device an f64
// division is ~12 dependent instructions on the LM's critical path, and the lane-distributed form in
// csrc/localize_kernels.cu (lm_refine_lanes) does the same operations in the same order, so host and device
// refinements stay bit-identical.
RLB_HD inline bool SolveSpd6(const double* A, const double* b, double* x, double* L, double* y) {
  double rinv[6];
  RLB_ROLLED
  for (int i = 0; i < 6; ++i) {
    RLB_ROLLED
    for (int j = 0; j <= i; ++j) {
      double s = A[6 * i + j];
      RLB_ROLLED
      for (int k = 0; k < j; ++k) s -= L[6 * i + k] * L[6 * j + k];
      if (i == j) {
        if (!(s > 0.0)) return false;
        L[6 * i + i] = std::sqrt(s);
        rinv[i] = 1.0 / L[6 * i + i];
      } else {
        L[6 * i + j] = s * rinv[j];
      }
    }
  }
  RLB_ROLLED
  for (int i = 0; i < 6; ++i) {
    double s = b[i];
    RLB_ROLLED
    for (int k = 0; k < i; ++k) s -= L[6 * i + k] * y[k];
    y[i] = s * rinv[i];
  }
  RLB_ROLLED
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    RLB_ROLLED
    for (int k = i + 1; k < 6; ++k) s -= L[6 * k + i] * x[k];
    x[i] = s * rinv[i];
  }
  return true;
}

// Vectors and matrices of one LM run (host: on the stack; device: one per warp in shared memory).
struct LmWorkspace {
  double x[6], xn[6], g[6], g2[6], rhs[6], step[6], y[6];
  double JtJ[36], JtJ2[36], A[36], L[36];
  double R[9], cam[6];
};

// Sequential lineariser (host): cost, J^T J and J^T r over the sample, points in sample order,
// per point the two residual rows in order, entries accumulated as acc += d[a] * d[b].
struct SequentialLineariser {
  static constexpr bool kSpeculativeFull = false;  // the host follows Ceres' evaluation order
  int k;
  const int* idx;
  const double* p2d;
  const double* p3d;
  double fx, fy;

  // Cost only: value arithmetic without derivatives (bit-identical residuals).
  RLB_HD double Cost(const double* c) const {
    Value cam[6];
    for (int i = 0; i < 6; ++i) cam[i] = Value::Constant(c[i]);
    double cost = 0.0;
    for (int s = 0; s < k; ++s) {
      const int i = idx[s];
      Value r[2];
      Residual(cam, p2d[2 * i], p2d[2 * i + 1], p3d[3 * i], p3d[3 * i + 1], p3d[3 * i + 2], fx, fy, r);
      cost += 0.5 * (r[0].a * r[0].a + r[1].a * r[1].a);
    }
    return cost;
  }

  RLB_HD double Full(const double* c, double* JtJ, double* Jtr) const {
    Dual cam[6];
    for (int i = 0; i < 6; ++i) cam[i] = Dual::Variable(c[i], i);
    double cost = 0.0;
    for (int i = 0; i < 36; ++i) JtJ[i] = 0.0;
    for (int i = 0; i < 6; ++i) Jtr[i] = 0.0;
    for (int s = 0; s < k; ++s) {
      const int i = idx[s];
      Dual r[2];
      Residual(cam, p2d[2 * i], p2d[2 * i + 1], p3d[3 * i], p3d[3 * i + 1], p3d[3 * i + 2], fx, fy, r);
      cost += 0.5 * (r[0].a * r[0].a + r[1].a * r[1].a);
      for (int e = 0; e < 2; ++e) {
        for (int a = 0; a < 6; ++a) {
          Jtr[a] += r[e].d[a] * r[e].a;
          for (int b = 0; b <= a; ++b) JtJ[6 * a + b] += r[e].d[a] * r[e].d[b];
        }
      }
    }
    for (int a = 0; a < 6; ++a)
      for (int b = a + 1; b < 6; ++b) JtJ[6 * a + b] = JtJ[6 * b + a];
    return cost;
  }
};

// Levenberg-Marquardt with Ceres' defaults: initial radius 1e4, at most 50 iterations,
// function tolerance 1e-6, gradient tolerance 1e-10, parameter tolerance 1e-8, minimum
// relative decrease 1e-3, diagonal scaling clamped to [1e-6, 1e32]. Returns false when the
// result is not usable (non-finite), like !summary.IsSolutionUsable() (.cc:202).
// Lin provides Cost(x) and Full(x, JtJ, g) (see SequentialLineariser).
template <class Lin>
RLB_HD inline bool RefinePoseWith(const Lin& lin, double* cam, LmWorkspace* w) {
  double* x = w->x;
  double* xn = w->xn;
  double* JtJ = w->JtJ;
  double* g = w->g;
  double* JtJ_next = w->JtJ2;  // linearisation at the trial point; becomes current when accepted
  double* g_next = w->g2;
  double* A = w->A;
  double* rhs = w->rhs;
  double* step = w->step;
  for (int i = 0; i < 6; ++i) x[i] = cam[i];
  RLB_WARP_SYNC();
  double cost = lin.Full(x, JtJ, g);
  if (!(cost == cost) || cost > DBL_MAX) return false;
  double radius = 1e4;
  double shrink = 2.0;
  for (int iter = 0; iter < 50; ++iter) {
    RLB_WARP_SYNC();
    double gmax = 0.0;
    for (int i = 0; i < 6; ++i) gmax = Max(gmax, std::fabs(g[i]));
    if (gmax <= 1e-10) break;
    // A = JtJ + diag(clamp(diag JtJ)) / radius; every entry is written exactly once
    RLB_ROLLED
    for (int i = 0; i < 36; ++i) {
      double v = JtJ[i];
      if (i % 7 == 0) {
        const double dgl = Min(Max(v, 1e-6), 1e32);
        v += dgl / radius;
      }
      A[i] = v;
    }
    for (int i = 0; i < 6; ++i) rhs[i] = -g[i];
    RLB_WARP_SYNC();
    bool ok = SolveSpd6(A, rhs, step, w->L, w->y);
    RLB_WARP_SYNC();
    double model_decrease = 0.0;
    if (ok) {
      double gs = 0.0, sJs = 0.0;
      RLB_ROLLED
      for (int i = 0; i < 6; ++i) {
        gs += g[i] * step[i];
        double t = 0.0;
        RLB_ROLLED
        for (int j = 0; j < 6; ++j) t += JtJ[6 * i + j] * step[j];
        sJs += step[i] * t;
      }
      model_decrease = -(gs + 0.5 * sJs);
      ok = model_decrease > 0.0;
    }
    if (!ok) {
      radius /= shrink;
      shrink *= 2.0;
      if (radius < 1e-32) break;
      continue;
    }
    double sn = 0.0, xnorm = 0.0;
    for (int i = 0; i < 6; ++i) {
      xn[i] = x[i] + step[i];
      sn += step[i] * step[i];
      xnorm += x[i] * x[i];
    }
    RLB_WARP_SYNC();
    sn = std::sqrt(sn);
    xnorm = std::sqrt(xnorm);
    if (sn <= 1e-8 * (xnorm + 1e-8)) break;
    // Ceres evaluates the cost at the trial point and, if the step is accepted, the Jacobian
    // there. Most steps are accepted, so the trial point is linearised right away: its cost is
    // the same number (same value arithmetic, same summation order), and an accepted step only
    // swaps the buffers instead of evaluating every residual a second time.
    const double cost_new = Lin::kSpeculativeFull ? lin.Full(xn, JtJ_next, g_next) : lin.Cost(xn);
    const double rho = (cost - cost_new) / model_decrease;
    if (cost_new == cost_new && rho > 1e-3) {
      const double delta = cost - cost_new;
      const double prev = cost;
      if (Lin::kSpeculativeFull) {
        double* t = x;
        x = xn;
        xn = t;
        t = JtJ;
        JtJ = JtJ_next;
        JtJ_next = t;
        t = g;
        g = g_next;
        g_next = t;
        cost = cost_new;
      } else {
        RLB_WARP_SYNC();
        for (int i = 0; i < 6; ++i) x[i] = xn[i];
        RLB_WARP_SYNC();
        cost = lin.Full(x, JtJ, g);
      }
      const double t = 2.0 * rho - 1.0;
      radius = radius / Max(1.0 / 3.0, 1.0 - t * t * t);
      radius = Min(radius, 1e16);
      shrink = 2.0;
      if (std::fabs(delta) <= 1e-6 * prev) break;
    } else {
      radius /= shrink;
      shrink *= 2.0;
      if (radius < 1e-32) break;
    }
  }
  RLB_WARP_SYNC();
  for (int i = 0; i < 6; ++i)
    if (!(x[i] - x[i] == 0.0)) return false;
  for (int i = 0; i < 6; ++i) cam[i] = x[i];
  return true;
}

inline bool RefinePose(int k, const int* idx, const double* p2d, const double* p3d, double fx,
                       double fy, double* cam) {
  const SequentialLineariser lin = {k, idx, p2d, p3d, fx, fy};
  LmWorkspace w;
  return RefinePoseWith(lin, cam, &w);
}

// CalibratedAbsolutePoseEstimator::LeastSquares around the LM (.cc:170-212): column-major
// [R | c] -> (angle-axis, c) -> refine -> back. The pose is left untouched when the refinement
// is not usable.
template <class Lin>
RLB_HD inline void LeastSquaresWith(const Lin& lin, double* pose /* col-major 3x4 [R|c] */,
                                    LmWorkspace* w) {
  double* R = w->R;
  double* cam = w->cam;
  RLB_WARP_SYNC();
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) R[3 * r + c] = pose[r + 3 * c];
  RLB_WARP_SYNC();
  RotationToAngleAxis(R, cam);
  cam[3] = pose[9];
  cam[4] = pose[10];
  cam[5] = pose[11];
  RLB_WARP_SYNC();
  if (!RefinePoseWith(lin, cam, w)) return;
  RLB_WARP_SYNC();
  AngleAxisToRotation(cam, R);
  RLB_WARP_SYNC();
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) pose[r + 3 * c] = R[3 * r + c];
  pose[9] = cam[3];
  pose[10] = cam[4];
  pose[11] = cam[5];
  RLB_WARP_SYNC();
}

}  // namespace refine
}  // namespace calibrated_absolute_pose
}  // namespace ransac_lib

#if defined(RLB_FP_CONTRACT_PUSHED_POSE_REF)
#pragma GCC pop_options
#undef RLB_FP_CONTRACT_PUSHED_POSE_REF
#endif

#endif  // RANSACLIB_B200_EXAMPLES_POSE_REFINEMENT_H_
