// Arithmetic of the reference's two line solvers, shared by the host callbacks of LineEstimator /
// HybridLineEstimator and by the device-resident line LO-MSAC (csrc/line_lomsac_kernel.cu, compiled
// with -fmad=false): one source, so host and device run the same sequence of IEEE operations.
#ifndef RANSACLIB_B200_EXAMPLES_LINE_FIT_H_
#define RANSACLIB_B200_EXAMPLES_LINE_FIT_H_

#include <cmath>

#ifdef __CUDACC__
#define RLB_LF_HD __host__ __device__
#else
#define RLB_LF_HD
#endif
// Device build: let the loads of the next points issue ahead of the (sequential) sums.
#ifdef __CUDA_ARCH__
#define RLB_LF_UNROLL _Pragma("unroll 4")
#else
#define RLB_LF_UNROLL
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
#define RLB_FP_CONTRACT_PUSHED_LINE_FIT 1
#endif

namespace ransac_lib {

namespace line_fit {
// Unit eigenvector of the LARGEST eigenvalue of [[a,b],[b,d]]: what the reference reads
// from SelfAdjointEigenSolver::eigenvectors().col(1) (line_estimator.cc:97-100). Note the
// reference uses it as the line NORMAL although it is the direction of largest spread;
// that behaviour is kept (the refit then loses against the minimal model in
// UpdateBestModel). Eigen's sign convention is internal; squared residuals ignore it.
RLB_LF_HD inline bool LargestEigenvector2x2(double a, double b, double d, double* ex, double* ey) {
  const double half_diff = 0.5 * (a - d);
  const double radius = sqrt(half_diff * half_diff + b * b);
  const double lambda = 0.5 * (a + d) + radius;
  const double c0x = b, c0y = lambda - a;   // from row 0 of (A - lambda I)
  const double c1x = lambda - d, c1y = b;   // from row 1
  const double n0 = c0x * c0x + c0y * c0y, n1 = c1x * c1x + c1y * c1y;
  double vx = c0x, vy = c0y, nn = n0;
  if (n1 > n0) {
    vx = c1x;
    vy = c1y;
    nn = n1;
  }
  if (!(nn > 0.0)) {
    *ex = 0.0;
    *ey = 1.0;
    return (a == a) && (d == d);
  }
  const double inv = 1.0 / sqrt(nn);
  *ex = vx * inv;
  *ey = vy * inv;
  return true;
}

// Covariance line fit over points gathered by `get(i, &x, &y)` (line_estimator.cc:74-107,
// hybrid_line_estimator.cc:58-97).
template <class Getter>
RLB_LF_HD inline bool FitCovariance(int count, Getter get, double* line) {
  double mx = 0.0, my = 0.0;
  RLB_LF_UNROLL
  for (int i = 0; i < count; ++i) {
    double x, y;
    get(i, &x, &y);
    mx += x;
    my += y;
  }
  mx /= static_cast<double>(count);
  my /= static_cast<double>(count);
  double cxx = 0.0, cxy = 0.0, cyy = 0.0;
  RLB_LF_UNROLL
  for (int i = 0; i < count; ++i) {
    double x, y;
    get(i, &x, &y);
    const double dx = x - mx, dy = y - my;
    cxx += dx * dx;
    cxy += dx * dy;
    cyy += dy * dy;
  }
  const double denom = static_cast<double>(count - 1);
  cxx /= denom;
  cxy /= denom;
  cyy /= denom;
  double ex, ey;
  if (!LargestEigenvector2x2(cxx, cxy, cyy, &ex, &ey)) return false;
  line[0] = ex;
  line[1] = ey;
  line[2] = -(ex * mx + ey * my);
  return true;
}

// (x1,y1,1) x (x2,y2,1), normal scaled to unit length (line_estimator.cc:53-72).
RLB_LF_HD inline bool ThroughTwoPoints(double x1, double y1, double x2, double y2, double* l) {
  const double a = y1 * 1.0 - 1.0 * y2;
  const double b = 1.0 * x2 - x1 * 1.0;
  const double c = x1 * y2 - y1 * x2;
  const double nrm = sqrt(a * a + b * b);
  if (nrm == 0.0) return false;
  l[0] = a / nrm;
  l[1] = b / nrm;
  l[2] = c / nrm;
  return true;
}
}  // namespace line_fit

}  // namespace ransac_lib

#if defined(RLB_FP_CONTRACT_PUSHED_LINE_FIT)
#pragma GCC pop_options
#undef RLB_FP_CONTRACT_PUSHED_LINE_FIT
#endif

#endif  // RANSACLIB_B200_EXAMPLES_LINE_FIT_H_
