// LineEstimator — the reference's 2-point line solver (examples/line_estimator.{h,cc})
// with its data set resident on the GPU. Host callbacks (NonMinimalSolver,
// LeastSquares, a host MinimalSolver for LO) keep the reference's arithmetic; scoring,
// inlier collection and the batched 2-point solver run on the device.
#ifndef RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_

#include <cmath>
#include <vector>

#include <RansacLib/gpu_scorer.h>

#include "model_types.h"

namespace ransac_lib {

namespace line_fit {
// Unit eigenvector of the LARGEST eigenvalue of [[a,b],[b,d]]: what the reference reads
// from SelfAdjointEigenSolver::eigenvectors().col(1) (line_estimator.cc:97-100). Note the
// reference uses it as the line NORMAL although it is the direction of largest spread;
// that behaviour is kept (the refit then loses against the minimal model in
// UpdateBestModel). Eigen's sign convention is internal; squared residuals ignore it.
inline bool LargestEigenvector2x2(double a, double b, double d, double* ex, double* ey) {
  const double half_diff = 0.5 * (a - d);
  const double radius = std::sqrt(half_diff * half_diff + b * b);
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
  const double inv = 1.0 / std::sqrt(nn);
  *ex = vx * inv;
  *ey = vy * inv;
  return true;
}

// Covariance line fit over points gathered by `get(i, &x, &y)` (line_estimator.cc:74-107,
// hybrid_line_estimator.cc:58-97).
template <class Getter>
inline bool FitCovariance(int count, Getter get, Line2D* line) {
  double mx = 0.0, my = 0.0;
  for (int i = 0; i < count; ++i) {
    double x, y;
    get(i, &x, &y);
    mx += x;
    my += y;
  }
  mx /= static_cast<double>(count);
  my /= static_cast<double>(count);
  double cxx = 0.0, cxy = 0.0, cyy = 0.0;
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
  (*line)[0] = ex;
  (*line)[1] = ey;
  (*line)[2] = -(ex * mx + ey * my);
  return true;
}

// (x1,y1,1) x (x2,y2,1), normal scaled to unit length (line_estimator.cc:53-72).
inline bool ThroughTwoPoints(double x1, double y1, double x2, double y2, Line2D* l) {
  const double a = y1 * 1.0 - 1.0 * y2;
  const double b = 1.0 * x2 - x1 * 1.0;
  const double c = x1 * y2 - y1 * x2;
  const double nrm = std::sqrt(a * a + b * b);
  if (nrm == 0.0) return false;
  (*l)[0] = a / nrm;
  (*l)[1] = b / nrm;
  (*l)[2] = c / nrm;
  return true;
}
}  // namespace line_fit

class LineEstimator {
 public:
  // data: n x 2 doubles (x0,y0,x1,y1,...) == column-major Eigen::Matrix2Xd. Copied to the
  // host member and uploaded once (line_estimator.cc:47-50).
  LineEstimator(int n, const double* data, gpu::Context* ctx = nullptr)
      : data_(data, data + 2 * static_cast<size_t>(n)), num_data_(n) {
    scorer_.InitLine(ctx ? *ctx : gpu::Context::ThreadDefault(), n, data_.data());
  }

  inline int min_sample_size() const { return 2; }
  inline int non_minimal_sample_size() const { return 6; }
  inline int num_data() const { return num_data_; }

  int MinimalSolver(const std::vector<int>& sample, Lines2D* lines) const {
    lines->clear();
    if (sample.size() < 2u) return 0;
    Line2D l;
    if (!line_fit::ThroughTwoPoints(x(sample[0]), y(sample[0]), x(sample[1]), y(sample[1]), &l)) return 0;
    lines->push_back(l);
    return 1;
  }

  int NonMinimalSolver(const std::vector<int>& sample, Line2D* line) const {
    if (sample.size() < 6u) return 0;
    return line_fit::FitCovariance(static_cast<int>(sample.size()),
                                   [&](int i, double* px, double* py) {
                                     *px = x(sample[i]);
                                     *py = y(sample[i]);
                                   },
                                   line)
               ? 1
               : 0;
  }

  // Host evaluation of ONE point (kept for API completeness; the drivers never call it).
  double EvaluateModelOnPoint(const Line2D& line, int i) const {
    const double r = line[0] * x(i) + line[1] * y(i) + line[2] * 1.0;
    return r * r;
  }

  inline void LeastSquares(const std::vector<int>& sample, Line2D* line) const {
    NonMinimalSolver(sample, line);
  }

  // ---- GPU opt-in (RansacLib/gpu_scorer.h) ----
  const gpu::DeviceScorer& device_scorer() const { return scorer_; }
  gpu::DeviceScorer& device_scorer() { return scorer_; }
  int device_minimal_solver() const { return 0; }
  static void PackModel(const Line2D& l, double* out) {
    out[0] = l[0];
    out[1] = l[1];
    out[2] = l[2];
  }
  static void UnpackModel(const double* in, Line2D* l) {
    (*l)[0] = in[0];
    (*l)[1] = in[1];
    (*l)[2] = in[2];
  }

 protected:
  double x(int i) const { return data_[2 * static_cast<size_t>(i)]; }
  double y(int i) const { return data_[2 * static_cast<size_t>(i) + 1]; }
  std::vector<double> data_;
  int num_data_;
  gpu::DeviceScorer scorer_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_
