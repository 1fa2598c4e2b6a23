// CircleEstimator — a solver that is NOT one of the reference's examples, written against the
// solver concept of the reference's README.md:105-154 to show how ANY residual opts in to the
// GPU scorer: the body of EvaluateModelOnPoint is handed over as CUDA source and compiled at
// construction (gpu::DeviceScorer::InitCustom -> rlb_scorer_create_custom, NVRTC for sm_100a).
// MinimalSolver / NonMinimalSolver / LeastSquares stay host callbacks; every pass over the data
// (ScoreModel, GetInliers, ransac.h:295-336) runs on the device.
#ifndef RANSACLIB_B200_EXAMPLES_CIRCLE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_CIRCLE_ESTIMATOR_H_

#include <cmath>
#include <vector>

#include <RansacLib/gpu_scorer.h>

namespace ransac_lib {

// (cx, cy, r)
struct Circle2D {
  double p[3];
  double& operator[](int i) { return p[i]; }
  const double& operator[](int i) const { return p[i]; }
};
typedef std::vector<Circle2D> Circles2D;

class CircleEstimator {
 public:
  // The device twin of EvaluateModelOnPoint below: same expression, same order of operations.
  static const char* DeviceEvalSource() {
    return R"(
__device__ double rlb_eval(const double* m, const double* p) {   // m = (cx, cy, r), p = (x, y)
  const double dx = p[0] - m[0], dy = p[1] - m[1];
  const double d = sqrt(dx * dx + dy * dy) - m[2];
  return d * d;
})";
  }

  // data: n x 2 doubles (x0,y0,x1,y1,...), copied and uploaded once.
  CircleEstimator(int n, const double* data, gpu::Context* ctx = nullptr)
      : pts_(data, data + 2 * static_cast<size_t>(n)), n_(n) {
    scorer_.InitCustom(ctx ? *ctx : gpu::Context::ThreadDefault(), n, /*dims=*/2, pts_.data(),
                       /*model_doubles=*/3, DeviceEvalSource());
    scorer_.set_precision(RLB_F64);  // user functors are compiled for f64 only
  }

  inline int min_sample_size() const { return 3; }
  inline int non_minimal_sample_size() const { return 6; }
  inline int num_data() const { return n_; }

  // Circumscribed circle: the centre solves the two perpendicular-bisector equations
  //   2 (q - p) . c = |q|^2 - |p|^2,   2 (s - p) . c = |s|^2 - |p|^2.
  int MinimalSolver(const std::vector<int>& sample, Circles2D* circles) const {
    circles->clear();
    if (sample.size() < 3u) return 0;
    const double px = x(sample[0]), py = y(sample[0]);
    const double qx = x(sample[1]), qy = y(sample[1]);
    const double sx = x(sample[2]), sy = y(sample[2]);
    const double ux = qx - px, uy = qy - py;
    const double vx = sx - px, vy = sy - py;
    const double rhs_u = ux * (px + qx) + uy * (py + qy);
    const double rhs_v = vx * (px + sx) + vy * (py + sy);
    const double det2 = 2.0 * (ux * (sy - qy) - uy * (sx - qx));
    if (det2 == 0.0) return 0;  // collinear
    Circle2D c;
    c[0] = (vy * rhs_u - uy * rhs_v) / det2;
    c[1] = (ux * rhs_v - vx * rhs_u) / det2;
    const double ex = px - c[0], ey = py - c[1];
    c[2] = std::sqrt(ex * ex + ey * ey);
    circles->push_back(c);
    return 1;
  }

  // Algebraic (Kasa) fit: x^2 + y^2 + A x + B y + C = 0 in the least-squares sense, 3 x 3 normal
  // equations solved by Cramer's rule.
  int NonMinimalSolver(const std::vector<int>& sample, Circle2D* circle) const {
    if (sample.size() < 6u) return 0;
    double Sx = 0, Sy = 0, Sxx = 0, Syy = 0, Sxy = 0, Sxz = 0, Syz = 0, Sz = 0;
    const double m = static_cast<double>(sample.size());
    for (size_t k = 0; k < sample.size(); ++k) {
      const double xi = x(sample[k]), yi = y(sample[k]);
      const double zi = xi * xi + yi * yi;
      Sx += xi;
      Sy += yi;
      Sxx += xi * xi;
      Syy += yi * yi;
      Sxy += xi * yi;
      Sxz += xi * zi;
      Syz += yi * zi;
      Sz += zi;
    }
    // | Sxx Sxy Sx | |A|     |Sxz|
    // | Sxy Syy Sy | |B| = - |Syz|
    // | Sx  Sy  m  | |C|     |Sz |
    const double c00 = Syy * m - Sy * Sy, c01 = Sxy * m - Sy * Sx, c02 = Sxy * Sy - Syy * Sx;
    const double det = Sxx * c00 - Sxy * c01 + Sx * c02;
    if (det == 0.0 || !(det == det)) return 0;
    const double b0 = -Sxz, b1 = -Syz, b2 = -Sz;
    const double A = (b0 * c00 - Sxy * (b1 * m - Sy * b2) + Sx * (b1 * Sy - Syy * b2)) / det;
    const double B = (Sxx * (b1 * m - b2 * Sy) - b0 * c01 + Sx * (Sxy * b2 - b1 * Sx)) / det;
    const double Cc = (Sxx * (Syy * b2 - Sy * b1) - Sxy * (Sxy * b2 - b1 * Sx) + b0 * c02) / det;
    const double cx = -0.5 * A, cy = -0.5 * B;
    const double rr = cx * cx + cy * cy - Cc;
    if (!(rr > 0.0)) return 0;
    (*circle)[0] = cx;
    (*circle)[1] = cy;
    (*circle)[2] = std::sqrt(rr);
    return 1;
  }

  // Host evaluation of ONE point (API completeness; the drivers never call it).
  double EvaluateModelOnPoint(const Circle2D& c, int i) const {
    const double dx = x(i) - c[0], dy = y(i) - c[1];
    const double d = std::sqrt(dx * dx + dy * dy) - c[2];
    return d * d;
  }

  inline void LeastSquares(const std::vector<int>& sample, Circle2D* circle) const {
    NonMinimalSolver(sample, circle);
  }

  // ---- GPU opt-in (RansacLib/gpu_scorer.h); no device_minimal_solver(): MinimalSolver runs on the host ----
  const gpu::DeviceScorer& device_scorer() const { return scorer_; }
  gpu::DeviceScorer& device_scorer() { return scorer_; }
  static void PackModel(const Circle2D& c, double* out) {
    for (int i = 0; i < 3; ++i) out[i] = c[i];
  }
  static void UnpackModel(const double* in, Circle2D* c) {
    for (int i = 0; i < 3; ++i) (*c)[i] = in[i];
  }

 private:
  double x(int i) const { return pts_[2 * static_cast<size_t>(i)]; }
  double y(int i) const { return pts_[2 * static_cast<size_t>(i) + 1]; }
  std::vector<double> pts_;
  int n_;
  gpu::DeviceScorer scorer_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_CIRCLE_ESTIMATOR_H_
