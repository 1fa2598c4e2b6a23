// HybridLineEstimator — the reference's two-solver / two-data-type line example
// (examples/hybrid_line_estimator.{h,cc}): solver 0 takes two plain points, solver 1 one
// point with a normal. Both data types live in one device scorer; scoring applies the
// per-type thresholds and weights of HybridLORansacOptions.
#ifndef RANSACLIB_B200_EXAMPLES_HYBRID_LINE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_HYBRID_LINE_ESTIMATOR_H_

#include <cmath>
#include <vector>

#include <RansacLib/gpu_scorer.h>

#include "line_estimator.h"
#include "model_types.h"

namespace ransac_lib {

class HybridLineEstimator {
 public:
  // points: n0 x 2; points_with_normals: n1 x 4 (x, y, nx, ny) == column-major Matrix4Xd.
  HybridLineEstimator(int n0, const double* points, int n1, const double* points_with_normals,
                      const std::vector<double>& prior_probabilities, gpu::Context* ctx = nullptr)
      : own_points_(points, points + 2 * static_cast<size_t>(n0)),
        own_points_with_normals_(points_with_normals, points_with_normals + 4 * static_cast<size_t>(n1)),
        points_(own_points_.data()),
        num_points_(n0),
        points_with_normals_(own_points_with_normals_.data()),
        num_points_with_normals_(n1),
        prior_probabilities_(prior_probabilities) {
    scorer_.InitHybridLine(ctx ? *ctx : gpu::Context::ThreadDefault(), n0, points_, n1, points_with_normals_);
  }
  // The same without the host copy: both arrays must outlive the estimator.
  HybridLineEstimator(gpu::BorrowData, int n0, const double* points, int n1, const double* points_with_normals,
                      const std::vector<double>& prior_probabilities, gpu::Context* ctx = nullptr)
      : points_(points),
        num_points_(n0),
        points_with_normals_(points_with_normals),
        num_points_with_normals_(n1),
        prior_probabilities_(prior_probabilities) {
    scorer_.InitHybridLine(ctx ? *ctx : gpu::Context::ThreadDefault(), n0, points_, n1, points_with_normals_);
  }

  inline int num_minimal_solvers() const { return 2; }
  inline void min_sample_sizes(std::vector<std::vector<int>>* min_sample_sizes) const {
    min_sample_sizes->resize(2);
    (*min_sample_sizes)[0] = {2, 0};
    (*min_sample_sizes)[1] = {0, 1};
  }
  inline int num_data_types() const { return 2; }
  inline void num_data(std::vector<int>* num_data) const {
    num_data->resize(2);
    (*num_data)[0] = num_points_;
    (*num_data)[1] = num_points_with_normals_;
  }
  inline void solver_probabilities(std::vector<double>* p) const { *p = prior_probabilities_; }

  // Optional per-point sampling weights for HybridBiasedSampling (reference
  // hybrid_sampling.h:146-160 reads them through get_weights); default: all ones.
  void set_weights(const std::vector<double>& w0, const std::vector<double>& w1) {
    weights_ = {w0, w1};
  }
  void get_weights(std::vector<std::vector<double>>& weights) const {
    if (!weights_.empty()) {
      weights = weights_;
      return;
    }
    weights.assign(2, std::vector<double>());
    weights[0].assign(static_cast<size_t>(num_points_), 1.0);
    weights[1].assign(static_cast<size_t>(num_points_with_normals_), 1.0);
  }

  inline int MinimalSolver(const std::vector<std::vector<int>>& sample, const int solver_idx,
                           Lines2D* lines) const {
    lines->clear();
    Line2D l;
    if (solver_idx == 0) {  // hybrid_line_estimator.cc:111-130
      const std::vector<int>& s = sample[0];
      if (s.size() < 2u) return 0;
      if (!line_fit::ThroughTwoPoints(points_[2 * s[0]], points_[2 * s[0] + 1], points_[2 * s[1]],
                                      points_[2 * s[1] + 1], l.coeff))
        return 0;
    } else {  // hybrid_line_estimator.cc:132-152
      const std::vector<int>& s = sample[1];
      if (s.size() < 1u) return 0;
      const double* p = &points_with_normals_[4 * static_cast<size_t>(s[0])];
      double nx = p[2], ny = p[3];
      const double n2 = nx * nx + ny * ny;
      if (n2 > 0.0) {
        const double n = std::sqrt(n2);
        nx /= n;
        ny /= n;
      }
      const double c = -(nx * p[0] + ny * p[1]);
      const double nrm = std::sqrt(nx * nx + ny * ny);
      if (nrm == 0.0) return 0;
      l[0] = nx / nrm;
      l[1] = ny / nrm;
      l[2] = c / nrm;
    }
    lines->push_back(l);
    return 1;
  }

  // Host evaluation of one point (hybrid_line_estimator.cc:100-109); unused by the drivers.
  double EvaluateModelOnPoint(const Line2D& line, int t, int i) const {
    const double* p = (t == 0) ? &points_[2 * static_cast<size_t>(i)]
                               : &points_with_normals_[4 * static_cast<size_t>(i)];
    const double r = line[0] * p[0] + line[1] * p[1] + line[2] * 1.0;
    return r * r;
  }

  // hybrid_line_estimator.cc:58-97
  void LeastSquares(const std::vector<std::vector<int>>& sample, Line2D* line) const {
    const int k0 = static_cast<int>(sample[0].size());
    const int k1 = static_cast<int>(sample[1].size());
    if (k0 + k1 < 6) return;
    line_fit::FitCovariance(k0 + k1,
                            [&](int i, double* x, double* y) {
                              if (i < k0) {
                                *x = points_[2 * static_cast<size_t>(sample[0][i])];
                                *y = points_[2 * static_cast<size_t>(sample[0][i]) + 1];
                              } else {
                                *x = points_with_normals_[4 * static_cast<size_t>(sample[1][i - k0])];
                                *y = points_with_normals_[4 * static_cast<size_t>(sample[1][i - k0]) + 1];
                              }
                            },
                            line->coeff);
  }

  // ---- GPU opt-in ----
  const gpu::DeviceScorer& device_scorer() const { return scorer_; }
  gpu::DeviceScorer& device_scorer() { return scorer_; }
  // solver_idx -> device solver id (0: 2-point on type 0, 1: point+normal on type 1)
  int device_minimal_solver(int solver_idx) const { return solver_idx; }
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
  std::vector<double> own_points_, own_points_with_normals_;  // empty when the data is borrowed
  const double* points_;
  int num_points_;
  const double* points_with_normals_;
  int num_points_with_normals_;
  std::vector<double> prior_probabilities_;
  std::vector<std::vector<double>> weights_;
  gpu::DeviceScorer scorer_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_HYBRID_LINE_ESTIMATOR_H_
