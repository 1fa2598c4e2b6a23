// LineEstimator — the reference's 2-point line solver (examples/line_estimator.{h,cc})
// with its data set resident on the GPU. Host callbacks (NonMinimalSolver,
// LeastSquares, a host MinimalSolver for LO) keep the reference's arithmetic; scoring,
// inlier collection and the batched 2-point solver run on the device.
#ifndef RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_

#include <cmath>
#include <vector>

#include <RansacLib/gpu_scorer.h>

#include "line_fit.h"
#include "model_types.h"

namespace ransac_lib {

class LineEstimator {
 public:
  // data: n x 2 doubles (x0,y0,x1,y1,...) == column-major Eigen::Matrix2Xd. Copied to the
  // host member and uploaded once (line_estimator.cc:47-50).
  LineEstimator(int n, const double* data, gpu::Context* ctx = nullptr)
      : own_(data, data + 2 * static_cast<size_t>(n)), data_(own_.data()), num_data_(n) {
    scorer_.InitLine(ctx ? *ctx : gpu::Context::ThreadDefault(), n, data_);
  }
  // The same without the host copy: the array must outlive the estimator.
  LineEstimator(gpu::BorrowData, int n, const double* data, gpu::Context* ctx = nullptr)
      : data_(data), num_data_(n) {
    scorer_.InitLine(ctx ? *ctx : gpu::Context::ThreadDefault(), n, data_);
  }

  static constexpr int kDeviceLomsacMaxPoints = 24576;  // measured crossover ~30k (tools/line_crossover.py)

  inline int min_sample_size() const { return 2; }
  inline int non_minimal_sample_size() const { return 6; }
  inline int num_data() const { return num_data_; }

  int MinimalSolver(const std::vector<int>& sample, Lines2D* lines) const {
    lines->clear();
    if (sample.size() < 2u) return 0;
    Line2D l;
    if (!line_fit::ThroughTwoPoints(x(sample[0]), y(sample[0]), x(sample[1]), y(sample[1]), l.coeff)) return 0;
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
                                   line->coeff)
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

  // ---- device-resident driver (csrc/line_lomsac_kernel.cu) ----
  // The whole of LocallyOptimizedMSAC::EstimateModel for this solver in one launch: sampling,
  // 2-point solver, scoring, local optimisation with the covariance fit above, same run as the
  // host-driven loop. false = options outside what the kernel supports (shuffle targets <= 32).
  bool device_lomsac(const rlb_lomsac_options& opt, uint32_t seed, double* packed_model,
                     rlb_lomsac_result* result, std::vector<int>* inliers) const {
    if (opt.min_sample_mult < 1 || opt.min_sample_mult * 2 > 32 || opt.non_min_sample_mult < 0 ||
        opt.non_min_sample_mult * 2 > 32 || opt.lo_steps < 0 || opt.lsq_iters < 0)
      return false;
    // One CTA = one SM works on the problem: beyond this size a pass over the points takes longer there
    // than a launch of the whole-GPU pass costs, and the host-driven loop is the faster driver.
    if (num_data_ > kDeviceLomsacMaxPoints) return false;
    inliers->resize(static_cast<size_t>(num_data_ > 0 ? num_data_ : 1));
    scorer_.LineLomsac(opt, seed, packed_model, result, reinterpret_cast<int32_t*>(inliers->data()));
    inliers->resize(static_cast<size_t>(result->num_inliers));
    return true;
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
  std::vector<double> own_;  // empty when the data is borrowed
  const double* data_;
  int num_data_;
  gpu::DeviceScorer scorer_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_LINE_ESTIMATOR_H_
