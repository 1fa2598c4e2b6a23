// CalibratedAbsolutePoseEstimator — the reference's P3P + 4th-point absolute pose solver
// (examples/calibrated_absolute_pose_estimator.{h,cc}) with the 2D-3D matches resident on
// the GPU. EvaluateModelOnPoint over all matches (scoring, inliers) and the batched P3P
// minimal solver run on the device; the non-linear refinement stays on the host.
#ifndef RANSACLIB_B200_EXAMPLES_CALIBRATED_ABSOLUTE_POSE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_CALIBRATED_ABSOLUTE_POSE_ESTIMATOR_H_

#include <cmath>
#include <cstdint>
#include <limits>
#include <stdexcept>
#include <vector>

#include <RansacLib/gpu_scorer.h>

#include "model_types.h"
#include "p3p.h"
#include "pose_refinement.h"

namespace ransac_lib {
namespace calibrated_absolute_pose {

class CalibratedAbsolutePoseEstimator {
 public:
  // points2D: n x 2 (centred on the principal point, undistorted); points3D: n x 3.
  // The viewing rays of the reference's constructor (.cc:99-109) are derived on the device
  // exactly as PixelsToViewingRays does (.cc:214-227).
  CalibratedAbsolutePoseEstimator(double f_x, double f_y, double squared_inlier_threshold, int n,
                                  const double* points2D, const double* points3D,
                                  gpu::Context* ctx = nullptr)
      : focal_x_(f_x),
        focal_y_(f_y),
        squared_inlier_threshold_(squared_inlier_threshold),
        own2D_(points2D, points2D + 2 * static_cast<size_t>(n)),
        own3D_(points3D, points3D + 3 * static_cast<size_t>(n)),
        points2D_(own2D_.data()),
        points3D_(own3D_.data()),
        num_data_(n) {
    InitScorer(ctx);
  }
  // The same without the host copy: both arrays must outlive the estimator.
  CalibratedAbsolutePoseEstimator(gpu::BorrowData, double f_x, double f_y, double squared_inlier_threshold, int n,
                                  const double* points2D, const double* points3D, gpu::Context* ctx = nullptr)
      : focal_x_(f_x),
        focal_y_(f_y),
        squared_inlier_threshold_(squared_inlier_threshold),
        points2D_(points2D),
        points3D_(points3D),
        num_data_(n) {
    InitScorer(ctx);
  }

  inline int min_sample_size() const { return 4; }          // reference .h:88
  inline int non_minimal_sample_size() const { return 6; }  // reference .h:90
  inline int num_data() const { return num_data_; }

  // One sample on the host (.cc:111-137): the same P3P as the device solver (examples/p3p.h, bit-identical
  // under -ffp-contract=off), without a launch - NonMinimalSolver calls this once per LO step. The batched
  // driver solves its hypothesis batches on the device.
  int MinimalSolver(const std::vector<int>& sample, CameraPoses* poses) const {
    poses->clear();
    if (sample.size() < 4u) return 0;
    for (int j = 0; j < 4; ++j)
      if (sample[j] < 0 || sample[j] >= num_data_)
        throw std::runtime_error("ransaclib_b200: MinimalSolver: sample index out of range");
    double y[9], x[9];
    for (int j = 0; j < 3; ++j) {
      const size_t i = static_cast<size_t>(sample[j]);
      // PixelsToViewingRays, .cc:214-227 (same operations as csrc/solver_kernels.cu: rays_kernel)
      const double rx = points2D_[2 * i] / focal_x_;
      const double ry = points2D_[2 * i + 1] / focal_y_;
      const double rz = 1.0;
      const double nrm = std::sqrt(rx * rx + ry * ry + rz * rz);
      y[3 * j] = rx / nrm;
      y[3 * j + 1] = ry / nrm;
      y[3 * j + 2] = rz / nrm;
      for (int c = 0; c < 3; ++c) x[3 * j + c] = points3D_[3 * i + c];
    }
    const size_t i4 = static_cast<size_t>(sample[3]);
    const double X4[3] = {points3D_[3 * i4], points3D_[3 * i4 + 1], points3D_[3 * i4 + 2]};
    const double uv4[2] = {points2D_[2 * i4], points2D_[2 * i4 + 1]};
    CameraPose P;
    if (rlb::p3pdev::p3p_pick(y, x, X4, uv4, focal_x_, focal_y_, squared_inlier_threshold_, P.m))
      poses->push_back(P);
    return static_cast<int>(poses->size());
  }

  // .cc:141-151
  int NonMinimalSolver(const std::vector<int>& sample, CameraPose* pose) const {
    CameraPoses poses;
    if (MinimalSolver(sample, &poses) == 1) {
      *pose = poses[0];
      LeastSquares(sample, pose);
      return 1;
    }
    return 0;
  }

  // Host evaluation of ONE match (.cc:154-167); the drivers never call it.
  double EvaluateModelOnPoint(const CameraPose& P, int i) const {
    const double d0 = points3D_[3 * i] - P.m[9];
    const double d1 = points3D_[3 * i + 1] - P.m[10];
    const double d2 = points3D_[3 * i + 2] - P.m[11];
    const double pz = P.m[2] * d0 + P.m[5] * d1 + P.m[8] * d2;
    if (pz < 0.0) return std::numeric_limits<double>::max();
    const double px = P.m[0] * d0 + P.m[3] * d1 + P.m[6] * d2;
    const double py = P.m[1] * d0 + P.m[4] * d1 + P.m[7] * d2;
    const double du = (px / pz) * focal_x_ - points2D_[2 * i];
    const double dv = (py / pz) * focal_y_ - points2D_[2 * i + 1];
    return du * du + dv * dv;
  }

  // .cc:170-212
  void LeastSquares(const std::vector<int>& sample, CameraPose* pose) const {
    const refine::SequentialLineariser lin = {static_cast<int>(sample.size()), sample.data(), points2D_,
                                              points3D_, focal_x_, focal_y_};
    refine::LmWorkspace w;
    refine::LeastSquaresWith(lin, pose->m, &w);
  }

  // ---- GPU opt-in ----
  const gpu::DeviceScorer& device_scorer() const { return scorer_; }
  gpu::DeviceScorer& device_scorer() { return scorer_; }
  int device_minimal_solver() const { return 0; }
  double device_minimal_solver_threshold() const { return squared_inlier_threshold_; }
  static void PackModel(const CameraPose& P, double* out) {
    for (int i = 0; i < 12; ++i) out[i] = P.m[i];
  }
  static void UnpackModel(const double* in, CameraPose* P) {
    for (int i = 0; i < 12; ++i) P->m[i] = in[i];
  }

  double focal_x() const { return focal_x_; }
  double focal_y() const { return focal_y_; }

 protected:
  // One GPU (the caller's context or the thread's default one), or with RLB_NUM_GPUS=G in the environment the
  // data set replicated on G GPUs and large hypothesis batches sharded over them (gpu::DeviceScorer::InitPoseMulti).
  void InitScorer(gpu::Context* ctx) {
    const int g = gpu::NumGpusFromEnv();
    if (g >= 1 && ctx == nullptr) {
      scorer_.InitPoseMulti(g, num_data_, points2D_, points3D_, focal_x_, focal_y_);
    } else {
      scorer_.InitPose(ctx ? *ctx : gpu::Context::ThreadDefault(), num_data_, points2D_, points3D_, focal_x_,
                       focal_y_);
    }
  }

  double focal_x_, focal_y_, squared_inlier_threshold_;
  std::vector<double> own2D_, own3D_;  // empty when the data is borrowed
  const double* points2D_;
  const double* points3D_;
  int num_data_;
  gpu::DeviceScorer scorer_;
};

}  // namespace calibrated_absolute_pose
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_CALIBRATED_ABSOLUTE_POSE_ESTIMATOR_H_
