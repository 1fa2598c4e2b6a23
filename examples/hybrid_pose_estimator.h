// HybridPoseEstimator - a HybridSolver (reference README.md:164-208, RansacLib/hybrid_ransac.h:126-137) for camera
// pose from two kinds of matches, the use of hybrid RANSAC the reference names but ships no example of:
//   data type 0: 2D-3D matches (u, v | X, Y, Z), squared reprojection error as CalibratedAbsolutePoseEstimator
//                (calibrated_absolute_pose_estimator.cc:154-167);
//   data type 1: 3D-3D matches (a point measured in the camera frame, e.g. by a depth sensor | X, Y, Z), squared
//                distance between R (X - c) and the measurement.
// Minimal solvers: 0 = P3P + 4th point on type 0 ({4, 0}), 1 = three 3D-3D matches ({0, 3}: the orthonormal frames of
// the two triangles give R, one vertex gives c). LeastSquares refines on the sample's 2D-3D matches (the LM of
// examples/pose_refinement.h) when it holds six or more.
// Both residuals reach the GPU as SOURCE through rlb_scorer_create_custom_hybrid (model = 12 doubles, rows of 5 and 6
// doubles), i.e. this is the path of any user-defined two-data-type solver; nothing in the library knows this model.
// oracle/solvers_oracle.h: HybridPoseSolver is the test-side counterpart (same operations in the same order), driven
// by the reference's own hybrid_ransac.h in oracle/_ref.
#ifndef RANSACLIB_B200_EXAMPLES_HYBRID_POSE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_HYBRID_POSE_ESTIMATOR_H_

#include <cmath>
#include <cstdio>
#include <limits>
#include <string>
#include <vector>

#include <RansacLib/gpu_scorer.h>

#include "model_types.h"
#include "p3p.h"
#include "pose_refinement.h"

namespace ransac_lib {
namespace calibrated_absolute_pose {

class HybridPoseEstimator {
 public:
  // points2D: n0 x 2 (centred on the principal point), points3D: n0 x 3; cam1 / world1: n1 x 3 each.
  // sq_thr_p3p: the 4th-point test of the P3P solver (the type-0 inlier threshold).
  HybridPoseEstimator(double f_x, double f_y, double sq_thr_p3p, int n0, const double* points2D,
                      const double* points3D, int n1, const double* cam1, const double* world1,
                      const std::vector<double>& prior_probabilities, gpu::Context* ctx = nullptr)
      : focal_x_(f_x),
        focal_y_(f_y),
        sq_thr_p3p_(sq_thr_p3p),
        points2D_(points2D, points2D + 2 * static_cast<size_t>(n0)),
        points3D_(points3D, points3D + 3 * static_cast<size_t>(n0)),
        cam1_(cam1, cam1 + 3 * static_cast<size_t>(n1)),
        world1_(world1, world1 + 3 * static_cast<size_t>(n1)),
        n0_(n0),
        n1_(n1),
        prior_probabilities_(prior_probabilities) {
    // rows as the functor reads them: type 0 (u, v, X, Y, Z), type 1 (xc, yc, zc, X, Y, Z)
    std::vector<double> rows0(5 * static_cast<size_t>(n0)), rows1(6 * static_cast<size_t>(n1));
    for (int i = 0; i < n0; ++i) {
      rows0[5 * static_cast<size_t>(i)] = points2D_[2 * static_cast<size_t>(i)];
      rows0[5 * static_cast<size_t>(i) + 1] = points2D_[2 * static_cast<size_t>(i) + 1];
      for (int c = 0; c < 3; ++c) rows0[5 * static_cast<size_t>(i) + 2 + c] = points3D_[3 * static_cast<size_t>(i) + c];
    }
    for (int i = 0; i < n1; ++i)
      for (int c = 0; c < 3; ++c) {
        rows1[6 * static_cast<size_t>(i) + c] = cam1_[3 * static_cast<size_t>(i) + c];
        rows1[6 * static_cast<size_t>(i) + 3 + c] = world1_[3 * static_cast<size_t>(i) + c];
      }
    // EvaluateModelOnPoint below, statement by statement (NVRTC compiles it with -fmad=false); the focal lengths
    // travel as 17-digit literals, which round-trip a double exactly.
    char src[2048];
    std::snprintf(src, sizeof(src),
                  "__device__ double rlb_eval_t(const double* m, int type, const double* p) {\n"
                  "  const double* X = (type == 0) ? p + 2 : p + 3;\n"
                  "  const double d0 = X[0] - m[9];\n"
                  "  const double d1 = X[1] - m[10];\n"
                  "  const double d2 = X[2] - m[11];\n"
                  "  const double pz = m[2] * d0 + m[5] * d1 + m[8] * d2;\n"
                  "  const double px = m[0] * d0 + m[3] * d1 + m[6] * d2;\n"
                  "  const double py = m[1] * d0 + m[4] * d1 + m[7] * d2;\n"
                  "  if (type == 0) {\n"
                  "    if (pz < 0.0) return 1.7976931348623157e308;\n"
                  "    const double du = (px / pz) * %.17g - p[0];\n"
                  "    const double dv = (py / pz) * %.17g - p[1];\n"
                  "    return du * du + dv * dv;\n"
                  "  }\n"
                  "  const double ex = px - p[0];\n"
                  "  const double ey = py - p[1];\n"
                  "  const double ez = pz - p[2];\n"
                  "  return ex * ex + ey * ey + ez * ez;\n"
                  "}\n",
                  focal_x_, focal_y_);
    scorer_.InitCustomHybrid(ctx ? *ctx : gpu::Context::ThreadDefault(), n0, 5, rows0.data(), n1, 6, rows1.data(), 12,
                             src);
  }

  inline int num_minimal_solvers() const { return 2; }
  inline void min_sample_sizes(std::vector<std::vector<int>>* min_sample_sizes) const {
    min_sample_sizes->resize(2);
    (*min_sample_sizes)[0] = {4, 0};
    (*min_sample_sizes)[1] = {0, 3};
  }
  inline int num_data_types() const { return 2; }
  inline void num_data(std::vector<int>* num_data) const {
    num_data->resize(2);
    (*num_data)[0] = n0_;
    (*num_data)[1] = n1_;
  }
  inline void solver_probabilities(std::vector<double>* p) const { *p = prior_probabilities_; }

  inline int MinimalSolver(const std::vector<std::vector<int>>& sample, const int solver_idx,
                           CameraPoses* poses) const {
    poses->clear();
    CameraPose P;
    if (solver_idx == 0) {
      const std::vector<int>& s = sample[0];
      if (s.size() < 4u) return 0;
      double y[9], x[9];
      for (int j = 0; j < 3; ++j) {
        const size_t i = static_cast<size_t>(s[j]);
        // PixelsToViewingRays, calibrated_absolute_pose_estimator.cc:214-227
        const double rx = points2D_[2 * i] / focal_x_;
        const double ry = points2D_[2 * i + 1] / focal_y_;
        const double rz = 1.0;
        const double nrm = std::sqrt(rx * rx + ry * ry + rz * rz);
        y[3 * j] = rx / nrm;
        y[3 * j + 1] = ry / nrm;
        y[3 * j + 2] = rz / nrm;
        for (int c = 0; c < 3; ++c) x[3 * j + c] = points3D_[3 * i + c];
      }
      const size_t i4 = static_cast<size_t>(s[3]);
      const double X4[3] = {points3D_[3 * i4], points3D_[3 * i4 + 1], points3D_[3 * i4 + 2]};
      const double uv4[2] = {points2D_[2 * i4], points2D_[2 * i4 + 1]};
      if (!rlb::p3pdev::p3p_pick(y, x, X4, uv4, focal_x_, focal_y_, sq_thr_p3p_, P.m)) return 0;
    } else {
      const std::vector<int>& s = sample[1];
      if (s.size() < 3u) return 0;
      if (!FromTriangles(&world1_[3 * static_cast<size_t>(s[0])], &world1_[3 * static_cast<size_t>(s[1])],
                         &world1_[3 * static_cast<size_t>(s[2])], &cam1_[3 * static_cast<size_t>(s[0])],
                         &cam1_[3 * static_cast<size_t>(s[1])], &cam1_[3 * static_cast<size_t>(s[2])], P.m))
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
  // m: R column-major, then c (CameraPose).
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

  // Host evaluation of one match of either type; the drivers never call it (the GPU functor above is its copy).
  double EvaluateModelOnPoint(const CameraPose& P, int t, int i) const {
    const double* X = (t == 0) ? &points3D_[3 * static_cast<size_t>(i)] : &world1_[3 * static_cast<size_t>(i)];
    const double d0 = X[0] - P.m[9];
    const double d1 = X[1] - P.m[10];
    const double d2 = X[2] - P.m[11];
    const double pz = P.m[2] * d0 + P.m[5] * d1 + P.m[8] * d2;
    const double px = P.m[0] * d0 + P.m[3] * d1 + P.m[6] * d2;
    const double py = P.m[1] * d0 + P.m[4] * d1 + P.m[7] * d2;
    if (t == 0) {
      if (pz < 0.0) return std::numeric_limits<double>::max();
      const double du = (px / pz) * focal_x_ - points2D_[2 * static_cast<size_t>(i)];
      const double dv = (py / pz) * focal_y_ - points2D_[2 * static_cast<size_t>(i) + 1];
      return du * du + dv * dv;
    }
    const double ex = px - cam1_[3 * static_cast<size_t>(i)];
    const double ey = py - cam1_[3 * static_cast<size_t>(i) + 1];
    const double ez = pz - cam1_[3 * static_cast<size_t>(i) + 2];
    return ex * ex + ey * ey + ez * ez;
  }

  // Non-linear refinement on the sample's 2D-3D matches (six or more; otherwise the model stays).
  void LeastSquares(const std::vector<std::vector<int>>& sample, CameraPose* pose) const {
    if (sample[0].size() < 6u) return;
    const refine::SequentialLineariser lin = {static_cast<int>(sample[0].size()), sample[0].data(), points2D_.data(),
                                              points3D_.data(), focal_x_, focal_y_};
    refine::LmWorkspace w;
    refine::LeastSquaresWith(lin, pose->m, &w);
  }

  // ---- GPU opt-in ----
  const gpu::DeviceScorer& device_scorer() const { return scorer_; }
  gpu::DeviceScorer& device_scorer() { return scorer_; }
  static void PackModel(const CameraPose& P, double* out) {
    for (int i = 0; i < 12; ++i) out[i] = P.m[i];
  }
  static void UnpackModel(const double* in, CameraPose* P) {
    for (int i = 0; i < 12; ++i) P->m[i] = in[i];
  }

 protected:
  double focal_x_, focal_y_, sq_thr_p3p_;
  std::vector<double> points2D_, points3D_, cam1_, world1_;
  int n0_, n1_;
  std::vector<double> prior_probabilities_;
  gpu::DeviceScorer scorer_;
};

}  // namespace calibrated_absolute_pose
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_HYBRID_POSE_ESTIMATOR_H_
