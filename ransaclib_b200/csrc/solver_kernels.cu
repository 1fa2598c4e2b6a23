// Batched minimal solvers, one sample per thread, f64. Compiled with -fmad=false
// (no FMA contraction) and using only + - * / sqrt frexp ldexp, so every
// hypothesis is bit-identical to a strict-IEEE host evaluation of the same
// formulas: the sequential replay of the reference's control flow then sees
// exactly the hypotheses a CPU run would.
//
// Replaces (reference): examples/line_estimator.cc:53-72,
// hybrid_line_estimator.cc:111-152, calibrated_absolute_pose_estimator.cc:111-137
// (+ pose_lib::p3p, external to the reference: the published Lambda Twist
// algorithm, Persson & Nordberg ECCV 2018, with a fixed solution order:
// '+' branch of the degenerate conic first, then ascending tau), and
// PixelsToViewingRays, .cc:214-227.
#include <cfloat>

#include "../../examples/p3p.h"
#include "rlb_internal.cuh"

namespace rlb {

namespace {

using namespace p3pdev;

__global__ void __launch_bounds__(128)
    solve_p3p_kernel(const DeviceData d, int k, const int32_t* __restrict__ samples, double sq_thr,
                     double* __restrict__ models, uint8_t* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  int id[4];
  for (int j = 0; j < 4; ++j) id[j] = samples[4 * static_cast<size_t>(i) + j];
  double y[9], x[9];
  for (int j = 0; j < 3; ++j)
    for (int c = 0; c < 3; ++c) {
      y[3 * j + c] = d.rays[3 * static_cast<size_t>(id[j]) + c];
      x[3 * j + c] = d.raw1[3 * static_cast<size_t>(id[j]) + c];
    }
  const double X4[3] = {d.raw1[3 * static_cast<size_t>(id[3])], d.raw1[3 * static_cast<size_t>(id[3]) + 1],
                        d.raw1[3 * static_cast<size_t>(id[3]) + 2]};
  const double uv4[2] = {d.raw0[2 * static_cast<size_t>(id[3])], d.raw0[2 * static_cast<size_t>(id[3]) + 1]};
  double P[12];
  const bool ok = p3p_pick(y, x, X4, uv4, d.fx, d.fy, sq_thr, P);
  valid[i] = ok ? 1 : 0;
  for (int j = 0; j < 12; ++j) models[12 * static_cast<size_t>(i) + j] = ok ? P[j] : 0.0;
}

// line_estimator.cc:53-72 / hybrid_line_estimator.cc:111-130
__global__ void solve_line2pt_kernel(const DeviceData d, int k, const int32_t* __restrict__ samples,
                                     double* __restrict__ models, uint8_t* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  const int i0 = samples[2 * static_cast<size_t>(i)], i1 = samples[2 * static_cast<size_t>(i) + 1];
  const double x1 = d.raw0[2 * static_cast<size_t>(i0)], y1 = d.raw0[2 * static_cast<size_t>(i0) + 1];
  const double x2 = d.raw0[2 * static_cast<size_t>(i1)], y2 = d.raw0[2 * static_cast<size_t>(i1) + 1];
  const double a = y1 * 1.0 - 1.0 * y2;
  const double b = 1.0 * x2 - x1 * 1.0;
  const double c = x1 * y2 - y1 * x2;
  const double nrm = sqrt(a * a + b * b);
  const bool ok = !(nrm == 0.0);
  valid[i] = ok ? 1 : 0;
  models[3 * static_cast<size_t>(i) + 0] = ok ? a / nrm : 0.0;
  models[3 * static_cast<size_t>(i) + 1] = ok ? b / nrm : 0.0;
  models[3 * static_cast<size_t>(i) + 2] = ok ? c / nrm : 0.0;
}

// hybrid_line_estimator.cc:132-152
__global__ void solve_point_normal_kernel(const DeviceData d, int k,
                                          const int32_t* __restrict__ samples,
                                          double* __restrict__ models,
                                          uint8_t* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  const size_t s = static_cast<size_t>(samples[i]);
  double nx = d.raw1[4 * s + 2], ny = d.raw1[4 * s + 3];
  const double n2 = nx * nx + ny * ny;
  if (n2 > 0.0) {  // Eigen normalize(): only divides when the norm is positive
    const double n = sqrt(n2);
    nx /= n;
    ny /= n;
  }
  const double c = -(nx * d.raw1[4 * s] + ny * d.raw1[4 * s + 1]);
  const double nrm = sqrt(nx * nx + ny * ny);
  const bool ok = !(nrm == 0.0);
  valid[i] = ok ? 1 : 0;
  models[3 * static_cast<size_t>(i) + 0] = ok ? nx / nrm : 0.0;
  models[3 * static_cast<size_t>(i) + 1] = ok ? ny / nrm : 0.0;
  models[3 * static_cast<size_t>(i) + 2] = ok ? c / nrm : 0.0;
}

// calibrated_absolute_pose_estimator.cc:214-227
__global__ void rays_kernel(const DeviceData d, double* __restrict__ rays) {
  const int n = d.n[0];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double x = d.raw0[2 * static_cast<size_t>(i)] / d.fx;
    const double y = d.raw0[2 * static_cast<size_t>(i) + 1] / d.fy;
    const double z = 1.0;
    const double nrm = sqrt(x * x + y * y + z * z);
    rays[3 * static_cast<size_t>(i)] = x / nrm;
    rays[3 * static_cast<size_t>(i) + 1] = y / nrm;
    rays[3 * static_cast<size_t>(i) + 2] = z / nrm;
  }
}

}  // namespace

cudaError_t launch_rays(const DeviceData& d, double* rays, cudaStream_t st, int* launches) {
  ++*launches;
  const int blocks = max(1, min(2048, (d.n[0] + 255) / 256));
  rays_kernel<<<blocks, 256, 0, st>>>(d, rays);
  return cudaGetLastError();
}
cudaError_t launch_solve_line2pt(const DeviceData& d, int k, const int32_t* samples, double* models,
                                 uint8_t* valid, cudaStream_t st, int* launches) {
  ++*launches;
  solve_line2pt_kernel<<<(k + 255) / 256, 256, 0, st>>>(d, k, samples, models, valid);
  return cudaGetLastError();
}
cudaError_t launch_solve_point_normal(const DeviceData& d, int k, const int32_t* samples,
                                      double* models, uint8_t* valid, cudaStream_t st,
                                      int* launches) {
  ++*launches;
  solve_point_normal_kernel<<<(k + 255) / 256, 256, 0, st>>>(d, k, samples, models, valid);
  return cudaGetLastError();
}
cudaError_t launch_solve_p3p(const DeviceData& d, int k, const int32_t* samples, double sq_thr,
                             double* models, uint8_t* valid, cudaStream_t st, int* launches) {
  ++*launches;
  solve_p3p_kernel<<<(k + 127) / 128, 128, 0, st>>>(d, k, samples, sq_thr, models, valid);
  return cudaGetLastError();
}

}  // namespace rlb
