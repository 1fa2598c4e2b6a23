// C entry points over the drop-in template layer (include/RansacLib/*.h) instantiated with
// the three example solvers. This is what the Python mirror (ransaclib_b200/host.py,
// ransaclib_b200/pyransaclib.py) binds; C++ users include the headers directly, exactly as
// they include the reference's (examples/line_estimation.cc:133-143,
// camera_pose_estimation.cc:172-192, hybrid_line_estimation.cc:175-186,
// ransaclib_localization.cc:58-165).
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>

#include <RansacLib/hybrid_ransac.h>
#include <RansacLib/ransac.h>

#include "circle_estimator.h"
#include "calibrated_absolute_pose_estimator.h"
#include "hybrid_custom_line_estimator.h"
#include "hybrid_line_estimator.h"
#include "hybrid_pose_estimator.h"
#include "line_estimator.h"
#include "localization_io.h"

using namespace ransac_lib;
using calibrated_absolute_pose::CalibratedAbsolutePoseEstimator;
using calibrated_absolute_pose::CameraPose;
using calibrated_absolute_pose::CameraPoses;
using calibrated_absolute_pose::HybridPoseEstimator;

extern "C" {

struct rlbh_options {
  uint32_t min_iters, max_iters;
  double success_prob;
  double sq_thr;
  uint32_t seed;
  int32_t lo_steps;
  double thr_mult;
  int32_t lsq_iters;
  int32_t min_sample_mult;
  int32_t non_min_sample_mult;
  uint32_t lo_start;
  int32_t final_lsq;
};

struct rlbh_stats {
  uint32_t num_iterations;
  int32_t best_num_inliers;
  double best_model_score;
  double inlier_ratio;
  int32_t number_lo_iterations;
  int32_t found;
  // driver counters (not in the reference)
  uint64_t batches, hypotheses_scored, single_passes;
  double seconds;  // wall time of EstimateModel
  uint64_t handbacks, handbacks_resumed;  // device-resident runs handed back / of those, resumed with their state
};

struct rlbh_hybrid_options {
  uint32_t min_iters, max_iters, max_iters_per_solver;
  double success_prob;
  double sq_thr[2];
  double weights[2];
  uint32_t seed;
  int32_t lo_steps;
  double thr_mult;
  int32_t lsq_iters;
  int32_t min_sample_mult;
  uint32_t lo_start;
  int32_t final_lsq;
};

struct rlbh_hybrid_stats {
  uint32_t num_iterations_total;
  uint32_t num_iterations_per_solver[2];
  int32_t best_num_inliers;
  int32_t best_solver_type;
  double best_model_score;
  double inlier_ratios[2];
  int32_t num_inliers_per_type[2];
  int32_t number_lo_iterations;
};

}  // extern "C"

namespace {

thread_local std::string g_error;

LORansacOptions ToOptions(const rlbh_options& o) {
  LORansacOptions r;
  r.min_num_iterations_ = o.min_iters;
  r.max_num_iterations_ = o.max_iters;
  r.success_probability_ = o.success_prob;
  r.squared_inlier_threshold_ = o.sq_thr;
  r.random_seed_ = o.seed;
  r.num_lo_steps_ = o.lo_steps;
  r.threshold_multiplier_ = o.thr_mult;
  r.num_lsq_iterations_ = o.lsq_iters;
  r.min_sample_multiplicator_ = o.min_sample_mult;
  r.non_min_sample_multiplier_ = o.non_min_sample_mult;
  r.lo_starting_iterations_ = o.lo_start;
  r.final_least_squares_ = o.final_lsq != 0;
  return r;
}

template <class Model, class ModelVector, class Solver>
void Run(const rlbh_options& o, Solver& solver, int precision, Model* model, rlbh_stats* st,
         int32_t* inliers) {
  if (precision >= 0) solver.device_scorer().set_precision(precision);
  LocallyOptimizedMSAC<Model, ModelVector, Solver> lomsac;
  GpuRunCounters counters;
  lomsac.counters = &counters;
  RansacStatistics rs;
  const auto t0 = std::chrono::steady_clock::now();
  lomsac.EstimateModel(ToOptions(o), solver, model, &rs);
  const auto t1 = std::chrono::steady_clock::now();
  st->num_iterations = rs.num_iterations;
  st->best_num_inliers = rs.best_num_inliers;
  st->best_model_score = rs.best_model_score;
  st->inlier_ratio = rs.inlier_ratio;
  st->number_lo_iterations = rs.number_lo_iterations;
  st->found = rs.best_model_score < std::numeric_limits<double>::max() ? 1 : 0;
  st->batches = counters.batches;
  st->hypotheses_scored = counters.hypotheses_scored;
  st->single_passes = counters.single_passes;
  st->seconds = std::chrono::duration<double>(t1 - t0).count();
  st->handbacks = counters.handbacks;
  st->handbacks_resumed = counters.handbacks_resumed;
  if (inliers)
    for (size_t i = 0; i < rs.inlier_indices.size(); ++i) inliers[i] = rs.inlier_indices[i];
}

using localization_io::PoseToQvecTvec;  // [R|c] -> (qvec w,x,y,z ; tvec = -R c), ransaclib_localization.cc:147-150

// What the samplers ask of a solver, and nothing else.
struct SizesOnly {
  std::vector<int> n;
  std::vector<std::vector<double>> weights;
  int sample_size;
  int num_data() const { return n[0]; }
  int min_sample_size() const { return sample_size; }
  void num_data(std::vector<int>* out) const { *out = n; }
  void get_weights(std::vector<std::vector<double>>& w) const { w = weights; }
};

template <class Sampler, class HSolver = HybridLineEstimator, class Model = Line2D, class ModelVector = Lines2D>
void RunHybrid(const rlbh_hybrid_options& o, HSolver& solver, int precision, double* model,
               rlbh_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  if (precision >= 0) solver.device_scorer().set_precision(precision);
  HybridLORansacOptions ho;
  ho.min_num_iterations_ = o.min_iters;
  ho.max_num_iterations_ = o.max_iters;
  ho.max_num_iterations_per_solver_ = o.max_iters_per_solver;
  ho.success_probability_ = o.success_prob;
  ho.squared_inlier_thresholds_ = {o.sq_thr[0], o.sq_thr[1]};
  ho.data_type_weights_ = {o.weights[0], o.weights[1]};
  ho.random_seed_ = o.seed;
  ho.num_lo_steps_ = o.lo_steps;
  ho.threshold_multiplier_ = o.thr_mult;
  ho.num_lsq_iterations_ = o.lsq_iters;
  ho.min_sample_multiplicator_ = o.min_sample_mult;
  ho.lo_starting_iterations_ = o.lo_start;
  ho.final_least_squares_ = o.final_lsq != 0;
  HybridLocallyOptimizedMSAC<Model, ModelVector, HSolver, Sampler> lomsac;
  HybridRansacStatistics hs;
  lomsac.EstimateModel(ho, solver, reinterpret_cast<Model*>(model), &hs);
  st->num_iterations_total = hs.num_iterations_total;
  for (int i = 0; i < 2; ++i) {
    st->num_iterations_per_solver[i] = hs.num_iterations_per_solver[i];
    st->inlier_ratios[i] = hs.inlier_ratios[i];
    st->num_inliers_per_type[i] = static_cast<int32_t>(hs.inlier_indices[i].size());
  }
  st->best_num_inliers = hs.best_num_inliers;
  st->best_solver_type = hs.best_solver_type;
  st->best_model_score = hs.best_model_score;
  st->number_lo_iterations = hs.number_lo_iterations;
  if (inl0) std::copy(hs.inlier_indices[0].begin(), hs.inlier_indices[0].end(), inl0);
  if (inl1) std::copy(hs.inlier_indices[1].begin(), hs.inlier_indices[1].end(), inl1);
}

// The callers' arrays outlive every call of this file, so the solvers borrow them (gpu::BorrowData);
// RLB_HOST_COPY=1 selects the copying constructors (A/B timing, and the tests run both).
bool HostCopy() {
  const char* copy = std::getenv("RLB_HOST_COPY");
  return copy && copy[0] == '1';
}

template <class F>
int Guard(F f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

}  // namespace

extern "C" {

const char* rlbh_last_error(void) { return g_error.c_str(); }

// precision: RLB_F64 / RLB_F32, or -1 for the default (env RLB_PRECISION, else f64).
int rlbh_lomsac_line(const rlbh_options* o, int device, int precision, int n, const double* xy,
                     double* model, rlbh_stats* st, int32_t* inliers) {
  return Guard([&] {
    if (HostCopy()) {
      LineEstimator solver(n, xy, &gpu::Context::ThreadDefault(device));
      Run<Line2D, Lines2D, LineEstimator>(*o, solver, precision, reinterpret_cast<Line2D*>(model), st, inliers);
      return;
    }
    LineEstimator solver(gpu::BorrowData(), n, xy, &gpu::Context::ThreadDefault(device));
    Run<Line2D, Lines2D, LineEstimator>(*o, solver, precision, reinterpret_cast<Line2D*>(model), st, inliers);
  });
}

// A user-defined solver (examples/circle_estimator.h): host minimal solver, NVRTC-compiled residual.
int rlbh_lomsac_circle(const rlbh_options* o, int device, int n, const double* xy, double* model,
                       rlbh_stats* st, int32_t* inliers) {
  return Guard([&] {
    CircleEstimator solver(n, xy, &gpu::Context::ThreadDefault(device));
    Run<Circle2D, Circles2D, CircleEstimator>(*o, solver, RLB_F64, reinterpret_cast<Circle2D*>(model), st,
                                              inliers);
  });
}

int rlbh_lomsac_pose(const rlbh_options* o, int device, int precision, double fx, double fy, int n,
                     const double* p2d, const double* p3d, double* model, rlbh_stats* st,
                     int32_t* inliers) {
  return Guard([&] {
    // RLB_NUM_GPUS=G: no context of our own - the estimator replicates its data over G GPUs (gpu_scorer.h)
    gpu::Context* cx = (gpu::NumGpusFromEnv() >= 1) ? nullptr : &gpu::Context::ThreadDefault(device);
    if (HostCopy()) {
      CalibratedAbsolutePoseEstimator solver(fx, fy, o->sq_thr, n, p2d, p3d, cx);
      Run<CameraPose, CameraPoses, CalibratedAbsolutePoseEstimator>(
          *o, solver, precision, reinterpret_cast<CameraPose*>(model), st, inliers);
      return;
    }
    CalibratedAbsolutePoseEstimator solver(gpu::BorrowData(), fx, fy, o->sq_thr, n, p2d, p3d, cx);
    Run<CameraPose, CameraPoses, CalibratedAbsolutePoseEstimator>(
        *o, solver, precision, reinterpret_cast<CameraPose*>(model), st, inliers);
  });
}

// CalibratedAbsolutePoseEstimator::MinimalSolver (host P3P, examples/p3p.h) on k samples of 4 indices.
int rlbh_pose_minimal_solver(int device, double fx, double fy, double sq_thr, int n, const double* p2d,
                             const double* p3d, int k, const int32_t* samples, double* models, uint8_t* valid) {
  return Guard([&] {
    CalibratedAbsolutePoseEstimator solver(fx, fy, sq_thr, n, p2d, p3d, &gpu::Context::ThreadDefault(device));
    CameraPoses poses;
    for (int i = 0; i < k; ++i) {
      const std::vector<int> sample(samples + 4 * static_cast<size_t>(i), samples + 4 * static_cast<size_t>(i + 1));
      valid[i] = solver.MinimalSolver(sample, &poses) == 1 ? 1 : 0;
      for (int j = 0; j < 12; ++j) models[12 * static_cast<size_t>(i) + j] = valid[i] ? poses[0].m[j] : 0.0;
    }
  });
}

int rlbh_hybrid_lomsac_line(const rlbh_hybrid_options* o, int device, int precision, int n0,
                            const double* xy0, int n1, const double* xynn1, const double* priors,
                            double* model, rlbh_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  return Guard([&] {
    if (HostCopy()) {
      HybridLineEstimator solver(n0, xy0, n1, xynn1, {priors[0], priors[1]},
                                 &gpu::Context::ThreadDefault(device));
      RunHybrid<HybridUniformSampling<HybridLineEstimator>>(*o, solver, precision, model, st, inl0, inl1);
      return;
    }
    HybridLineEstimator solver(gpu::BorrowData(), n0, xy0, n1, xynn1, {priors[0], priors[1]},
                               &gpu::Context::ThreadDefault(device));
    RunHybrid<HybridUniformSampling<HybridLineEstimator>>(*o, solver, precision, model, st, inl0, inl1);
  });
}

// The same run with the solver's EvaluateModelOnPoint handed over as source (rlb_scorer_create_custom_hybrid):
// the path of any user-defined HybridSolver.
int rlbh_hybrid_lomsac_line_custom(const rlbh_hybrid_options* o, int device, int precision, int n0,
                                   const double* xy0, int n1, const double* xynn1, const double* priors,
                                   double* model, rlbh_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  return Guard([&] {
    HybridCustomLineEstimator solver(n0, xy0, n1, xynn1, {priors[0], priors[1]}, &gpu::Context::ThreadDefault(device));
    RunHybrid<HybridUniformSampling<HybridCustomLineEstimator>, HybridCustomLineEstimator>(*o, solver, precision, model,
                                                                                          st, inl0, inl1);
  });
}

// Hybrid camera pose (examples/hybrid_pose_estimator.h): 2D-3D matches (type 0: p2d n0 x 2, p3d n0 x 3) and 3D-3D
// matches (type 1: cam1 / world1 n1 x 3), P3P and three-point solvers, both residuals on the GPU as source.
// model: 12 doubles, R column-major then the camera centre.
int rlbh_hybrid_lomsac_pose(const rlbh_hybrid_options* o, int device, int precision, double fx, double fy, int n0,
                            const double* p2d, const double* p3d, int n1, const double* cam1, const double* world1,
                            const double* priors, double* model, rlbh_hybrid_stats* st, int32_t* inl0,
                            int32_t* inl1) {
  return Guard([&] {
    HybridPoseEstimator solver(fx, fy, o->sq_thr[0], n0, p2d, p3d, n1, cam1, world1, {priors[0], priors[1]},
                               &gpu::Context::ThreadDefault(device));
    RunHybrid<HybridUniformSampling<HybridPoseEstimator>, HybridPoseEstimator, CameraPose, CameraPoses>(
        *o, solver, precision, model, st, inl0, inl1);
  });
}

// Same with the reference's weighted sampler (hybrid_sampling.h:146-223): w0 / w1 are the per-point
// sampling weights of the two data types (<= 0: never sampled, still scored).
int rlbh_hybrid_lomsac_line_biased(const rlbh_hybrid_options* o, int device, int precision, int n0,
                                   const double* xy0, const double* w0, int n1, const double* xynn1,
                                   const double* w1, const double* priors, double* model,
                                   rlbh_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  return Guard([&] {
    HybridLineEstimator solver(n0, xy0, n1, xynn1, {priors[0], priors[1]},
                               &gpu::Context::ThreadDefault(device));
    solver.set_weights(std::vector<double>(w0, w0 + n0), std::vector<double>(w1, w1 + n1));
    RunHybrid<HybridBiasedSampling<HybridLineEstimator>>(*o, solver, precision, model, st, inl0, inl1);
  });
}

// ---- the samplers of the template layer on their own (host only; no device is touched) ----------
// k samples of size s from n data (include/RansacLib/sampling.h) -> out[k][s].
int rlbh_sample_uniform(uint32_t seed, int n, int s, int k, int32_t* out) {
  return Guard([&] {
    SizesOnly solver{{n}, {}, s};
    UniformSampling<SizesOnly> sampler(seed, solver);
    // first half one by one, second half as the drivers draw them (SampleBatch): one stream
    std::vector<int> one, flat;
    const int singles = k / 2;
    for (int i = 0; i < singles; ++i) {
      sampler.Sample(&one);
      std::copy(one.begin(), one.end(), out + static_cast<size_t>(i) * s);
    }
    sampler.SampleBatch(k - singles, &flat);
    std::copy(flat.begin(), flat.end(), out + static_cast<size_t>(singles) * s);
  });
}

// The drivers' local-optimisation stream on its own (host only): `reps` consecutive
// utils::RandomShuffleAndResize(target, Mt19937*, iota(n)) on one engine -> kept prefix of the last one,
// the engine's next raw output (checks the number of consumed draws) and, in first_words[0..8), the
// engine's first outputs after seeding.
int rlbh_shuffle_and_resize(uint32_t seed, int n, int target, int reps, int32_t* out, uint32_t* next_draw,
                            uint32_t* first_words) {
  return Guard([&] {
    utils::Mt19937 probe(seed);
    for (int i = 0; i < 8; ++i) first_words[i] = probe();
    utils::Mt19937 rng;
    rng.seed(seed);
    std::vector<int> v;
    for (int r = 0; r < reps; ++r) {
      v.resize(n);
      for (int i = 0; i < n; ++i) v[i] = i;
      utils::RandomShuffleAndResize(target, &rng, &v);
    }
    for (int i = 0; i < target; ++i) out[i] = v[i];
    *next_draw = rng();
  });
}

// The same on a read-only list (list[i] = 3 i + 1): utils::ShuffleSampleFromList.
int rlbh_shuffle_from_list(uint32_t seed, int n, int target, int reps, int32_t* out, uint32_t* next_draw) {
  return Guard([&] {
    utils::Mt19937 rng;
    rng.seed(seed);
    std::vector<int> list(n), v;
    for (int i = 0; i < n; ++i) list[i] = 3 * i + 1;
    for (int r = 0; r < reps; ++r) utils::ShuffleSampleFromList(target, &rng, list.data(), n, &v);
    for (int i = 0; i < target; ++i) out[i] = v[i];
    *next_draw = rng();
  });
}

// k hybrid samples over T data types; sizes[k][T] per draw; weights == NULL: HybridUniformSampling,
// else HybridBiasedSampling with the concatenated per-type weights. out[k][T][max_size], -1 padded.
int rlbh_sample_hybrid(uint32_t seed, int T, const int32_t* num_data, const double* weights, int k,
                       const int32_t* sizes, int max_size, int32_t* out) {
  return Guard([&] {
    SizesOnly solver{std::vector<int>(num_data, num_data + T), {}, 0};
    if (weights) {
      const double* w = weights;
      for (int t = 0; t < T; ++t) {
        solver.weights.emplace_back(w, w + num_data[t]);
        w += num_data[t];
      }
    }
    std::vector<std::vector<int>> one;
    auto draw = [&](auto& sampler) {
      for (int i = 0; i < k; ++i) {
        std::vector<int> sz(sizes + static_cast<size_t>(i) * T, sizes + static_cast<size_t>(i + 1) * T);
        sampler.Sample(sz, &one);
        for (int t = 0; t < T; ++t) {
          int32_t* row = out + (static_cast<size_t>(i) * T + t) * max_size;
          std::fill(row, row + max_size, -1);
          std::copy(one[t].begin(), one[t].end(), row);
        }
      }
    };
    if (weights) {
      HybridBiasedSampling<SizesOnly> sampler(seed, solver);
      draw(sampler);
    } else {
      HybridUniformSampling<SizesOnly> sampler(seed, solver);
      draw(sampler);
    }
  });
}

// pyransaclib.ransaclib_localization (reference examples/ransaclib_localization.cc:58-165):
// float focals / threshold, fixed options (:100-107), [R|c] -> (qvec w,x,y,z ; tvec = -R c)
// (:147-150). The reference seeds from std::random_device (:109-110); here the seed is an
// argument so runs are reproducible. Returns 1 on success, 0 for <= 3 matches (:86-92).
int rlbh_localize(int device, int precision, float focal_x, float focal_y, int n, const double* p2d,
                  const double* p3d, float inlier_threshold, int number_lo_steps,
                  uint32_t min_num_iterations, uint32_t max_num_iterations, uint32_t seed,
                  double* qvec, double* tvec, int32_t* num_inliers, int32_t* inliers,
                  rlbh_stats* st) {
  if (n <= 3) return 0;
  int success = 0;
  const int rc = Guard([&] {
    rlbh_options o;
    o.min_iters = min_num_iterations;
    o.max_iters = max_num_iterations;
    o.success_prob = 0.9999;
    const double thr_px = static_cast<double>(inlier_threshold);
    o.sq_thr = thr_px * thr_px;
    o.seed = seed;
    o.lo_steps = number_lo_steps;
    o.thr_mult = std::sqrt(2.0);
    o.lsq_iters = 4;
    o.min_sample_mult = 7;
    o.non_min_sample_mult = 3;
    o.lo_start = 60;
    o.final_lsq = 1;
    CalibratedAbsolutePoseEstimator solver(focal_x, focal_y, o.sq_thr, n, p2d, p3d,
                                           &gpu::Context::ThreadDefault(device));
    CameraPose best;
    std::memset(&best, 0, sizeof(best));
    rlbh_stats local;
    Run<CameraPose, CameraPoses, CalibratedAbsolutePoseEstimator>(o, solver, precision, &best,
                                                                  st ? st : &local, inliers);
    const rlbh_stats& s = st ? *st : local;
    *num_inliers = s.best_num_inliers;
    PoseToQvecTvec(best.m, qvec, tvec);
    success = 1;
  });
  return rc == 0 ? success : -1;
}

// Batched ransaclib_localization: Q independent queries in ONE launch of the device-resident
// LO-MSAC (rlb_localize_batch); per-query semantics are those of rlbh_localize above (float
// focals / threshold, fixed options, <= 3 matches -> success 0). Queries are concatenated:
// query q owns matches [offsets[q], offsets[q+1]). inliers (optional): `total` int32, query q's
// ascending inlier indices start at offsets[q].
int rlbh_localize_batch(int device, int num_queries, const int64_t* offsets, const double* p2d,
                        const double* p3d, const float* focal_x, const float* focal_y,
                        float inlier_threshold, int number_lo_steps, uint32_t min_num_iterations,
                        uint32_t max_num_iterations, const uint32_t* seeds, int32_t* success,
                        double* qvecs, double* tvecs, int32_t* num_inliers, int32_t* inliers,
                        rlb_localize_result* results_out) {
  return Guard([&] {
    gpu::Context& ctx = gpu::Context::ThreadDefault(device);
    rlb_localize_options o;
    rlb_localize_default_options(&o, static_cast<double>(inlier_threshold), number_lo_steps,
                                 min_num_iterations, max_num_iterations);
    std::vector<double> fx(num_queries), fy(num_queries), poses(static_cast<size_t>(num_queries) * 12);
    for (int q = 0; q < num_queries; ++q) {
      fx[q] = static_cast<double>(focal_x[q]);
      fy[q] = static_cast<double>(focal_y[q]);
    }
    std::vector<rlb_localize_result> res(num_queries);
    gpu::Check(rlb_localize_batch(ctx.get(), num_queries, offsets, p2d, p3d, fx.data(), fy.data(), seeds, &o,
                                  poses.data(), res.data(), inliers),
               "rlb_localize_batch");
    for (int q = 0; q < num_queries; ++q) {
      success[q] = res[q].success;
      num_inliers[q] = res[q].num_inliers;
      if (res[q].success) {
        PoseToQvecTvec(&poses[static_cast<size_t>(q) * 12], qvecs + 4 * static_cast<size_t>(q),
                       tvecs + 3 * static_cast<size_t>(q));
      } else {
        for (int i = 0; i < 4; ++i) qvecs[4 * static_cast<size_t>(q) + i] = 0.0;
        for (int i = 0; i < 3; ++i) tvecs[3 * static_cast<size_t>(q) + i] = 0.0;
      }
      if (results_out) results_out[q] = res[q];
    }
  });
}

}  // extern "C"
