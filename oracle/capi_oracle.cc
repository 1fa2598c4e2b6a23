// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// One C entry-point file, compiled twice by oracle/Makefile:
//   liboracle.so        (no flag)        : the restated LO-MSAC of lomsac_oracle.h
//   _ref/libransac_ref.so (-DRLO_USE_REFERENCE -I/root/reference)
//                                         : the reference's OWN unmodified
//                                           RansacLib/*.h templates.
// Both drive the same Eigen-free solver restatements (solvers_oracle.h), so a
// difference between the two libraries isolates the restated control flow /
// scoring loops. Symbols are prefixed oracle_ / ref_ accordingly.
#include <chrono>
#include <cstring>
#include <vector>

#include "lomsac_oracle.h"
#include "solvers_oracle.h"

#ifdef RLO_USE_REFERENCE
#include <RansacLib/hybrid_ransac.h>
#include <RansacLib/ransac.h>
#define RLO(name) ref_##name
#else
#define RLO(name) oracle_##name
#endif

using namespace rlb_oracle;

extern "C" {

struct rlo_options {
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

struct rlo_stats {
  uint32_t num_iterations;
  int32_t best_num_inliers;
  double best_model_score;
  double inlier_ratio;
  int32_t number_lo_iterations;
  int32_t found;  // 1 if best_model was written
};

struct rlo_hybrid_options {
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

struct rlo_hybrid_stats {
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

#ifdef RLO_USE_REFERENCE
// Exposes the protected hot-path members of the reference class (ransac.h:277).
template <class Model, class ModelVector, class Solver>
struct OpenMsac : public ransac_lib::LocallyOptimizedMSAC<Model, ModelVector, Solver> {
  typedef ransac_lib::LocallyOptimizedMSAC<Model, ModelVector, Solver> B;
  using B::GetBestEstimatedModelId;
  using B::GetInliers;
  using B::ScoreModel;
};
template <class Model, class ModelVector, class Solver>
struct OpenHybrid
    : public ransac_lib::HybridLocallyOptimizedMSAC<Model, ModelVector, Solver> {
  typedef ransac_lib::HybridLocallyOptimizedMSAC<Model, ModelVector, Solver> B;
  using B::GetInliers;
  using B::ScoreModel;
};

ransac_lib::LORansacOptions to_ref(const rlo_options& o) {
  ransac_lib::LORansacOptions r;
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
ransac_lib::HybridLORansacOptions to_ref(const rlo_hybrid_options& o) {
  ransac_lib::HybridLORansacOptions r;
  r.min_num_iterations_ = o.min_iters;
  r.max_num_iterations_ = o.max_iters;
  r.max_num_iterations_per_solver_ = o.max_iters_per_solver;
  r.success_probability_ = o.success_prob;
  r.squared_inlier_thresholds_ = {o.sq_thr[0], o.sq_thr[1]};
  r.data_type_weights_ = {o.weights[0], o.weights[1]};
  r.random_seed_ = o.seed;
  r.num_lo_steps_ = o.lo_steps;
  r.threshold_multiplier_ = o.thr_mult;
  r.num_lsq_iterations_ = o.lsq_iters;
  r.min_sample_multiplicator_ = o.min_sample_mult;
  r.lo_starting_iterations_ = o.lo_start;
  r.final_least_squares_ = o.final_lsq != 0;
  return r;
}
#endif

LoMsacOptions to_oracle(const rlo_options& o) {
  LoMsacOptions r;
  r.min_iters = o.min_iters;
  r.max_iters = o.max_iters;
  r.success_prob = o.success_prob;
  r.sq_thr = o.sq_thr;
  r.seed = o.seed;
  r.lo_steps = o.lo_steps;
  r.thr_mult = o.thr_mult;
  r.lsq_iters = o.lsq_iters;
  r.min_sample_mult = o.min_sample_mult;
  r.non_min_sample_mult = o.non_min_sample_mult;
  r.lo_start = o.lo_start;
  r.final_lsq = o.final_lsq != 0;
  return r;
}
HybridOptions to_oracle(const rlo_hybrid_options& o) {
  HybridOptions r;
  r.min_iters = o.min_iters;
  r.max_iters = o.max_iters;
  r.max_iters_per_solver = o.max_iters_per_solver;
  r.success_prob = o.success_prob;
  r.sq_thr = {o.sq_thr[0], o.sq_thr[1]};
  r.weights = {o.weights[0], o.weights[1]};
  r.seed = o.seed;
  r.lo_steps = o.lo_steps;
  r.thr_mult = o.thr_mult;
  r.lsq_iters = o.lsq_iters;
  r.min_sample_mult = o.min_sample_mult;
  r.lo_start = o.lo_start;
  r.final_lsq = o.final_lsq != 0;
  return r;
}

// A solver stub that only carries sizes, for the samplers.
struct SizeOnly {
  int n, s;
  int num_data() const { return n; }
  int min_sample_size() const { return s; }
};
struct HybridSizeOnly {
  std::vector<int> n;
  std::vector<std::vector<double>> w;
  void num_data(std::vector<int>* nd) const { *nd = n; }
  void get_weights(std::vector<std::vector<double>>& out) const { out = w; }
};

// Generic single-type batch scoring through the scoring loops under test.
template <class Model, class ModelVector, class Solver>
void score_batch(const Solver& s, int k, const Model* models, double thr,
                 double* scores, int32_t* counts) {
#ifdef RLO_USE_REFERENCE
  OpenMsac<Model, ModelVector, Solver> r;
  for (int m = 0; m < k; ++m) {
    r.ScoreModel(s, models[m], thr, &scores[m]);
    if (counts) counts[m] = r.GetInliers(s, models[m], thr, nullptr);
  }
#else
  for (int m = 0; m < k; ++m) {
    scores[m] = score_model(s, models[m], thr);
    if (counts) counts[m] = get_inliers(s, models[m], thr, nullptr);
  }
#endif
}

template <class Model, class ModelVector, class Solver>
int inliers_one(const Solver& s, const Model& m, double thr, int32_t* idx) {
  std::vector<int> inl;
#ifdef RLO_USE_REFERENCE
  OpenMsac<Model, ModelVector, Solver> r;
  const int c = r.GetInliers(s, m, thr, &inl);
#else
  const int c = get_inliers(s, m, thr, &inl);
#endif
  for (int i = 0; i < c; ++i) idx[i] = inl[i];
  return c;
}

template <class Model, class ModelVector, class Solver>
void run_lomsac(const rlo_options& o, const Solver& s, Model* model, rlo_stats* st,
                int32_t* inliers) {
  // Sentinel so that "best_model left untouched" (ransac.h:135-139) is observable.
  st->found = 0;
#ifdef RLO_USE_REFERENCE
  ransac_lib::LocallyOptimizedMSAC<Model, ModelVector, Solver> r;
  ransac_lib::RansacStatistics rs;
  r.EstimateModel(to_ref(o), s, model, &rs);
  st->num_iterations = rs.num_iterations;
  st->best_num_inliers = rs.best_num_inliers;
  st->best_model_score = rs.best_model_score;
  st->inlier_ratio = rs.inlier_ratio;
  st->number_lo_iterations = rs.number_lo_iterations;
  const std::vector<int>& ii = rs.inlier_indices;
#else
  LoMsacStats rs;
  lomsac_estimate<Model, ModelVector, Solver>(to_oracle(o), s, model, &rs);
  st->num_iterations = rs.num_iterations;
  st->best_num_inliers = rs.best_num_inliers;
  st->best_model_score = rs.best_model_score;
  st->inlier_ratio = rs.inlier_ratio;
  st->number_lo_iterations = rs.number_lo_iterations;
  const std::vector<int>& ii = rs.inlier_indices;
#endif
  st->found = st->best_model_score < std::numeric_limits<double>::max() ? 1 : 0;
  if (inliers)
    for (size_t i = 0; i < ii.size(); ++i) inliers[i] = ii[i];
}

}  // namespace

extern "C" {

// ---- samplers / utils --------------------------------------------------------
// K consecutive minimal samples of size s from N (sampling.h:58-64).
void RLO(sample_uniform)(uint32_t seed, int n, int s, int k, int32_t* out) {
  std::vector<int> v;
#ifdef RLO_USE_REFERENCE
  SizeOnly so{n, s};
  ransac_lib::UniformSampling<SizeOnly> smp(seed, so);
#else
  UniformSampler smp(seed, n, s);
#endif
  for (int i = 0; i < k; ++i) {
    smp.Sample(&v);
    for (int j = 0; j < s; ++j) out[i * s + j] = v[j];
  }
}

// Hybrid sampler for T data types; sizes is k x T; out is k x stride (stride =
// max row sum), padded with -1. hybrid_sampling.h:64-87.
void RLO(sample_hybrid)(uint32_t seed, int T, const int32_t* nd, int k,
                        const int32_t* sizes, int stride, int32_t* out) {
  std::vector<int> ndv(nd, nd + T);
#ifdef RLO_USE_REFERENCE
  HybridSizeOnly so{ndv};
  ransac_lib::HybridUniformSampling<HybridSizeOnly> smp(seed, so);
#else
  HybridUniformSampler smp(seed, ndv);
#endif
  std::vector<std::vector<int>> v;
  for (int i = 0; i < k; ++i) {
    std::vector<int> sz(sizes + i * T, sizes + (i + 1) * T);
    smp.Sample(sz, &v);
    int p = 0;
    for (int t = 0; t < T; ++t)
      for (size_t j = 0; j < v[t].size(); ++j) out[i * stride + p++] = v[t][j];
    for (; p < stride; ++p) out[i * stride + p] = -1;
  }
}

// Weighted hybrid sampler (hybrid_sampling.h:146-223); weights = the T per-type weight arrays
// concatenated; same output layout as sample_hybrid.
void RLO(sample_hybrid_biased)(uint32_t seed, int T, const int32_t* nd, const double* weights,
                               int k, const int32_t* sizes, int stride, int32_t* out) {
  std::vector<std::vector<double>> w(T);
  for (int t = 0; t < T; ++t) {
    w[t].assign(weights, weights + nd[t]);
    weights += nd[t];
  }
#ifdef RLO_USE_REFERENCE
  HybridSizeOnly so{std::vector<int>(nd, nd + T), w};
  ransac_lib::HybridBiasedSampling<HybridSizeOnly> smp(seed, so);
#else
  HybridBiasedSampler smp(seed, w);
#endif
  std::vector<std::vector<int>> v;
  for (int i = 0; i < k; ++i) {
    std::vector<int> sz(sizes + i * T, sizes + (i + 1) * T);
    smp.Sample(sz, &v);
    int p = 0;
    for (int t = 0; t < T; ++t)
      for (size_t j = 0; j < v[t].size(); ++j) out[i * stride + p++] = v[t][j];
    for (; p < stride; ++p) out[i * stride + p] = -1;
  }
}

uint32_t RLO(num_required_iterations)(double ratio, double prob_miss, int s,
                                      uint32_t mn, uint32_t mx) {
#ifdef RLO_USE_REFERENCE
  return ransac_lib::utils::NumRequiredIterations(ratio, prob_miss, s, mn, mx);
#else
  return num_required_iterations(ratio, prob_miss, s, mn, mx);
#endif
}

uint32_t RLO(num_required_iterations_hybrid)(int T, const double* ratios,
                                             double prob_miss, const int32_t* sizes,
                                             uint32_t mn, uint32_t mx) {
  std::vector<double> r(ratios, ratios + T);
  std::vector<int> s(sizes, sizes + T);
#ifdef RLO_USE_REFERENCE
  return ransac_lib::utils::NumRequiredIterations(r, prob_miss, s, mn, mx);
#else
  return num_required_iterations(r, prob_miss, s, mn, mx);
#endif
}

// iota(n) shuffled by mt19937(seed), first `target` kept; *next_draw receives the
// next raw rng output so the number of consumed draws is checked too (F7).
void RLO(shuffle_and_resize)(uint32_t seed, int n, int target, int32_t* out,
                             uint32_t* next_draw) {
  std::mt19937 rng;
  rng.seed(seed);
  std::vector<int> v(n);
  std::iota(v.begin(), v.end(), 0);
#ifdef RLO_USE_REFERENCE
  ransac_lib::utils::RandomShuffleAndResize(target, &rng, &v);
#else
  shuffle_and_resize(target, &rng, &v);
#endif
  for (int i = 0; i < target; ++i) out[i] = v[i];
  *next_draw = static_cast<uint32_t>(rng());
}

// `reps` consecutive shuffles of iota(n) on ONE engine (the local-optimisation stream shuffles list
// after list); out = the kept prefix of the last one.
void RLO(shuffle_and_resize_reps)(uint32_t seed, int n, int target, int reps, int32_t* out,
                                  uint32_t* next_draw) {
  std::mt19937 rng;
  rng.seed(seed);
  std::vector<int> v;
  for (int r = 0; r < reps; ++r) {
    v.resize(n);
    std::iota(v.begin(), v.end(), 0);
#ifdef RLO_USE_REFERENCE
    ransac_lib::utils::RandomShuffleAndResize(target, &rng, &v);
#else
    shuffle_and_resize(target, &rng, &v);
#endif
  }
  for (int i = 0; i < target; ++i) out[i] = v[i];
  *next_draw = static_cast<uint32_t>(rng());
}

// ---- pose ----------------------------------------------------------------------
void RLO(pixels_to_rays)(double fx, double fy, int n, const double* p2d, double* rays) {
  PoseSolver::PixelsToViewingRays(fx, fy, n, p2d, rays);
}

void RLO(pose_errors)(const double* model, double fx, double fy, int n,
                      const double* p2d, const double* p3d, double* err) {
  PoseSolver s(fx, fy, 0.0, n, p2d, nullptr, p3d);
  Pose34 P;
  std::memcpy(P.m, model, sizeof(P.m));
  for (int i = 0; i < n; ++i) err[i] = s.EvaluateModelOnPoint(P, i);
}

void RLO(pose_score_models)(double fx, double fy, int n, const double* p2d,
                            const double* p3d, int k, const double* models,
                            double thr, double* scores, int32_t* counts) {
  PoseSolver s(fx, fy, thr, n, p2d, nullptr, p3d);
  score_batch<Pose34, Poses, PoseSolver>(s, k, reinterpret_cast<const Pose34*>(models),
                                         thr, scores, counts);
}

int RLO(pose_inliers)(double fx, double fy, int n, const double* p2d, const double* p3d,
                      const double* model, double thr, int32_t* idx) {
  PoseSolver s(fx, fy, thr, n, p2d, nullptr, p3d);
  return inliers_one<Pose34, Poses, PoseSolver>(
      s, *reinterpret_cast<const Pose34*>(model), thr, idx);
}

int RLO(p3p)(const double* y, const double* x, double* R, double* t) {
  P3PSolutions sols;
  const int n = p3p_lambda_twist(y, x, &sols);
  for (int i = 0; i < n; ++i) {
    std::memcpy(R + 9 * i, sols.R[i], 9 * sizeof(double));
    std::memcpy(t + 3 * i, sols.t[i], 3 * sizeof(double));
  }
  return n;
}

// MinimalSolver over K 4-samples (calibrated_absolute_pose_estimator.cc:111-137).
void RLO(pose_minimal_batch)(double fx, double fy, double thr, int n, const double* p2d,
                             const double* p3d, int k, const int32_t* samples,
                             double* models, uint8_t* valid) {
  std::vector<double> rays(3 * static_cast<size_t>(n));
  PoseSolver::PixelsToViewingRays(fx, fy, n, p2d, rays.data());
  PoseSolver s(fx, fy, thr, n, p2d, rays.data(), p3d);
  Poses out;
  std::vector<int> smp(4);
  for (int i = 0; i < k; ++i) {
    for (int j = 0; j < 4; ++j) smp[j] = samples[4 * i + j];
    valid[i] = s.MinimalSolver(smp, &out) == 1 ? 1 : 0;
    if (valid[i]) std::memcpy(models + 12 * i, out[0].m, 12 * sizeof(double));
    else for (int j = 0; j < 12; ++j) models[12 * i + j] = 0.0;
  }
}

void RLO(pose_least_squares)(double fx, double fy, int n, const double* p2d,
                             const double* p3d, int k, const int32_t* sample,
                             double* model) {
  PoseSolver s(fx, fy, 0.0, n, p2d, nullptr, p3d);
  std::vector<int> smp(sample, sample + k);
  Pose34 P;
  std::memcpy(P.m, model, sizeof(P.m));
  s.LeastSquares(smp, &P);
  std::memcpy(model, P.m, sizeof(P.m));
}

void RLO(lomsac_pose)(const rlo_options* o, double fx, double fy, int n,
                      const double* p2d, const double* p3d, double* model,
                      rlo_stats* st, int32_t* inliers) {
  std::vector<double> rays(3 * static_cast<size_t>(n));
  PoseSolver::PixelsToViewingRays(fx, fy, n, p2d, rays.data());
  PoseSolver s(fx, fy, o->sq_thr, n, p2d, rays.data(), p3d);
  run_lomsac<Pose34, Poses, PoseSolver>(*o, s, reinterpret_cast<Pose34*>(model), st,
                                        inliers);
}

// Times `reps` ScoreModel passes over k models; returns seconds (steady_clock).
// Used only by bench.py's cpu_baseline / --impl reference legs.
double RLO(time_pose_scoring)(double fx, double fy, int n, const double* p2d,
                              const double* p3d, int k, const double* models, double thr,
                              int reps, double* checksum) {
  PoseSolver s(fx, fy, thr, n, p2d, nullptr, p3d);
  std::vector<double> scores(k);
  const auto t0 = std::chrono::steady_clock::now();
  double acc = 0.0;
  for (int r = 0; r < reps; ++r) {
    score_batch<Pose34, Poses, PoseSolver>(
        s, k, reinterpret_cast<const Pose34*>(models), thr, scores.data(), nullptr);
    for (int i = 0; i < k; ++i) acc += scores[i];
  }
  const auto t1 = std::chrono::steady_clock::now();
  if (checksum) *checksum = acc;
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---- line ----------------------------------------------------------------------
void RLO(line_score_models)(int n, const double* xy, int k, const double* models,
                            double thr, double* scores, int32_t* counts) {
  LineSolver s(n, xy);
  score_batch<Line, Lines, LineSolver>(s, k, reinterpret_cast<const Line*>(models), thr,
                                       scores, counts);
}

int RLO(line_inliers)(int n, const double* xy, const double* model, double thr,
                      int32_t* idx) {
  LineSolver s(n, xy);
  return inliers_one<Line, Lines, LineSolver>(s, *reinterpret_cast<const Line*>(model),
                                              thr, idx);
}

void RLO(line_minimal_batch)(int n, const double* xy, int k, const int32_t* samples,
                             double* models, uint8_t* valid) {
  LineSolver s(n, xy);
  Lines out;
  std::vector<int> smp(2);
  for (int i = 0; i < k; ++i) {
    smp[0] = samples[2 * i];
    smp[1] = samples[2 * i + 1];
    valid[i] = s.MinimalSolver(smp, &out) == 1 ? 1 : 0;
    for (int j = 0; j < 3; ++j) models[3 * i + j] = valid[i] ? out[0].v[j] : 0.0;
  }
}

void RLO(line_least_squares)(int n, const double* xy, int k, const int32_t* sample,
                             double* model, int32_t* ok) {
  LineSolver s(n, xy);
  std::vector<int> smp(sample, sample + k);
  Line l;
  std::memcpy(l.v, model, sizeof(l.v));
  *ok = s.NonMinimalSolver(smp, &l);
  std::memcpy(model, l.v, sizeof(l.v));
}

void RLO(lomsac_line)(const rlo_options* o, int n, const double* xy, double* model,
                      rlo_stats* st, int32_t* inliers) {
  LineSolver s(n, xy);
  run_lomsac<Line, Lines, LineSolver>(*o, s, reinterpret_cast<Line*>(model), st, inliers);
}

// ---- circle: a user-defined solver through the same LO-MSAC (user-functor parity test) ----------
void RLO(lomsac_circle)(const rlo_options* o, int n, const double* xy, double* model,
                        rlo_stats* st, int32_t* inliers) {
  CircleSolver s(n, xy);
  run_lomsac<Circle, Circles, CircleSolver>(*o, s, reinterpret_cast<Circle*>(model), st, inliers);
}

// ---- hybrid line ---------------------------------------------------------------
void RLO(hybrid_line_score_models)(int n0, const double* xy0, int n1, const double* p1,
                                   int k, const double* models, const double* thr,
                                   const double* w, double* scores, int32_t* counts) {
  HybridLineSolver s(n0, xy0, n1, p1, {0.5, 0.5});
  std::vector<double> thrv = {thr[0], thr[1]};
  std::vector<int> nd = {n0, n1};
  const Line* L = reinterpret_cast<const Line*>(models);
  std::vector<std::vector<int>> inl;
#ifdef RLO_USE_REFERENCE
  OpenHybrid<Line, Lines, HybridLineSolver> r;
  ransac_lib::HybridLORansacOptions ho;
  ho.squared_inlier_thresholds_ = thrv;
  ho.data_type_weights_ = {w[0], w[1]};
  for (int m = 0; m < k; ++m) {
    r.ScoreModel(ho, s, L[m], thrv, 2, nd, &scores[m]);
    if (counts) counts[m] = r.GetInliers(s, L[m], thrv, &inl);
  }
#else
  HybridOptions ho;
  ho.sq_thr = thrv;
  ho.weights = {w[0], w[1]};
  for (int m = 0; m < k; ++m) {
    scores[m] = hybrid_score(ho, s, L[m], thrv, nd);
    if (counts) counts[m] = hybrid_inliers(s, L[m], thrv, &inl);
  }
#endif
}

// idx0 / idx1 receive the per-type ascending lists; returns total, writes per-type counts.
int RLO(hybrid_line_inliers)(int n0, const double* xy0, int n1, const double* p1,
                             const double* model, const double* thr, int32_t* idx0,
                             int32_t* idx1, int32_t* counts) {
  HybridLineSolver s(n0, xy0, n1, p1, {0.5, 0.5});
  std::vector<double> thrv = {thr[0], thr[1]};
  std::vector<std::vector<int>> inl;
  const Line& L = *reinterpret_cast<const Line*>(model);
#ifdef RLO_USE_REFERENCE
  OpenHybrid<Line, Lines, HybridLineSolver> r;
  const int c = r.GetInliers(s, L, thrv, &inl);
#else
  const int c = hybrid_inliers(s, L, thrv, &inl);
#endif
  for (size_t i = 0; i < inl[0].size(); ++i) idx0[i] = inl[0][i];
  for (size_t i = 0; i < inl[1].size(); ++i) idx1[i] = inl[1][i];
  counts[0] = static_cast<int>(inl[0].size());
  counts[1] = static_cast<int>(inl[1].size());
  return c;
}

void RLO(point_normal_batch)(int n1, const double* p1, int k, const int32_t* samples,
                             double* models, uint8_t* valid) {
  for (int i = 0; i < k; ++i) {
    Line l;
    valid[i] = HybridLineSolver::PointNormal(p1, samples[i], &l) ? 1 : 0;
    for (int j = 0; j < 3; ++j) models[3 * i + j] = valid[i] ? l.v[j] : 0.0;
  }
  (void)n1;
}

}  // extern "C"

template <class RefSampler, class OracleSampler>
static void run_hybrid_line(const rlo_hybrid_options* o, const HybridLineSolver& s, double* model,
                            rlo_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  Line* L = reinterpret_cast<Line*>(model);
#ifdef RLO_USE_REFERENCE
  ransac_lib::HybridLocallyOptimizedMSAC<Line, Lines, HybridLineSolver, RefSampler> r;
  ransac_lib::HybridRansacStatistics hs;
  r.EstimateModel(to_ref(*o), s, L, &hs);
#else
  HybridStats hs;
  hybrid_lomsac_estimate<Line, Lines, HybridLineSolver, OracleSampler>(to_oracle(*o), s, L, &hs);
#endif
  st->num_iterations_total = hs.num_iterations_total;
  for (int i = 0; i < 2; ++i) {
    st->num_iterations_per_solver[i] =
        i < (int)hs.num_iterations_per_solver.size() ? hs.num_iterations_per_solver[i] : 0;
    st->inlier_ratios[i] = i < (int)hs.inlier_ratios.size() ? hs.inlier_ratios[i] : 0.0;
  }
  st->best_num_inliers = hs.best_num_inliers;
  st->best_solver_type = hs.best_solver_type;
  st->best_model_score = hs.best_model_score;
  st->number_lo_iterations = hs.number_lo_iterations;
  const std::vector<std::vector<int>>& ii = hs.inlier_indices;
  st->num_inliers_per_type[0] = ii.size() > 0 ? (int)ii[0].size() : 0;
  st->num_inliers_per_type[1] = ii.size() > 1 ? (int)ii[1].size() : 0;
  if (inl0 && ii.size() > 0)
    for (size_t i = 0; i < ii[0].size(); ++i) inl0[i] = ii[0][i];
  if (inl1 && ii.size() > 1)
    for (size_t i = 0; i < ii[1].size(); ++i) inl1[i] = ii[1][i];
}

// The same for the hybrid pose solver (solvers_oracle.h: HybridPoseSolver).
static void run_hybrid_pose(const rlo_hybrid_options* o, const HybridPoseSolver& s, double* model,
                            rlo_hybrid_stats* st, int32_t* inl0, int32_t* inl1) {
  Pose34* P = reinterpret_cast<Pose34*>(model);
#ifdef RLO_USE_REFERENCE
  ransac_lib::HybridLocallyOptimizedMSAC<Pose34, Poses, HybridPoseSolver> r;
  ransac_lib::HybridRansacStatistics hs;
  r.EstimateModel(to_ref(*o), s, P, &hs);
#else
  HybridStats hs;
  hybrid_lomsac_estimate<Pose34, Poses, HybridPoseSolver, HybridUniformSampler>(to_oracle(*o), s, P, &hs);
#endif
  st->num_iterations_total = hs.num_iterations_total;
  for (int i = 0; i < 2; ++i) {
    st->num_iterations_per_solver[i] =
        i < (int)hs.num_iterations_per_solver.size() ? hs.num_iterations_per_solver[i] : 0;
    st->inlier_ratios[i] = i < (int)hs.inlier_ratios.size() ? hs.inlier_ratios[i] : 0.0;
  }
  st->best_num_inliers = hs.best_num_inliers;
  st->best_solver_type = hs.best_solver_type;
  st->best_model_score = hs.best_model_score;
  st->number_lo_iterations = hs.number_lo_iterations;
  const std::vector<std::vector<int>>& ii = hs.inlier_indices;
  st->num_inliers_per_type[0] = ii.size() > 0 ? (int)ii[0].size() : 0;
  st->num_inliers_per_type[1] = ii.size() > 1 ? (int)ii[1].size() : 0;
  if (inl0 && ii.size() > 0)
    for (size_t i = 0; i < ii[0].size(); ++i) inl0[i] = ii[0][i];
  if (inl1 && ii.size() > 1)
    for (size_t i = 0; i < ii[1].size(); ++i) inl1[i] = ii[1][i];
}

#ifdef RLO_USE_REFERENCE
typedef ransac_lib::HybridUniformSampling<HybridLineSolver> RefUniformS;
typedef ransac_lib::HybridBiasedSampling<HybridLineSolver> RefBiasedS;
#else
typedef void RefUniformS;
typedef void RefBiasedS;
#endif

extern "C" {

void RLO(hybrid_lomsac_line)(const rlo_hybrid_options* o, int n0, const double* xy0,
                             int n1, const double* p1, const double* priors,
                             double* model, rlo_hybrid_stats* st, int32_t* inl0,
                             int32_t* inl1) {
  HybridLineSolver s(n0, xy0, n1, p1, {priors[0], priors[1]});
  run_hybrid_line<RefUniformS, HybridUniformSampler>(o, s, model, st, inl0, inl1);
}

// Hybrid pose: 2D-3D matches (type 0) and 3D-3D matches (type 1), P3P and three-point solvers.
void RLO(hybrid_lomsac_pose)(const rlo_hybrid_options* o, double fx, double fy, int n0, const double* p2d,
                             const double* p3d, int n1, const double* cam1, const double* world1,
                             const double* priors, double* model, rlo_hybrid_stats* st, int32_t* inl0,
                             int32_t* inl1) {
  std::vector<double> rays(3 * static_cast<size_t>(n0));
  PoseSolver::PixelsToViewingRays(fx, fy, n0, p2d, rays.data());
  HybridPoseSolver s(fx, fy, o->sq_thr[0], n0, p2d, rays.data(), p3d, n1, cam1, world1, {priors[0], priors[1]});
  run_hybrid_pose(o, s, model, st, inl0, inl1);
}

// One model against both data types (scores with the per-type thresholds / weights, inlier counts per type).
void RLO(hybrid_pose_errors)(const double* model, double fx, double fy, int n0, const double* p2d,
                             const double* p3d, int n1, const double* cam1, const double* world1, double* e0,
                             double* e1) {
  std::vector<double> rays(3 * static_cast<size_t>(std::max(n0, 1)));
  HybridPoseSolver s(fx, fy, 0.0, n0, p2d, rays.data(), p3d, n1, cam1, world1, {0.5, 0.5});
  Pose34 P;
  std::memcpy(P.m, model, sizeof(P.m));
  for (int i = 0; i < n0; ++i) e0[i] = s.EvaluateModelOnPoint(P, 0, i);
  for (int i = 0; i < n1; ++i) e1[i] = s.EvaluateModelOnPoint(P, 1, i);
}

// HybridBiasedSampling as the Sampler template argument; w0 / w1: per-point sampling weights.
void RLO(hybrid_lomsac_line_biased)(const rlo_hybrid_options* o, int n0, const double* xy0,
                                    const double* w0, int n1, const double* p1, const double* w1,
                                    const double* priors, double* model, rlo_hybrid_stats* st,
                                    int32_t* inl0, int32_t* inl1) {
  HybridLineSolver s(n0, xy0, n1, p1, {priors[0], priors[1]});
  s.set_weights(std::vector<double>(w0, w0 + n0), std::vector<double>(w1, w1 + n1));
  run_hybrid_line<RefBiasedS, HybridBiasedSampler>(o, s, model, st, inl0, inl1);
}

}  // extern "C"
