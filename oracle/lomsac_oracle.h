// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// CPU restatement of the reference's hot path and of the sequential control
// flow around it, written as free functions over the Solver concept. Each
// function cites the reference lines it follows (paths relative to
// /root/reference). Pinned against oracle/_ref (the reference's own unmodified
// headers compiled here) through tests/golden/*.json.
#ifndef RLB_ORACLE_LOMSAC_ORACLE_H_
#define RLB_ORACLE_LOMSAC_ORACLE_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <numeric>
#include <random>
#include <utility>
#include <vector>

namespace rlb_oracle {

// ---- options / statistics (values of RansacLib/ransac.h:47-101) -------------
struct LoMsacOptions {
  uint32_t min_iters = 100u;
  uint32_t max_iters = 10000u;
  double success_prob = 0.9999;
  double sq_thr = 1.0;
  unsigned seed = 0u;
  int lo_steps = 10;
  double thr_mult = 1.4142135623730951;  // std::sqrt(2.0)
  int lsq_iters = 4;
  int min_sample_mult = 7;
  int non_min_sample_mult = 3;
  uint32_t lo_start = 50u;
  bool final_lsq = false;
};

struct LoMsacStats {
  uint32_t num_iterations = 0;
  int best_num_inliers = 0;
  double best_model_score = std::numeric_limits<double>::max();
  double inlier_ratio = 0.0;
  std::vector<int> inlier_indices;
  int number_lo_iterations = 0;
};

// ---- RansacLib/utils.h -------------------------------------------------------
// utils.h:48-56 Fisher-Yates with a fresh distribution per position.
template <class T>
inline void fisher_yates(std::mt19937* rng, std::vector<T>* v) {
  const int n = static_cast<int>(v->size());
  for (int i = 0; i + 1 < n; ++i) {
    std::uniform_int_distribution<int> d(i, n - 1);
    std::swap((*v)[i], (*v)[d(*rng)]);
  }
}
// utils.h:69-73
inline void shuffle_and_resize(int target, std::mt19937* rng, std::vector<int>* v) {
  fisher_yates(rng, v);
  v->resize(target);
}
// utils.h:96-105 (per-type sizes)
inline void shuffle_and_resize(const std::vector<int>& sizes, std::mt19937* rng,
                               std::vector<std::vector<int>>* v) {
  for (size_t t = 0; t < v->size(); ++t) {
    const int have = static_cast<int>((*v)[t].size());
    shuffle_and_resize(std::min(have, sizes[t]), rng, &(*v)[t]);
  }
}
// utils.h:110-140
inline uint32_t num_required_iterations(double inlier_ratio, double prob_miss,
                                        int sample_size, uint32_t min_it,
                                        uint32_t max_it) {
  if (inlier_ratio <= 0.0) return max_it;
  if (inlier_ratio >= 1.0) return min_it;
  const double p_bad = 1.0 - std::pow(inlier_ratio, static_cast<double>(sample_size));
  if (p_bad >= 0.99999999999999) return max_it;
  const double it = std::ceil(std::log(prob_miss) / std::log(p_bad) + 0.5);
  uint32_t r = std::min(static_cast<uint32_t>(it), max_it);
  return std::max(min_it, r);
}
// utils.h:142-178
inline uint32_t num_required_iterations(const std::vector<double>& ratios,
                                        double prob_miss,
                                        const std::vector<int>& sizes,
                                        uint32_t min_it, uint32_t max_it) {
  double p_good = 1.0;
  for (size_t i = 0; i < sizes.size(); ++i)
    p_good *= std::pow(ratios[i], static_cast<double>(sizes[i]));
  if (p_good <= 0.0) return max_it;
  if (p_good >= 1.0) return min_it;
  const double p_bad = 1.0 - p_good;
  if (p_bad >= 0.99999999999999) return max_it;
  const double it = std::ceil(std::log(prob_miss) / std::log(p_bad) + 0.5);
  uint32_t r = std::min(static_cast<uint32_t>(it), max_it);
  return std::max(min_it, r);
}

// ---- RansacLib/sampling.h:47-135 ----------------------------------------------
class UniformSampler {
 public:
  UniformSampler(unsigned seed, int num_data, int sample_size)
      : n_(num_data), s_(sample_size), dist_(0, num_data - 1) {
    rng_.seed(seed);
    // sampling.h:69-78: rejection sampling iff N/(N-s) < e
    draw_ = (static_cast<double>(n_) / static_cast<double>(n_ - s_)) < M_E;
  }
  void Sample(std::vector<int>* out) {
    if (draw_) {
      out->resize(s_);
      for (int i = 0; i < s_; ++i) {  // sampling.h:81-97
        for (;;) {
          (*out)[i] = dist_(rng_);
          bool dup = false;
          for (int j = 0; j < i; ++j) dup = dup || ((*out)[j] == (*out)[i]);
          if (!dup) break;
        }
      }
    } else {  // sampling.h:102-110
      out->resize(n_);
      std::iota(out->begin(), out->end(), 0);
      if (s_ == n_) return;
      fisher_yates(&rng_, out);
      out->resize(s_);
    }
  }

 private:
  std::mt19937 rng_;
  int n_, s_;
  bool draw_;
  std::uniform_int_distribution<int> dist_;
};

// ---- RansacLib/hybrid_sampling.h:47-140 ---------------------------------------
class HybridUniformSampler {
 public:
  HybridUniformSampler(unsigned seed, const std::vector<int>& num_data)
      : n_(num_data) {
    rng_.seed(seed);
    dist_.resize(n_.size());
    for (size_t t = 0; t < n_.size(); ++t) {
      if (n_[t] < 2) continue;  // hybrid_sampling.h:58
      dist_[t].param(std::uniform_int_distribution<int>::param_type(0, n_[t] - 1));
    }
  }
  void Sample(const std::vector<int>& sizes, std::vector<std::vector<int>>* out) {
    out->resize(n_.size());
    for (size_t t = 0; t < n_.size(); ++t) {
      std::vector<int>& s = (*out)[t];
      s.clear();
      if (sizes[t] <= 0) continue;
      const double coeff = static_cast<double>(n_[t]) /
                           static_cast<double>(n_[t] - sizes[t]);
      if (coeff < M_E) {  // hybrid_sampling.h:87-104
        s.resize(sizes[t]);
        for (int i = 0; i < sizes[t]; ++i) {
          for (;;) {
            s[i] = dist_[t](rng_);
            bool dup = false;
            for (int j = 0; j < i; ++j) dup = dup || (s[j] == s[i]);
            if (!dup) break;
          }
        }
      } else {  // hybrid_sampling.h:109-120
        s.resize(n_[t]);
        std::iota(s.begin(), s.end(), 0);
        if (sizes[t] == n_[t]) continue;
        fisher_yates(&rng_, &s);
        s.resize(sizes[t]);
      }
    }
  }

 private:
  std::mt19937 rng_;
  std::vector<int> n_;
  std::vector<std::uniform_int_distribution<int>> dist_;
};

// ---- RansacLib/hybrid_sampling.h:146-223 (HybridBiasedSampling) ----------------
// Positive-weight points only; std::discrete_distribution on the sampler's own mt19937;
// duplicates inside one sample re-drawn; "all candidates" short-cut without a draw.
class HybridBiasedSampler {
 public:
  HybridBiasedSampler(unsigned seed, const std::vector<std::vector<double>>& weights) {
    rng_.seed(seed);
    ids_.resize(weights.size());
    dist_.resize(weights.size());
    for (size_t t = 0; t < weights.size(); ++t) {
      std::vector<double> kept;
      for (size_t j = 0; j < weights[t].size(); ++j) {
        if (weights[t][j] > 0.0) {
          kept.push_back(weights[t][j]);
          ids_[t].push_back(static_cast<int>(j));
        }
      }
      std::discrete_distribution<int> d(kept.begin(), kept.end());
      dist_[t].param(d.param());
    }
  }
  void Sample(const std::vector<int>& sizes, std::vector<std::vector<int>>* out) {
    out->resize(ids_.size());
    for (size_t t = 0; t < ids_.size(); ++t) {
      std::vector<int>& s = (*out)[t];
      s.clear();
      if (sizes[t] <= 0) continue;
      if (sizes[t] == static_cast<int>(ids_[t].size())) {
        s = ids_[t];
        continue;
      }
      s.resize(sizes[t]);
      for (int i = 0; i < sizes[t]; ++i) {
        for (;;) {
          s[i] = ids_[t][dist_[t](rng_)];
          bool dup = false;
          for (int j = 0; j < i; ++j) dup = dup || (s[j] == s[i]);
          if (!dup) break;
        }
      }
    }
  }

 private:
  std::mt19937 rng_;
  std::vector<std::discrete_distribution<int>> dist_;
  std::vector<std::vector<int>> ids_;
};

// ---- the hot path --------------------------------------------------------------
// ransac.h:306-309. std::min(a,b) == (b < a) ? b : a : NaN errors propagate.
inline double msac_cost(double err, double thr) { return (thr < err) ? thr : err; }

// ransac.h:295-303: sequential f64 sum in index order.
template <class Solver, class Model>
inline double score_model(const Solver& s, const Model& m, double thr) {
  const int n = s.num_data();
  double score = 0.0;
  for (int i = 0; i < n; ++i) score += msac_cost(s.EvaluateModelOnPoint(m, i), thr);
  return score;
}

// ransac.h:311-336: strict '<', ascending list.
template <class Solver, class Model>
inline int get_inliers(const Solver& s, const Model& m, double thr,
                       std::vector<int>* inl) {
  const int n = s.num_data();
  if (inl) inl->clear();
  int c = 0;
  for (int i = 0; i < n; ++i) {
    if (s.EvaluateModelOnPoint(m, i) < thr) {
      ++c;
      if (inl) inl->push_back(i);
    }
  }
  return c;
}

// ransac.h:278-293: argmin with strict '<' (lowest id wins ties).
template <class Solver, class ModelVector>
inline void best_model_id(const Solver& s, const ModelVector& models, int num,
                          double thr, double* best_score, int* best_id) {
  *best_score = std::numeric_limits<double>::max();
  *best_id = 0;
  for (int m = 0; m < num; ++m) {
    const double sc = score_model(s, models[m], thr);
    if (sc < *best_score) {
      *best_score = sc;
      *best_id = m;
    }
  }
}

// ransac.h:422-428
template <class Model>
inline void keep_better(double sc, const Model& m, double* best_sc, Model* best) {
  if (sc < *best_sc) {
    *best_sc = sc;
    *best = m;
  }
}

// ransac.h:409-420
template <class Solver, class Model>
inline void lsq_fit(const LoMsacOptions& o, double thr, const Solver& s,
                    std::mt19937* rng, Model* m) {
  const int cap = o.min_sample_mult * s.min_sample_size();
  std::vector<int> inl;
  const int ni = get_inliers(s, *m, thr, &inl);
  if (ni < s.min_sample_size()) return;
  shuffle_and_resize(std::min(cap, ni), rng, &inl);
  s.LeastSquares(inl, m);
}

// ransac.h:341-407
template <class Solver, class Model>
inline void local_optimization(const LoMsacOptions& o, const Solver& s,
                               std::mt19937* rng, Model* best, double* best_sc) {
  const int n = s.num_data();
  const int min_non_min = s.non_minimal_sample_size();
  if (min_non_min > n) return;
  const int min_size = s.min_sample_size();
  const double thr = o.sq_thr;
  const double mult = o.thr_mult;

  Model m_init = *best;
  lsq_fit(o, thr * mult, s, rng, &m_init);
  double sc = score_model(s, m_init, thr);
  keep_better(sc, m_init, best_sc, best);

  std::vector<int> base;
  get_inliers(s, m_init, thr * mult, &base);

  const int non_min_size =
      std::max(min_non_min, std::min(min_size * o.non_min_sample_mult,
                                     static_cast<int>(base.size()) / 2));
  std::vector<int> sample;
  for (int r = 0; r < o.lo_steps; ++r) {
    sample = base;
    shuffle_and_resize(non_min_size, rng, &sample);
    Model m;
    if (!s.NonMinimalSolver(sample, &m)) continue;
    sc = score_model(s, m, thr);
    keep_better(sc, m, best_sc, best);
    lsq_fit(o, thr, s, rng, &m);
    double t = mult * thr;
    // ransac.h:395-397: divides by (lsq_iters - 1) as an int (inf if 1).
    const double dt = (mult - 1.0) * thr / static_cast<int>(o.lsq_iters - 1);
    for (int i = 0; i < o.lsq_iters; ++i) {
      lsq_fit(o, t, s, rng, &m);
      sc = score_model(s, m, thr);
      keep_better(sc, m, best_sc, best);
      t -= dt;
    }
  }
}

// ransac.h:128-275
template <class Model, class ModelVector, class Solver>
inline int lomsac_estimate(const LoMsacOptions& o, const Solver& s, Model* best,
                           LoMsacStats* st) {
  *st = LoMsacStats();
  const int ss = s.min_sample_size();
  const int n = s.num_data();
  if (ss > n || ss <= 0) return 0;

  UniformSampler sampler(o.seed, n, ss);
  std::mt19937 rng;
  rng.seed(o.seed);
  uint32_t max_it = std::max(o.max_iters, o.min_iters);
  const double thr = o.sq_thr;
  const double kMax = std::numeric_limits<double>::max();

  Model best_min;
  double best_min_sc = kMax;
  std::vector<int> sample(ss);
  ModelVector models;

  auto refresh_termination = [&]() {
    st->best_num_inliers = get_inliers(s, *best, thr, &st->inlier_indices);
    st->inlier_ratio = static_cast<double>(st->best_num_inliers) / static_cast<double>(n);
    max_it = num_required_iterations(st->inlier_ratio, 1.0 - o.success_prob, ss,
                                     o.min_iters, o.max_iters);
  };

  for (st->num_iterations = 0u; st->num_iterations < max_it; ++st->num_iterations) {
    if (st->num_iterations == o.lo_start && best_min_sc < kMax) {  // 163-178
      ++st->number_lo_iterations;
      local_optimization(o, s, &rng, best, &st->best_model_score);
      refresh_termination();
    }
    sampler.Sample(&sample);
    const int nm = s.MinimalSolver(sample, &models);
    if (nm <= 0) continue;
    double loc_sc;
    int loc_id;
    best_model_id(s, models, nm, thr, &loc_sc, &loc_id);
    if (loc_sc < best_min_sc || st->num_iterations == o.lo_start) {  // 195-196
      const bool new_best = loc_sc < best_min_sc;
      if (new_best) {
        best_min_sc = loc_sc;
        best_min = models[loc_id];
        keep_better(best_min_sc, best_min, &st->best_model_score, best);
      }
      const bool run_lo = st->num_iterations >= o.lo_start && best_min_sc < kMax;
      if (!new_best && !run_lo) continue;
      if (run_lo) {
        ++st->number_lo_iterations;
        double sc = best_min_sc;
        local_optimization(o, s, &rng, &best_min, &sc);
        keep_better(sc, best_min, &st->best_model_score, best);
      }
      refresh_termination();
    }
  }
  if (st->num_iterations <= o.lo_start && st->best_model_score < kMax) {  // 245-255
    ++st->number_lo_iterations;
    local_optimization(o, s, &rng, best, &st->best_model_score);
    st->best_num_inliers = get_inliers(s, *best, thr, &st->inlier_indices);
    st->inlier_ratio = static_cast<double>(st->best_num_inliers) / static_cast<double>(n);
  }
  if (o.final_lsq) {  // 257-272
    Model refined = *best;
    s.LeastSquares(st->inlier_indices, &refined);
    const double sc = score_model(s, refined, thr);
    if (sc < st->best_model_score) {
      st->best_model_score = sc;
      *best = refined;
      st->best_num_inliers = get_inliers(s, *best, thr, &st->inlier_indices);
      st->inlier_ratio = static_cast<double>(st->best_num_inliers) / static_cast<double>(n);
    }
  }
  return st->best_num_inliers;
}

// ================================================================================
// Hybrid LO-MSAC: RansacLib/hybrid_ransac.h
struct HybridOptions {
  uint32_t min_iters = 100u;
  uint32_t max_iters = 10000u;
  uint32_t max_iters_per_solver = 10000u;
  double success_prob = 0.9999;
  std::vector<double> sq_thr;
  std::vector<double> weights;
  unsigned seed = 0u;
  int lo_steps = 10;
  double thr_mult = 1.4142135623730951;
  int lsq_iters = 4;
  int min_sample_mult = 7;
  uint32_t lo_start = 50u;
  bool final_lsq = false;
};

struct HybridStats {
  uint32_t num_iterations_total = 0;
  std::vector<uint32_t> num_iterations_per_solver;
  int best_num_inliers = 0;
  int best_solver_type = -1;
  double best_model_score = std::numeric_limits<double>::max();
  std::vector<double> inlier_ratios;
  std::vector<std::vector<int>> inlier_indices;
  int number_lo_iterations = 0;
};

// hybrid_ransac.h:407-421: sum_t w_t sum_i min(err, thr_t), weight applied per term.
template <class Solver, class Model>
inline double hybrid_score(const HybridOptions& o, const Solver& s, const Model& m,
                           const std::vector<double>& thr,
                           const std::vector<int>& nd) {
  double score = 0.0;
  for (size_t t = 0; t < nd.size(); ++t)
    for (int i = 0; i < nd[t]; ++i)
      score += msac_cost(s.EvaluateModelOnPoint(m, static_cast<int>(t), i), thr[t]) *
               o.weights[t];
  return score;
}

// hybrid_ransac.h:429-462 (non-null branch; the nullptr branch is dead code).
template <class Solver, class Model>
inline int hybrid_inliers(const Solver& s, const Model& m,
                          const std::vector<double>& thr,
                          std::vector<std::vector<int>>* inl) {
  std::vector<int> nd;
  s.num_data(&nd);
  inl->clear();
  inl->resize(nd.size());
  int c = 0;
  for (size_t t = 0; t < nd.size(); ++t) {
    for (int i = 0; i < nd[t]; ++i) {
      if (s.EvaluateModelOnPoint(m, static_cast<int>(t), i) < thr[t]) {
        ++c;
        (*inl)[t].push_back(i);
      }
    }
  }
  return c;
}

template <class Model>
inline void hybrid_keep_better(double sc, const Model& m, int type, double* best_sc,
                               Model* best, int* best_type) {  // :597-605
  if (sc < *best_sc) {
    *best_sc = sc;
    *best = m;
    *best_type = type;
  }
}

// hybrid_ransac.h:464-495
template <class Solver, class Model>
inline void hybrid_update_termination(const HybridOptions& o, const Solver& s,
                                      const Model& m, HybridStats* st,
                                      std::vector<uint32_t>* max_per_solver) {
  st->best_num_inliers = hybrid_inliers(s, m, o.sq_thr, &st->inlier_indices);
  std::vector<int> nd;
  s.num_data(&nd);
  for (size_t d = 0; d < nd.size(); ++d) {
    st->inlier_ratios[d] =
        nd[d] > 0 ? static_cast<double>(st->inlier_indices[d].size()) /
                        static_cast<double>(nd[d])
                  : 0.0;
  }
  std::vector<std::vector<int>> sizes;
  s.min_sample_sizes(&sizes);
  for (int k = 0; k < s.num_minimal_solvers(); ++k) {
    (*max_per_solver)[k] = num_required_iterations(
        st->inlier_ratios, 1.0 - o.success_prob, sizes[k], o.min_iters,
        o.max_iters_per_solver);
  }
}

// hybrid_ransac.h:569-595
template <class Solver, class Model>
inline void hybrid_lsq_fit(const HybridOptions& o, const std::vector<double>& thr,
                           int solver_type, const Solver& s, std::mt19937* rng,
                           Model* m) {
  std::vector<std::vector<int>> sizes;
  s.min_sample_sizes(&sizes);
  std::vector<std::vector<int>> inl;
  hybrid_inliers(s, *m, thr, &inl);
  for (int t = 0; t < s.num_data_types(); ++t) {
    const int have = static_cast<int>(inl[t].size());
    if (have < sizes[solver_type][t]) return;
    sizes[solver_type][t] = std::min(sizes[solver_type][t] * o.min_sample_mult, have);
  }
  shuffle_and_resize(sizes[solver_type], rng, &inl);
  s.LeastSquares(inl, m);
}

// hybrid_ransac.h:500-567. Keeps the quirk that the relaxed thresholds are
// initialised once and keep shrinking over ALL lo_steps (they turn negative).
template <class Solver, class Model>
inline void hybrid_local_optimization(const HybridOptions& o, const Solver& s,
                                      int solver_type, std::mt19937* rng,
                                      Model* best, double* best_sc, int* best_type) {
  std::vector<int> nd;
  s.num_data(&nd);
  const int T = static_cast<int>(nd.size());
  std::vector<double> thr = o.sq_thr;
  std::vector<double> dthr(thr.size());
  for (int t = 0; t < T; ++t) {
    dthr[t] = (o.thr_mult - 1.0) * thr[t] / static_cast<int>(o.lsq_iters - 1);
    thr[t] *= o.thr_mult;
  }
  Model m_init = *best;
  hybrid_lsq_fit(o, thr, solver_type, s, rng, &m_init);
  double sc = hybrid_score(o, s, m_init, o.sq_thr, nd);
  hybrid_keep_better(sc, m_init, solver_type, best_sc, best, best_type);
  std::vector<std::vector<int>> base;
  hybrid_inliers(s, m_init, o.sq_thr, &base);  // :534-536 (result unused, pass kept)
  for (int r = 0; r < o.lo_steps; ++r) {
    Model m = m_init;
    hybrid_lsq_fit(o, o.sq_thr, solver_type, s, rng, &m);
    for (int i = 0; i < o.lsq_iters; ++i) {
      hybrid_lsq_fit(o, thr, solver_type, s, rng, &m);
      sc = hybrid_score(o, s, m, o.sq_thr, nd);
      hybrid_keep_better(sc, m, solver_type, best_sc, best, best_type);
      for (int t = 0; t < T; ++t) thr[t] -= dthr[t];
    }
  }
}

// hybrid_ransac.h:324-384
template <class Solver>
inline int hybrid_select_solver(const Solver& s, const std::vector<double>& priors,
                                const HybridStats& st, uint32_t min_iters,
                                std::mt19937* rng) {
  const int S = static_cast<int>(priors.size());
  std::vector<std::vector<int>> sizes;
  s.min_sample_sizes(&sizes);
  std::vector<double> p(S, 0.0);
  double sum = 0.0;
  const int T = s.num_data_types();
  const double sum_ratios =
      std::accumulate(st.inlier_ratios.begin(), st.inlier_ratios.end(), 0.0);
  if (sum_ratios == 0.0) {
    for (int i = 0; i < S; ++i) {
      p[i] = priors[i];
      sum += p[i];
    }
  } else {
    for (int i = 0; i < S; ++i) {
      double it = static_cast<double>(st.num_iterations_per_solver[i]);
      if (it > 0.0) it -= 1.0;
      double all_in = 1.0;
      for (int j = 0; j < T; ++j)
        all_in *= std::pow(st.inlier_ratios[j], static_cast<double>(sizes[i][j]));
      if (it < static_cast<double>(min_iters)) {
        p[i] = all_in * priors[i];
      } else {
        p[i] = all_in * std::pow(1.0 - all_in, it) * priors[i];
      }
      sum += p[i];
    }
  }
  std::uniform_real_distribution<double> d(0.0, sum);
  const double x = d(*rng);
  double acc = 0.0;
  for (int i = 0; i < S; ++i) {
    if (priors[i] == 0.0) continue;
    acc += p[i];
    if (x <= acc) return i;
  }
  return -1;
}

// hybrid_ransac.h:135-320
// Sampler construction differs per sampler type (the reference's take (seed, solver)).
template <class Sampler>
struct MakeHybridSampler;
template <>
struct MakeHybridSampler<HybridUniformSampler> {
  template <class Solver>
  static HybridUniformSampler make(unsigned seed, const Solver& s) {
    std::vector<int> nd;
    s.num_data(&nd);
    return HybridUniformSampler(seed, nd);
  }
};
template <>
struct MakeHybridSampler<HybridBiasedSampler> {
  template <class Solver>
  static HybridBiasedSampler make(unsigned seed, const Solver& s) {
    std::vector<std::vector<double>> w;
    s.get_weights(w);
    return HybridBiasedSampler(seed, w);
  }
};

template <class Model, class ModelVector, class Solver, class Sampler = HybridUniformSampler>
inline int hybrid_lomsac_estimate(const HybridOptions& o, const Solver& s, Model* best,
                                  HybridStats* st) {
  *st = HybridStats();
  const int S = s.num_minimal_solvers();
  const int T = s.num_data_types();
  st->num_iterations_per_solver.assign(S, 0u);
  st->inlier_ratios.assign(T, 0.0);
  st->inlier_indices.resize(T);
  std::vector<double> priors;
  s.solver_probabilities(&priors);
  std::vector<std::vector<int>> sizes;
  s.min_sample_sizes(&sizes);
  std::vector<int> nd;
  s.num_data(&nd);
  {  // VerifyData, :610-638
    for (int i = 0; i < S; ++i)
      for (int j = 0; j < T; ++j)
        if (sizes[i][j] > nd[j] || sizes[i][j] < 0) {
          priors[i] = 0.0;
          break;
        }
    bool any = false;
    for (int i = 0; i < S; ++i) any = any || priors[i] > 0.0;
    if (!any) return 0;
  }
  Sampler sampler = MakeHybridSampler<Sampler>::make(o.seed, s);
  uint32_t max_it = std::max(o.max_iters, o.min_iters);
  std::vector<uint32_t> max_per(S, std::max(o.max_iters_per_solver, o.min_iters));
  const double kMax = std::numeric_limits<double>::max();
  Model best_min;
  double best_min_sc = kMax;
  std::vector<std::vector<int>> sample(T);
  ModelVector models;
  std::mt19937 rng;
  rng.seed(o.seed);

  for (st->num_iterations_total = 0u; st->num_iterations_total < max_it;
       ++st->num_iterations_total) {
    if (st->num_iterations_total == o.lo_start && best_min_sc < kMax) {
      ++st->number_lo_iterations;
      hybrid_local_optimization(o, s, st->best_solver_type, &rng, best,
                                &st->best_model_score, &st->best_solver_type);
      hybrid_update_termination(o, s, *best, st, &max_per);
    }
    const int type = hybrid_select_solver(s, priors, *st, o.min_iters, &rng);
    if (type < 0) break;  // reference: `< -1` can never fire and -1 is UB; we stop.
    st->num_iterations_per_solver[type] += 1;
    sampler.Sample(sizes[type], &sample);
    const int nm = s.MinimalSolver(sample, type, &models);
    if (nm > 0) {
      double loc_sc = kMax;
      int loc_id = 0;
      for (int m = 0; m < nm; ++m) {  // :386-405
        const double sc = hybrid_score(o, s, models[m], o.sq_thr, nd);
        if (sc < loc_sc) {
          loc_sc = sc;
          loc_id = m;
        }
      }
      if (loc_sc < best_min_sc || st->num_iterations_total == o.lo_start) {
        const bool new_best = loc_sc < best_min_sc;
        if (new_best) {
          best_min_sc = loc_sc;
          best_min = models[loc_id];
          hybrid_keep_better(best_min_sc, best_min, type, &st->best_model_score, best,
                             &st->best_solver_type);
        }
        const bool run_lo = st->num_iterations_total >= o.lo_start && best_min_sc < kMax;
        if (!new_best && !run_lo) continue;  // NB: skips the per-solver cap check (:247)
        if (run_lo) {
          ++st->number_lo_iterations;
          double sc = best_min_sc;
          hybrid_local_optimization(o, s, st->best_solver_type, &rng, &best_min, &sc,
                                    &st->best_solver_type);
          hybrid_keep_better(sc, best_min, type, &st->best_model_score, best,
                             &st->best_solver_type);
        }
        hybrid_update_termination(o, s, *best, st, &max_per);
      }
    }
    if (st->num_iterations_per_solver[type] >= max_per[type]) break;  // :279-282
  }
  if (st->num_iterations_total <= o.lo_start && st->best_model_score < kMax) {
    ++st->number_lo_iterations;
    hybrid_local_optimization(o, s, st->best_solver_type, &rng, best,
                              &st->best_model_score, &st->best_solver_type);
    hybrid_update_termination(o, s, *best, st, &max_per);
  }
  if (o.final_lsq) {
    Model refined = *best;
    s.LeastSquares(st->inlier_indices, &refined);
    const double sc = hybrid_score(o, s, refined, o.sq_thr, nd);
    if (sc < st->best_model_score) {
      st->best_model_score = sc;
      *best = refined;
      hybrid_update_termination(o, s, *best, st, &max_per);
    }
  }
  return st->best_num_inliers;
}

}  // namespace rlb_oracle

#endif  // RLB_ORACLE_LOMSAC_ORACLE_H_
