// ransaclib_b200 — drop-in for RansacLib/hybrid_ransac.h (Hybrid LO-MSAC, Camposeco et al.
// CVPR 2018 + Lebeda et al. BMVC 2012). Same public surface as reference
// hybrid_ransac.h:47-137; all passes over the data run on the GPU.
//
// Speculation is harder than in ransac.h because the solver choice of every iteration is
// drawn from the SAME RNG that local optimisation consumes and depends on the current
// inlier ratios (reference hybrid_ransac.h:181-205, 324-384). The driver therefore
// speculates on COPIES of the RNG, the sampler and the per-solver counters: it simulates
// the next B iterations assuming no model update happens, sweeps all their hypotheses in
// one launch, and replays the reference's decisions with the real objects. A model update
// (the only thing that touches the RNG or the ratios) discards the rest of the batch and
// the next iteration re-speculates from the real state. Updates are O(log iterations).
#ifndef RANSACLIB_B200_RANSACLIB_HYBRID_RANSAC_H_
#define RANSACLIB_B200_RANSACLIB_HYBRID_RANSAC_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <numeric>
#include <random>
#include <vector>

#include "gpu_scorer.h"
#include "hybrid_sampling.h"
#include "utils.h"

namespace ransac_lib {

// reference hybrid_ransac.h:47-66
class HybridRansacOptions {
 public:
  HybridRansacOptions()
      : min_num_iterations_(100u),
        max_num_iterations_(10000u),
        max_num_iterations_per_solver_(10000u),
        success_probability_(0.9999),
        random_seed_(0u) {}
  uint32_t min_num_iterations_;
  uint32_t max_num_iterations_;
  uint32_t max_num_iterations_per_solver_;
  double success_probability_;
  std::vector<double> squared_inlier_thresholds_;
  std::vector<double> data_type_weights_;
  unsigned int random_seed_;
};

// reference hybrid_ransac.h:70-93
class HybridLORansacOptions : public HybridRansacOptions {
 public:
  HybridLORansacOptions()
      : num_lo_steps_(10),
        threshold_multiplier_(std::sqrt(2.0)),
        num_lsq_iterations_(4),
        min_sample_multiplicator_(7),
        lo_starting_iterations_(50u),
        final_least_squares_(false) {}
  int num_lo_steps_;
  double threshold_multiplier_;
  int num_lsq_iterations_;
  int min_sample_multiplicator_;
  uint32_t lo_starting_iterations_;
  bool final_least_squares_;
};

// reference hybrid_ransac.h:95-104
struct HybridRansacStatistics {
  uint32_t num_iterations_total;
  std::vector<uint32_t> num_iterations_per_solver;
  int best_num_inliers;
  int best_solver_type;
  double best_model_score;
  std::vector<double> inlier_ratios;
  std::vector<std::vector<int>> inlier_indices;
  int number_lo_iterations;
};

class HybridRansacBase {
 protected:
  void ResetStatistics(HybridRansacStatistics* s) const {
    s->num_iterations_total = 0u;
    s->num_iterations_per_solver.clear();
    s->best_num_inliers = 0;
    s->best_solver_type = -1;
    s->best_model_score = std::numeric_limits<double>::max();
    s->inlier_ratios.clear();
    s->inlier_indices.clear();
    s->number_lo_iterations = 0;
  }
};

template <class Model, class ModelVector, class HybridSolver,
          class Sampler = HybridUniformSampling<HybridSolver> >
class HybridLocallyOptimizedMSAC : public HybridRansacBase {
  static_assert(gpu::HasDeviceScorer<HybridSolver>::value,
                "ransaclib_b200 has no CPU scoring path: the HybridSolver must expose "
                "device_scorer(), PackModel() and UnpackModel() (see RansacLib/gpu_scorer.h)");

 public:
  int EstimateModel(const HybridLORansacOptions& options, const HybridSolver& solver,
                    Model* best_model, HybridRansacStatistics* statistics) const {
    ResetStatistics(statistics);
    HybridRansacStatistics& stats = *statistics;
    const int num_solvers = solver.num_minimal_solvers();
    const int num_types = solver.num_data_types();
    stats.num_iterations_per_solver.assign(num_solvers, 0u);
    stats.inlier_ratios.assign(num_types, 0.0);
    stats.inlier_indices.resize(num_types);

    std::vector<double> priors;
    solver.solver_probabilities(&priors);
    std::vector<std::vector<int>> sample_sizes;
    solver.min_sample_sizes(&sample_sizes);
    std::vector<int> num_data;
    solver.num_data(&num_data);
    if (!VerifyData(sample_sizes, num_data, num_solvers, num_types, &priors)) return 0;

    Sampler sampler(options.random_seed_, solver);
    uint32_t max_iters = std::max(options.max_num_iterations_, options.min_num_iterations_);
    std::vector<uint32_t> max_per_solver(
        num_solvers, std::max(options.max_num_iterations_per_solver_, options.min_num_iterations_));
    const double kNever = std::numeric_limits<double>::max();
    Model best_minimal_model;
    double best_minimal_score = kNever;
    utils::Mt19937 rng;
    rng.seed(options.random_seed_);
    const int precision = solver.device_scorer().precision();

    Batch batch;
    bool lists_stale = false;  // stats.inlier_indices lag behind *best_model
    uint32_t next_batch_size = 128u;
    std::vector<std::vector<int>> sample(num_types);

    for (stats.num_iterations_total = 0u; stats.num_iterations_total < max_iters;
         ++stats.num_iterations_total) {
      const uint32_t it = stats.num_iterations_total;
      if (it == options.lo_starting_iterations_ && best_minimal_score < kNever) {  // :191-201
        ++stats.number_lo_iterations;
        LocalOptimization(options, solver, stats.best_solver_type, &rng, best_model,
                          &stats.best_model_score, &stats.best_solver_type);
        UpdateRANSACTerminationCriteria(options, solver, *best_model, statistics, &max_per_solver, &lists_stale);
        batch.count = 0;  // RNG and ratios changed: speculation is stale
      }

      if (batch.count == 0 || it >= batch.first + batch.count) {
        const uint32_t want = std::min(next_batch_size, std::max(max_iters, it + 1u) - it);
        Speculate(options, solver, priors, sample_sizes, stats, max_per_solver, rng, sampler, it, want,
                  precision, &batch);
        next_batch_size = std::min(next_batch_size * 2u, 8192u);
      }
      const uint32_t slot = it - batch.first;

      // The real draw; by construction it equals the speculated one.
      const int type = SelectMinimalSolver(solver, priors, stats, options.min_num_iterations_, &rng);
      if (type < 0) break;  // reference: `< -1` can never fire and -1 would index out of range
      stats.num_iterations_per_solver[type] += 1;
      sampler.Sample(sample_sizes[type], &sample);
      if (type != batch.type[slot])
        throw std::logic_error("ransaclib_b200: hybrid speculation diverged from the replay");

      if (batch.num_models[slot] > 0) {
        double local_score = kNever;
        int local_id = -1;
        for (int m = 0; m < batch.num_models[slot]; ++m) {  // :386-405
          const double s = batch.scores[batch.model_offset[slot] + m];
          if (s < local_score) {
            local_score = s;
            local_id = batch.model_offset[slot] + m;
          }
        }
        const bool forced = (it == options.lo_starting_iterations_);
        if (precision == RLB_F32 && local_id >= 0 &&
            (forced || best_minimal_score == kNever || local_score <= best_minimal_score * (1.0 + 4e-4))) {
          Model cand;
          HybridSolver::UnpackModel(&batch.models[static_cast<size_t>(local_id) * batch.model_size], &cand);
          local_score = ScoreOnly(options, solver, cand);
        }
        if (local_score < best_minimal_score || forced) {  // :225-226
          const bool new_best = local_score < best_minimal_score;
          if (new_best) {
            best_minimal_score = local_score;
            HybridSolver::UnpackModel(&batch.models[static_cast<size_t>(local_id) * batch.model_size],
                                      &best_minimal_model);
            UpdateBestModel(best_minimal_score, best_minimal_model, type, &stats.best_model_score,
                            best_model, &stats.best_solver_type);
          }
          const bool run_lo = (it >= options.lo_starting_iterations_ && best_minimal_score < kNever);
          if (!new_best && !run_lo) continue;  // NB: skips the per-solver cap test (:247)
          if (run_lo) {
            ++stats.number_lo_iterations;
            double score = best_minimal_score;
            LocalOptimization(options, solver, stats.best_solver_type, &rng, &best_minimal_model, &score,
                              &stats.best_solver_type);
            UpdateBestModel(score, best_minimal_model, type, &stats.best_model_score, best_model,
                            &stats.best_solver_type);
          }
          UpdateRANSACTerminationCriteria(options, solver, *best_model, statistics, &max_per_solver,
                                          &lists_stale);
          batch.count = 0;
        }
      }
      if (stats.num_iterations_per_solver[type] >= max_per_solver[type]) break;  // :279-282
    }

    if (stats.num_iterations_total <= options.lo_starting_iterations_ &&
        stats.best_model_score < kNever) {  // :288-297
      ++stats.number_lo_iterations;
      LocalOptimization(options, solver, stats.best_solver_type, &rng, best_model,
                        &stats.best_model_score, &stats.best_solver_type);
      UpdateRANSACTerminationCriteria(options, solver, *best_model, statistics, &max_per_solver);
      lists_stale = false;
    }
    if (lists_stale)  // the lists of the last update inside the loop
      FusedPass(options, solver, *best_model, options.squared_inlier_thresholds_, false,
                &stats.inlier_indices, nullptr);

    if (options.final_least_squares_) {  // :299-317
      Model refined = *best_model;
      solver.LeastSquares(stats.inlier_indices, &refined);
      const double score = ScoreOnly(options, solver, refined);
      if (score < stats.best_model_score) {
        stats.best_model_score = score;
        *best_model = refined;
        UpdateRANSACTerminationCriteria(options, solver, *best_model, statistics, &max_per_solver);
      }
    }
    return stats.best_num_inliers;
  }

 protected:
  struct Batch {
    uint32_t first = 0, count = 0;
    int model_size = 0;
    std::vector<int> type;
    std::vector<int> num_models;
    std::vector<int> model_offset;
    std::vector<double> models;
    std::vector<double> scores;
  };

  // Simulates the next `want` iterations on copies of the RNG / sampler / counters, solves
  // their minimal problems and sweeps all resulting models in one launch.
  void Speculate(const HybridLORansacOptions& options, const HybridSolver& solver,
                 const std::vector<double>& priors, const std::vector<std::vector<int>>& sample_sizes,
                 const HybridRansacStatistics& stats, const std::vector<uint32_t>& max_per_solver,
                 utils::Mt19937 rng_copy, Sampler sampler_copy, const uint32_t first, const uint32_t want,
                 const int precision, Batch* b) const {
    const gpu::DeviceScorer& dev = solver.device_scorer();
    b->first = first;
    b->model_size = dev.model_size();
    b->type.clear();
    b->num_models.clear();
    b->model_offset.clear();
    b->models.clear();
    HybridRansacStatistics sim;
    sim.inlier_ratios = stats.inlier_ratios;
    sim.num_iterations_per_solver = stats.num_iterations_per_solver;
    std::vector<std::vector<int>> sample;
    ModelVector found;
    for (uint32_t j = 0; j < want; ++j) {
      const int type = SelectMinimalSolver(solver, priors, sim, options.min_num_iterations_, &rng_copy);
      if (type < 0) break;
      sim.num_iterations_per_solver[type] += 1;
      sampler_copy.Sample(sample_sizes[type], &sample);
      const int nm = std::max(0, solver.MinimalSolver(sample, type, &found));
      b->type.push_back(type);
      b->num_models.push_back(nm);
      b->model_offset.push_back(static_cast<int>(b->models.size() / b->model_size));
      for (int m = 0; m < nm; ++m) {
        const size_t at = b->models.size();
        b->models.resize(at + b->model_size);
        HybridSolver::PackModel(found[m], &b->models[at]);
      }
      if (sim.num_iterations_per_solver[type] >= max_per_solver[type]) break;  // loop would end here
    }
    b->count = static_cast<uint32_t>(b->type.size());
    const int total = static_cast<int>(b->models.size() / b->model_size);
    b->scores.assign(total, std::numeric_limits<double>::max());
    if (total > 0) {
      dev.ScoreModels(total, b->models.data(), nullptr, options.squared_inlier_thresholds_.data(),
                      options.data_type_weights_.data(), b->scores.data(), precision);
    }
    if (b->count == 0) {  // keep the replay loop well-defined
      b->type.push_back(-1);
      b->num_models.push_back(0);
      b->model_offset.push_back(0);
      b->count = 1;
    }
  }

  // Eq. 1 of Camposeco et al. (reference hybrid_ransac.h:324-384).
  int SelectMinimalSolver(const HybridSolver& solver, const std::vector<double>& priors,
                          const HybridRansacStatistics& stats, const uint32_t min_num_iterations,
                          utils::Mt19937* rng) const {
    const int num_solvers = static_cast<int>(priors.size());
    std::vector<std::vector<int>> sizes;
    solver.min_sample_sizes(&sizes);
    std::vector<double> prob(num_solvers, 0.0);
    double total = 0.0;
    const int num_types = solver.num_data_types();
    const double ratio_sum = std::accumulate(stats.inlier_ratios.begin(), stats.inlier_ratios.end(), 0.0);
    if (ratio_sum == 0.0) {
      for (int i = 0; i < num_solvers; ++i) {
        prob[i] = priors[i];
        total += prob[i];
      }
    } else {
      for (int i = 0; i < num_solvers; ++i) {
        double iters = static_cast<double>(stats.num_iterations_per_solver[i]);
        if (iters > 0.0) iters -= 1.0;
        double all_inlier = 1.0;
        for (int t = 0; t < num_types; ++t)
          all_inlier *= std::pow(stats.inlier_ratios[t], static_cast<double>(sizes[i][t]));
        if (iters < static_cast<double>(min_num_iterations)) {
          prob[i] = all_inlier * priors[i];
        } else {
          prob[i] = all_inlier * std::pow(1.0 - all_inlier, iters) * priors[i];
        }
        total += prob[i];
      }
    }
    std::uniform_real_distribution<double> pick(0.0, total);
    const double x = pick(*rng);
    double acc = 0.0;
    for (int i = 0; i < num_solvers; ++i) {
      if (priors[i] == 0.0) continue;
      acc += prob[i];
      if (x <= acc) return i;
    }
    return -1;
  }

  // Per-type inlier lists somebody else owns: the scorer's pinned result buffer (valid until the next
  // pass) or a std::vector<std::vector<int>>.
  struct Lists {
    std::vector<const int*> list;
    std::vector<int> count;
    Lists() {}
    explicit Lists(const std::vector<std::vector<int>>& v) {
      for (const auto& l : v) {
        list.push_back(l.data());
        count.push_back(static_cast<int>(l.size()));
      }
    }
  };

  // The same pass with the lists left where the scorer wrote them.
  Lists FusedPassView(const HybridLORansacOptions& options, const HybridSolver& solver, const Model& model,
                      const std::vector<double>& thr_inl, const bool want_score, double* score) const {
    double packed[16];
    HybridSolver::PackModel(model, packed);
    int32_t counts[2] = {0, 0};
    const int* at = reinterpret_cast<const int*>(solver.device_scorer().InliersViewAndScore(
        packed, thr_inl.data(), want_score ? options.squared_inlier_thresholds_.data() : nullptr,
        options.data_type_weights_.data(), counts, score));
    Lists out;
    for (int t = 0; t < solver.num_data_types(); ++t) {
      out.list.push_back(at);
      out.count.push_back(counts[t]);
      at += counts[t];
    }
    return out;
  }

  // Fused K=1 pass with per-type thresholds: per-type ascending inlier lists at thr_inl and
  // (optionally) the weighted MSAC score at the options' thresholds (:407-462).
  int FusedPass(const HybridLORansacOptions& options, const HybridSolver& solver, const Model& model,
                const std::vector<double>& thr_inl, const bool want_score,
                std::vector<std::vector<int>>* inliers, double* score) const {
    if (inliers) {
      const Lists v = FusedPassView(options, solver, model, thr_inl, want_score, score);
      inliers->resize(v.list.size());
      int total = 0;
      for (size_t t = 0; t < v.list.size(); ++t) {
        (*inliers)[t].assign(v.list[t], v.list[t] + v.count[t]);
        total += v.count[t];
      }
      return total;
    }
    double packed[16];
    HybridSolver::PackModel(model, packed);
    int32_t counts[2] = {0, 0};
    solver.device_scorer().InliersAndScore(
        packed, thr_inl.data(), want_score ? options.squared_inlier_thresholds_.data() : nullptr,
        options.data_type_weights_.data(), nullptr, counts, score);
    return counts[0] + (solver.num_data_types() > 1 ? counts[1] : 0);
  }

  double ScoreOnly(const HybridLORansacOptions& options, const HybridSolver& solver,
                   const Model& model) const {
    double score = std::numeric_limits<double>::max();
    FusedPass(options, solver, model, options.squared_inlier_thresholds_, true, nullptr, &score);
    return score;
  }

  // :464-495
  void UpdateRANSACTerminationCriteria(const HybridLORansacOptions& options,
                                       const HybridSolver& solver, const Model& model,
                                       HybridRansacStatistics* statistics,
                                       std::vector<uint32_t>* max_iterations,
                                       bool* lists_stale = nullptr) const {
    // The ratios need the per-type counts only. Inside the main loop the lists themselves are left for the
    // end of EstimateModel (*lists_stale): every new best model would otherwise fetch and copy them.
    std::vector<int> nd;
    solver.num_data(&nd);
    int32_t counts[2] = {0, 0};
    if (lists_stale) {
      double packed[16];
      HybridSolver::PackModel(model, packed);
      solver.device_scorer().InliersAndScore(packed, options.squared_inlier_thresholds_.data(), nullptr,
                                             options.data_type_weights_.data(), nullptr, counts, nullptr);
      *lists_stale = true;
    } else {
      FusedPass(options, solver, model, options.squared_inlier_thresholds_, false,
                &statistics->inlier_indices, nullptr);
      for (size_t d = 0; d < nd.size() && d < 2; ++d)
        counts[d] = static_cast<int32_t>(statistics->inlier_indices[d].size());
    }
    statistics->best_num_inliers = 0;
    for (size_t d = 0; d < nd.size(); ++d) {
      const int c = d < 2 ? counts[d] : 0;
      statistics->best_num_inliers += c;
      statistics->inlier_ratios[d] = nd[d] > 0 ? static_cast<double>(c) / static_cast<double>(nd[d]) : 0.0;
    }
    std::vector<std::vector<int>> sizes;
    solver.min_sample_sizes(&sizes);
    for (int s = 0; s < solver.num_minimal_solvers(); ++s) {
      (*max_iterations)[s] = utils::NumRequiredIterations(
          statistics->inlier_ratios, 1.0 - options.success_probability_, sizes[s],
          options.min_num_iterations_, options.max_num_iterations_per_solver_);
    }
  }

  // Second half of LeastSquaresFit (:578-594) once the inlier lists are known.
  // The samples are copied out of the lists before the solver runs. Returns whether the solver ran (and
  // *m may have changed).
  bool FitOnInliers(const HybridLORansacOptions& options, const HybridSolver& solver,
                    const int solver_type, utils::Mt19937* rng, const Lists& inliers, Model* m) const {
    std::vector<std::vector<int>> sizes;
    solver.min_sample_sizes(&sizes);
    const int num_types = solver.num_data_types();
    for (int t = 0; t < num_types; ++t) {
      const int have = inliers.count[t];
      if (have < sizes[solver_type][t]) return false;
      sizes[solver_type][t] = std::min(sizes[solver_type][t] * options.min_sample_multiplicator_, have);
    }
    std::vector<std::vector<int>> sample(num_types);
    for (int t = 0; t < num_types; ++t)  // utils.h:96-105, one type after the other on one engine
      utils::ShuffleSampleFromList(sizes[solver_type][t], rng, inliers.list[t], inliers.count[t], &sample[t]);
    solver.LeastSquares(sample, m);
    return true;
  }

  // :500-567 with ScoreModel + the following GetInliers on the same model fused. Keeps the
  // reference's behaviour that the relaxed thresholds are initialised once and keep
  // shrinking over all LO steps (they become negative).
  void LocalOptimization(const HybridLORansacOptions& options, const HybridSolver& solver,
                         const int solver_type, utils::Mt19937* rng, Model* best_minimal_model,
                         double* score_best_minimal_model, int* best_solver_type) const {
    const int num_types = solver.num_data_types();
    std::vector<double> relaxed = options.squared_inlier_thresholds_;
    std::vector<double> step(relaxed.size());
    for (int t = 0; t < num_types; ++t) {
      step[t] = (options.threshold_multiplier_ - 1.0) * relaxed[t] /
                static_cast<int>(options.num_lsq_iterations_ - 1);
      relaxed[t] *= options.threshold_multiplier_;
    }
    double score = std::numeric_limits<double>::max();

    Model m_init = *best_minimal_model;
    FitOnInliers(options, solver, solver_type, rng,
                 FusedPassView(options, solver, m_init, relaxed, false, nullptr), &m_init);

    // ScoreModel(m_init) + GetInliers(m_init, thresholds): the list is what every LO step's
    // first LeastSquaresFit(m_non_min = m_init, thresholds) recomputes in the reference.
    std::vector<std::vector<int>> base;
    FusedPass(options, solver, m_init, options.squared_inlier_thresholds_, true, &base, &score);
    UpdateBestModel(score, m_init, solver_type, score_best_minimal_model, best_minimal_model,
                    best_solver_type);

    const Lists base_lists(base);
    // The relaxed thresholds keep shrinking across the LO steps (reference quirk, see above) and are <= 0
    // from some step on: no squared error is below them, so those inlier lists are empty without a pass
    // over the data. `scored` tracks whether m has changed since its score was last offered to
    // UpdateBestModel - offering the same model again cannot change the best (strict <), so that pass is
    // left out as well. Same sequence of solver calls, RNG draws and best-model updates as the reference.
    Lists none;
    none.list.assign(num_types, nullptr);
    none.count.assign(num_types, 0);
    auto all_lists_empty = [&]() {
      for (int t = 0; t < num_types; ++t)
        if (relaxed[t] > 0.0) return false;
      return true;
    };
    auto offer = [&](const Model& m) {
      UpdateBestModel(score, m, solver_type, score_best_minimal_model, best_minimal_model, best_solver_type);
    };
    // Once every relaxed threshold is <= 0 and cannot grow again, a step is: shuffle + fit on the base lists
    // (the only RNG use), then the score of that ONE model. The models of all remaining steps are therefore
    // known up front; they are scored in one sweep, and only those within 1e-9 of the best so far (the sweep
    // sums its partials in another order than a K = 1 pass) get the exact pass and the offer, in step order.
    bool thresholds_only_shrink = options.num_lsq_iterations_ > 0;
    for (int t = 0; t < num_types; ++t) thresholds_only_shrink = thresholds_only_shrink && step[t] >= 0.0;
    bool empty_fit_returns = false;  // a fit on empty lists returns before it shuffles
    {
      std::vector<std::vector<int>> sizes;
      solver.min_sample_sizes(&sizes);
      for (int t = 0; t < num_types; ++t) empty_fit_returns = empty_fit_returns || sizes[solver_type][t] > 0;
    }
    for (int r = 0; r < options.num_lo_steps_; ++r) {
      if (thresholds_only_shrink && empty_fit_returns && all_lists_empty() && options.num_lo_steps_ - r >= 3) {
        const gpu::DeviceScorer& dev = solver.device_scorer();
        const int md = dev.model_size();
        std::vector<Model> tail;
        std::vector<double> packed;
        for (; r < options.num_lo_steps_; ++r) {
          Model m = m_init;
          if (!FitOnInliers(options, solver, solver_type, rng, base_lists, &m)) continue;  // m_init: scored
          tail.push_back(m);
          packed.resize(packed.size() + md);
          HybridSolver::PackModel(m, &packed[packed.size() - md]);
        }
        if (tail.empty()) break;
        std::vector<double> approx(tail.size());
        dev.ScoreModels(static_cast<int>(tail.size()), packed.data(), nullptr,
                        options.squared_inlier_thresholds_.data(), options.data_type_weights_.data(),
                        approx.data(), RLB_F64);
        for (size_t j = 0; j < tail.size(); ++j) {
          if (!(approx[j] <= *score_best_minimal_model * (1.0 + 1e-9))) continue;  // cannot be strictly better
          score = ScoreOnly(options, solver, tail[j]);
          offer(tail[j]);
        }
        break;
      }
      Model m = m_init;
      bool scored = !FitOnInliers(options, solver, solver_type, rng, base_lists, &m);  // m_init was scored
      if (options.num_lsq_iterations_ <= 0) continue;
      if (FitOnInliers(options, solver, solver_type, rng,
                       all_lists_empty() ? none : FusedPassView(options, solver, m, relaxed, false, nullptr),
                       &m))
        scored = false;
      for (int i = 0; i < options.num_lsq_iterations_; ++i) {
        for (int t = 0; t < num_types; ++t) relaxed[t] -= step[t];
        const bool want_fit = i + 1 < options.num_lsq_iterations_;
        if (want_fit && !all_lists_empty()) {
          const Lists inliers = FusedPassView(options, solver, m, relaxed, true, &score);
          offer(m);
          scored = !FitOnInliers(options, solver, solver_type, rng, inliers, &m);
          continue;
        }
        if (!scored) {
          score = ScoreOnly(options, solver, m);
          offer(m);
          scored = true;
        }
        if (want_fit && FitOnInliers(options, solver, solver_type, rng, none, &m)) scored = false;
      }
    }
  }

  inline void UpdateBestModel(const double score_curr, const Model& m_curr, const int solver_type,
                              double* score_best, Model* m_best, int* best_solver_type) const {
    if (score_curr < *score_best) {
      *score_best = score_curr;
      *m_best = m_curr;
      *best_solver_type = solver_type;
    }
  }

  // :610-638
  bool VerifyData(const std::vector<std::vector<int>>& min_sample_sizes, const std::vector<int>& num_data,
                  const int num_solvers, const int num_data_types, std::vector<double>* priors) const {
    for (int i = 0; i < num_solvers; ++i) {
      for (int j = 0; j < num_data_types; ++j) {
        if (min_sample_sizes[i][j] > num_data[j] || min_sample_sizes[i][j] < 0) {
          (*priors)[i] = 0.0;
          break;
        }
      }
    }
    for (int i = 0; i < num_solvers; ++i)
      if ((*priors)[i] > 0.0) return true;
    return false;
  }
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_HYBRID_RANSAC_H_
