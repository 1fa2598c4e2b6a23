// ransaclib_b200 — drop-in for RansacLib/ransac.h.
//
// Same public names, option/statistics structs and EstimateModel signature as the
// reference (ransac.h:47-129), but every pass over the data runs on the GPU through
// the C ABI (include/ransaclib_b200.h); there is no CPU scoring path. The control flow
// is NOT a translation of the reference's sample-score-compare loop:
//
//  * Hypotheses are produced and scored in speculative BATCHES: the sampler's stream is
//    a pure function of (seed, N, sample size) (reference sampling.h:49-55: private RNG),
//    so B samples are pre-drawn, solved (device-side minimal solver when the solver has
//    one) and scored in one K x N sweep; the host then REPLAYS the reference's sequential
//    decisions (ransac.h:158-240) over the returned scores. Because hypotheses do not
//    depend on the local-optimisation RNG, a best-model update never invalidates the rest
//    of a batch; speculation past the final iteration only wastes GPU work.
//  * Inside local optimisation each ScoreModel(m) that is followed by GetInliers(m, thr')
//    on the same model (ransac.h:364->369, 386->391, 401->399) is ONE fused K=1 pass that
//    returns the score at thr and the ascending inlier list at thr'; an LO call needs
//    2 + lo_steps * (2 + lsq_iters) passes instead of the reference's ~105.
//  * With RLB_F32 sweeps, any hypothesis whose f32 score could beat the current best
//    (within 4e-4 relative) is re-scored in f64 before the comparison, so the sequence of
//    accepted models is the f64 one.
//
// Results (iterations, inliers, LO count, model) are those of the reference run with the
// same seed, up to points whose squared error is within rounding of the threshold.
#ifndef RANSACLIB_B200_RANSACLIB_RANSAC_H_
#define RANSACLIB_B200_RANSACLIB_RANSAC_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <random>
#include <vector>

#include "gpu_scorer.h"
#include "sampling.h"
#include "utils.h"

namespace ransac_lib {

// Members and defaults as reference ransac.h:47-60.
class RansacOptions {
 public:
  RansacOptions()
      : min_num_iterations_(100u),
        max_num_iterations_(10000u),
        success_probability_(0.9999),
        squared_inlier_threshold_(1.0),
        random_seed_(0u) {}
  uint32_t min_num_iterations_;
  uint32_t max_num_iterations_;
  double success_probability_;
  double squared_inlier_threshold_;
  unsigned int random_seed_;
};

// Members and defaults as reference ransac.h:64-92 (Lebeda et al., BMVC 2012, Table 1).
class LORansacOptions : public RansacOptions {
 public:
  LORansacOptions()
      : num_lo_steps_(10),
        threshold_multiplier_(std::sqrt(2.0)),
        num_lsq_iterations_(4),
        min_sample_multiplicator_(7),
        non_min_sample_multiplier_(3),
        lo_starting_iterations_(50u),
        final_least_squares_(false) {}
  int num_lo_steps_;
  double threshold_multiplier_;
  int num_lsq_iterations_;
  int min_sample_multiplicator_;
  int non_min_sample_multiplier_;
  uint32_t lo_starting_iterations_;
  bool final_least_squares_;
};

// reference ransac.h:94-101.
struct RansacStatistics {
  uint32_t num_iterations;
  int best_num_inliers;
  double best_model_score;
  double inlier_ratio;
  std::vector<int> inlier_indices;
  int number_lo_iterations;
};

// Out-of-band counters of the GPU driver (not part of the reference API).
struct GpuRunCounters {
  uint64_t batches = 0;           // speculative hypothesis batches
  uint64_t hypotheses_scored = 0; // K summed over batches (includes speculation waste)
  uint64_t single_passes = 0;     // fused K=1 passes (LO, inlier collection, f64 re-scores)
  uint64_t handbacks = 0;         // device-resident runs handed back to the host-driven loop (LORansacGpuOptions)
  uint64_t handback_iterations = 0;  // iterations such runs had done on the device when they came back WITHOUT their state (work done twice)
  uint64_t handbacks_resumed = 0;    // handed-back runs that came with their state and were resumed, not restarted
};

// When the device-resident driver of a solver (one CTA runs the whole problem) may hand a run back to the
// host-driven loop, whose sweeps use the whole GPU: a run that after `handback_after` iterations still needs
// more than `handback_min_remaining` hypotheses (very low inlier ratios). One SM sweeps ~7 hypothesis x point
// evaluations per cycle, so a few thousand hypotheses over 10k points are where the whole-GPU sweep wins
// (DESIGN.md section 3c). handback_after = 0 never hands back. The run is the same either way.
struct LORansacGpuOptions {
  uint32_t handback_after = 384u;
  uint32_t handback_min_remaining = 3000u;
};

class RansacBase {
 protected:
  void ResetStatistics(RansacStatistics* statistics) const {
    statistics->num_iterations = 0u;
    statistics->best_num_inliers = 0;
    statistics->best_model_score = std::numeric_limits<double>::max();
    statistics->inlier_ratio = 0.0;
    statistics->inlier_indices.clear();
    statistics->number_lo_iterations = 0;
  }
};

template <class Model, class ModelVector, class Solver,
          class Sampler = UniformSampling<Solver> >
class LocallyOptimizedMSAC : public RansacBase {
  static_assert(gpu::HasDeviceScorer<Solver>::value,
                "ransaclib_b200 has no CPU scoring path: the Solver must expose "
                "device_scorer(), PackModel() and UnpackModel() (see RansacLib/gpu_scorer.h)");

 public:
  // Optional: where to accumulate driver counters (may stay null).
  mutable GpuRunCounters* counters = nullptr;
  // GPU-side policy knobs (not part of the reference API; the defaults are what the tests and benches use).
  LORansacGpuOptions gpu_options;

  // Same contract as reference ransac.h:128-275: returns the number of inliers; leaves
  // *best_model untouched when there is too little data or no model was found.
  int EstimateModel(const LORansacOptions& options, const Solver& solver, Model* best_model,
                    RansacStatistics* statistics) const {
    ResetStatistics(statistics);
    RansacStatistics& stats = *statistics;
    const int sample_size = solver.min_sample_size();
    const int num_data = solver.num_data();
    if (sample_size > num_data || sample_size <= 0) return 0;

    // A solver may run the whole of this function on the device (gpu_scorer.h: device_lomsac). A run that is
    // handed back (very low inlier ratios) comes with its state and is resumed below.
    Resume resume;
    if (DeviceResident(options, solver, best_model, statistics, &resume,
                       std::integral_constant<bool, gpu::HasDeviceLomsac<Solver>::value &&
                                                        std::is_same<Sampler, UniformSampling<Solver> >::value>()))
      return stats.best_num_inliers;

    Sampler sampler(options.random_seed_, solver);
    utils::Mt19937 rng;  // local optimisation's own stream (ransac.h:142-144)
    rng.seed(options.random_seed_);

    uint32_t max_iters = std::max(options.max_num_iterations_, options.min_num_iterations_);
    const double thr = options.squared_inlier_threshold_;
    const double kNever = std::numeric_limits<double>::max();
    const gpu::DeviceScorer& dev = solver.device_scorer();
    const int precision = dev.precision();

    Model best_minimal_model;
    double best_minimal_score = kNever;

    Batch batch;
    uint32_t next_batch_size = 256u;
    uint32_t first_iteration = 0u;

    auto refresh_termination = [&]() {  // ransac.h:170-177, 231-238
      stats.best_num_inliers = CollectInliers(solver, *best_model, thr, &stats.inlier_indices);
      stats.inlier_ratio = static_cast<double>(stats.best_num_inliers) / static_cast<double>(num_data);
      max_iters = utils::NumRequiredIterations(stats.inlier_ratio, 1.0 - options.success_probability_,
                                               sample_size, options.min_num_iterations_,
                                               options.max_num_iterations_);
    };

    if (resume.valid) {
      // Continue the device-resident run: the sampler's stream is a function of (seed, iteration), so the
      // samples it already consumed are drawn again and dropped (32 ns each); the local optimisation's engine,
      // the best models and the LO count come from the record.
      std::vector<int> dropped;
      for (uint32_t i = 0; i < resume.iterations; ++i) sampler.Sample(&dropped);
      rng.SetState(resume.lo_block, resume.lo_pos);
      best_minimal_score = resume.best_minimal_score;
      if (best_minimal_score < kNever) Solver::UnpackModel(resume.best_minimal_model, &best_minimal_model);
      stats.best_model_score = resume.best_score;
      stats.number_lo_iterations = resume.num_lo;
      if (resume.best_score < kNever) {
        Solver::UnpackModel(resume.best_model, best_model);
        refresh_termination();  // inlier list, ratio and iteration bound of the best model (ransac.h:231-238)
      }
      first_iteration = resume.iterations;
      next_batch_size = 4096u;  // a handed-back run needs thousands of hypotheses
    }

    for (stats.num_iterations = first_iteration; stats.num_iterations < max_iters; ++stats.num_iterations) {
      const uint32_t it = stats.num_iterations;
      // LO on the best model so far once the start iteration is reached (ransac.h:163-178).
      // (a resumed run has already been through this check for its first iteration on the device)
      if (it == options.lo_starting_iterations_ && best_minimal_score < kNever &&
          !(resume.valid && it == first_iteration)) {
        ++stats.number_lo_iterations;
        LocalOptimization(options, solver, &rng, best_model, &stats.best_model_score);
        refresh_termination();
      }

      if (it >= batch.first + batch.count) {
        // Speculate: everything up to the current iteration bound, in growing chunks.
        const uint32_t bound = std::max(max_iters, it + 1u);
        const uint32_t want = std::min(next_batch_size, bound - it);
        FillBatch(solver, dev, &sampler, thr, precision, it, want, &batch);
        next_batch_size = std::min(next_batch_size * 2u, 16384u);
      }
      const uint32_t slot = it - batch.first;
      if (batch.num_models[slot] <= 0) continue;  // ransac.h:185

      // Best model of this sample (ransac.h:278-293) from the swept scores.
      double local_score = kNever;
      int local_id = -1;
      for (int m = 0; m < batch.num_models[slot]; ++m) {
        const double s = batch.scores[batch.model_offset[slot] + m];
        if (s < local_score) {
          local_score = s;
          local_id = batch.model_offset[slot] + m;
        }
      }
      // local_id < 0: every score was NaN -> local_score stays DBL_MAX and can never win, but
      // the forced entry at the LO start iteration below still applies (ransac.h:195-196).
      const bool forced = (it == options.lo_starting_iterations_);
      if (precision == RLB_F32 && local_id >= 0 &&
          (forced || local_score <= best_minimal_score * (1.0 + 4e-4) || best_minimal_score == kNever)) {
        // f32 sweep says "maybe better": settle it with the exact f64 score.
        Model cand;
        Solver::UnpackModel(&batch.models[static_cast<size_t>(local_id) * batch.model_size], &cand);
        local_score = ScoreOnly(solver, cand, thr);
      }

      if (local_score < best_minimal_score || forced) {  // ransac.h:195-196
        const bool new_best = local_score < best_minimal_score;
        if (new_best) {
          best_minimal_score = local_score;
          Solver::UnpackModel(&batch.models[static_cast<size_t>(local_id) * batch.model_size],
                              &best_minimal_model);
          UpdateBestModel(best_minimal_score, best_minimal_model, &stats.best_model_score, best_model);
        }
        const bool run_lo = (it >= options.lo_starting_iterations_ && best_minimal_score < kNever);
        if (!new_best && !run_lo) continue;
        if (run_lo) {
          ++stats.number_lo_iterations;
          double score = best_minimal_score;
          LocalOptimization(options, solver, &rng, &best_minimal_model, &score);
          UpdateBestModel(score, best_minimal_model, &stats.best_model_score, best_model);
        }
        refresh_termination();
      }
    }

    // LO for runs that ended before the start iteration (ransac.h:245-255).
    if (stats.num_iterations <= options.lo_starting_iterations_ && stats.best_model_score < kNever) {
      ++stats.number_lo_iterations;
      LocalOptimization(options, solver, &rng, best_model, &stats.best_model_score);
      stats.best_num_inliers = CollectInliers(solver, *best_model, thr, &stats.inlier_indices);
      stats.inlier_ratio = static_cast<double>(stats.best_num_inliers) / static_cast<double>(num_data);
    }

    if (options.final_least_squares_) {  // ransac.h:257-272
      Model refined = *best_model;
      solver.LeastSquares(stats.inlier_indices, &refined);
      std::vector<int> inl;
      double score = kNever;
      const int n_inl = FusedPass(solver, refined, thr, &thr, &inl, &score);
      if (score < stats.best_model_score) {
        stats.best_model_score = score;
        *best_model = refined;
        stats.best_num_inliers = n_inl;
        stats.inlier_indices.swap(inl);
        stats.inlier_ratio = static_cast<double>(n_inl) / static_cast<double>(num_data);
      }
    }
    return stats.best_num_inliers;
  }

 protected:
  // State of a device-resident run that was handed back (rlb_lomsac_resume_record, ransaclib_b200.h).
  struct Resume {
    bool valid = false;
    uint32_t iterations = 0u;
    int num_lo = 0;
    int lo_pos = 624;
    double best_score = std::numeric_limits<double>::max();
    double best_minimal_score = std::numeric_limits<double>::max();
    double best_model[16] = {0.0};
    double best_minimal_model[16] = {0.0};
    uint32_t lo_block[624];
  };

  bool DeviceResident(const LORansacOptions&, const Solver&, Model*, RansacStatistics*, Resume*,
                      std::false_type) const {
    return false;
  }
  static int dev_model_doubles(const Solver& solver) { return solver.device_scorer().model_size(); }
  // One launch: sampler, minimal solver, scoring, local optimisation and refits all run on the
  // device (same sequence of decisions as the loop above; csrc/line_lomsac_kernel.cu).
  bool DeviceResident(const LORansacOptions& options, const Solver& solver, Model* best_model,
                      RansacStatistics* statistics, Resume* resume, std::true_type) const {
    if (gpu::HostDriverForced()) return false;
    rlb_lomsac_options o;
    o.min_iters = options.min_num_iterations_;
    o.max_iters = options.max_num_iterations_;
    o.success_prob = options.success_probability_;
    o.sq_thr = options.squared_inlier_threshold_;
    o.lo_steps = options.num_lo_steps_;
    o.thr_mult = options.threshold_multiplier_;
    o.lsq_iters = options.num_lsq_iterations_;
    o.min_sample_mult = options.min_sample_multiplicator_;
    o.non_min_sample_mult = options.non_min_sample_multiplier_;
    o.lo_start = options.lo_starting_iterations_;
    o.final_lsq = options.final_least_squares_ ? 1 : 0;
    o.debug_flags = 0;
    // One CTA runs the whole problem; a run that turns out to need thousands of hypotheses (very low
    // inlier ratios) is handed back and restarts in the loop below, whose sweeps use the whole GPU
    // (gpu_options; counted in GpuRunCounters). Same run either way.
    o.handback_after = gpu_options.handback_after;
    o.handback_min_remaining = gpu_options.handback_min_remaining;
    double packed[16] = {0.0};
    Solver::PackModel(*best_model, packed);  // read by the final least-squares step (ransac.h:257-259)
    rlb_lomsac_result r;
    std::vector<int> inliers;
    if (!solver.device_lomsac(o, options.random_seed_, packed, &r, &inliers)) return false;
    RansacStatistics& stats = *statistics;
    if (r.success == 2) {
      // handed back: with its state (resumed by the caller) or, for tiny problems, without (the run restarts)
      const bool has_state = r.num_inliers == RLB_LOMSAC_RESUME_WORDS &&
                             static_cast<int>(inliers.size()) >= RLB_LOMSAC_RESUME_WORDS &&
                             inliers[0] == RLB_LOMSAC_RESUME_MAGIC && dev_model_doubles(solver) <= 3;
      if (has_state) {
        const int* rec = inliers.data();
        auto dbl = [&](int word) {
          const uint64_t bits = static_cast<uint64_t>(static_cast<uint32_t>(rec[word])) |
                                (static_cast<uint64_t>(static_cast<uint32_t>(rec[word + 1])) << 32);
          double v;
          std::memcpy(&v, &bits, sizeof(v));
          return v;
        };
        resume->valid = true;
        resume->iterations = static_cast<uint32_t>(rec[1]);
        resume->num_lo = rec[2];
        resume->lo_pos = rec[3];
        resume->best_score = dbl(4);
        resume->best_minimal_score = dbl(12);
        for (int i = 0; i < 3; ++i) {
          resume->best_model[i] = dbl(6 + 2 * i);
          resume->best_minimal_model[i] = dbl(14 + 2 * i);
        }
        for (int i = 0; i < 624; ++i) resume->lo_block[i] = static_cast<uint32_t>(rec[20 + i]);
      }
      if (counters) {
        counters->handbacks += 1;
        if (has_state) {
          counters->handbacks_resumed += 1;
        } else {
          counters->handback_iterations += r.num_iterations;
        }
      }
      return false;
    }
    if (!r.success) return true;
    stats.num_iterations = r.num_iterations;
    stats.best_num_inliers = r.num_inliers;
    stats.best_model_score = r.score;
    stats.inlier_ratio = r.inlier_ratio;
    stats.number_lo_iterations = r.num_lo;
    stats.inlier_indices.swap(inliers);
    if (r.score < std::numeric_limits<double>::max()) Solver::UnpackModel(packed, best_model);
    if (counters) counters->single_passes += static_cast<uint64_t>(r.passes);
    return true;
  }

  // One speculative batch of hypotheses with their scores, on the host.
  struct Batch {
    uint32_t first = 0, count = 0;
    int model_size = 0;
    std::vector<int> num_models;     // per sample: models returned by the minimal solver
    std::vector<int> model_offset;   // per sample: first model index
    std::vector<double> models;      // packed
    std::vector<double> scores;      // per model
    std::vector<int32_t> samples;
    std::vector<uint8_t> valid;
  };

  void FillBatch(const Solver& solver, const gpu::DeviceScorer& dev, Sampler* sampler,
                 const double thr, const int precision, const uint32_t first, const uint32_t count,
                 Batch* b) const {
    b->first = first;
    b->count = count;
    b->model_size = dev.model_size();
    b->num_models.assign(count, 0);
    b->model_offset.assign(count, 0);
    const double thr_v[2] = {thr, thr};
    FillBatchImpl(solver, dev, sampler, thr_v, precision, b, gpu::HasDeviceMinimalSolver<Solver>());
    if (counters) {
      counters->batches += 1;
      counters->hypotheses_scored += b->scores.size();
    }
  }

  // Device-side minimal solver: samples go up, models + scores come back.
  void FillBatchImpl(const Solver& solver, const gpu::DeviceScorer& dev, Sampler* sampler,
                     const double* thr_v, const int precision, Batch* b, std::true_type) const {
    const int count = static_cast<int>(b->count);
    const int ss = solver.min_sample_size();
    std::vector<int> flat;
    std::vector<int> one;
    b->samples.resize(static_cast<size_t>(count) * ss);
    for (int k = 0; k < count; ++k) {
      sampler->Sample(&one);
      for (int j = 0; j < ss; ++j) b->samples[static_cast<size_t>(k) * ss + j] = one[j];
    }
    b->models.resize(static_cast<size_t>(count) * b->model_size);
    b->valid.resize(count);
    b->scores.resize(count);
    const double solver_thr = SolverThreshold(solver, thr_v[0]);
    if (solver_thr == thr_v[0]) {
      dev.HypothesizeAndScore(solver.device_minimal_solver(), count, b->samples.data(), thr_v, nullptr,
                              b->models.data(), b->valid.data(), b->scores.data(), precision);
    } else {
      // The solver's own consistency threshold differs from the scoring threshold.
      const double st[2] = {solver_thr, solver_thr};
      std::vector<double> tmp(count);
      dev.HypothesizeAndScore(solver.device_minimal_solver(), count, b->samples.data(), st, nullptr,
                              b->models.data(), b->valid.data(), tmp.data(), precision);
      dev.ScoreModels(count, b->models.data(), b->valid.data(), thr_v, nullptr, b->scores.data(),
                      precision);
    }
    for (int k = 0; k < count; ++k) {
      b->num_models[k] = b->valid[k] ? 1 : 0;
      b->model_offset[k] = k;
    }
  }

  // Host minimal solver (any Solver): models are packed and swept on the device.
  void FillBatchImpl(const Solver& solver, const gpu::DeviceScorer& dev, Sampler* sampler,
                     const double* thr_v, const int precision, Batch* b, std::false_type) const {
    const int count = static_cast<int>(b->count);
    std::vector<int> one;
    ModelVector found;
    b->models.clear();
    for (int k = 0; k < count; ++k) {
      sampler->Sample(&one);
      const int nm = solver.MinimalSolver(one, &found);
      b->num_models[k] = std::max(nm, 0);
      b->model_offset[k] = static_cast<int>(b->models.size() / b->model_size);
      for (int m = 0; m < nm; ++m) {
        const size_t at = b->models.size();
        b->models.resize(at + b->model_size);
        Solver::PackModel(found[m], &b->models[at]);
      }
    }
    const int total = static_cast<int>(b->models.size() / b->model_size);
    b->scores.assign(total, std::numeric_limits<double>::max());
    if (total > 0)
      dev.ScoreModels(total, b->models.data(), nullptr, thr_v, nullptr, b->scores.data(), precision);
  }

  template <class S>
  static auto SolverThresholdImpl(const S& s, double, int) -> decltype(s.device_minimal_solver_threshold()) {
    return s.device_minimal_solver_threshold();
  }
  template <class S>
  static double SolverThresholdImpl(const S&, double fallback, long) {
    return fallback;
  }
  static double SolverThreshold(const Solver& s, double fallback) {
    return SolverThresholdImpl(s, fallback, 0);
  }

  // Fused K=1 pass: ascending inliers at thr_inl and (optionally) the score at *thr_score.
  int FusedPass(const Solver& solver, const Model& model, const double thr_inl,
                const double* thr_score, std::vector<int>* inliers, double* score) const {
    double packed[16];
    Solver::PackModel(model, packed);
    const double ti[2] = {thr_inl, thr_inl};
    double ts[2] = {0.0, 0.0};
    if (thr_score) ts[0] = ts[1] = *thr_score;
    int32_t counts[2] = {0, 0};
    static_assert(sizeof(int) == sizeof(int32_t), "int must be 32 bits");
    if (inliers) inliers->resize(solver.num_data());
    solver.device_scorer().InliersAndScore(packed, ti, thr_score ? ts : nullptr, nullptr,
                                           inliers ? reinterpret_cast<int32_t*>(inliers->data()) : nullptr,
                                           counts, score);
    if (inliers) inliers->resize(counts[0]);
    if (counters) counters->single_passes += 1;
    return counts[0];
  }

  // The same pass with the list left in the scorer's pinned result buffer: *list is valid until the next
  // pass. For the LeastSquaresFit steps, which keep a short sample of a long list.
  int FusedPassView(const Solver& solver, const Model& model, const double thr_inl, const double* thr_score,
                    const int** list, double* score) const {
    double packed[16];
    Solver::PackModel(model, packed);
    const double ti[2] = {thr_inl, thr_inl};
    double ts[2] = {0.0, 0.0};
    if (thr_score) ts[0] = ts[1] = *thr_score;
    int32_t counts[2] = {0, 0};
    *list = reinterpret_cast<const int*>(solver.device_scorer().InliersViewAndScore(
        packed, ti, thr_score ? ts : nullptr, nullptr, counts, score));
    if (counters) counters->single_passes += 1;
    return counts[0];
  }

  double ScoreOnly(const Solver& solver, const Model& model, const double thr) const {
    double score = std::numeric_limits<double>::max();
    FusedPass(solver, model, thr, &thr, nullptr, &score);
    return score;
  }

  // GetInliers (ransac.h:311-336).
  int CollectInliers(const Solver& solver, const Model& model, const double thr,
                     std::vector<int>* inliers) const {
    return FusedPass(solver, model, thr, nullptr, inliers, nullptr);
  }

  // One UpdateBestModel call of the reference's LO, in the reference's order. Nothing inside an LO call reads
  // the best model, so the calls are recorded and applied when the LO call ends; `deferred` ones (the
  // ScoreModel of the last LSQ iteration of each step, which no fit follows) are scored together then.
  struct Offer {
    double score;
    Model model;
    bool deferred;
  };

  // ScoreModel(m) fused with the LeastSquaresFit(fit_thr, m) that follows it in the reference
  // (ransac.h:409-420 for the fit); the score is recorded as an offer.
  void ScoreThenFit(const LORansacOptions& options, const Solver& solver, utils::Mt19937* rng,
                    const bool want_score, const double fit_thr, Model* m, std::vector<Offer>* offers) const {
    const double thr = options.squared_inlier_threshold_;
    const int* list = nullptr;
    double score = std::numeric_limits<double>::max();
    const int n_inl = FusedPassView(solver, *m, fit_thr, want_score ? &thr : nullptr, &list, &score);
    if (want_score) offers->push_back(Offer{score, *m, false});
    if (n_inl < solver.min_sample_size()) return;
    const int cap = options.min_sample_multiplicator_ * solver.min_sample_size();
    std::vector<int> sample;  // copied out of the result buffer before the solver runs
    utils::ShuffleSampleFromList(std::min(cap, n_inl), rng, list, n_inl, &sample);
    solver.LeastSquares(sample, m);
  }

  // Algorithms 2 and 3 of Lebeda et al.; sequence of passes, shuffles and solver calls as
  // reference ransac.h:341-407, with score/inlier passes on the same model fused.
  void LocalOptimization(const LORansacOptions& options, const Solver& solver, utils::Mt19937* rng,
                         Model* best_minimal_model, double* score_best_minimal_model) const {
    const int num_data = solver.num_data();
    const int min_non_min = solver.non_minimal_sample_size();
    if (min_non_min > num_data) return;
    const int min_size = solver.min_sample_size();
    const double thr = options.squared_inlier_threshold_;
    const double mult = options.threshold_multiplier_;
    std::vector<Offer> offers;

    Model m_init = *best_minimal_model;
    ScoreThenFit(options, solver, rng, false, thr * mult, &m_init, &offers);

    std::vector<int> inliers_base;
    double score = std::numeric_limits<double>::max();
    FusedPass(solver, m_init, thr * mult, &thr, &inliers_base, &score);
    offers.push_back(Offer{score, m_init, false});

    const int non_min_size =
        std::max(min_non_min, std::min(min_size * options.non_min_sample_multiplier_,
                                       static_cast<int>(inliers_base.size()) / 2));
    std::vector<int> sample;
    for (int r = 0; r < options.num_lo_steps_; ++r) {
      utils::ShuffleSampleFromList(non_min_size, rng, inliers_base.data(),
                                   static_cast<int>(inliers_base.size()), &sample);
      Model m;
      if (!solver.NonMinimalSolver(sample, &m)) continue;
      // ScoreModel(m) + LeastSquaresFit(thr, m)   (ransac.h:386-391)
      ScoreThenFit(options, solver, rng, true, thr, &m, &offers);

      double fit_thr = mult * thr;
      // Reference divides by (num_lsq_iterations_ - 1) as an int (ransac.h:395-397).
      const double step = (mult - 1.0) * thr / static_cast<int>(options.num_lsq_iterations_ - 1);
      if (options.num_lsq_iterations_ <= 0) continue;
      ScoreThenFit(options, solver, rng, false, fit_thr, &m, &offers);  // fit of i = 0
      for (int i = 0; i < options.num_lsq_iterations_; ++i) {
        fit_thr -= step;
        if (i + 1 < options.num_lsq_iterations_) {
          // ScoreModel of iteration i fused with the fit of iteration i + 1.
          ScoreThenFit(options, solver, rng, true, fit_thr, &m, &offers);
        } else {
          offers.push_back(Offer{std::numeric_limits<double>::max(), m, true});
        }
      }
    }
    ApplyOffers(options, solver, &offers, best_minimal_model, score_best_minimal_model);
  }

  // The recorded UpdateBestModel calls, in order. The deferred models are scored in one f64 sweep first; the
  // sweep sums its partials in another order than a K = 1 pass, so its score only filters: a model within
  // 1e-9 of the best so far gets the exact pass before it is offered, any other cannot be strictly better.
  void ApplyOffers(const LORansacOptions& options, const Solver& solver, std::vector<Offer>* offers,
                   Model* best, double* best_score) const {
    const double thr = options.squared_inlier_threshold_;
    std::vector<double> packed, approx;
    int num_deferred = 0;
    for (const Offer& o : *offers) num_deferred += o.deferred ? 1 : 0;
    if (num_deferred >= 3) {
      const gpu::DeviceScorer& dev = solver.device_scorer();
      const int md = dev.model_size();
      packed.resize(static_cast<size_t>(num_deferred) * md);
      size_t at = 0;
      for (const Offer& o : *offers) {
        if (!o.deferred) continue;
        Solver::PackModel(o.model, &packed[at]);
        at += md;
      }
      approx.resize(num_deferred);
      const double thr2[2] = {thr, thr};
      dev.ScoreModels(num_deferred, packed.data(), nullptr, thr2, nullptr, approx.data(), RLB_F64);
    }
    int k = 0;
    for (Offer& o : *offers) {
      if (o.deferred) {
        const int id = k++;
        if (!approx.empty() && !(approx[id] <= *best_score * (1.0 + 1e-9))) continue;
        o.score = ScoreOnly(solver, o.model, thr);
      }
      UpdateBestModel(o.score, o.model, best_score, best);
    }
  }

  inline void UpdateBestModel(const double score_curr, const Model& m_curr, double* score_best,
                              Model* m_best) const {
    if (score_curr < *score_best) {  // strict: the earlier model keeps ties (ransac.h:422-428)
      *score_best = score_curr;
      *m_best = m_curr;
    }
  }
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_RANSAC_H_
