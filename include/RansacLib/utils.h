// ransaclib_b200 — drop-in for RansacLib/utils.h (reference lines cited per function).
// Host-side helpers of the LO-MSAC control flow. Written fresh; semantics (including
// the exact number of RNG draws each call consumes) follow the reference because the
// local-optimisation RNG stream must stay aligned with a CPU run.
#ifndef RANSACLIB_B200_RANSACLIB_UTILS_H_
#define RANSACLIB_B200_RANSACLIB_UTILS_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <random>
#include <utility>
#include <vector>

namespace ransac_lib {
namespace utils {

// Fisher-Yates, one freshly constructed uniform_int_distribution(i, n-1) per position
// (reference utils.h:48-67). Consumes n-1 draws (plus libstdc++'s rare rejections).
template <class T>
inline void RandomShuffle(std::mt19937* rng, std::vector<T>* v) {
  const int n = static_cast<int>(v->size());
  for (int i = 0; i + 1 < n; ++i) {
    std::uniform_int_distribution<int> pick(i, n - 1);
    const int j = pick(*rng);
    std::swap((*v)[i], (*v)[j]);
  }
}

// utils.h:69-73. The whole list is shuffled even though only `target_size` survive.
inline void RandomShuffleAndResize(const int target_size, std::mt19937* rng,
                                   std::vector<int>* v) {
  RandomShuffle<int>(rng, v);
  v->resize(target_size);
}

// utils.h:76-94: joint shuffle over (type, index) pairs of all data types.
inline void RandomShuffleAndResize(const int target_size, std::mt19937* rng,
                                   std::vector<std::vector<int>>* v) {
  std::vector<std::pair<int, int>> all;
  for (size_t t = 0; t < v->size(); ++t) {
    for (int idx : (*v)[t]) all.emplace_back(static_cast<int>(t), idx);
    (*v)[t].clear();
  }
  RandomShuffle<std::pair<int, int>>(rng, &all);
  all.resize(target_size);
  for (const auto& p : all) (*v)[p.first].push_back(p.second);
}

// utils.h:96-105: independent per-type shuffles, sizes clamped to what is available.
inline void RandomShuffleAndResize(const std::vector<int> sample_sizes, std::mt19937* rng,
                                   std::vector<std::vector<int>>* v) {
  for (size_t t = 0; t < v->size(); ++t) {
    const int have = static_cast<int>((*v)[t].size());
    RandomShuffleAndResize(std::min(have, sample_sizes[t]), rng, &(*v)[t]);
  }
}

namespace detail {
inline uint32_t IterationsFor(const double prob_all_inlier_sample, const double prob_missing,
                              const uint32_t min_it, const uint32_t max_it) {
  if (prob_all_inlier_sample <= 0.0) return max_it;
  if (prob_all_inlier_sample >= 1.0) return min_it;
  const double prob_bad_sample = 1.0 - prob_all_inlier_sample;
  if (prob_bad_sample >= 0.99999999999999) return max_it;
  const double needed = std::ceil(std::log(prob_missing) / std::log(prob_bad_sample) + 0.5);
  return std::max(min_it, std::min(static_cast<uint32_t>(needed), max_it));
}
}  // namespace detail

// utils.h:110-140.
inline uint32_t NumRequiredIterations(const double inlier_ratio,
                                      const double prob_missing_best_model,
                                      const int sample_size, const uint32_t min_iterations,
                                      const uint32_t max_iterations) {
  if (inlier_ratio <= 0.0) return max_iterations;
  if (inlier_ratio >= 1.0) return min_iterations;
  // The reference forms 1 - pow(.) first; going through IterationsFor(1 - (1 - pow)) would
  // round differently, so the complement is computed the same way here.
  const double prob_bad_sample = 1.0 - std::pow(inlier_ratio, static_cast<double>(sample_size));
  if (prob_bad_sample >= 0.99999999999999) return max_iterations;
  const double needed =
      std::ceil(std::log(prob_missing_best_model) / std::log(prob_bad_sample) + 0.5);
  return std::max(min_iterations, std::min(static_cast<uint32_t>(needed), max_iterations));
}

// utils.h:142-178 (hybrid: product over data types).
inline uint32_t NumRequiredIterations(const std::vector<double> inlier_ratios,
                                      const double prob_missing_best_model,
                                      const std::vector<int> sample_sizes,
                                      const uint32_t min_iterations,
                                      const uint32_t max_iterations) {
  double all_inlier = 1.0;
  for (size_t t = 0; t < sample_sizes.size(); ++t)
    all_inlier *= std::pow(inlier_ratios[t], static_cast<double>(sample_sizes[t]));
  return detail::IterationsFor(all_inlier, prob_missing_best_model, min_iterations,
                               max_iterations);
}

}  // namespace utils
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_UTILS_H_
