// ransaclib_b200 — drop-in for RansacLib/hybrid_sampling.h: per-data-type uniform
// sampling with one shared private RNG (reference hybrid_sampling.h:47-140).
// HybridBiasedSampling (hybrid_sampling.h:146-223): per-point weights from
// solver.get_weights(), zero-weight points never drawn.
#ifndef RANSACLIB_B200_RANSACLIB_HYBRID_SAMPLING_H_
#define RANSACLIB_B200_RANSACLIB_HYBRID_SAMPLING_H_

#include <cmath>
#include <numeric>
#include <random>
#include <vector>

namespace ransac_lib {

template <class Solver>
class HybridUniformSampling {
 public:
  HybridUniformSampling(const unsigned int random_seed, const Solver& solver) {
    solver.num_data(&num_data_);
    rng_.seed(random_seed);
    index_.resize(num_data_.size());
    for (size_t t = 0; t < num_data_.size(); ++t) {
      if (num_data_[t] < 2) continue;  // hybrid_sampling.h:58
      index_[t].param(std::uniform_int_distribution<int>::param_type(0, num_data_[t] - 1));
    }
  }

  // hybrid_sampling.h:64-87.
  void Sample(const std::vector<int>& num_samples_per_data_type,
              std::vector<std::vector<int>>* random_sample) {
    const size_t T = num_data_.size();
    random_sample->resize(T);
    for (size_t t = 0; t < T; ++t) {
      std::vector<int>& s = (*random_sample)[t];
      s.clear();
      const int want = num_samples_per_data_type[t];
      if (want <= 0) continue;
      const double ratio = static_cast<double>(num_data_[t]) /
                           static_cast<double>(num_data_[t] - want);
      if (ratio < M_E) {
        s.resize(want);
        for (int i = 0; i < want; ++i) {
          bool again = true;
          while (again) {
            s[i] = index_[t](rng_);
            again = false;
            for (int j = 0; j < i && !again; ++j) again = (s[j] == s[i]);
          }
        }
      } else {
        s.resize(num_data_[t]);
        std::iota(s.begin(), s.end(), 0);
        if (want == num_data_[t]) continue;
        for (int i = 0; i + 1 < num_data_[t]; ++i) {
          std::uniform_int_distribution<int> pick(i, num_data_[t] - 1);
          std::swap(s[i], s[pick(rng_)]);
        }
        s.resize(want);
      }
    }
  }

 protected:
  std::mt19937 rng_;
  std::vector<std::uniform_int_distribution<int>> index_;
  std::vector<int> num_data_;
};

// Weighted sampling (reference hybrid_sampling.h:146-223). The solver additionally provides
//     void get_weights(std::vector<std::vector<double>>& weights) const;   // one weight per datum and type
// Points with weight <= 0 are never drawn; the others are drawn with probability proportional
// to their weight (std::discrete_distribution over the shared mt19937, so the stream of
// samples for a given seed is the reference's), without repetition inside one sample.
template <class Solver>
class HybridBiasedSampling {
 public:
  HybridBiasedSampling(const unsigned int random_seed, const Solver& solver) {
    std::vector<std::vector<double>> w;
    solver.get_weights(w);
    rng_.seed(random_seed);
    const size_t T = w.size();
    candidates_.resize(T);
    pick_.resize(T);
    for (size_t t = 0; t < T; ++t) {
      std::vector<double> positive;
      positive.reserve(w[t].size());
      candidates_[t].reserve(w[t].size());
      for (size_t j = 0; j < w[t].size(); ++j) {
        if (!(w[t][j] > 0.0)) continue;
        positive.push_back(w[t][j]);
        candidates_[t].push_back(static_cast<int>(j));
      }
      pick_[t] = std::discrete_distribution<int>(positive.begin(), positive.end());
    }
  }

  void Sample(const std::vector<int>& num_samples_per_data_type,
              std::vector<std::vector<int>>* random_sample) {
    const size_t T = candidates_.size();
    random_sample->resize(T);
    for (size_t t = 0; t < T; ++t) {
      std::vector<int>& s = (*random_sample)[t];
      s.clear();
      const int want = num_samples_per_data_type[t];
      if (want <= 0) continue;
      if (want == static_cast<int>(candidates_[t].size())) {  // every candidate, no draw (:189-191)
        s = candidates_[t];
        continue;
      }
      s.resize(want);
      for (int i = 0; i < want; ++i) {
        bool again = true;
        while (again) {
          s[i] = candidates_[t][pick_[t](rng_)];
          again = false;
          for (int j = 0; j < i && !again; ++j) again = (s[j] == s[i]);
        }
      }
    }
  }

 protected:
  std::mt19937 rng_;
  std::vector<std::discrete_distribution<int>> pick_;
  std::vector<std::vector<int>> candidates_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_HYBRID_SAMPLING_H_
