// ransaclib_b200 — drop-in for RansacLib/hybrid_sampling.h: per-data-type uniform
// sampling with one shared private RNG (reference hybrid_sampling.h:47-140).
// HybridBiasedSampling (hybrid_sampling.h:146-223) is not provided: no example uses
// it and it needs a solver.get_weights() that is outside the documented concept.
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

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_HYBRID_SAMPLING_H_
