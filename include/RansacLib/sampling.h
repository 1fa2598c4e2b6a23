// ransaclib_b200 — drop-in for RansacLib/sampling.h (reference sampling.h:47-135).
// The minimal-sample stream stays on the host and stays a pure function of
// (seed, num_data, sample_size): the GPU path pre-draws BATCHES of samples with this
// class, so both implementations score identical sample sets.
#ifndef RANSACLIB_B200_RANSACLIB_SAMPLING_H_
#define RANSACLIB_B200_RANSACLIB_SAMPLING_H_

#include <cmath>
#include <numeric>
#include <random>
#include <vector>

namespace ransac_lib {

template <class Solver>
class UniformSampling {
 public:
  UniformSampling(const unsigned int random_seed, const Solver& solver)
      : num_data_(solver.num_data()), sample_size_(solver.min_sample_size()) {
    rng_.seed(random_seed);
    // Expected cost of rejection sampling vs a full shuffle (sampling.h:69-78).
    const double ratio = static_cast<double>(num_data_) /
                         static_cast<double>(num_data_ - sample_size_);
    rejection_ = ratio < M_E;
    index_.param(std::uniform_int_distribution<int>::param_type(0, num_data_ - 1));
  }

  // One minimal sample (sampling.h:58-64).
  void Sample(std::vector<int>* random_sample) {
    std::vector<int>& s = *random_sample;
    if (rejection_) {
      s.resize(sample_size_);
      for (int i = 0; i < sample_size_; ++i) {
        bool again = true;
        while (again) {  // redraw until distinct from the earlier picks (sampling.h:81-97)
          s[i] = index_(rng_);
          again = false;
          for (int j = 0; j < i && !again; ++j) again = (s[j] == s[i]);
        }
      }
      return;
    }
    // Shuffle path (sampling.h:102-135).
    s.resize(num_data_);
    std::iota(s.begin(), s.end(), 0);
    if (sample_size_ == num_data_) return;
    for (int i = 0; i + 1 < num_data_; ++i) {
      std::uniform_int_distribution<int> pick(i, num_data_ - 1);
      std::swap(s[i], s[pick(rng_)]);
    }
    s.resize(sample_size_);
  }

  // Batched form used by the GPU driver: `count` consecutive samples, row-major.
  void SampleBatch(const int count, std::vector<int>* flat) {
    flat->resize(static_cast<size_t>(count) * sample_size_);
    std::vector<int> one;
    for (int k = 0; k < count; ++k) {
      Sample(&one);
      std::copy(one.begin(), one.end(), flat->begin() + static_cast<size_t>(k) * sample_size_);
    }
  }

  int sample_size() const { return sample_size_; }

 protected:
  std::mt19937 rng_;
  std::uniform_int_distribution<int> index_;
  int num_data_;
  int sample_size_;
  bool rejection_;
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_SAMPLING_H_
