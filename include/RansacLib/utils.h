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

// ---- the local-optimisation stream of the GPU drivers ------------------------------------------
// The reference's LO spends most of its host time here: every LeastSquaresFit / LO step shuffles the
// WHOLE inlier list (n - 1 draws and n - 1 random swaps, utils.h:48-73) to keep <= 28 entries. The
// drivers in ransac.h / hybrid_ransac.h only need (a) the kept prefix and (b) the engine left exactly
// where std::mt19937 would be, so they use an engine with the same output stream whose block of 624
// tempered words is accessible: the draws of the steps that cannot touch the prefix are consumed by a
// branch-free (vectorisable) Lemire test over the block, and only a rejected draw (probability
// range / 2^32) is replayed one by one. Bit-identical to the functions above on std::mt19937
// (tests: golden vectors of the reference + repeated shuffles on one engine, CPU only).
class Mt19937 {
 public:
  typedef uint32_t result_type;
  static constexpr result_type min() { return 0u; }
  static constexpr result_type max() { return 0xffffffffu; }
  static constexpr result_type default_seed = 5489u;

  Mt19937() { seed(default_seed); }
  explicit Mt19937(const uint32_t s) { seed(s); }

  void seed(const uint32_t s) {  // std::mt19937::seed(value)
    uint32_t x = s;
    state_[0] = x;
    for (uint32_t i = 1; i < 624u; ++i) {
      x = 1812433253u * (x ^ (x >> 30)) + i;
      state_[i] = x;
    }
    idx_ = 624;
  }

  result_type operator()() {
    if (idx_ >= 624) Refill();
    return out_[idx_++];
  }

  void discard(unsigned long long z) {
    for (; z > 0; --z) (*this)();
  }

  // Continue another engine's stream: `block` = the 624 state words of its current block (after the twist that
  // produced it), `pos` = how many of them have been drawn (624: the next draw twists first; the state right
  // after seed() is pos = 624 as well). What a device-resident run hands back (rlb_lomsac_resume_record).
  void SetState(const uint32_t* block, const int pos) {
    for (int k = 0; k < 624; ++k) state_[k] = block[k];
    TemperBlock();
    idx_ = pos;
  }

  // std::uniform_int_distribution<int>(0, range - 1) of libstdc++ (GCC >= 11) on this engine:
  // Lemire's multiply-shift with rejection.
  uint32_t Below(const uint32_t range) {
    uint64_t prod = static_cast<uint64_t>((*this)()) * range;
    uint32_t low = static_cast<uint32_t>(prod);
    if (low < range) {
      const uint32_t thresh = (0u - range) % range;
      while (low < thresh) {
        prod = static_cast<uint64_t>((*this)()) * range;
        low = static_cast<uint32_t>(prod);
      }
    }
    return static_cast<uint32_t>(prod >> 32);
  }

  // Consumes the draws of Fisher-Yates steps i = first .. first + count - 1 over a list of n entries
  // (step i draws from [i, n - 1], i.e. range n - i) without using their values.
  void ConsumeShuffleDraws(const int n, int first, int count) {
    while (count > 0) {
      if (idx_ >= 624) Refill();
      const int chunk = std::min(count, 624 - idx_);
      const uint32_t* w = out_ + idx_;
      const uint32_t range0 = static_cast<uint32_t>(n - first);
      uint32_t maybe_rejected = 0u;
      for (int k = 0; k < chunk; ++k) {
        const uint32_t range = range0 - static_cast<uint32_t>(k);
        const uint32_t low = w[k] * range;  // low 32 bits of the 64-bit product
        maybe_rejected |= (low < range) ? 1u : 0u;
      }
      if (!maybe_rejected) {
        idx_ += chunk;
      } else {
        for (int k = 0; k < chunk; ++k) (void)Below(static_cast<uint32_t>(n - (first + k)));
      }
      first += chunk;
      count -= chunk;
    }
  }

 private:
  void Refill() {
    uint32_t* s = state_;
    for (int k = 0; k < 227; ++k) s[k] = Twist(s[k], s[k + 1], s[k + 397]);
    for (int k = 227; k < 623; ++k) s[k] = Twist(s[k], s[k + 1], s[k - 227]);
    s[623] = Twist(s[623], s[0], s[396]);
    TemperBlock();
    idx_ = 0;
  }
  void TemperBlock() {
    for (int k = 0; k < 624; ++k) {
      uint32_t y = state_[k];
      y ^= y >> 11;
      y ^= (y << 7) & 0x9d2c5680u;
      y ^= (y << 15) & 0xefc60000u;
      y ^= y >> 18;
      out_[k] = y;
    }
  }
  static uint32_t Twist(const uint32_t a, const uint32_t b, const uint32_t src) {
    const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return src ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
  }

  uint32_t state_[624];
  uint32_t out_[624];  // the current block, tempered
  int idx_;            // next word of out_; 624 = exhausted
};

// utils.h:69-73 on Mt19937: same kept prefix, same engine position afterwards. Only the first
// min(target_size, n - 1) steps can move an entry into the prefix; later steps swap positions >= i.
inline void RandomShuffleAndResize(const int target_size, Mt19937* rng, std::vector<int>* v) {
  const int n = static_cast<int>(v->size());
  const int steps = n > 0 ? n - 1 : 0;
  const int t = std::max(0, std::min(target_size, steps));
  for (int i = 0; i < t; ++i) {
    const int j = i + static_cast<int>(rng->Below(static_cast<uint32_t>(n - i)));
    std::swap((*v)[i], (*v)[j]);
  }
  rng->ConsumeShuffleDraws(n, t, steps - t);
  v->resize(target_size);
}

// utils.h:96-105 on Mt19937.
inline void RandomShuffleAndResize(const std::vector<int> sample_sizes, Mt19937* rng,
                                   std::vector<std::vector<int>>* v) {
  for (size_t t = 0; t < v->size(); ++t) {
    const int have = static_cast<int>((*v)[t].size());
    RandomShuffleAndResize(std::min(have, sample_sizes[t]), rng, &(*v)[t]);
  }
}

// The same shuffle over a list the caller only has a read-only view of (the pinned result buffer of a
// fused pass): *out receives what RandomShuffleAndResize(target_size, rng, &copy_of_list) would leave in
// the copy, and the engine ends at the same position - without copying the n entries. The few entries the
// first steps displace beyond the prefix are remembered in a short side list.
inline void ShuffleSampleFromList(const int target_size, Mt19937* rng, const int* list, const int n,
                                  std::vector<int>* out) {
  const int steps = n > 0 ? n - 1 : 0;
  const int t = std::max(0, std::min(target_size, steps));
  if (t > 64) {  // long samples: the side list would be searched too often
    out->assign(list, list + n);
    RandomShuffleAndResize(target_size, rng, out);
    return;
  }
  const int head = std::max(0, std::min(target_size, n));
  out->assign(static_cast<size_t>(std::max(target_size, 0)), 0);
  std::copy(list, list + head, out->begin());
  int far_pos[64], far_val[64], num_far = 0;
  for (int i = 0; i < t; ++i) {
    const int j = i + static_cast<int>(rng->Below(static_cast<uint32_t>(n - i)));
    if (j < head) {
      std::swap((*out)[i], (*out)[j]);
      continue;
    }
    int k = 0;
    while (k < num_far && far_pos[k] != j) ++k;
    if (k == num_far) {
      far_pos[k] = j;
      far_val[k] = list[j];
      ++num_far;
    }
    std::swap((*out)[i], far_val[k]);
  }
  rng->ConsumeShuffleDraws(n, t, steps - t);
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
