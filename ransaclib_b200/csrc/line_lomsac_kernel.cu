// Device-resident LO-MSAC for the 2D line solver: whole EstimateModel runs, one problem per CTA.
//
// SURVEY.md section 8(f) rank 1 names two refinements to move onto the device so that "a whole LO
// chain runs without host round trips": the pose LM (localize_kernels.cu) and the 2x2 covariance
// line fit (line_estimator.cc:74-107). This file is the second one, for BASELINE configs[0]
// (examples/line_estimation.cc: N = 10,000 points, 100 .. 100,000 iterations): driven from the
// host, that run is a chain of 130 .. 470 dependent K = 1 passes (one launch + one sync each) with
// a full-list Fisher-Yates shuffle and a tiny fit between them; here one CTA of 512 threads runs
// the complete control flow of LocallyOptimizedMSAC::EstimateModel (ransac.h:128-275) in ONE launch.
//
//   reference                                              here (all 512 threads, CTA-uniform)
//   ------------------------------------------------------ ------------------------------------
//   UniformSampling (sampling.h:47-135), std::mt19937      Mt: double-buffered state in shared
//                                                          memory, the twist is three 227-wide rounds;
//                                                          fill_samples: Lemire test of a whole 624-
//                                                          word block in parallel, one thread pairs
//   utils::RandomShuffleAndResize (utils.h:48-73)          shuffle_resize: the <= 32 steps that can
//                                                          touch the kept prefix on a sparse view, the
//                                                          other n-1-32 draws checked 624 at a time
//   MinimalSolver (line_estimator.cc:53-72)                line_fit::ThroughTwoPoints, thread = sample
//   ScoreModel of a batch of hypotheses (ransac.h:295-309) sweep: thread = hypothesis, broadcast loads
//   ScoreModel / GetInliers that decide (ransac.h:295-336) cta_pass: reference formulation, ordered
//                                                          compaction with one barrier per 16k points
//   NonMinimalSolver / LeastSquares (.cc:74-107)           line_fit::FitCovariance (one thread, the
//                                                          host's order of operations)
//   LocalOptimization / LeastSquaresFit (ransac.h:341-420) local_optimization / score_then_fit
//   NumRequiredIterations (utils.h:110-140)                num_required_iterations
//
// Everything every thread knows alike - options, data pointers, RNG positions, counters - lives ONCE per
// CTA in shared memory and models travel by value in registers: per-thread local memory would be 512
// copies, and with the points staged in shared memory only ~30 KB of L1 are left to hold them.
//
// The sequence of decisions is that of include/RansacLib/ransac.h (speculative batches, every
// "maybe better" hypothesis settled by an exact pass, sequential replay of ransac.h:158-240).
// Compiled with -fmad=false: samples, shuffles, hypotheses, per-point errors and refits are those
// of a strict-IEEE CPU run (log/pow come from the CUDA math library: <= 2 ulp from glibc). Only
// the speculative sweep uses explicit FMAs; it never decides anything by itself.
#include <algorithm>
#include <cfloat>
#include <climits>

#include "../../examples/line_fit.h"
#include "rlb_internal.cuh"

namespace rlb {

namespace {

namespace line_fit = ransac_lib::line_fit;

constexpr int kT = 512;  // threads per CTA = largest speculative batch
constexpr int kWarps = kT / 32;
constexpr unsigned kFull = 0xffffffffu;
constexpr double kNever = DBL_MAX;
constexpr int kFirstBatch = 128;
constexpr int kSuper = kT * 32;  // points per ordered-compaction round (one 32-bit mask per lane)
// hand-back record (include/ransaclib_b200.h: rlb_lomsac_resume_record), in int32 words
constexpr int kResumeWords = RLB_LOMSAC_RESUME_WORDS;
constexpr int32_t kResumeMagic = RLB_LOMSAC_RESUME_MAGIC;

struct Line {
  double a, b, c;
};

struct __align__(16) Shared {
  uint32_t mt[2][2 * 624];  // [sampler, local optimisation][double buffer]
  double part[kT];          // sweep partial sums [group][hypothesis]; staging area of the fits
  double models[kT * 3];    // hypotheses of the current batch
  double scores[kT];
  int samples[kT * 2];
  int vals[624];            // sampler: accepted draw of each state word, -1 = Lemire rejection
  unsigned char valid[kT];
  double wscore[kWarps];
  int wcount[kWarps];
  int samp[32];  // output of shuffle_resize
  double fit[3];
  int fit_ok;
  int have_first, first_val, walk_idx, walk_produced;  // sampler pairing state
  int rejpos;
  // ---- state of the current problem (CTA-uniform; written by thread 0 between barriers) ----
  const double2* xy;  // n x (x, y), the reference's Matrix2Xd layout (global or the staged copy)
  int n;
  rlb_localize_options opt;
  int32_t* list_best;  // stats.inlier_indices
  int list_best_is_output;  // list_best is the caller's inlier buffer (it comes back to the host)
  int32_t* list_base;  // LocalOptimization's inliers_base
  int32_t* list_fit;   // inliers of the current fit / scratch
  int mt_cur[2], mt_idx[2];  // per engine: which half holds the current state, next word to temper
  int passes, lsq_calls;
  // debug_flags bit 1: cycles per phase (0 sampler, 1 minimal solver + sweep, 2 exact passes, 3 shuffles,
  // 4 fits, 5 whole problem, 6 seeding)
  long long prof[16];  // 8.. : inside cta_pass (evaluation, reduction + barrier, list, closing barrier)
};
#define RLB_LL_PROF_BEGIN() const long long prof_t0__ = clock64()
#define RLB_LL_PROF_END(sh, slot) \
  if (threadIdx.x == 0) (sh).prof[slot] += clock64() - prof_t0__

// ----------------------------------------------------------------- std::mt19937
// Register view of one engine for the duration of a call; load() / store() bracket it.
struct Mt {
  uint32_t* base;  // 2 x 624 words
  int cur;         // which half holds the current state
  int idx;         // next word to temper; 624 = exhausted (state right after seed())
  __device__ __forceinline__ uint32_t* state() const { return base + cur * 624; }
};
__device__ __forceinline__ Mt mt_load(Shared& sh, int engine) {
  Mt r;
  r.base = sh.mt[engine];
  r.cur = sh.mt_cur[engine];
  r.idx = sh.mt_idx[engine];
  return r;
}
// Call right before the barrier that ends the function.
__device__ __forceinline__ void mt_store(Shared& sh, int engine, const Mt& r) {
  if (threadIdx.x == 0) {
    sh.mt_cur[engine] = r.cur;
    sh.mt_idx[engine] = r.idx;
  }
}

__device__ __forceinline__ uint32_t twist_word(uint32_t a, uint32_t b, uint32_t src) {
  const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
  return src ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}

// New state into the other half: word k needs the OLD k, k+1 and either the old k+397 (k < 227) or
// the NEW k-227, so the 624 words fall into three rounds of <= 227 independent words. Writing the
// other half means no round has to separate its reads from its writes. Collective (whole CTA).
__device__ __forceinline__ void mt_twist(Mt& r) {
  const int tid = threadIdx.x;
  const uint32_t* o = r.state();
  uint32_t* n = r.base + (r.cur ^ 1) * 624;
  if (tid < 227) n[tid] = twist_word(o[tid], o[tid + 1], o[tid + 397]);
  __syncthreads();
  if (tid < 227) {
    const int k = tid + 227;
    n[k] = twist_word(o[k], o[k + 1], n[k - 227]);
  }
  __syncthreads();
  if (tid < 170) {
    const int k = tid + 454;
    const uint32_t next = (k == 623) ? n[0] : o[k + 1];
    n[k] = twist_word(o[k], next, n[k - 227]);
  }
  __syncthreads();
  r.cur ^= 1;
  r.idx = 0;
}

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// Next 32-bit draw; CTA-uniform (every thread reads the same word and keeps the same idx).
__device__ __forceinline__ uint32_t mt_next(Mt& r) {
  if (r.idx >= 624) mt_twist(r);
  return mt_temper(r.state()[r.idx++]);
}

// libstdc++ std::uniform_int_distribution<int> on a 32-bit engine (bits/uniform_int_dist.h,
// GCC >= 11): Lemire's multiply-shift with rejection. Returns a value in [0, range).
__device__ __forceinline__ uint32_t lemire(Mt& r, uint32_t range) {
  unsigned long long prod = static_cast<unsigned long long>(mt_next(r)) * range;
  uint32_t low = static_cast<uint32_t>(prod);
  if (low < range) {
    const uint32_t thresh = (0u - range) % range;
    while (low < thresh) {
      prod = static_cast<unsigned long long>(mt_next(r)) * range;
      low = static_cast<uint32_t>(prod);
    }
  }
  return static_cast<uint32_t>(prod >> 32);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// `count` minimal samples of two distinct indices into sh.samples (sampling.h:58-121).
__device__ __noinline__ void fill_samples(Shared& sh, int count) {
  Mt r = mt_load(sh, 0);
  const int tid = threadIdx.x;
  const int n = sh.n;
  RLB_LL_PROF_BEGIN();
  if (n <= 3) {
    // n / (n - 2) >= e: ShuffleSample. n == 2 returns iota without a draw (sampling.h:114).
    for (int k = 0; k < count; ++k) {
      int a0 = 0, a1 = 1, a2 = 2;
      if (n == 3) {
        const int j0 = static_cast<int>(lemire(r, 3u));  // i = 0: swap(a[0], a[j0])
        if (j0 == 1) {
          const int t = a0;
          a0 = a1;
          a1 = t;
        } else if (j0 == 2) {
          const int t = a0;
          a0 = a2;
          a2 = t;
        }
        const int j1 = 1 + static_cast<int>(lemire(r, 2u));  // i = 1: swap(a[1], a[j1])
        if (j1 == 2) a1 = a2;
      }
      if (tid == 0) {
        sh.samples[2 * k] = a0;
        sh.samples[2 * k + 1] = a1;
      }
    }
    mt_store(sh, 0, r);
    __syncthreads();
    return;
  }
  // DrawSample: every draw has the same range n, so the Lemire test of a whole block of state
  // words is independent per word; one thread then walks the accepted values and pairs them
  // (second index redrawn while it equals the first).
  const uint32_t range = static_cast<uint32_t>(n);
  const uint32_t thresh = (0u - range) % range;
  int produced = 0;
  while (produced < count) {
    if (r.idx >= 624) mt_twist(r);
    const uint32_t* s = r.state();
    for (int w = r.idx + tid; w < 624; w += kT) {
      const unsigned long long prod = static_cast<unsigned long long>(mt_temper(s[w])) * range;
      const uint32_t low = static_cast<uint32_t>(prod);
      sh.vals[w] = (low < thresh) ? -1 : static_cast<int>(prod >> 32);
    }
    __syncthreads();
    if (tid == 0) {
      int w = r.idx, p = produced, hf = sh.have_first, a = sh.first_val;
      while (w < 624 && p < count) {
        const int v = sh.vals[w++];
        if (v < 0) continue;
        if (!hf) {
          a = v;
          hf = 1;
        } else if (v != a) {
          sh.samples[2 * p] = a;
          sh.samples[2 * p + 1] = v;
          ++p;
          hf = 0;
        }
      }
      sh.have_first = hf;
      sh.first_val = a;
      sh.walk_idx = w;
      sh.walk_produced = p;
    }
    __syncthreads();
    r.idx = sh.walk_idx;
    produced = sh.walk_produced;
    __syncthreads();
  }
  mt_store(sh, 0, r);
  RLB_LL_PROF_END(sh, 0);
  __syncthreads();
}

// utils::RandomShuffleAndResize(target, rng, &copy_of(list[0..n))) -> sh.samp[0..target).
// The reference shuffles the WHOLE list (n-1 draws, Fisher-Yates from the front) and keeps the
// first `target` entries; only the first min(target, n-1) steps can touch those. Their draws are
// taken first (one lane each when the state block has enough words left and none is rejected), the
// list entries they name are requested, then the remaining steps consume their draws - the Lemire
// test of up to 624 of them at a time, one word per thread; a rejection (probability range / 2^32)
// drops that one step to the sequential path - and with the entries now arrived the kept prefix is
// resolved on a sparse view of the list (lane l holds head position l; positions beyond the head
// live in at most one (pos, val) entry per step; every warp does this redundantly).
// target <= 32. Entries past n (resize growing the vector) are 0 like std::vector::resize.
__device__ __noinline__ void shuffle_resize(Shared& sh, const int32_t* list, int n, int target) {
  Mt r = mt_load(sh, 1);
  const int tid = threadIdx.x, lane = tid & 31;
  const bool force_slow = (sh.opt.debug_flags & 1) != 0;
  RLB_LL_PROF_BEGIN();
  const int steps = (n > 0) ? n - 1 : 0;
  const int head_n = min(target, n);
  const int t_steps = min(target, steps);
  int my_j = -1;
  bool drawn = false;
  if (!force_slow && t_steps > 0 && r.idx + t_steps <= 624) {
    // step `lane` takes state word idx + lane; valid unless some step's draw is rejected
    bool rej = false;
    if (lane < t_steps) {
      const uint32_t range = static_cast<uint32_t>(n - lane);
      const unsigned long long prod = static_cast<unsigned long long>(mt_temper(r.state()[r.idx + lane])) * range;
      const uint32_t low = static_cast<uint32_t>(prod);
      if (low < range) rej = low < ((0u - range) % range);
      my_j = lane + static_cast<int>(prod >> 32);
    }
    if (__ballot_sync(kFull, rej) == 0u) {  // same words in every warp: same verdict
      drawn = true;
      r.idx += t_steps;
    } else {
      my_j = -1;
    }
  }
  if (!drawn) {
    for (int i = 0; i < t_steps; ++i) {
      const int j = i + static_cast<int>(lemire(r, static_cast<uint32_t>(n - i)));
      if (lane == i) my_j = j;
    }
  }
  int head = (lane < head_n) ? list[lane] : 0;
  const int pre = (lane < t_steps && my_j >= head_n) ? list[my_j] : 0;

  // draws of the steps that cannot reach the kept prefix
  int i = t_steps;
  int remaining = steps - t_steps;
  while (remaining > 0) {
    if (r.idx >= 624) mt_twist(r);
    const int chunk = min(remaining, 624 - r.idx);
    const uint32_t* s = r.state();
    int rej = INT_MAX;
    for (int w = tid; w < chunk; w += kT) {
      const uint32_t range = static_cast<uint32_t>(n - (i + w));
      const unsigned long long prod = static_cast<unsigned long long>(mt_temper(s[r.idx + w])) * range;
      const uint32_t low = static_cast<uint32_t>(prod);
      if (low < range && low < ((0u - range) % range)) rej = min(rej, w);
    }
    int any = __syncthreads_or(rej != INT_MAX);
    if (force_slow) any = 1;
    if (!any) {
      r.idx += chunk;
      i += chunk;
      remaining -= chunk;
    } else {
      if (tid == 0) sh.rejpos = force_slow ? 0 : INT_MAX;
      __syncthreads();
      if (rej != INT_MAX) atomicMin(&sh.rejpos, rej);
      __syncthreads();
      const int f = sh.rejpos;
      __syncthreads();
      r.idx += f;
      i += f;
      remaining -= f;
      (void)lemire(r, static_cast<uint32_t>(n - i));  // this step, exactly
      i += 1;
      remaining -= 1;
    }
  }

  // A step whose j lies beyond the head and is hit by no other step ("simple") leaves head[step] = list[j] and a
  // displaced entry nobody reads; it reads nothing an earlier step wrote and no later step reads its head entry. So
  // only the other steps (j inside the head, or a j shared by several steps) go through the sequential resolution,
  // in order; the simple ones are applied at the end (csrc/localize_kernels.cu: shuffle_resize, same argument).
  const bool is_step = lane < t_steps;
  const unsigned same_j = __match_any_sync(kFull, is_step ? my_j : -1 - lane);
  const bool simple = !force_slow && is_step && my_j >= head_n && __popc(same_j) == 1;
  unsigned todo = __ballot_sync(kFull, is_step && !simple);
  int pos = -1, val = 0;
  while (todo) {
    const int k = __ffs(todo) - 1;
    todo &= todo - 1u;
    const int j = __shfl_sync(kFull, my_j, k);
    const int a = __shfl_sync(kFull, head, k);
    int b;
    if (j < head_n) {
      b = __shfl_sync(kFull, head, j);
      if (lane == j) head = a;
    } else {
      const unsigned match = __ballot_sync(kFull, pos == j);
      if (match) {
        const int m = __ffs(match) - 1;
        b = __shfl_sync(kFull, val, m);
        if (lane == m) val = a;
      } else {
        b = __shfl_sync(kFull, pre, k);
        if (lane == k) {
          pos = j;
          val = a;
        }
      }
    }
    if (lane == k) head = b;
  }
  if (simple) head = pre;
  if (tid < target) sh.samp[tid] = (tid < head_n) ? head : 0;
  mt_store(sh, 1, r);
  RLB_LL_PROF_END(sh, 3);
  __syncthreads();
}

struct PassResult {
  int count;     // points with error < thr_inl
  double score;  // MSAC score at thr_score (0 unless want_score)
};

// ScoreModel fused with GetInliers for ONE model (ransac.h:295-336), reference formulation
// (line_estimator.cc:110-114). Warp w owns a contiguous segment of 32 * per points of the round and
// lane l its points l, l + 32, ... (coalesced 16-byte loads, conflict-free from shared memory); the
// lane keeps one mask bit per point, so the ascending list needs one barrier (for the warps' counts)
// per round of kSuper points and is written with coalesced stores, one ballot per 32 points. (Giving
// every thread a contiguous run instead saves the ballots but scatters the stores: measured slower.)
// Sums: lane (five interleaved partial sums over its points, added at the end) -> warp butterfly ->
// warps in order -> rounds in order.
__device__ __noinline__ PassResult cta_pass(Shared& sh, const Line m, double thr_inl, bool want_score,
                                            double thr_score, int32_t* __restrict__ list) {
  constexpr int U = 5;  // points in flight per lane; 20 = 4 x 5 points per lane at N = 10,000
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double2* __restrict__ xy = sh.xy;
  const int n = sh.n;
  const unsigned lt = (1u << lane) - 1u;
  RLB_LL_PROF_BEGIN();
  int off_base = 0;
  double total = 0.0;
  for (int sb = 0; sb < n; sb += kSuper) {
    const int nb = min(n - sb, kSuper);
    const int per = (nb + kT - 1) / kT;  // <= 32
    const int seg = sb + warp * 32 * per;
    const int end = sb + nb;
    unsigned mask = 0u;
    double sc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) sc[u] = 0.0;
    for (int j0 = 0; j0 < per; j0 += U) {
      double2 p[U];
      if (j0 + U <= per && seg + (j0 + U) * 32 <= end) {
        // all 32 x U points of this step exist (warp-uniform test): no per-point predicates
#pragma unroll
        for (int u = 0; u < U; ++u) p[u] = xy[seg + (j0 + u) * 32 + lane];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const double r = m.a * p[u].x + m.b * p[u].y + m.c * 1.0;
          const double e = r * r;
          if (want_score) sc[u] += (thr_score < e) ? thr_score : e;  // std::min(e, thr): NaN propagates
          mask |= (e < thr_inl ? 1u : 0u) << (j0 + u);
        }
      } else {
        bool live[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int i = seg + (j0 + u) * 32 + lane;
          live[u] = (j0 + u < per) && (i < end);
          p[u] = live[u] ? xy[i] : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const double r = m.a * p[u].x + m.b * p[u].y + m.c * 1.0;
          const double e = r * r;
          if (live[u]) {
            if (want_score) sc[u] += (thr_score < e) ? thr_score : e;
            if (e < thr_inl) mask |= 1u << (j0 + u);
          }
        }
      }
    }
    const long long t1 = clock64();
    int cnt = __popc(mask);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(kFull, cnt, o);
    double ws = 0.0;
    if (want_score) ws = warp_sum(((sc[0] + sc[1]) + (sc[2] + sc[3])) + sc[4]);
    if (lane == 0) {
      sh.wcount[warp] = cnt;
      sh.wscore[warp] = ws;
    }
    __syncthreads();
    const long long t2 = clock64();
    int off = off_base, tot = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const int cw = sh.wcount[w];
      if (w < warp) off += cw;
      tot += cw;
      if (want_score) total += sh.wscore[w];
    }
    if (list != nullptr) {
      int32_t* out = list + off;
      int idx = seg + lane;
#pragma unroll 4
      for (int j = 0; j < per; ++j) {
        const bool in = ((mask >> j) & 1u) != 0u;
        const unsigned bl = __ballot_sync(kFull, in);
        if (in) out[__popc(bl & lt)] = idx;
        out += __popc(bl);
        idx += 32;
      }
    }
    off_base += tot;
    const long long t3 = clock64();
    __syncthreads();
    if (tid == 0) {
      sh.prof[8] += t1 - prof_t0__;
      sh.prof[9] += t2 - t1;
      sh.prof[10] += t3 - t2;
      sh.prof[11] += clock64() - t3;
    }
  }
  if (tid == 0) ++sh.passes;
  RLB_LL_PROF_END(sh, 2);
  PassResult out;
  out.count = off_base;
  out.score = total;
  return out;
}

// Minimal solver for the `count` samples of the batch + their MSAC scores (thread = hypothesis;
// when the batch is smaller than the CTA, kT / kpad groups of threads split the points).
__device__ __noinline__ void hypothesize_and_sweep(Shared& sh, int count) {
  const int tid = threadIdx.x;
  const double2* __restrict__ xy = sh.xy;
  const int n = sh.n;
  const double thr = sh.opt.sq_thr;
  RLB_LL_PROF_BEGIN();
  if (tid < count) {
    const double2 p0 = xy[sh.samples[2 * tid]], p1 = xy[sh.samples[2 * tid + 1]];
    double l[3] = {0.0, 0.0, 0.0};
    const bool ok = line_fit::ThroughTwoPoints(p0.x, p0.y, p1.x, p1.y, l);
    sh.models[3 * tid] = l[0];
    sh.models[3 * tid + 1] = l[1];
    sh.models[3 * tid + 2] = l[2];
    sh.valid[tid] = ok ? 1 : 0;
  }
  __syncthreads();
  int kpad = 32;
  while (kpad < count) kpad <<= 1;
  const int groups = kT / kpad;
  const int h = tid & (kpad - 1), g = tid / kpad;
  const int per = (n + groups - 1) / groups;
  const int i0 = min(n, g * per), i1 = min(n, i0 + per);
  double a = 0.0, b = 0.0, c = 0.0;
  if (h < count) {
    a = sh.models[3 * h];
    b = sh.models[3 * h + 1];
    c = sh.models[3 * h + 2];
  }
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  int i = i0;
  for (; i + 3 < i1; i += 4) {  // every lane of a warp reads the same four points (broadcast)
    const double2 p = xy[i], q = xy[i + 1], u = xy[i + 2], v = xy[i + 3];
    const double rp = __fma_rn(a, p.x, __fma_rn(b, p.y, c));
    const double rq = __fma_rn(a, q.x, __fma_rn(b, q.y, c));
    const double ru = __fma_rn(a, u.x, __fma_rn(b, u.y, c));
    const double rv = __fma_rn(a, v.x, __fma_rn(b, v.y, c));
    const double ep = rp * rp, eq = rq * rq, eu = ru * ru, ev = rv * rv;
    s0 += (thr < ep) ? thr : ep;
    s1 += (thr < eq) ? thr : eq;
    s2 += (thr < eu) ? thr : eu;
    s3 += (thr < ev) ? thr : ev;
  }
  for (; i < i1; ++i) {
    const double2 p = xy[i];
    const double rp = __fma_rn(a, p.x, __fma_rn(b, p.y, c));
    const double ep = rp * rp;
    s0 += (thr < ep) ? thr : ep;
  }
  sh.part[tid] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (tid < count) {
    double t = 0.0;
    for (int gg = 0; gg < groups; ++gg) t += sh.part[gg * kpad + tid];
    // MinimalSolver returned nothing (ransac.h:185): never scored, never wins
    sh.scores[tid] = sh.valid[tid] ? t : kNever;
  }
  RLB_LL_PROF_END(sh, 1);
  __syncthreads();
}

// utils.h:110-140.
__device__ __noinline__ uint32_t num_required_iterations(double inlier_ratio, double prob_missing, int sample_size,
                                                         uint32_t min_it, uint32_t max_it) {
  if (inlier_ratio <= 0.0) return max_it;
  if (inlier_ratio >= 1.0) return min_it;
  const double prob_bad = 1.0 - pow(inlier_ratio, static_cast<double>(sample_size));
  if (prob_bad >= 0.99999999999999) return max_it;
  const double needed = ceil(log(prob_missing) / log(prob_bad) + 0.5);
  const uint32_t nd = static_cast<uint32_t>(needed);
  return max(min_it, min(nd, max_it));
}

struct Scored {
  Line m;
  double score;
};

// UpdateBestModel (ransac.h:422-428): strict, the earlier model keeps ties.
__device__ __forceinline__ void update_best(double score, const Line& m, Scored& best) {
  if (score < best.score) {
    best.score = score;
    best.m = m;
  }
}

struct FitResult {
  bool ok;
  Line m;
};

// LineEstimator::NonMinimalSolver on sample idx[0..k) (line_estimator.cc:74-107): sequential sums,
// so one thread runs the host's own code; the sample's points are gathered into shared memory by k
// threads first (thread 0 alone would wait for 2k dependent loads). ok = false (model to be left
// untouched) for k < 6 or a failed fit.
__device__ __noinline__ FitResult fit_covariance(Shared& sh, const int32_t* idx, int k) {
  FitResult out;
  out.ok = false;
  out.m.a = out.m.b = out.m.c = 0.0;
  if (threadIdx.x == 0) ++sh.lsq_calls;
  if (k < 6) return out;
  RLB_LL_PROF_BEGIN();
  constexpr int kStaged = kT / 2;  // double2 slots in sh.part
  double2* staged = reinterpret_cast<double2*>(sh.part);
  const bool use_staged = k <= kStaged;
  const double2* xy = sh.xy;
  if (use_staged && static_cast<int>(threadIdx.x) < k) staged[threadIdx.x] = xy[idx[threadIdx.x]];
  __syncthreads();
  if (threadIdx.x == 0) {
    double l[3];
    bool ok;
    if (use_staged) {
      ok = line_fit::FitCovariance(
          k,
          [&](int i, double* x, double* y) {
            const double2 p = staged[i];
            *x = p.x;
            *y = p.y;
          },
          l);
    } else {  // final least squares over all inliers
      ok = line_fit::FitCovariance(
          k,
          [&](int i, double* x, double* y) {
            const double2 p = xy[idx[i]];
            *x = p.x;
            *y = p.y;
          },
          l);
    }
    sh.fit_ok = ok ? 1 : 0;
    sh.fit[0] = l[0];
    sh.fit[1] = l[1];
    sh.fit[2] = l[2];
  }
  __syncthreads();
  out.ok = sh.fit_ok != 0;
  out.m.a = sh.fit[0];
  out.m.b = sh.fit[1];
  out.m.c = sh.fit[2];
  RLB_LL_PROF_END(sh, 4);
  __syncthreads();
  return out;
}

// [ScoreModel(m) + UpdateBestModel] fused with the LeastSquaresFit(fit_thr, m) that follows it
// (ransac.h:409-420; include/RansacLib/ransac.h ScoreThenFit). Returns the (possibly refitted) model.
__device__ __noinline__ Line score_then_fit(Shared& sh, bool want_score, double fit_thr, Line m, Scored& best) {
  const double thr = sh.opt.sq_thr;
  const PassResult pr = cta_pass(sh, m, fit_thr, want_score, thr, sh.list_fit);
  if (want_score) update_best(pr.score, m, best);
  if (pr.count < 2) return m;
  const int cap = sh.opt.min_sample_mult * 2;
  const int target = min(cap, pr.count);
  shuffle_resize(sh, sh.list_fit, pr.count, target);
  const FitResult f = fit_covariance(sh, sh.samp, target);  // LeastSquares (line_estimator.h:69-72)
  return f.ok ? f.m : m;
}

// Algorithms 2 and 3 of Lebeda et al. in the order of ransac.h:341-407. `best` = (best minimal
// model, its score), updated in place.
__device__ __noinline__ void local_optimization(Shared& sh, Scored& best) {
  if (6 > sh.n) return;  // non_minimal_sample_size() (line_estimator.h:64)
  const double thr = sh.opt.sq_thr;
  const double mult = sh.opt.thr_mult;
  const int lo_steps = sh.opt.lo_steps, lsq_iters = sh.opt.lsq_iters;

  Scored unused;
  unused.score = -kNever;
  const Line m_init = score_then_fit(sh, false, thr * mult, best.m, unused);

  const PassResult base = cta_pass(sh, m_init, thr * mult, true, thr, sh.list_base);
  update_best(base.score, m_init, best);
  const int n_base = base.count;

  const int non_min_size = max(6, min(2 * sh.opt.non_min_sample_mult, n_base / 2));
  for (int r = 0; r < lo_steps; ++r) {
    shuffle_resize(sh, sh.list_base, n_base, non_min_size);
    const FitResult nm = fit_covariance(sh, sh.samp, non_min_size);  // NonMinimalSolver
    if (!nm.ok) continue;
    Line m = score_then_fit(sh, true, thr, nm.m, best);  // ransac.h:386-391
    double fit_thr = mult * thr;
    // Reference divides by (num_lsq_iterations_ - 1) as an int (ransac.h:395-397).
    const double step = (mult - 1.0) * thr / static_cast<double>(lsq_iters - 1);
    if (lsq_iters <= 0) continue;
    m = score_then_fit(sh, false, fit_thr, m, unused);
    for (int i = 0; i < lsq_iters; ++i) {
      fit_thr -= step;
      if (i + 1 < lsq_iters) {
        m = score_then_fit(sh, true, fit_thr, m, best);
      } else {
        const PassResult last = cta_pass(sh, m, thr, true, thr, nullptr);
        update_best(last.score, m, best);
      }
    }
  }
}

// LocallyOptimizedMSAC::EstimateModel (ransac.h:128-275) for one problem. `initial` is the
// caller's model (the reference reads it in the final least-squares step even when no hypothesis
// was ever found, ransac.h:257-259). Returns the best model.
__device__ Line run_problem(Shared& sh, const Line initial, rlb_localize_result* res) {
  const rlb_localize_options o = sh.opt;
  const int n = sh.n;
  const double thr = o.sq_thr;

  uint32_t num_iterations = 0u;
  int best_num_inliers = 0;
  double inlier_ratio = 0.0;
  int num_lo = 0;

  uint32_t max_iters = max(o.max_iters, o.min_iters);
  Scored best;  // *best_model, stats.best_model_score
  best.m = initial;
  best.score = kNever;
  Scored best_minimal;
  best_minimal.m.a = best_minimal.m.b = best_minimal.m.c = 0.0;
  best_minimal.score = kNever;

  uint32_t batch_first = 0u, batch_count = 0u, next_batch = kFirstBatch;

  auto refresh_termination = [&]() {  // ransac.h:170-177, 231-238
    best_num_inliers = cta_pass(sh, best.m, thr, false, thr, sh.list_best).count;
    inlier_ratio = static_cast<double>(best_num_inliers) / static_cast<double>(n);
    max_iters = num_required_iterations(inlier_ratio, 1.0 - o.success_prob, 2, o.min_iters, o.max_iters);
  };

  for (num_iterations = 0u; num_iterations < max_iters; ++num_iterations) {
    const uint32_t it = num_iterations;
    if (it == o.lo_start && best_minimal.score < kNever) {  // ransac.h:163-178
      ++num_lo;
      local_optimization(sh, best);
      refresh_termination();
    }

    if (it >= batch_first + batch_count) {
      // Speculate on the next iterations: samples are a pure function of the sampler stream,
      // hypotheses do not depend on the LO stream.
      const uint32_t bound = max(max_iters, it + 1u);
      if (o.handback_after > 0u && it >= o.handback_after && bound - it > o.handback_min_remaining) {
        // One SM sweeps ~7 hypothesis x point evaluations per cycle; a run that still needs thousands
        // of hypotheses is better served by the whole-GPU sweep of the host-driven loop. The caller
        // allowed handing it back (same run from scratch there: results do not depend on the driver).
        // The state goes with it (rlb_lomsac_resume_record in the caller's inlier buffer), so that the caller
        // resumes at iteration `it` instead of starting over: the local optimisation's engine (position included),
        // both best models and their scores, the LO count. The sampler is a pure function of (seed, it): replayed.
        int rec_words = 0;
        if (sh.list_best_is_output && n >= kResumeWords) {
          rec_words = kResumeWords;
          int32_t* rec = sh.list_best;
          const uint32_t* st = sh.mt[1] + sh.mt_cur[1] * 624;
          for (int i = threadIdx.x; i < 624; i += kT) rec[20 + i] = static_cast<int32_t>(st[i]);
          if (threadIdx.x == 0) {
            rec[0] = kResumeMagic;
            rec[1] = static_cast<int32_t>(it);
            rec[2] = num_lo;
            rec[3] = sh.mt_idx[1];
            const double dv[8] = {best.score, best.m.a, best.m.b, best.m.c,
                                  best_minimal.score, best_minimal.m.a, best_minimal.m.b, best_minimal.m.c};
            for (int i = 0; i < 8; ++i) {
              const long long bits = __double_as_longlong(dv[i]);
              rec[4 + 2 * i] = static_cast<int32_t>(bits & 0xffffffffLL);
              rec[5 + 2 * i] = static_cast<int32_t>(bits >> 32);
            }
          }
        }
        if (threadIdx.x == 0) {
          res->success = 2;
          res->num_inliers = rec_words;
          res->num_iterations = it;
          res->num_lo = num_lo;
          res->score = kNever;
          res->inlier_ratio = 0.0;
          res->passes = sh.passes;
          res->lsq_calls = sh.lsq_calls;
        }
        return initial;
      }
      batch_first = it;
      batch_count = min(next_batch, bound - it);
      next_batch = min(next_batch * 2u, static_cast<uint32_t>(kT));
      fill_samples(sh, static_cast<int>(batch_count));
      hypothesize_and_sweep(sh, static_cast<int>(batch_count));
    }

    const int slot = static_cast<int>(it - batch_first);
    if (!sh.valid[slot]) continue;  // ransac.h:185

    double local_score = kNever;
    const double swept = sh.scores[slot];
    const bool have = swept < kNever;  // GetBestEstimatedModelId: NaN never becomes the local best
    if (have) local_score = swept;
    const bool forced = (it == o.lo_start);
    Line cand;
    cand.a = sh.models[3 * slot];
    cand.b = sh.models[3 * slot + 1];
    cand.c = sh.models[3 * slot + 2];
    if (have && (forced || local_score <= best_minimal.score * (1.0 + 1e-9) || best_minimal.score == kNever)) {
      // the sweep says "maybe better": settle it with the exact pass
      local_score = cta_pass(sh, cand, thr, true, thr, nullptr).score;
    }

    if (local_score < best_minimal.score || forced) {  // ransac.h:195-196
      const bool new_best = local_score < best_minimal.score;
      if (new_best) {
        best_minimal.score = local_score;
        best_minimal.m = cand;
        update_best(best_minimal.score, best_minimal.m, best);
      }
      const bool run_lo = (it >= o.lo_start && best_minimal.score < kNever);
      if (!new_best && !run_lo) continue;
      if (run_lo) {
        ++num_lo;
        Scored lo = best_minimal;
        local_optimization(sh, lo);
        best_minimal.m = lo.m;  // the reference optimises best_minimal_model in place, its score by copy
        update_best(lo.score, lo.m, best);
      }
      refresh_termination();
    }
  }

  // LO for runs that ended before the start iteration (ransac.h:245-255).
  if (num_iterations <= o.lo_start && best.score < kNever) {
    ++num_lo;
    local_optimization(sh, best);
    best_num_inliers = cta_pass(sh, best.m, thr, false, thr, sh.list_best).count;
    inlier_ratio = static_cast<double>(best_num_inliers) / static_cast<double>(n);
  }

  if (o.final_lsq) {  // ransac.h:257-272
    const FitResult f = fit_covariance(sh, sh.list_best, best_num_inliers);
    const Line refined = f.ok ? f.m : best.m;
    const PassResult pr = cta_pass(sh, refined, thr, true, thr, sh.list_fit);
    if (pr.score < best.score) {
      best.score = pr.score;
      best.m = refined;
      best_num_inliers = pr.count;
      for (int i = threadIdx.x; i < pr.count; i += kT) sh.list_best[i] = sh.list_fit[i];
      inlier_ratio = static_cast<double>(pr.count) / static_cast<double>(n);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    res->success = 1;
    res->num_inliers = best_num_inliers;
    res->num_iterations = num_iterations;
    res->num_lo = num_lo;
    res->score = best.score;
    res->inlier_ratio = inlier_ratio;
    res->passes = sh.passes;
    res->lsq_calls = sh.lsq_calls;
  }
  return best.m;
}

}  // namespace

// Problems of up to A.smem_points points are copied into (dynamic) shared memory once and every pass,
// sweep and fit reads them from there; larger ones are read from global memory (L1 / L2).
__global__ void __launch_bounds__(kT, 1) line_lomsac_kernel(const LineLomsacArgs A) {
  extern __shared__ __align__(16) double2 s_xy[];
  __shared__ Shared sh;
  __shared__ int s_problem;
  const int tid = threadIdx.x;
  int32_t* lists = A.lists + static_cast<size_t>(blockIdx.x) * 3 * A.nmax;

  for (;;) {
    if (tid == 0) s_problem = static_cast<int>(atomicAdd(A.ticket, 1u));
    __syncthreads();
    const int pi = s_problem;
    __syncthreads();
    if (pi >= A.num_problems) break;

    const long long o0 = A.offsets ? A.offsets[pi] : 0;
    const long long o1 = A.offsets ? A.offsets[pi + 1] : A.n_single;
    const int n = static_cast<int>(o1 - o0);
    rlb_localize_result* res = A.results + pi;
    double* line_io = A.lines + 3 * static_cast<size_t>(pi);
    if (n < 2) {  // sample_size > num_data: nothing to estimate, model untouched (ransac.h:135-139)
      if (tid == 0) {
        res->success = 0;
        res->num_inliers = 0;
        res->num_iterations = 0u;
        res->num_lo = 0;
        res->score = kNever;
        res->inlier_ratio = 0.0;
        res->passes = 0;
        res->lsq_calls = 0;
      }
      continue;
    }
    const long long t_problem = clock64();
    const double2* xy = reinterpret_cast<const double2*>(A.xy) + o0;
    if (n <= A.smem_points) {
      for (int i = tid; i < n; i += kT) s_xy[i] = xy[i];
      xy = s_xy;
    }
    // Sampler and local optimisation own separate engines seeded alike (ransac.h:142-144,
    // sampling.h:49-55): s[i] = 1812433253 * (s[i-1] ^ (s[i-1] >> 30)) + i.
    if (tid == 0) {
      sh.xy = xy;
      sh.n = n;
      sh.opt = A.opt;
      sh.list_base = lists;
      sh.list_fit = lists + A.nmax;
      sh.list_best = (A.inliers != nullptr) ? A.inliers + o0 : lists + 2 * A.nmax;
      sh.list_best_is_output = (A.inliers != nullptr) ? 1 : 0;
      sh.passes = 0;
      sh.lsq_calls = 0;
      for (int i = 0; i < 16; ++i) sh.prof[i] = 0;
      sh.mt_cur[0] = sh.mt_cur[1] = 0;
      sh.mt_idx[0] = sh.mt_idx[1] = 624;
      sh.have_first = 0;
      sh.first_val = 0;
      uint32_t x = A.seeds ? A.seeds[pi] : A.seed_single;
      sh.mt[0][0] = x;
#pragma unroll 1
      for (int i = 1; i < 624; ++i) {
        x = 1812433253u * (x ^ (x >> 30)) + static_cast<uint32_t>(i);
        sh.mt[0][i] = x;
      }
    }
    __syncthreads();
    for (int i = tid; i < 624; i += kT) sh.mt[1][i] = sh.mt[0][i];
    __syncthreads();
    if (tid == 0) sh.prof[6] = clock64() - t_problem;

    Line initial;
    initial.a = line_io[0];
    initial.b = line_io[1];
    initial.c = line_io[2];
    const Line best = run_problem(sh, initial, res);
    __syncthreads();
    if (tid == 0) {
      line_io[0] = best.a;
      line_io[1] = best.b;
      line_io[2] = best.c;
      sh.prof[5] = clock64() - t_problem;
      if ((A.opt.debug_flags & 2) && A.inliers != nullptr && n >= 64) {
        // debug: the cycle counters replace the tail of this problem's inlier region
        long long* out = reinterpret_cast<long long*>(A.inliers + ((o0 + n - 32) & ~1LL));
        for (int i = 0; i < 16; ++i) out[i] = sh.prof[i];
      }
    }
    __syncthreads();
  }
}

cudaError_t launch_line_lomsac(const LineLomsacArgs& a, int ctas, cudaStream_t st, int* launches) {
  ++*launches;
  const size_t dyn = static_cast<size_t>(a.smem_points) * sizeof(double2);
  if (dyn > 0) {
    // per device and cheap next to a launch: set every time rather than remember per device
    cudaError_t e = cudaFuncSetAttribute(line_lomsac_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(dyn));
    if (e != cudaSuccess) return e;
  }
  line_lomsac_kernel<<<ctas, kT, dyn, st>>>(a);
  return cudaGetLastError();
}

// Largest problem that is staged in shared memory (227 KB per CTA minus the static part).
int line_lomsac_smem_points() { return 11900; }
// 512 threads x up to 128 registers: one CTA per SM.
int line_lomsac_ctas_per_sm(int) { return 1; }

}  // namespace rlb
