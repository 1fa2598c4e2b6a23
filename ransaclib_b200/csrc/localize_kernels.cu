// Device-resident batched localization: whole P3P LO-MSAC runs on the device, one query per CTA
// (localize_cta_kernel, the default: the passes over the matches are shared by the CTA's four warps,
// the f32 copy of the matches lives in shared memory) or one query per WARP (localize_kernel).
//
// SURVEY.md section 8(f) ranks 1 and 2 (device-side refinement, bit-exact device sampler) for the
// batched ransaclib_localization workload (BASELINE configs[3]): 4096 independent queries of
// ~2k matches each. A per-query host loop is a chain of ~100 dependent K=1 passes and ~50
// tiny least-squares problems per local optimisation and is bound by launch + sync latency;
// here each warp of a persistent grid takes queries from a ticket counter and runs the
// complete control flow of LocallyOptimizedMSAC::EstimateModel on the device:
//
//   reference                                              here
//   ------------------------------------------------------ --------------------------------
//   UniformSampling (sampling.h:47-135), std::mt19937      Mt (warp-parallel twist), lemire()
//   utils::RandomShuffleAndResize (utils.h:48-73)          shuffle_resize (sparse Fisher-Yates)
//   MinimalSolver = P3P + 4th point (.cc:111-137)          p3pdev::p3p_pick, one sample per lane
//   ScoreModel / GetInliers (ransac.h:295-336)             sweep32 (32 hypotheses x N, f32) +
//                                                          fused_pass (f64, reference formula)
//   LocalOptimization / LeastSquaresFit (ransac.h:341-420) local_optimization / score_then_fit
//   LeastSquares (Ceres LM, .cc:170-212)                   refine::LeastSquaresWith<WarpLineariser>
//   NumRequiredIterations (utils.h:110-140)                num_required_iterations
//   ransaclib_localization.cc:58-165                       localize_kernel (options, <=3 matches)
//
// The sequence of decisions is the one of include/RansacLib/ransac.h (the host template layer):
// speculative batches of 32 hypotheses scored in f32, every "maybe better" hypothesis settled
// by an exact f64 pass, sequential replay of ransac.h:158-240. All f64 arithmetic is compiled
// with -fmad=false and written in the reference's formulation, so per-point errors, inlier
// sets, samples, shuffles and refinements are those of a strict-IEEE CPU run (sin/cos/atan2/
// log/pow come from the CUDA math library instead of glibc: <= 2 ulp apart).
#include <cfloat>
#include <climits>

// A/B switches of the LM inner code (see DESIGN.md 3b for the measurements)
#ifndef RLB_LM_SMEM
#define RLB_LM_SMEM 0   // LM vectors/matrices: one copy per warp in shared memory (else per lane, local)
#endif
#ifndef RLB_ROLL_LM
#define RLB_ROLL_LM 0   // keep the 6x6 loops rolled
#endif
#ifndef RLB_LOC_CTA_SMEM
#define RLB_LOC_CTA_SMEM 0  // one query per CTA: 1 = stage the f32 copy of the matches in shared memory. Default 0:
                            // the sweeps that read it are 2 % of a query, and 4 x 40 KB of shared memory per SM
                            // leave ~40 KB of L1 for the warps' stack frames (profiles/r02_localize_ab.txt: +7 %)
#endif
#ifndef RLB_LOC_STREAM
#define RLB_LOC_STREAM 1    // one query per CTA: the passes read the matches with L1::no_allocate loads
#endif
#ifndef RLB_FULL_SLICES
#define RLB_FULL_SLICES 0  // form the 6 partials in two slices of 3 (smaller body, more instructions)
#endif

#include "../../examples/pose_refinement.h"
#include "../../examples/p3p.h"
#include "rlb_internal.cuh"

namespace rlb {

namespace {

namespace refine = ransac_lib::calibrated_absolute_pose::refine;

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWarpsPerCta = 4;
constexpr int kSlabDoubles = 32 * 14;  // residual rows of one 32-point chunk: 2 x (value + 6 partials)

// ----------------------------------------------------------------- std::mt19937
struct Mt {
  uint32_t* s;  // 624 words of shared memory
  int idx;      // next word to temper; 624 = exhausted (state right after seed())
};

// std::mt19937::seed(value): s[i] = 1812433253 * (s[i-1] ^ (s[i-1] >> 30)) + i.
__device__ __noinline__ void mt_seed(uint32_t* s, uint32_t seed, int lane) {
  if (lane == 0) {
    uint32_t x = seed;
    s[0] = x;
#pragma unroll 1
    for (int i = 1; i < 624; ++i) {
      x = 1812433253u * (x ^ (x >> 30)) + static_cast<uint32_t>(i);
      s[i] = x;
    }
  }
  __syncwarp();
}

// Regenerates the 624 words in place, 128 per round (four per lane: four independent load -> twist chains in
// flight): word k needs the OLD k, k+1 and either the old k+397 (k < 227) or the NEW k-227, which an earlier round
// has already written (k - 227 < base for every k of a round; k + 1 and k + 397 are never in the round's own write
// set before they are read: all reads of a round precede its writes).
__device__ __noinline__ void mt_twist(uint32_t* s, int lane) {
#pragma unroll 1
  for (int base = 0; base < 623; base += 128) {
    uint32_t v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = base + lane + 32 * q;
      v[q] = 0u;
      if (k < 623) {
        const uint32_t y = (s[k] & 0x80000000u) | (s[k + 1] & 0x7fffffffu);
        const uint32_t src = (k < 227) ? s[k + 397] : s[k - 227];
        v[q] = src ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
      }
    }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = base + lane + 32 * q;
      if (k < 623) s[k] = v[q];
    }
    __syncwarp();
  }
  if (lane == 0) {
    const uint32_t y = (s[623] & 0x80000000u) | (s[0] & 0x7fffffffu);
    s[623] = s[396] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
  }
  __syncwarp();
}

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// Next 32-bit draw; warp-uniform (every lane reads the same word).
__device__ uint32_t mt_next(Mt& r, int lane) {
  if (r.idx >= 624) {
    mt_twist(r.s, lane);
    r.idx = 0;
  }
  return mt_temper(r.s[r.idx++]);
}

// libstdc++ std::uniform_int_distribution<int> on a 32-bit engine (bits/uniform_int_dist.h,
// GCC >= 11): Lemire's multiply-shift with rejection. Returns a value in [0, range).
__device__ uint32_t lemire(Mt& r, uint32_t range, int lane) {
  unsigned long long prod = static_cast<unsigned long long>(mt_next(r, lane)) * range;
  uint32_t low = static_cast<uint32_t>(prod);
  if (low < range) {
    const uint32_t thresh = (0u - range) % range;
    while (low < thresh) {
      prod = static_cast<unsigned long long>(mt_next(r, lane)) * range;
      low = static_cast<uint32_t>(prod);
    }
  }
  return static_cast<uint32_t>(prod >> 32);
}

// One minimal sample of 4 distinct indices (sampling.h:58-110), warp-uniform.
__device__ __noinline__ void draw_sample4(Mt& r, int n, int* s, int lane) {
  const double ratio = static_cast<double>(n) / static_cast<double>(n - 4);
  if (ratio < 2.718281828459045235360287471352662498 /* M_E */) {
    for (int i = 0; i < 4; ++i) {
      bool again = true;
      while (again) {
        s[i] = static_cast<int>(lemire(r, static_cast<uint32_t>(n), lane));
        again = false;
        for (int j = 0; j < i; ++j) again = again || (s[j] == s[i]);
      }
    }
    return;
  }
  // Shuffle path: only n in {4, 5, 6} gets here (n / (n - 4) >= e).
  int a[8];
  for (int i = 0; i < 8; ++i) a[i] = i;
  if (n != 4) {
    for (int i = 0; i + 1 < n; ++i) {
      const int j = i + static_cast<int>(lemire(r, static_cast<uint32_t>(n - i), lane));
      int ai = 0, aj = 0;
      for (int q = 0; q < 8; ++q) {  // register-resident: no dynamic indexing
        ai = (q == i) ? a[q] : ai;
        aj = (q == j) ? a[q] : aj;
      }
      for (int q = 0; q < 8; ++q) a[q] = (q == i) ? aj : ((q == j) ? ai : a[q]);
    }
  }
  for (int i = 0; i < 4; ++i) s[i] = a[i];
}

// utils::RandomShuffleAndResize(target, rng, &copy_of(list[0..n))) -> samp[0..target) (shared).
// The reference shuffles the WHOLE list (n-1 draws, Fisher-Yates from the front) and keeps the
// first `target` entries; only the first min(target, n-1) steps can touch those, so they are
// resolved on a sparse view of the list (lane l holds head position l; positions beyond the
// head live in at most one (pos,val) entry per step), and the remaining steps only consume
// their draws: 32 steps per round, each lane testing its own draw for a Lemire rejection
// (probability range / 2^32); a rejection drops that one step to the sequential path.
// target <= 32. Entries past n (resize growing the vector) are 0 like std::vector::resize.
__device__ __noinline__ void shuffle_resize(Mt& r, const int32_t* list, int n, int target, int32_t* samp,
                               int lane, bool force_slow) {
  const int steps = (n > 0) ? n - 1 : 0;
  const int head_n = min(target, n);
  const int t_steps = min(target, steps);
  // The draws of the first t_steps steps, one per lane when they are the next t_steps words of the current block
  // and none of them is a Lemire rejection (p ~ n / 2^32 each); else one after the other.
  int my_j = -1;
  bool drawn = false;
  if (t_steps > 0 && !force_slow) {
    if (r.idx >= 624) {
      mt_twist(r.s, lane);
      r.idx = 0;
    }
    if (r.idx + t_steps <= 624) {
      bool rej = false;
      int j = -1;
      if (lane < t_steps) {
        const uint32_t range = static_cast<uint32_t>(n - lane);
        const unsigned long long prod = static_cast<unsigned long long>(mt_temper(r.s[r.idx + lane])) * range;
        const uint32_t low = static_cast<uint32_t>(prod);
        if (low < range) rej = low < ((0u - range) % range);
        j = lane + static_cast<int>(prod >> 32);
      }
      if (__ballot_sync(kFull, rej) == 0u) {
        my_j = j;
        r.idx += t_steps;
        drawn = true;
      }
    }
  }
  if (!drawn) {
    for (int i = 0; i < t_steps; ++i) {
      const int j = i + static_cast<int>(lemire(r, static_cast<uint32_t>(n - i), lane));
      if (lane == i) my_j = j;
    }
  }
  int head = (lane < head_n) ? list[lane] : 0;
  const int pre = (lane < t_steps && my_j >= head_n) ? list[my_j] : 0;
  // A step i whose j lies beyond the head and is hit by no other step ("simple") leaves head[i] = list[j] and a
  // displaced entry nobody reads; it reads nothing a step before it wrote and nothing after it reads head[i]
  // (later steps touch positions >= their own index). So only the other steps - j inside the head, or a j shared
  // by several steps - go through the sequential resolution, in order; the simple ones are applied at the end.
  const bool is_step = lane < t_steps;
  const unsigned same_j = __match_any_sync(kFull, is_step ? my_j : -1 - lane);
  const bool simple = is_step && my_j >= head_n && __popc(same_j) == 1;
  unsigned todo = __ballot_sync(kFull, is_step && !simple);
  if (force_slow) todo = __ballot_sync(kFull, is_step);
  int pos = -1, val = 0;
  while (todo) {
    const int i = __ffs(todo) - 1;
    todo &= todo - 1u;
    const int j = __shfl_sync(kFull, my_j, i);
    const int a = __shfl_sync(kFull, head, i);
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
        b = __shfl_sync(kFull, pre, i);
        if (lane == i) {
          pos = j;
          val = a;
        }
      }
    }
    if (lane == i) head = b;
  }
  if (simple && !force_slow) head = pre;
  // draws of the steps that cannot reach the kept prefix: 128 per round, every lane testing four of them for a
  // Lemire rejection (four independent load -> temper -> multiply chains in flight)
  int i = t_steps;
  int remaining = steps - t_steps;
  while (remaining > 0) {
    if (r.idx >= 624) {
      mt_twist(r.s, lane);
      r.idx = 0;
    }
    const int chunk = min(min(128, remaining), 624 - r.idx);
    bool rej[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int p = lane + 32 * q;
      rej[q] = false;
      if (p < chunk) {
        const uint32_t range = static_cast<uint32_t>(n - (i + p));
        const unsigned long long prod = static_cast<unsigned long long>(mt_temper(r.s[r.idx + p])) * range;
        const uint32_t low = static_cast<uint32_t>(prod);
        if (low < range) rej[q] = low < ((0u - range) % range);
      }
    }
    const unsigned m0 = __ballot_sync(kFull, rej[0]), m1 = __ballot_sync(kFull, rej[1]);
    const unsigned m2 = __ballot_sync(kFull, rej[2]), m3 = __ballot_sync(kFull, rej[3]);
    if (!force_slow && (m0 | m1 | m2 | m3) == 0u) {
      r.idx += chunk;
      i += chunk;
      remaining -= chunk;
    } else {
      // the first rejected draw of the round (position 0 when the slow path is forced)
      int f = 0;
      if (!force_slow)
        f = m0 ? __ffs(m0) - 1 : (m1 ? 32 + __ffs(m1) - 1 : (m2 ? 64 + __ffs(m2) - 1 : 96 + __ffs(m3) - 1));
      r.idx += f;
      i += f;
      remaining -= f;
      (void)lemire(r, static_cast<uint32_t>(n - i), lane);  // this step, exactly
      i += 1;
      remaining -= 1;
    }
  }
  __syncwarp();
  if (lane < target) samp[lane] = (lane < head_n) ? head : 0;
  __syncwarp();
}

// ------------------------------------------------------------------ one query
struct Query {
  int n;
  const double* p2;  // n x 2 (reference layout)
  const double* p3;  // n x 3
  double fx, fy;
  float* soa32;      // [5][npad]: X-cx, Y-cy, Z-cz, -u, -v; padding 0,0,0,NaN,NaN
  int npad;          // multiple of 4
  double centre[3];
};

struct WarpShared {
  uint32_t mt_sampler[624];
  uint32_t mt_lo[624];
  double slab[kSlabDoubles];
  double sums[32];
  // Warp-uniform state lives here ONCE per warp instead of 32 times in local memory (every lane
  // executes the same scalar code and writes identical values): the LM vectors / matrices and
  // the models handed between the phases.
#if RLB_LM_SMEM
  refine::LmWorkspace lm;
#endif
  double models[5][12];  // 0 best, 1 best minimal, 2 candidate, 3 LO initial / final refit, 4 LO step
  int32_t samp[32];
};

constexpr double kNever = DBL_MAX;

// Warp-wide sum, identical bits in every lane (commutative butterfly).
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}


// Recentred f32 copy of the query's matches for the 32-hypothesis sweeps.
__device__ __noinline__ void prepare_soa32(Query& q, int lane) {
  double c[3] = {0.0, 0.0, 0.0};
  const int take = min(q.n, 32);
  if (lane < take) {
    const int i = static_cast<int>(static_cast<long long>(lane) * q.n / take);
    c[0] = q.p3[3 * static_cast<size_t>(i)];
    c[1] = q.p3[3 * static_cast<size_t>(i) + 1];
    c[2] = q.p3[3 * static_cast<size_t>(i) + 2];
  }
  for (int k = 0; k < 3; ++k) q.centre[k] = warp_sum(c[k]) / static_cast<double>(take);
  for (int i = lane; i < q.npad; i += 32) {
    const bool live = i < q.n;
    const float nanv = __int_as_float(0x7fc00000);
    q.soa32[0 * q.npad + i] = live ? static_cast<float>(q.p3[3 * static_cast<size_t>(i)] - q.centre[0]) : 0.0f;
    q.soa32[1 * q.npad + i] = live ? static_cast<float>(q.p3[3 * static_cast<size_t>(i) + 1] - q.centre[1]) : 0.0f;
    q.soa32[2 * q.npad + i] = live ? static_cast<float>(q.p3[3 * static_cast<size_t>(i) + 2] - q.centre[2]) : 0.0f;
    q.soa32[3 * q.npad + i] = live ? -static_cast<float>(q.p2[2 * static_cast<size_t>(i)]) : nanv;
    q.soa32[4 * q.npad + i] = live ? -static_cast<float>(q.p2[2 * static_cast<size_t>(i) + 1]) : nanv;
  }
  __syncwarp();
}

// MSAC score of THIS LANE's hypothesis over all matches in f32 (ransac.h:295-309 with the
// f32 deviations documented in DESIGN.md section 3): every lane reads the same match
// (broadcast), 4 matches per vector load; f32 sums are flushed into f64 every 512 matches.
// f32 coefficients of one hypothesis for the sweeps: p = R (X - c) = R (X - centre) + R (centre - c),
// rows 0/1 pre-scaled by fx/fy. cf[0..8] = rows of R, cf[9..11] = t.
__device__ __forceinline__ void sweep_coefficients(const Query& q, const double* P, float* cf) {
  const double dc[3] = {q.centre[0] - P[9], q.centre[1] - P[10], q.centre[2] - P[11]};
  const double sc[3] = {q.fx, q.fy, 1.0};
#pragma unroll
  for (int row = 0; row < 3; ++row) {
    const double r0 = P[row], r1 = P[row + 3], r2 = P[row + 6];
    cf[3 * row + 0] = static_cast<float>(r0 * sc[row]);
    cf[3 * row + 1] = static_cast<float>(r1 * sc[row]);
    cf[3 * row + 2] = static_cast<float>(r2 * sc[row]);
    cf[9 + row] = static_cast<float>((r0 * dc[0] + r1 * dc[1] + r2 * dc[2]) * sc[row]);
  }
}

__device__ __noinline__ double sweep32(const Query& q, const double* P, double thr64) {
  float cf[12];
  sweep_coefficients(q, P, cf);
  const float* r = cf;
  const float* t = cf + 9;
  const float thr = static_cast<float>(thr64);
  const int groups = q.npad >> 2;
  const float4* X = reinterpret_cast<const float4*>(q.soa32);
  const float4* Y = reinterpret_cast<const float4*>(q.soa32 + q.npad);
  const float4* Z = reinterpret_cast<const float4*>(q.soa32 + 2 * q.npad);
  const float4* U = reinterpret_cast<const float4*>(q.soa32 + 3 * q.npad);
  const float4* V = reinterpret_cast<const float4*>(q.soa32 + 4 * q.npad);
  double total = 0.0;
  for (int g0 = 0; g0 < groups; g0 += 128) {
    const int g1 = min(groups, g0 + 128);
    float s = 0.0f;
    int c = 0;
    for (int g = g0; g < g1; ++g) {
      const float4 x = X[g], y = Y[g], z = Z[g], u = U[g], v = V[g];
      const float xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
      const float zs[4] = {z.x, z.y, z.z, z.w}, us[4] = {u.x, u.y, u.z, u.w};
      const float vs[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float pz = fmaf(r[6], xs[e], fmaf(r[7], ys[e], fmaf(r[8], zs[e], t[2])));
        const float px = fmaf(r[0], xs[e], fmaf(r[1], ys[e], fmaf(r[2], zs[e], t[0])));
        const float py = fmaf(r[3], xs[e], fmaf(r[4], ys[e], fmaf(r[5], zs[e], t[1])));
        float inv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(fmaxf(pz, 0.0f)));
        const float du = fmaf(px, inv, us[e]);
        const float dv = fmaf(py, inv, vs[e]);
        const float err = fmaf(du, du, dv * dv);
        if (err < thr) {
          s += err;
          c += 1;
        }
      }
    }
    const int live = max(0, min(q.n, 4 * g1) - 4 * g0);
    total += static_cast<double>(s) + thr64 * static_cast<double>(live - c);
  }
  return total;
}

// ScoreModel fused with GetInliers for ONE model (ransac.h:295-336), reference formulation
// (calibrated_absolute_pose_estimator.cc:154-167): lane l evaluates matches l, l+32, ...;
// the ascending inlier list is compacted with a ballot per 32 matches.
__device__ __noinline__ int fused_pass(const Query& qr, const double* __restrict__ Pin, double thr_inl, bool want_score,
                          double thr_score, int32_t* __restrict__ list, double* score, int lane) {
  // Query fields live in the caller's local memory: read them once (stores to `list` could alias)
  const int qn = qr.n;
  const double* __restrict__ qp2 = qr.p2;
  const double* __restrict__ qp3 = qr.p3;
  const double qfx = qr.fx, qfy = qr.fy;
  int off = 0;
  double sc = 0.0;
  const unsigned lt = (1u << lane) - 1u;
  double P[12];  // the model lives in registers for the whole pass (the caller's copy is in local memory)
#pragma unroll
  for (int i = 0; i < 12; ++i) P[i] = Pin[i];
  // Two matches per lane and iteration (i and i + 32): f64 results take ~18 cycles to come
  // back, so the two independent error chains overlap; the loads of the next pair are in flight
  // while the current pair is evaluated. Sums and lists keep the order of the 32-wide loop
  // (lane l adds matches l, l + 32, l + 64, ... in that order; the list is compacted per 32).
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, au = 0.0, av = 0.0;
  double b0 = 0.0, b1 = 0.0, b2 = 0.0, bu = 0.0, bv = 0.0;
  auto load = [&](int i, double& x0, double& x1, double& x2, double& u, double& v) {
    if (i < qn) {
      const double* X = qp3 + 3 * static_cast<size_t>(i);
      const double* uv = qp2 + 2 * static_cast<size_t>(i);
      x0 = X[0];
      x1 = X[1];
      x2 = X[2];
      u = uv[0];
      v = uv[1];
    }
  };
  load(lane, a0, a1, a2, au, av);
  load(lane + 32, b0, b1, b2, bu, bv);
#pragma unroll 1
  for (int base = 0; base < qn; base += 64) {
    const int ia = base + lane, ib = ia + 32;
    const bool live_a = ia < qn, live_b = ib < qn;
    double na0 = 0.0, na1 = 0.0, na2 = 0.0, nau = 0.0, nav = 0.0;
    double nb0 = 0.0, nb1 = 0.0, nb2 = 0.0, nbu = 0.0, nbv = 0.0;
    load(ia + 64, na0, na1, na2, nau, nav);
    load(ib + 64, nb0, nb1, nb2, nbu, nbv);
    double ea = DBL_MAX, eb = DBL_MAX;
    if (live_a) ea = p3pdev::reproj_sq_error_v(P, a0, a1, a2, au, av, qfx, qfy);
    if (live_b) eb = p3pdev::reproj_sq_error_v(P, b0, b1, b2, bu, bv, qfx, qfy);
    if (want_score) {  // std::min(e, thr): NaN propagates
      if (live_a) sc += (thr_score < ea) ? thr_score : ea;
      if (live_b) sc += (thr_score < eb) ? thr_score : eb;
    }
    const bool in_a = live_a && (ea < thr_inl);
    const bool in_b = live_b && (eb < thr_inl);
    const unsigned mask_a = __ballot_sync(kFull, in_a);
    const unsigned mask_b = __ballot_sync(kFull, in_b);
    if (list != nullptr) {
      if (in_a) list[off + __popc(mask_a & lt)] = ia;
      if (in_b) list[off + __popc(mask_a) + __popc(mask_b & lt)] = ib;
    }
    off += __popc(mask_a) + __popc(mask_b);
    a0 = na0; a1 = na1; a2 = na2; au = nau; av = nav;
    b0 = nb0; b1 = nb1; b2 = nb2; bu = nbu; bv = nbv;
  }
  if (want_score) *score = warp_sum(sc);
  __syncwarp();
  return off;
}

// ------------------------------------------------- one query per CTA: the shared passes
// The master warp runs the control flow of the query (run_query, exactly the code of the
// one-query-per-warp form); the other three warps wait at a barrier for a command and share every
// pass over the matches with it: the K = 1 f64 pass (cta_pass), the 32-hypothesis f32 sweep
// (cta_sweep) and the recentred f32 copy (cta_prepare). Results cross through shared memory.
enum { kCmdDone = 0, kCmdPass = 1, kCmdSweep = 2, kCmdFull = 3 };
constexpr int kTilePoints = 4096;  // inlier masks of one tile of matches: 128 words

struct CtaComm {
  int cmd;
  int qi;
  // the query (a copy of Run::q for the workers)
  int n, npad;
  const double* p2;
  const double* p3;
  double fx, fy;
  float* soa32;
  double centre[3];
  // K = 1 pass
  double P[12];
  double thr_inl, thr_score;
  int want_score;
  int32_t* list;
  double part_score[kWarpsPerCta];
  unsigned mask[kTilePoints / 32];
  // 32-hypothesis sweep: coefficient c of hypothesis h at coef[c][h]
  float coef[12][32];
  double thr64;
  double part_sweep[kWarpsPerCta][32];
  // linearisation of the pose LM at lm_c over the sample lm_idx[0..lm_k) (cta_full)
  double lm_c[6];
  int lm_k;
  const int32_t* lm_idx;
  double* slab;  // [32][14] residual rows of one chunk of 32 matches
};

// A pass streams 80 KB of matches that the next pass reads again only after every other resident CTA has
// streamed its own: letting them allocate in L1 only evicts what does live there - the warps' stack frames
// (local memory) and the LM samples. RLB_LOC_STREAM = 1 reads them through the non-allocating path.
__device__ __forceinline__ double ldg_pass(const double* p) {
#if RLB_LOC_STREAM
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
#else
  return *p;
#endif
}

// ScoreModel fused with GetInliers for ONE model, all four warps (lw = logical warp, 0 = master).
// Per-point arithmetic is fused_pass's (the reference formulation); warp lw evaluates the 32-match
// words lw, lw + 4, ... of a tile (two words in flight per lane), the inlier bits of a tile go to
// shared memory, and after one barrier every warp scans the 128 word counts itself and writes the
// list entries of its own words: ascending order, two barriers per 4096 matches.
__device__ __noinline__ int cta_pass(CtaComm* C, int lw, int lane) {
  const int n = C->n;
  const double* __restrict__ qp2 = C->p2;
  const double* __restrict__ qp3 = C->p3;
  const double qfx = C->fx, qfy = C->fy;
  const double thr_inl = C->thr_inl, thr_score = C->thr_score;
  const bool want_score = C->want_score != 0;
  int32_t* __restrict__ list = C->list;
  double P[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) P[i] = C->P[i];
  const unsigned lt = (1u << lane) - 1u;
  double sc = 0.0;
  int total = 0;
  auto load = [&](int i, bool live, double& x0, double& x1, double& x2, double& u, double& v) {
    if (live) {
      const double* X = qp3 + 3 * static_cast<size_t>(i);
      const double* uv = qp2 + 2 * static_cast<size_t>(i);
      x0 = ldg_pass(X);
      x1 = ldg_pass(X + 1);
      x2 = ldg_pass(X + 2);
      u = ldg_pass(uv);
      v = ldg_pass(uv + 1);
    }
  };
#pragma unroll 1
  for (int t0 = 0; t0 < n; t0 += kTilePoints) {
    const int tn = min(kTilePoints, n - t0);
    const int words = (tn + 31) >> 5;
    if (t0 > 0) __syncthreads();  // the previous tile's masks have been consumed
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, au = 0.0, av = 0.0;
    double b0 = 0.0, b1 = 0.0, b2 = 0.0, bu = 0.0, bv = 0.0;
    {
      const int ia = t0 + 32 * lw + lane, ib = ia + 128;
      load(ia, lw < words && ia < n, a0, a1, a2, au, av);
      load(ib, lw + 4 < words && ib < n, b0, b1, b2, bu, bv);
    }
#pragma unroll 1
    for (int j = lw; j < words; j += 8) {
      const int ia = t0 + 32 * j + lane, ib = ia + 128;
      const bool live_a = ia < n, live_b = (j + 4 < words) && (ib < n);
      double na0 = 0.0, na1 = 0.0, na2 = 0.0, nau = 0.0, nav = 0.0;
      double nb0 = 0.0, nb1 = 0.0, nb2 = 0.0, nbu = 0.0, nbv = 0.0;
      load(ia + 256, (j + 8 < words) && (ia + 256 < n), na0, na1, na2, nau, nav);
      load(ib + 256, (j + 12 < words) && (ib + 256 < n), nb0, nb1, nb2, nbu, nbv);
      double ea = DBL_MAX, eb = DBL_MAX;
      if (live_a) ea = p3pdev::reproj_sq_error_v(P, a0, a1, a2, au, av, qfx, qfy);
      if (live_b) eb = p3pdev::reproj_sq_error_v(P, b0, b1, b2, bu, bv, qfx, qfy);
      if (want_score) {  // std::min(e, thr): NaN propagates
        if (live_a) sc += (thr_score < ea) ? thr_score : ea;
        if (live_b) sc += (thr_score < eb) ? thr_score : eb;
      }
      const unsigned mask_a = __ballot_sync(kFull, live_a && (ea < thr_inl));
      const unsigned mask_b = __ballot_sync(kFull, live_b && (eb < thr_inl));
      if (lane == 0) {
        C->mask[j] = mask_a;
        if (j + 4 < words) C->mask[j + 4] = mask_b;
      }
      a0 = na0; a1 = na1; a2 = na2; au = nau; av = nav;
      b0 = nb0; b1 = nb1; b2 = nb2; bu = nbu; bv = nbv;
    }
    __syncthreads();
    // lane l owns the counts of words 4l .. 4l+3; `mine` = list offset of word 4l + lw (this warp's)
    unsigned m[4];
    int cnt[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      m[e] = (4 * lane + e < words) ? C->mask[4 * lane + e] : 0u;
      cnt[e] = __popc(m[e]);
    }
    const int s4 = cnt[0] + cnt[1] + cnt[2] + cnt[3];
    int incl = s4;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += t;
    }
    const int tile_total = __shfl_sync(kFull, incl, 31);
    const int mine = total + (incl - s4) + (lw > 0 ? cnt[0] : 0) + (lw > 1 ? cnt[1] : 0) + (lw > 2 ? cnt[2] : 0);
    const unsigned my_word = (lw == 0) ? m[0] : (lw == 1) ? m[1] : (lw == 2) ? m[2] : m[3];
    if (list != nullptr) {
#pragma unroll 1
      for (int i = 0; 4 * i + lw < words; ++i) {
        const unsigned w = __shfl_sync(kFull, my_word, i);
        const int base = __shfl_sync(kFull, mine, i);
        if ((w >> lane) & 1u) list[base + __popc(w & lt)] = t0 + 32 * (4 * i + lw) + lane;
      }
    }
    total += tile_total;
  }
  if (want_score) {
    const double s = warp_sum(sc);
    if (lane == 0) C->part_score[lw] = s;
  }
  return total;
}

// The 32 hypotheses of a speculative batch against all matches in f32 (sweep32's arithmetic): lane =
// hypothesis, warp lw takes the 512-match groups lw, lw + 4, ...
__device__ __noinline__ void cta_sweep(CtaComm* C, int lw, int lane) {
  float r[9], t[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) r[i] = C->coef[i][lane];
#pragma unroll
  for (int i = 0; i < 3; ++i) t[i] = C->coef[9 + i][lane];
  const double thr64 = C->thr64;
  const float thr = static_cast<float>(thr64);
  const int n = C->n, npad = C->npad;
  const int groups = npad >> 2;
  const float* soa = C->soa32;
  const float4* X = reinterpret_cast<const float4*>(soa);
  const float4* Y = reinterpret_cast<const float4*>(soa + npad);
  const float4* Z = reinterpret_cast<const float4*>(soa + 2 * npad);
  const float4* U = reinterpret_cast<const float4*>(soa + 3 * npad);
  const float4* V = reinterpret_cast<const float4*>(soa + 4 * npad);
  double total = 0.0;
#pragma unroll 1
  for (int g0 = 128 * lw; g0 < groups; g0 += 128 * kWarpsPerCta) {
    const int g1 = min(groups, g0 + 128);
    float s = 0.0f;
    int c = 0;
    for (int g = g0; g < g1; ++g) {
      const float4 x = X[g], y = Y[g], z = Z[g], u = U[g], v = V[g];
      const float xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
      const float zs[4] = {z.x, z.y, z.z, z.w}, us[4] = {u.x, u.y, u.z, u.w};
      const float vs[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float pz = fmaf(r[6], xs[e], fmaf(r[7], ys[e], fmaf(r[8], zs[e], t[2])));
        const float px = fmaf(r[0], xs[e], fmaf(r[1], ys[e], fmaf(r[2], zs[e], t[0])));
        const float py = fmaf(r[3], xs[e], fmaf(r[4], ys[e], fmaf(r[5], zs[e], t[1])));
        float inv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(fmaxf(pz, 0.0f)));
        const float du = fmaf(px, inv, us[e]);
        const float dv = fmaf(py, inv, vs[e]);
        const float err = fmaf(du, du, dv * dv);
        if (err < thr) {
          s += err;
          c += 1;
        }
      }
    }
    const int live = max(0, min(n, 4 * g1) - 4 * g0);
    total += static_cast<double>(s) + thr64 * static_cast<double>(live - c);
  }
  C->part_sweep[lw][lane] = total;
}

// Recentred f32 copy of the query's matches (prepare_soa32's arithmetic), all four warps.
__device__ __noinline__ void cta_prepare(CtaComm* C, int lw, int lane) {
  const int n = C->n, npad = C->npad;
  const double* p2 = C->p2;
  const double* p3 = C->p3;
  float* dst = C->soa32;
  double c[3] = {0.0, 0.0, 0.0};
  const int take = min(n, 32);
  if (lane < take) {
    const int i = static_cast<int>(static_cast<long long>(lane) * n / take);
    c[0] = p3[3 * static_cast<size_t>(i)];
    c[1] = p3[3 * static_cast<size_t>(i) + 1];
    c[2] = p3[3 * static_cast<size_t>(i) + 2];
  }
  double centre[3];
  for (int k = 0; k < 3; ++k) centre[k] = warp_sum(c[k]) / static_cast<double>(take);
  if (lw == 0 && lane < 3) C->centre[lane] = (lane == 0) ? centre[0] : (lane == 1) ? centre[1] : centre[2];
  for (int i = lw * 32 + lane; i < npad; i += 32 * kWarpsPerCta) {
    const bool live = i < n;
    const float nanv = __int_as_float(0x7fc00000);
    dst[0 * npad + i] = live ? static_cast<float>(p3[3 * static_cast<size_t>(i)] - centre[0]) : 0.0f;
    dst[1 * npad + i] = live ? static_cast<float>(p3[3 * static_cast<size_t>(i) + 1] - centre[1]) : 0.0f;
    dst[2 * npad + i] = live ? static_cast<float>(p3[3 * static_cast<size_t>(i) + 2] - centre[2]) : 0.0f;
    dst[3 * npad + i] = live ? -static_cast<float>(p2[2 * static_cast<size_t>(i)]) : nanv;
    dst[4 * npad + i] = live ? -static_cast<float>(p2[2 * static_cast<size_t>(i) + 1]) : nanv;
  }
}

// Lineariser of the pose LM with one match per lane (see refine::SequentialLineariser for the
// arithmetic it reproduces): residual rows go to the shared slab, then lane e < 27 owns one
// entry of (J^T r, lower triangle of J^T J) and lane 27 the cost, each adding the chunk's
// contributions in match order, so every sum has the host's order of operations.
struct WarpLineariser {
  static constexpr bool kSpeculativeFull = true;  // see refine::RefinePoseWith
  int k;
  const int32_t* idx;  // sample (shared or global), k entries
  const double* p2d;
  const double* p3d;
  double fx, fy;
  double* slab;  // [32][14]
  double* sums;  // [28]
  int lane;
  long long* prof;

  RLB_HD __noinline__ double Cost(const double* c) const {
#ifdef __CUDA_ARCH__
    const long long t0 = clock64();
    refine::Value cam[6];
    for (int i = 0; i < 6; ++i) cam[i] = refine::Value::Constant(c[i]);
    double cost = 0.0;
    for (int base = 0; base < k; base += 32) {
      const int s = base + lane;
      double mine = 0.0;
      if (s < k) {
        const int i = idx[s];
        refine::Value r[2];
        refine::Residual(cam, p2d[2 * static_cast<size_t>(i)], p2d[2 * static_cast<size_t>(i) + 1],
                         p3d[3 * static_cast<size_t>(i)], p3d[3 * static_cast<size_t>(i) + 1],
                         p3d[3 * static_cast<size_t>(i) + 2], fx, fy, r);
        mine = 0.5 * (r[0].a * r[0].a + r[1].a * r[1].a);
      }
      const int m = min(32, k - base);
#pragma unroll 4
      for (int j = 0; j < m; ++j) cost += __shfl_sync(kFull, mine, j);
    }
    prof[5] += clock64() - t0;
    prof[11] += 1;
    return cost;
#else
    (void)c;
    return 0.0;
#endif
  }

  RLB_HD __noinline__ double Full(const double* c, double* JtJ, double* Jtr) const {
#ifdef __CUDA_ARCH__
    const long long t0 = clock64();
    // entry owned by this lane: 0..5 = Jtr[a]; 6..26 = JtJ[a][b], b <= a; 27 = cost
    int ea = 0, eb = 0;
    if (lane < 6) {
      ea = lane;
    } else if (lane < 27) {
      int t = lane - 6;
      while (t >= ea + 1) {
        t -= ea + 1;
        ++ea;
      }
      eb = t;
    }
    // slab columns multiplied by this lane: Jtr[a] = d[a] * value, JtJ[a][b] = d[a] * d[b]
    const int ia = (lane < 27) ? 1 + ea : 0;
    const int ib = (lane < 6) ? 0 : ((lane < 27) ? 1 + eb : 0);
    double acc = 0.0;
    for (int base = 0; base < k; base += 32) {
      const int s = base + lane;
      __syncwarp();
      if (s < k) {
        const int i = idx[s];
        const double u = p2d[2 * static_cast<size_t>(i)], v = p2d[2 * static_cast<size_t>(i) + 1];
        const double X = p3d[3 * static_cast<size_t>(i)], Y = p3d[3 * static_cast<size_t>(i) + 1],
                     Z = p3d[3 * static_cast<size_t>(i) + 2];
        double* row = slab + 14 * lane;
#if RLB_FULL_SLICES
        // The six partials in two slices of three (rotation, centre): each partial is the same
        // sequence of operations as in the host's 6-wide dual numbers.
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          typedef refine::DualN<3> D3;
          D3 cam[6];
#pragma unroll
          for (int a = 0; a < 6; ++a) cam[a] = D3::Variable(c[a], a - 3 * half);
          D3 r[2];
          refine::Residual(cam, u, v, X, Y, Z, fx, fy, r);
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            row[7 * e] = r[e].a;
#pragma unroll
            for (int a = 0; a < 3; ++a) row[7 * e + 1 + 3 * half + a] = r[e].d[a];
          }
        }
#else
        refine::Dual cam[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) cam[a] = refine::Dual::Variable(c[a], a);
        refine::Dual r[2];
        refine::Residual(cam, u, v, X, Y, Z, fx, fy, r);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          row[7 * e] = r[e].a;
#pragma unroll
          for (int a = 0; a < 6; ++a) row[7 * e + 1 + a] = r[e].d[a];
        }
#endif
      }
      __syncwarp();
      const int m = min(32, k - base);
      if (lane < 28) {
        // one loop for all 28 entries: acc += row[ia] * row[ib] for residual row 0, then row 1
        // (the cost lane has ia = ib = 0 and adds 0.5 * (r0^2 + r1^2) per match)
#pragma unroll 4
        for (int j = 0; j < m; ++j) {
          const double* row = slab + 14 * j;
          const double t0 = row[ia] * row[ib];
          const double t1 = row[7 + ia] * row[7 + ib];
          if (lane == 27) {
            acc += 0.5 * (t0 + t1);
          } else {
            acc += t0;
            acc += t1;
          }
        }
      }
    }
    __syncwarp();
    if (lane < 28) sums[lane] = acc;
    __syncwarp();
    for (int a = 0; a < 6; ++a) Jtr[a] = sums[a];
    int t = 6;
    for (int a = 0; a < 6; ++a)
      for (int b = 0; b <= a; ++b) {
        const double v = sums[t++];
        JtJ[6 * a + b] = v;
        JtJ[6 * b + a] = v;
      }
    const double cost = sums[27];
    __syncwarp();
    prof[4] += clock64() - t0;
    prof[10] += 1;
    return cost;
#else
    (void)c;
    (void)JtJ;
    (void)Jtr;
    return 0.0;
#endif
  }
};

// ------------------------------------------------- the pose LM, one query per CTA
// refine::RefinePoseWith (examples/pose_refinement.h) re-expressed for a warp, with the SAME
// floating-point operations in the SAME order per result (so poses, and with them every later
// decision, are those of the host / one-query-per-warp form), but
//  * no LM state in local memory: the 6 parameters, the gradient and the lower triangle of J^T J
//    live one entry per LANE (lane e < 6: vector element e; lane 6 + a(a+1)/2 + b: matrix entry
//    (a, b), b <= a; lane 27: the cost) - exactly where the linearisation's accumulation leaves them;
//  * the 6 x 6 Cholesky factorisation and the two substitutions run across those lanes (every entry
//    subtracts its products in the host's order k = 0, 1, ...), so the dependent chain is six
//    sqrt -> div -> mul -> sub links instead of ~130 operations in sequence;
//  * the linearisation (value + 6 partials of 2 residuals per match, forward-mode) is shared by
//    the CTA's four warps: the master forms the residual values, each worker two of the six
//    partials (DualN<2>: each partial is the same sequence of operations as in the 6-wide dual
//    numbers), then 28 lanes of the master add the chunk's contributions in match order.
__device__ __forceinline__ int mat_lane(int i, int j) { return 6 + (i * (i + 1)) / 2 + j; }  // i >= j

struct LmLanes {
  int i, j;    // matrix entry of this lane (i >= j), or -1
  int ia, ib;  // slab columns this lane multiplies in the accumulation
};

__device__ __forceinline__ LmLanes lm_lanes(int lane) {
  LmLanes E;
  E.i = -1;
  E.j = -1;
  int ea = 0, eb = 0;
  if (lane < 6) {
    ea = lane;
  } else if (lane < 27) {
    int t = lane - 6;
    while (t >= ea + 1) {
      t -= ea + 1;
      ++ea;
    }
    eb = t;
    E.i = ea;
    E.j = eb;
  }
  E.ia = (lane < 27) ? 1 + ea : 0;
  E.ib = (lane < 6) ? 0 : ((lane < 27) ? 1 + eb : 0);
  return E;
}

// Rows of the chunks (all warps) + accumulation (master, lw == 0). Returns the master lanes' entries.
__device__ __noinline__ double cta_full(CtaComm* C, int lw, int lane, int ia, int ib) {
  const int k = C->lm_k;
  const int32_t* idx = C->lm_idx;
  const double* __restrict__ p2d = C->p2;
  const double* __restrict__ p3d = C->p3;
  const double fx = C->fx, fy = C->fy;
  double* slab = C->slab;
  double c[6];
#pragma unroll
  for (int a = 0; a < 6; ++a) c[a] = C->lm_c[a];
  double acc = 0.0;
#pragma unroll 1
  for (int base = 0; base < k; base += 32) {
    const int s = base + lane;
    if (base > 0) __syncthreads();  // the previous chunk's rows have been added up
    if (s < k) {
      const int i = idx[s];
      const double u = p2d[2 * static_cast<size_t>(i)], v = p2d[2 * static_cast<size_t>(i) + 1];
      const double X = p3d[3 * static_cast<size_t>(i)], Y = p3d[3 * static_cast<size_t>(i) + 1],
                   Z = p3d[3 * static_cast<size_t>(i) + 2];
      double* row = slab + 14 * lane;
      if (lw == 0) {
        refine::Value cam[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) cam[a] = refine::Value::Constant(c[a]);
        refine::Value r[2];
        refine::Residual(cam, u, v, X, Y, Z, fx, fy, r);
        row[0] = r[0].a;
        row[7] = r[1].a;
      } else {
        typedef refine::DualN<2> D2;
        const int first = 2 * (lw - 1);  // this warp's two partials
        D2 cam[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) cam[a] = D2::Variable(c[a], a - first);
        D2 r[2];
        refine::Residual(cam, u, v, X, Y, Z, fx, fy, r);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          row[7 * e + 1 + first] = r[e].d[0];
          row[7 * e + 2 + first] = r[e].d[1];
        }
      }
    }
    __syncthreads();  // the chunk's rows are complete
    if (lw == 0 && lane < 28) {
      const int m = min(32, k - base);
#pragma unroll 4
      for (int j = 0; j < m; ++j) {
        const double* row = slab + 14 * j;
        const double t0 = row[ia] * row[ib];
        const double t1 = row[7 + ia] * row[7 + ib];
        if (lane == 27) {
          acc += 0.5 * (t0 + t1);
        } else {
          acc += t0;
          acc += t1;
        }
      }
    }
  }
  return acc;
}

// In-order sum of the values lanes 0..5 hold (0.0 + v0 + v1 + ...), identical in every lane.
__device__ __forceinline__ double sum6(double v) {
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i) s += __shfl_sync(kFull, v, i);
  return s;
}

struct Run;
__device__ double lm_linearise(Run& R, double xl, const int32_t* idx, int k, const LmLanes& E);

// RefinePoseWith with Lin::kSpeculativeFull == true. cam: 6 doubles in shared memory (in/out).
__device__ __noinline__ bool lm_refine_lanes(Run& R, double* cam, const int32_t* idx, int k, int lane) {
  const LmLanes E = lm_lanes(lane);
  const bool is_vec = lane < 6;
  const bool is_mat = E.i >= 0;
  double xl = is_vec ? cam[lane] : 0.0;  // x
  double acc = lm_linearise(R, xl, idx, k, E);  // g on lanes 0..5, J^T J on 6..26, cost on 27
  double cost = __shfl_sync(kFull, acc, 27);
  if (!(cost == cost) || cost > DBL_MAX) return false;
  double radius = 1e4;
  double shrink = 2.0;
#pragma unroll 1
  for (int iter = 0; iter < 50; ++iter) {
    {  // gmax = max |g_i| with refine::Max's NaN behaviour (a NaN entry is skipped)
      double ag = is_vec ? fabs(acc) : 0.0;
      if (!(ag == ag)) ag = 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double t = __shfl_xor_sync(kFull, ag, o);
        ag = (ag < t) ? t : ag;
      }
      if (ag <= 1e-10) break;
    }
    // A = JtJ + diag(clamp(diag JtJ)) / radius, then A = L L^T across the lanes
    double v = acc;
    if (is_mat && E.i == E.j) {
      const double dgl = refine::Min(refine::Max(v, 1e-6), 1e32);
      v += dgl / radius;
    }
    bool ok = true;
    double rinv[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // 1 / L[c][c], as in refine::SolveSpd6
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const double scc = __shfl_sync(kFull, v, mat_lane(c, c));
      if (!(scc > 0.0)) {
        ok = false;
        break;
      }
      const double lcc = sqrt(scc);
      rinv[c] = 1.0 / lcc;
      if (is_mat && E.j == c) v = (E.i == c) ? lcc : v * rinv[c];
      const bool below = is_mat && E.j > c;
      const double lic = __shfl_sync(kFull, v, below ? mat_lane(E.i, c) : 0);
      const double ljc = __shfl_sync(kFull, v, below ? mat_lane(E.j, c) : 0);
      if (below) v = v - lic * ljc;
    }
    double step = 0.0;
    double model_decrease = 0.0;
    if (ok) {
      // L y = -g
      double yv = is_vec ? -acc : 0.0;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        if (lane == c) yv = yv * rinv[c];
        const double yc = __shfl_sync(kFull, yv, c);
        const bool below = is_vec && lane > c;
        const double llc = __shfl_sync(kFull, v, below ? mat_lane(lane, c) : 0);
        if (below) yv = yv - llc * yc;
      }
      // L^T x = y: x_i = (y_i - sum_{c > i, ascending} L[c][i] x_c) / L[i][i]
      double prod[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int i = 5; i >= 0; --i) {
        double sv = yv;
#pragma unroll
        for (int c = i + 1; c < 6; ++c) sv = sv - prod[c];
        if (lane == i) step = sv * rinv[i];
        const double xi = __shfl_sync(kFull, step, i);
        const bool left = is_vec && lane < i;
        const double lil = __shfl_sync(kFull, v, left ? mat_lane(i, lane) : 0);
        prod[i] = lil * xi;
      }
      // model decrease -(g.s + s.JtJ.s / 2)
      const double gs = sum6(is_vec ? acc * step : 0.0);
      double t = 0.0;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const int hi = (lane > c) ? lane : c, lo = (lane > c) ? c : lane;
        const double a = __shfl_sync(kFull, acc, is_vec ? mat_lane(hi, lo) : 0);
        const double sc = __shfl_sync(kFull, step, c);
        t += a * sc;
      }
      const double sJs = sum6(is_vec ? step * t : 0.0);
      model_decrease = -(gs + 0.5 * sJs);
      ok = model_decrease > 0.0;
    }
    if (!ok) {
      radius /= shrink;
      shrink *= 2.0;
      if (radius < 1e-32) break;
      continue;
    }
    const double xn = is_vec ? xl + step : 0.0;
    double sn = sum6(is_vec ? step * step : 0.0);
    double xnorm = sum6(is_vec ? xl * xl : 0.0);
    sn = sqrt(sn);
    xnorm = sqrt(xnorm);
    if (sn <= 1e-8 * (xnorm + 1e-8)) break;
    // linearise the trial point right away (see RefinePoseWith): its cost is Ceres' trial cost
    const double acc_next = lm_linearise(R, xn, idx, k, E);
    const double cost_new = __shfl_sync(kFull, acc_next, 27);
    const double rho = (cost - cost_new) / model_decrease;
    if (cost_new == cost_new && rho > 1e-3) {
      const double delta = cost - cost_new;
      const double prev = cost;
      xl = xn;
      acc = acc_next;
      cost = cost_new;
      const double t = 2.0 * rho - 1.0;
      radius = radius / refine::Max(1.0 / 3.0, 1.0 - t * t * t);
      radius = refine::Min(radius, 1e16);
      shrink = 2.0;
      if (fabs(delta) <= 1e-6 * prev) break;
    } else {
      radius /= shrink;
      shrink *= 2.0;
      if (radius < 1e-32) break;
    }
  }
  const bool bad = is_vec && !(xl - xl == 0.0);
  if (__any_sync(kFull, bad)) return false;
  __syncwarp();
  if (is_vec) cam[lane] = xl;
  __syncwarp();
  return true;
}

// --------------------------------------------------------- LO-MSAC control flow
struct Run {
  Query q;
  rlb_localize_options opt;
  Mt sampler;
  Mt lo_rng;
  WarpShared* sh;
  CtaComm* cta;        // one query per CTA: the mailbox shared with the worker warps (else nullptr)
  int32_t* list_best;  // stats.inlier_indices
  int32_t* list_base;  // LocalOptimization's inliers_base
  int32_t* list_fit;   // inliers of the current fit / scratch
  int lane;
  bool force_slow;
  // counters (reported per query)
  int passes;
  int lsq_calls;
  // debug_flags bit 1: cycles per phase (0 hypothesis batches, 1 sweep32, 2 fused_pass, 3 shuffles,
  // 4 Full, 5 Cost, 6 least_squares, 7 whole query, 8 LO minimal solver, 9 setup, 10 #Full, 11 #Cost)
  long long prof[12];
};
#define RLB_PROF_BEGIN(R) const long long prof_t0__ = clock64()
#define RLB_PROF_END(R, slot) (R).prof[slot] += clock64() - prof_t0__

__device__ __forceinline__ int timed_pass(Run& R, const double* P, double thr_inl, bool want_score,
                                          double thr_score, int32_t* list, double* score) {
  RLB_PROF_BEGIN(R);
  int n;
  if (R.cta != nullptr) {
    CtaComm* C = R.cta;
    if (R.lane < 12) C->P[R.lane] = P[R.lane];
    if (R.lane == 0) {
      C->thr_inl = thr_inl;
      C->thr_score = thr_score;
      C->want_score = want_score ? 1 : 0;
      C->list = list;
      C->cmd = kCmdPass;
    }
    __syncthreads();  // releases the workers
    n = cta_pass(C, 0, R.lane);
    __syncthreads();  // joins them
    if (want_score) *score = ((C->part_score[0] + C->part_score[1]) + C->part_score[2]) + C->part_score[3];
  } else {
    n = fused_pass(R.q, P, thr_inl, want_score, thr_score, list, score, R.lane);
  }
  RLB_PROF_END(R, 2);
  ++R.passes;
  return n;
}
__device__ __forceinline__ void timed_shuffle(Run& R, const int32_t* list, int n, int target) {
  RLB_PROF_BEGIN(R);
  shuffle_resize(R.lo_rng, list, n, target, R.sh->samp, R.lane, R.force_slow);
  RLB_PROF_END(R, 3);
}

// The master's side of cta_full: the trial point goes to the mailbox, the workers are released, and the
// accumulated entries come back in the lanes (no join barrier: the workers' next stop is their command
// barrier, and the slab is not written again before the master's next command).
__device__ double lm_linearise(Run& R, double xl, const int32_t* idx, int k, const LmLanes& E) {
  const long long t0 = clock64();
  CtaComm* C = R.cta;
  if (R.lane < 6) C->lm_c[R.lane] = xl;
  if (R.lane == 0) {
    C->lm_k = k;
    C->lm_idx = idx;
    C->cmd = kCmdFull;
  }
  __syncthreads();
  const double acc = cta_full(C, 0, R.lane, E.ia, E.ib);
  R.prof[4] += clock64() - t0;
  R.prof[10] += 1;
  return acc;
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

__device__ __forceinline__ void update_best(double score, const double* m, double* best_score,
                                            double* best) {
  if (score < *best_score) {  // strict: the earlier model keeps ties (ransac.h:422-428)
    *best_score = score;
#pragma unroll
    for (int i = 0; i < 12; ++i) best[i] = m[i];
  }
  __syncwarp();
}

// CalibratedAbsolutePoseEstimator::LeastSquares on sample idx[0..k).
__device__ __noinline__ void least_squares(Run& R, const int32_t* idx, int k, double* pose) {
  if (R.cta != nullptr) {
    // refine::LeastSquaresWith around the lane-distributed LM: [R | c] -> (angle-axis, c) -> refine -> back
    RLB_PROF_BEGIN(R);
    double* cam = R.sh->sums;  // 6 doubles of shared scratch
    double Rm[9], aa[3];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) Rm[3 * r + c] = pose[r + 3 * c];
    refine::RotationToAngleAxis(Rm, aa);
    const double c3 = pose[9], c4 = pose[10], c5 = pose[11];
    __syncwarp();
    if (R.lane == 0) {
      cam[0] = aa[0];
      cam[1] = aa[1];
      cam[2] = aa[2];
      cam[3] = c3;
      cam[4] = c4;
      cam[5] = c5;
    }
    __syncwarp();
    const bool usable = lm_refine_lanes(R, cam, idx, k, R.lane);
    if (usable) {
      double cm[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) cm[i] = cam[i];
      refine::AngleAxisToRotation(cm, Rm);
      __syncwarp();
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) pose[r + 3 * c] = Rm[3 * r + c];
      pose[9] = cm[3];
      pose[10] = cm[4];
      pose[11] = cm[5];
      __syncwarp();
    }
    RLB_PROF_END(R, 6);
    ++R.lsq_calls;
    return;
  }
  WarpLineariser lin;
  lin.k = k;
  lin.idx = idx;
  lin.p2d = R.q.p2;
  lin.p3d = R.q.p3;
  lin.fx = R.q.fx;
  lin.fy = R.q.fy;
  lin.slab = R.sh->slab;
  lin.sums = R.sh->sums;
  lin.lane = R.lane;
  lin.prof = R.prof;
  RLB_PROF_BEGIN(R);
#if RLB_LM_SMEM
  refine::LeastSquaresWith(lin, pose, &R.sh->lm);
#else
  refine::LmWorkspace w;
  refine::LeastSquaresWith(lin, pose, &w);
#endif
  RLB_PROF_END(R, 6);
  ++R.lsq_calls;
}

// P3P on sample[0..2] + consistency with sample[3] (calibrated_absolute_pose_estimator.cc:111-137).
__device__ __noinline__ bool minimal_solver(const Query& q, const int* s, double sq_thr, double* P) {
  double y[9], x[9];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    // PixelsToViewingRays (.cc:214-227)
    const double rx = q.p2[2 * static_cast<size_t>(s[j])] / q.fx;
    const double ry = q.p2[2 * static_cast<size_t>(s[j]) + 1] / q.fy;
    const double rz = 1.0;
    const double nrm = sqrt(rx * rx + ry * ry + rz * rz);
    y[3 * j] = rx / nrm;
    y[3 * j + 1] = ry / nrm;
    y[3 * j + 2] = rz / nrm;
#pragma unroll
    for (int c = 0; c < 3; ++c) x[3 * j + c] = q.p3[3 * static_cast<size_t>(s[j]) + c];
  }
  const double* X4 = q.p3 + 3 * static_cast<size_t>(s[3]);
  const double* uv4 = q.p2 + 2 * static_cast<size_t>(s[3]);
  return p3pdev::p3p_pick(y, x, X4, uv4, q.fx, q.fy, sq_thr, P);
}

// [ScoreModel(m) + UpdateBestModel] fused with the LeastSquaresFit(fit_thr, m) that follows it
// (ransac.h:409-420; include/RansacLib/ransac.h ScoreThenFit).
__device__ __noinline__ void score_then_fit(Run& R, bool want_score, double fit_thr, double* m, double* best_score,
                               double* best) {
  const double thr = R.opt.sq_thr;
  double score = kNever;
  const int n_inl = timed_pass(R, m, fit_thr, want_score, thr, R.list_fit, &score);
  if (want_score) update_best(score, m, best_score, best);
  if (n_inl < 4) return;
  const int cap = R.opt.min_sample_mult * 4;
  const int target = min(cap, n_inl);
  timed_shuffle(R, R.list_fit, n_inl, target);
  least_squares(R, R.sh->samp, target, m);
}

// Algorithms 2 and 3 of Lebeda et al. in the order of ransac.h:341-407.
__device__ __noinline__ void local_optimization(Run& R, double* best_minimal_model, double* score_best_minimal) {
  const int n = R.q.n;
  if (6 > n) return;  // non_minimal_sample_size() (.h:90)
  const double thr = R.opt.sq_thr;
  const double mult = R.opt.thr_mult;

  double* m_init = R.sh->models[3];
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 12; ++i) m_init[i] = best_minimal_model[i];
  __syncwarp();
  score_then_fit(R, false, thr * mult, m_init, nullptr, nullptr);

  double score = kNever;
  const int n_base = timed_pass(R, m_init, thr * mult, true, thr, R.list_base, &score);
  update_best(score, m_init, score_best_minimal, best_minimal_model);

  const int non_min_size = max(6, min(4 * R.opt.non_min_sample_mult, n_base / 2));
  for (int r = 0; r < R.opt.lo_steps; ++r) {
    timed_shuffle(R, R.list_base, n_base, non_min_size);
    // NonMinimalSolver (.cc:141-151): minimal solver on the first four, then LeastSquares.
    int s4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) s4[i] = R.sh->samp[i];
    double* m = R.sh->models[4];
    __syncwarp();
    bool solved;
    {
      RLB_PROF_BEGIN(R);
      solved = minimal_solver(R.q, s4, thr, m);
      RLB_PROF_END(R, 8);
    }
    if (!solved) continue;
    __syncwarp();
    least_squares(R, R.sh->samp, non_min_size, m);

    score_then_fit(R, true, thr, m, score_best_minimal, best_minimal_model);  // ransac.h:386-391
    double fit_thr = mult * thr;
    // Reference divides by (num_lsq_iterations_ - 1) as an int (ransac.h:395-397).
    const double step = (mult - 1.0) * thr / static_cast<double>(R.opt.lsq_iters - 1);
    if (R.opt.lsq_iters <= 0) continue;
    score_then_fit(R, false, fit_thr, m, nullptr, nullptr);
    for (int i = 0; i < R.opt.lsq_iters; ++i) {
      fit_thr -= step;
      if (i + 1 < R.opt.lsq_iters) {
        score_then_fit(R, true, fit_thr, m, score_best_minimal, best_minimal_model);
      } else {
        double sc = kNever;
        timed_pass(R, m, thr, true, thr, nullptr, &sc);
        update_best(sc, m, score_best_minimal, best_minimal_model);
      }
    }
  }
}

// LocallyOptimizedMSAC::EstimateModel (ransac.h:128-275) as driven by
// ransaclib_localization (examples/ransaclib_localization.cc:58-165), for one query.
__device__ void run_query(Run& R, double* best_model, rlb_localize_result* res) {
  const int lane = R.lane;
  const rlb_localize_options& o = R.opt;
  const int n = R.q.n;
  const double thr = o.sq_thr;

  uint32_t num_iterations = 0u;
  int best_num_inliers = 0;
  double best_model_score = kNever;
  double inlier_ratio = 0.0;
  int num_lo = 0;

  uint32_t max_iters = max(o.max_iters, o.min_iters);
  double* best_minimal_model = R.sh->models[1];
  if (lane < 12) best_minimal_model[lane] = 0.0;
  __syncwarp();
  double best_minimal_score = kNever;

  // this lane's hypothesis of the current speculative batch
  uint32_t batch_first = 0u, batch_count = 0u;
  double my_model[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) my_model[i] = 0.0;
  bool my_valid = false;
  double my_score = kNever;

  auto refresh_termination = [&]() {  // ransac.h:170-177, 231-238
    best_num_inliers = timed_pass(R, best_model, thr, false, thr, R.list_best, nullptr);
    inlier_ratio = static_cast<double>(best_num_inliers) / static_cast<double>(n);
    max_iters = num_required_iterations(inlier_ratio, 1.0 - o.success_prob, 4, o.min_iters, o.max_iters);
  };

  for (num_iterations = 0u; num_iterations < max_iters; ++num_iterations) {
    const uint32_t it = num_iterations;
    if (it == o.lo_start && best_minimal_score < kNever) {  // ransac.h:163-178
      ++num_lo;
      local_optimization(R, best_model, &best_model_score);
      refresh_termination();
    }

    if (it >= batch_first + batch_count) {
      // Speculate on the next (up to) 32 iterations: samples are a pure function of the
      // sampler stream, hypotheses do not depend on the LO stream.
      const uint32_t bound = max(max_iters, it + 1u);
      RLB_PROF_BEGIN(R);
      batch_first = it;
      batch_count = min(32u, bound - it);
      int mine[4] = {0, 0, 0, 0};
      for (uint32_t k = 0; k < batch_count; ++k) {
        int s[4];
        draw_sample4(R.sampler, n, s, lane);
        if (static_cast<uint32_t>(lane) == k) {
#pragma unroll
          for (int j = 0; j < 4; ++j) mine[j] = s[j];
        }
      }
      my_valid = false;
      my_score = kNever;
      bool my_finite = false;
      if (static_cast<uint32_t>(lane) < batch_count) {
        my_valid = minimal_solver(R.q, mine, thr, my_model);
        if (my_valid) {
          my_finite = true;
#pragma unroll
          for (int i = 0; i < 12; ++i) my_finite = my_finite && (my_model[i] - my_model[i] == 0.0);
        }
      }
      // a model with a non-finite entry scores NaN and never wins (finalize_kernel)
      const long long ts = clock64();
      if (R.cta != nullptr) {
        CtaComm* C = R.cta;
        float cf[12];
        sweep_coefficients(R.q, my_model, cf);
#pragma unroll
        for (int i = 0; i < 12; ++i) C->coef[i][lane] = (my_valid && my_finite) ? cf[i] : 0.0f;
        if (lane == 0) {
          C->thr64 = thr;
          C->cmd = kCmdSweep;
        }
        __syncthreads();
        cta_sweep(C, 0, lane);
        __syncthreads();
        const double tot = ((C->part_sweep[0][lane] + C->part_sweep[1][lane]) + C->part_sweep[2][lane]) +
                           C->part_sweep[3][lane];
        if (my_valid) my_score = my_finite ? tot : __longlong_as_double(0x7ff8000000000000LL);
      } else if (my_valid) {
        my_score = my_finite ? sweep32(R.q, my_model, thr) : __longlong_as_double(0x7ff8000000000000LL);
      }
      R.prof[1] += clock64() - ts;
      __syncwarp();
      RLB_PROF_END(R, 0);
    }

    const int slot = static_cast<int>(it - batch_first);
    if (!__shfl_sync(kFull, my_valid ? 1 : 0, slot)) continue;  // ransac.h:185

    double local_score = kNever;
    const double s32 = __shfl_sync(kFull, my_score, slot);
    const bool have = s32 < kNever;  // GetBestEstimatedModelId: NaN never becomes the local best
    if (have) local_score = s32;
    const bool forced = (it == o.lo_start);
    double* cand = R.sh->models[2];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      const double v = __shfl_sync(kFull, my_model[i], slot);
      cand[i] = v;
    }
    __syncwarp();
    if (have && (forced || local_score <= best_minimal_score * (1.0 + 4e-4) ||
                 best_minimal_score == kNever)) {
      // the f32 sweep says "maybe better": settle it with the exact f64 score
      timed_pass(R, cand, thr, true, thr, nullptr, &local_score);
    }

    if (local_score < best_minimal_score || forced) {  // ransac.h:195-196
      const bool new_best = local_score < best_minimal_score;
      if (new_best) {
        best_minimal_score = local_score;
#pragma unroll
        for (int i = 0; i < 12; ++i) best_minimal_model[i] = cand[i];
        __syncwarp();
        update_best(best_minimal_score, best_minimal_model, &best_model_score, best_model);
      }
      const bool run_lo = (it >= o.lo_start && best_minimal_score < kNever);
      if (!new_best && !run_lo) continue;
      if (run_lo) {
        ++num_lo;
        double score = best_minimal_score;
        local_optimization(R, best_minimal_model, &score);
        update_best(score, best_minimal_model, &best_model_score, best_model);
      }
      refresh_termination();
    }
  }

  // LO for runs that ended before the start iteration (ransac.h:245-255).
  if (num_iterations <= o.lo_start && best_model_score < kNever) {
    ++num_lo;
    local_optimization(R, best_model, &best_model_score);
    best_num_inliers = timed_pass(R, best_model, thr, false, thr, R.list_best, nullptr);
    inlier_ratio = static_cast<double>(best_num_inliers) / static_cast<double>(n);
  }

  if (o.final_lsq) {  // ransac.h:257-272
    double* refined = R.sh->models[3];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 12; ++i) refined[i] = best_model[i];
    __syncwarp();
    least_squares(R, R.list_best, best_num_inliers, refined);
    double score = kNever;
    const int n_inl = timed_pass(R, refined, thr, true, thr, R.list_fit, &score);
    if (score < best_model_score) {
      best_model_score = score;
#pragma unroll
      for (int i = 0; i < 12; ++i) best_model[i] = refined[i];
      __syncwarp();
      best_num_inliers = n_inl;
      for (int i = lane; i < n_inl; i += 32) R.list_best[i] = R.list_fit[i];
      inlier_ratio = static_cast<double>(n_inl) / static_cast<double>(n);
    }
  }
  __syncwarp();
  if (lane == 0) {
    res->success = 1;
    res->num_inliers = best_num_inliers;
    res->num_iterations = num_iterations;
    res->num_lo = num_lo;
    res->score = best_model_score;
    res->inlier_ratio = inlier_ratio;
    res->passes = R.passes;
    res->lsq_calls = R.lsq_calls;
  }
}

// Writes the "nothing to estimate" record of a query with <= 3 matches (ransaclib_localization.cc:86-92).
__device__ __forceinline__ void write_empty_result(rlb_localize_result* res, double* pose_out, int lane) {
  if (lane == 0) {
    res->success = 0;
    res->num_inliers = 0;
    res->num_iterations = 0u;
    res->num_lo = 0;
    res->score = kNever;
    res->inlier_ratio = 0.0;
    res->passes = 0;
    res->lsq_calls = 0;
  }
  if (lane < 12) pose_out[lane] = 0.0;
}

// Seeds the two engines, runs the query and writes pose + debug counters (R.q and the lists are set).
__device__ __forceinline__ void seed_and_run(Run& R, const LocalizeArgs& A, int qi, long long o0, long long t_query) {
  const int lane = R.lane;
  const int n = R.q.n;
  // Sampler and local optimisation own separate engines seeded alike (ransac.h:142-144,
  // sampling.h:49-55).
  mt_seed(R.sh->mt_sampler, A.seeds[qi], lane);
  for (int i = lane; i < 624; i += 32) R.sh->mt_lo[i] = R.sh->mt_sampler[i];
  __syncwarp();
  R.sampler.s = R.sh->mt_sampler;
  R.sampler.idx = 624;
  R.lo_rng.s = R.sh->mt_lo;
  R.lo_rng.idx = 624;

  double* best_model = R.sh->models[0];
  if (lane < 12) best_model[lane] = 0.0;
  __syncwarp();
  R.prof[9] = clock64() - t_query;
  unsigned long long ns0 = 0, ns1 = 0;
  unsigned smid = 0;
  if (A.opt.debug_flags & 2) {
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns0));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  }
  run_query(R, best_model, A.results + qi);
  __syncwarp();
  R.prof[7] = clock64() - t_query;
  if ((A.opt.debug_flags & 2) && A.inliers != nullptr && n >= 64 && lane == 0) {
    // debug: lane 0's cycle counters replace the tail of this query's inlier region, followed by the
    // query's begin / end on the global nanosecond timer, the SM and the CTA it ran on
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    long long* out = reinterpret_cast<long long*>(A.inliers + ((o0 + n - 32) & ~1LL));
    for (int i = 0; i < 12; ++i) out[i] = R.prof[i];
    out[12] = static_cast<long long>(ns0);
    out[13] = static_cast<long long>(ns1);
    out[14] = static_cast<long long>(smid);
    out[15] = static_cast<long long>(blockIdx.x);
  }
  double* pose_out = A.poses + 12 * static_cast<size_t>(qi);
  if (lane < 12) pose_out[lane] = best_model[lane];
  __syncwarp();
}

}  // namespace

#ifndef RLB_LOC_MINB
#define RLB_LOC_MINB 3
#endif
// One query per warp.
__global__ void __launch_bounds__(kWarpsPerCta * 32, RLB_LOC_MINB) localize_kernel(const LocalizeArgs A) {
  __shared__ WarpShared shared[kWarpsPerCta];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * kWarpsPerCta + warp;

  Run R;
  R.opt = A.opt;
  R.sh = &shared[warp];
  R.cta = nullptr;
  R.lane = lane;
  R.force_slow = (A.opt.debug_flags & 1) != 0;
  R.q.soa32 = A.soa32 + static_cast<size_t>(slot) * 5 * A.npad_max;
  int32_t* lists = A.lists + static_cast<size_t>(slot) * 3 * A.nmax;
  R.list_base = lists;
  R.list_fit = lists + A.nmax;

  for (;;) {
    int qi = 0;
    if (lane == 0) qi = A.q_first + static_cast<int>(atomicAdd(A.ticket, 1u));
    qi = __shfl_sync(kFull, qi, 0);
    if (qi >= A.q_end) break;

    const long long o0 = A.offsets[qi], o1 = A.offsets[qi + 1];
    const int n = static_cast<int>(o1 - o0);
    if (n <= 3) {
      write_empty_result(A.results + qi, A.poses + 12 * static_cast<size_t>(qi), lane);
      continue;
    }
    R.q.n = n;
    R.q.p2 = A.p2d + 2 * static_cast<size_t>(o0);
    R.q.p3 = A.p3d + 3 * static_cast<size_t>(o0);
    R.q.fx = A.fx[qi];
    R.q.fy = A.fy[qi];
    R.q.npad = (n + 3) & ~3;
    R.list_best = (A.inliers != nullptr) ? A.inliers + o0 : lists + 2 * A.nmax;
    R.passes = 0;
    R.lsq_calls = 0;
    for (int i = 0; i < 12; ++i) R.prof[i] = 0;
    const long long t_query = clock64();
    prepare_soa32(R.q, lane);
    seed_and_run(R, A, qi, o0, t_query);
  }
}

#ifndef RLB_LOC_CTA_MINB
#define RLB_LOC_CTA_MINB 4
#endif
#if RLB_LOC_CTA_SMEM
constexpr int kCtaSmemPoints = 2048;  // f32 copy of the matches in shared memory: 5 x 2048 x 4 B = 40 KB
#else
constexpr int kCtaSmemPoints = 0;     // the f32 copy stays in global memory (L2): shared memory left to L1
#endif

// One query per CTA: master warp = control flow, all four warps = the passes over the matches.
__global__ void __launch_bounds__(kWarpsPerCta * 32, RLB_LOC_CTA_MINB) localize_cta_kernel(const LocalizeArgs A) {
  __shared__ WarpShared shared;
  __shared__ CtaComm comm;
  extern __shared__ float4 dyn_soa4[];
  float* dyn_soa = reinterpret_cast<float*>(dyn_soa4);
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  // The four master warps of an SM's four CTAs should sit on four different schedulers: CTAs of one
  // wave land on different SMs, so the master rotates with the wave index.
  const int master = (blockIdx.x / A.wave) & (kWarpsPerCta - 1);
  const int lw = (warp - master) & (kWarpsPerCta - 1);
  CtaComm* C = &comm;
  int32_t* lists = A.lists + static_cast<size_t>(blockIdx.x) * 3 * A.nmax;
  float* soa_global = (A.soa32 != nullptr) ? A.soa32 + static_cast<size_t>(blockIdx.x) * 5 * A.npad_max : nullptr;

  for (;;) {
    __syncthreads();  // everybody is done with the previous query's mailbox
    if (lw == 0 && lane == 0) C->qi = A.q_first + static_cast<int>(atomicAdd(A.ticket, 1u));
    __syncthreads();
    const int qi = C->qi;
    if (qi >= A.q_end) break;
    const long long o0 = A.offsets[qi], o1 = A.offsets[qi + 1];
    const int n = static_cast<int>(o1 - o0);
    if (n <= 3) {
      if (lw == 0) write_empty_result(A.results + qi, A.poses + 12 * static_cast<size_t>(qi), lane);
      continue;
    }
    const long long t_query = clock64();
    if (lw == 0 && lane == 0) {
      C->n = n;
      C->npad = (n + 3) & ~3;
      C->p2 = A.p2d + 2 * static_cast<size_t>(o0);
      C->p3 = A.p3d + 3 * static_cast<size_t>(o0);
      C->fx = A.fx[qi];
      C->fy = A.fy[qi];
      C->soa32 = (n <= A.smem_points) ? dyn_soa : soa_global;
      C->slab = shared.slab;
    }
    __syncthreads();
    cta_prepare(C, lw, lane);
    __syncthreads();

    if (lw != 0) {
      // worker: share the passes the master asks for until it reports the query done
      for (;;) {
        __syncthreads();
        const int cmd = C->cmd;
        if (cmd == kCmdDone) break;
        if (cmd == kCmdFull) {
          (void)cta_full(C, lw, lane, 0, 0);
          continue;  // no join: the next stop is the command barrier
        }
        if (cmd == kCmdPass) {
          (void)cta_pass(C, lw, lane);
        } else {
          cta_sweep(C, lw, lane);
        }
        __syncthreads();
      }
      continue;
    }

    Run R;
    R.opt = A.opt;
    R.sh = &shared;
    R.cta = C;
    R.lane = lane;
    R.force_slow = (A.opt.debug_flags & 1) != 0;
    R.list_base = lists;
    R.list_fit = lists + A.nmax;
    R.q.n = n;
    R.q.p2 = C->p2;
    R.q.p3 = C->p3;
    R.q.fx = C->fx;
    R.q.fy = C->fy;
    R.q.npad = C->npad;
    R.q.soa32 = C->soa32;
    for (int k = 0; k < 3; ++k) R.q.centre[k] = C->centre[k];
    R.list_best = (A.inliers != nullptr) ? A.inliers + o0 : lists + 2 * A.nmax;
    R.passes = 0;
    R.lsq_calls = 0;
    for (int i = 0; i < 12; ++i) R.prof[i] = 0;
    seed_and_run(R, A, qi, o0, t_query);
    if (lane == 0) C->cmd = kCmdDone;
    __syncthreads();  // the workers' command barrier: they leave their loop
  }
}

cudaError_t launch_localize(const LocalizeArgs& a, int mode, int ctas, cudaStream_t st, int* launches) {
  ++*launches;
  if (mode == 0) {
    localize_kernel<<<ctas, kWarpsPerCta * 32, 0, st>>>(a);
    return cudaGetLastError();
  }
  const int dyn = 5 * kCtaSmemPoints * static_cast<int>(sizeof(float));
  static PerDeviceOnce configured;
  const int dev = PerDeviceOnce::device();
  if (!configured.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(localize_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
    if (e != cudaSuccess) return e;
    configured.set(dev);
  }
  localize_cta_kernel<<<ctas, kWarpsPerCta * 32, dyn, st>>>(a);
  return cudaGetLastError();
}

int localize_warps_per_cta() { return kWarpsPerCta; }
int localize_ctas_per_sm(int mode) { return mode == 0 ? RLB_LOC_MINB : RLB_LOC_CTA_MINB; }
int localize_smem_points() { return kCtaSmemPoints; }

}  // namespace rlb
