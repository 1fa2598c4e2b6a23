// K x N pose MSAC sweep with the 3x4 projection on the 5th-generation tensor cores.
//
// Why: ncu on the FFMA2 sweep (profiles/history/r01f_ncu_sweep_h8.csv) shows the FMA pipe 78 % busy with
// math-pipe throttle as the top stall, and 9 of the 14 FMA-pipe operations of an evaluation are the
// projection p = R' X' + t' (BASELINE north_star: "tensor cores only if ncu shows the 3x4 projection
// contraction to be FMA-issue-bound"). Here the projection of 128 hypotheses x 32 points is six
// tcgen05.mma (kind::tf32, M = 128, N = 32, K = 8, A from TMEM) into TMEM, and the CUDA cores keep
// the part that is not a contraction: reciprocal, residual, MSAC test and the per-hypothesis sums.
// Replaces, for f32 pose sweeps, the N x K loop of LocallyOptimizedMSAC::ScoreModel (ransac.h:295-303)
// over EvaluateModelOnPoint (calibrated_absolute_pose_estimator.cc:154-167).
//
// Precision: TF32 keeps 11 significant bits, so both operands are split, x = hi + lo with hi = x
// with the low 13 mantissa bits cleared (exactly what the tensor core reads) and lo = x - hi (exact
// in f32). K = 16 carries all four partial products,
//     A row (hypothesis): [ rh0 rh1 rh2 th | rl0 rl1 rl2 tl ]   (the same 8 TMEM columns for both K steps)
//     B col (point)     : [ Xh  Yh  Zh  1  | Xh  Yh  Zh  1  | Xl  Yl  Zl  0 | Xl  Yl  Zl  0 ]
// accumulated in f32 by the tensor core: the result differs from the f32 FMA chain by the second
// truncation of the lo parts (<= 2^-22 relative per term), i.e. it is an f32-class projection.
//
// Operand layout: UMMA K-major, SWIZZLE_NONE canonical layout (core matrix = 8 rows x 16 bytes,
// rows 16 B apart; SBO = byte distance between core matrices along N, LBO = along K). The points
// are pre-packed in global memory in exactly the shared-memory image of a tile (32 points: 4 K-chunks
// x 4 cores x 128 B = 2 KB, followed by 32 x (-u) and 32 x (-v)), so one 1-D bulk TMA copy per tile
// brings everything in. The hypotheses of a CTA are written to TMEM once (tcgen05.st): the A operand
// is re-read by every MMA, and the TS form runs at the N/2-cycle floor where the SS form streams 4 KB
// of A from shared memory per MMA (tools/probes/umma_probe.cu). What was measured on the way, and what
// bounds the kernel now, is in profiles/r01r_tc_experiments.md and DESIGN.md section 3.
#include <cfloat>

#include "rlb_internal.cuh"

namespace rlb {

namespace {

constexpr int kTcThreads = 128;            // 4 warps = the 128 TMEM lanes = 128 hypotheses
constexpr int kTcTile = 32;                // points per MMA tile (N)
constexpr int kTcTileFloats = 512 + 64;    // 16 x 32 operand floats + 32 x (-u) + 32 x (-v)
constexpr int kTcTileBytes = kTcTileFloats * 4;
constexpr int kTcRing = 8;                 // TMA ring depth (tiles)
constexpr int kTcCols = 128;               // TMEM columns: A operand [0,24), px [32,64) py [64,96) pz [96,128)
constexpr int kTcFlush = 16;               // tiles (512 evaluations per thread) per f32 -> f64 flush

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Bounded wait: a protocol error must fault, not hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  uint32_t done = 0;
  for (long long spins = 0; !done; ++spins) {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(done)
        : "r"(a), "r"(parity)
        : "memory");
    if (spins > 200000000LL) __trap();
  }
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_NONE, sm_100 (version 1).
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3ffffu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;  // descriptor version for Blackwell
  return d;         // base offset 0, LBO mode 0, layout type SWIZZLE_NONE
}

// Instruction descriptor: D = f32, A = B = tf32, both K-major, M = 128, N = 32, dense.
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((kTcTile >> 3) << 17) | ((128u >> 4) << 24);

// A operand from TMEM (lane = hypothesis, 8 consecutive 32-bit columns = K), B from shared memory.
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(kIdesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
               "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
               "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
               "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
               : "memory");
}
// One lane of a converged warp. tcgen05.mma / commit issued under elect.sync compile to a single UTCHMMA;
// under `if (tid == 0)` ptxas wraps every MMA in a per-lane serialisation loop (~109 cycles per MMA
// measured with tools/probes/umma_probe.cu against ~40 for the elected form).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
#ifdef RLB_TC_PROBE_NOLDTM  // timing probe only (wrong results): is the TMEM read path the bound?
#pragma unroll
  for (int i = 0; i < 8; ++i) asm volatile("mov.b32 %0, %1;" : "=f"(v[i]) : "r"(taddr + i));
  return;
#endif
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
// tcgen05.wait::ld with the loaded registers as in/out operands: nothing that reads them can be
// scheduled above the wait.
__device__ __forceinline__ void tmem_wait_ld8x3(float* a, float* b, float* c) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7]),
                 "+f"(b[0]), "+f"(b[1]), "+f"(b[2]), "+f"(b[3]), "+f"(b[4]), "+f"(b[5]), "+f"(b[6]), "+f"(b[7]),
                 "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]), "+f"(c[4]), "+f"(c[5]), "+f"(c[6]), "+f"(c[7])
               :
               : "memory");
}

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) {
  u64 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) {
  u64 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float rsq_approx(float x) {
#ifdef RLB_TC_PROBE_NOMUFU  // timing probe only (wrong results): is the XU pipe the bound?
  return x;
#else
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}
// MSAC accumulation of one evaluation: s += e and c += 1 when e < thr (one FSETP + two predicated adds).
__device__ __forceinline__ void msac_add(float e, float thr, float& s, int& c) {
  asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %2, %3;\n\t@p add.f32 %0, %0, %2;\n\t@p add.s32 %1, %1, 1;\n\t}"
      : "+f"(s), "+r"(c)
      : "f"(e), "f"(thr));
}

__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

}  // namespace

// ---------------------------------------------------------------- point pre-pack
// tiles: [npad / 32][kTcTileFloats]; inputs are the recentred f32 stage values (the same numbers the
// FFMA2 sweep reads), padding = X 0, -u/-v NaN.
//
// bounds (two floats behind the tiles, zeroed before the launch): [0] = max |X'| (Euclidean norm of a recentred
// point), [1] = max(|u|, |v|) over the live points. The filtered f64 sweep derives its per-hypothesis error
// bound of the f32 classification from them.
__global__ void tc_pack_kernel(const DeviceData d, float* __restrict__ tiles, unsigned int* __restrict__ bounds) {
  const int total = d.npad;
  float rmax = 0.0f, umax = 0.0f;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    const float x = d.soa32[0 * static_cast<size_t>(total) + g];
    const float y = d.soa32[1 * static_cast<size_t>(total) + g];
    const float z = d.soa32[2 * static_cast<size_t>(total) + g];
    const float nu = d.soa32[3 * static_cast<size_t>(total) + g];
    const float nv = d.soa32[4 * static_cast<size_t>(total) + g];
    if (g < d.n[0]) {
      rmax = fmaxf(rmax, sqrtf(x * x + y * y + z * z));
      umax = fmaxf(umax, fmaxf(fabsf(nu), fabsf(nv)));
    }
    const float xh = tf32_hi(x), yh = tf32_hi(y), zh = tf32_hi(z);
    const float xl = x - xh, yl = y - yh, zl = z - zh;
    float* t = tiles + static_cast<size_t>(g / kTcTile) * kTcTileFloats;
    const int n = g % kTcTile;
    const int core = n >> 3, row = n & 7;
    const float ch[4][4] = {{xh, yh, zh, 1.0f}, {xh, yh, zh, 1.0f}, {xl, yl, zl, 0.0f}, {xl, yl, zl, 0.0f}};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float4* dst = reinterpret_cast<float4*>(t + ((c * 4 + core) * 8 + row) * 4);
      *dst = make_float4(ch[c][0], ch[c][1], ch[c][2], ch[c][3]);
    }
    t[512 + n] = nu;
    t[512 + 32 + n] = nv;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    rmax = fmaxf(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
    umax = fmaxf(umax, __shfl_xor_sync(0xffffffffu, umax, o));
  }
  if ((threadIdx.x & 31) == 0) {  // non-negative floats order like their bit patterns; NaN / inf stay what they are
    atomicMax(bounds + 0, __float_as_uint(rmax));
    atomicMax(bounds + 1, __float_as_uint(umax));
  }
}

static float* tc_bounds_of(float* tiles, int npad) {
  return tiles + static_cast<size_t>(npad / kTcTile) * kTcTileFloats;
}

cudaError_t launch_tc_pack(const DeviceData& d, float* tiles, cudaStream_t st, int* launches) {
  ++*launches;
  const int blocks = max(1, min(2048, (d.npad + 255) / 256));
  unsigned int* bounds = reinterpret_cast<unsigned int*>(tc_bounds_of(tiles, d.npad));
  cudaError_t e = cudaMemsetAsync(bounds, 0, 256, st);
  if (e != cudaSuccess) return e;
  tc_pack_kernel<<<blocks, 256, 0, st>>>(d, tiles, bounds);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ sweep kernel
struct TcSmem {
  float ring[kTcRing][kTcTileFloats];     // point tiles
  uint64_t full[kTcRing];                 // TMA -> MMA
  uint64_t mma_done;                      // MMA -> epilogue
  uint32_t tmem_base;
};

// COUNT = false (the caller did not ask for per-hypothesis inlier counts, as ScoreModel in ransac.h
// never does): the score is accumulated as sum min(e, thr), two instructions fewer per evaluation.
//
// HALVES = 2: 256 threads. A warp can only read the TMEM lanes of its own quarter (warp % 4), so warps w and
// w + 4 serve the SAME 32 hypotheses and split the tile's points (columns 0-15 / 16-31): every hypothesis'
// sums live in two threads and are joined once at the end. TMEM allows four 128-column CTAs per SM, registers
// three of these: 24 instead of 16 warps per SM. Measured (profiles/r02_tc_halves_ab.txt): SLOWER, 33.0 against
// 27.1 ms on C5 - both forms settle at ~60 % of the issue slots, and three CTAs keep only three tiles' MMA round
// trips in flight per SM instead of four. Kept as an A/B switch (-DRLB_TC_HALVES=2); the default is 1.
struct TcJoin {
  double s[128];
  int c[128];
};

template <bool COUNT, int HALVES>
__global__ void __launch_bounds__(128 * HALVES, HALVES == 1 ? 4 : 3)
    sweep_pose_tc_kernel(const DeviceData d, const SweepArgs a, const float* __restrict__ tiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcSmem& S = *reinterpret_cast<TcSmem*>(smem_raw);
  constexpr int kPts = kTcTile / HALVES;  // points of a tile this thread evaluates
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int quarter = warp & 3;            // TMEM lane quarter = hypotheses 32 quarter ... of the CTA
  const int half = warp >> 2;              // which points of the tile (HALVES = 1: always 0)
  const int hyp = quarter * 32 + (tid & 31);
  const int k = blockIdx.x * 128 + hyp;
  const int split = blockIdx.y;
  const int n_live = d.n[0];

  const int total_tiles = d.npad / kTcTile;
  const int t_begin = split * a.stages_per_split;  // here: tiles per split
  const int t_end = min(total_tiles, t_begin + a.stages_per_split);
  const int ntiles = max(0, t_end - t_begin);

  // ---- one-time setup -----------------------------------------------------------------
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)),
                 "r"(static_cast<uint32_t>(kTcCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 32) {
#pragma unroll
    for (int s = 0; s < kTcRing; ++s) mbar_init(&S.full[s], 1);
    mbar_init(&S.mma_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // TMEM base address must be visible before anyone uses it
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem_base;
  const uint32_t lane_addr = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
  if (half == 0) {
    // this thread's hypothesis -> its TMEM lane of each operand matrix (PoseF<float>::load arithmetic).
    // Columns [8 mat, 8 mat + 8) = [rh0 rh1 rh2 th | rl0 rl1 rl2 tl]; the same eight columns serve both
    // K steps: the second multiplies the points' lo parts, whose fourth component is 0.
    float r[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float t[3] = {0.f, 0.f, 1.f};
    if (k < a.k) {
      const double* m = a.models + static_cast<size_t>(k) * 12;
      const double dc[3] = {d.centre[0] - m[9], d.centre[1] - m[10], d.centre[2] - m[11]};
      const double sc[3] = {d.fx, d.fy, 1.0};
#pragma unroll
      for (int row = 0; row < 3; ++row) {
        const double r0 = m[row], r1 = m[row + 3], r2 = m[row + 6];
        r[3 * row + 0] = static_cast<float>(r0 * sc[row]);
        r[3 * row + 1] = static_cast<float>(r1 * sc[row]);
        r[3 * row + 2] = static_cast<float>(r2 * sc[row]);
        t[row] = static_cast<float>((r0 * dc[0] + r1 * dc[1] + r2 * dc[2]) * sc[row]);
      }
    }
#pragma unroll
    for (int mat = 0; mat < 3; ++mat) {
      const float v[4] = {r[3 * mat], r[3 * mat + 1], r[3 * mat + 2], t[mat]};
      float hl[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        hl[j] = tf32_hi(v[j]);
        hl[4 + j] = v[j] - hl[j];
      }
      tmem_st8(lane_addr + 8 * mat, hl);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");


  auto issue_tma = [&](int it) {
    const int slot = it % kTcRing;
    mbar_expect_tx(&S.full[slot], kTcTileBytes);
    tma_load_1d(S.ring[slot], tiles + static_cast<size_t>(t_begin + it) * kTcTileFloats, kTcTileBytes,
                &S.full[slot]);
  };
  if (tid == 0) {
    for (int it = 0; it < kTcRing && it < ntiles; ++it) issue_tma(it);
  }

  const double thr64 = a.thr[0];
  const float thr = static_cast<float>(thr64);
  const double w = a.w[0];
  double dacc = 0.0;
  int dcnt = 0;
  float s = 0.0f;
  int c = 0;
  u64 s2 = 0ull;  // COUNT = false: two running sums of min(e, thr)
  int live_in_chunk = 0;
  const int col0 = half * kPts;  // first TMEM column / tile point of this thread's share

  for (int it = 0; it < ntiles; ++it) {
    const int slot = it % kTcRing;
    if (warp == 0) {
      mbar_wait(&S.full[slot], (it / kTcRing) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (elect_one()) {
        const uint32_t b_addr = smem_u32(S.ring[slot]);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          // one MMA covers K = 8 = two 16-byte chunks of the points (hi parts, then lo parts), LBO apart
          const uint64_t db = umma_desc(b_addr + kk * 2 * 512, 512, 128);
#pragma unroll
          for (int mat = 0; mat < 3; ++mat) umma_tf32_ts(tmem + 32 + mat * kTcTile, tmem + 8 * mat, db, kk);
        }
        umma_commit(&S.mma_done);
      }
      __syncwarp();
    }
    mbar_wait(&S.mma_done, it & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    const float* nu = S.ring[slot] + 512 + col0;
    const float* nv = nu + 32;
    // Groups of 8 points: the TMEM loads of the next group are in flight while this one is
    // reduced (tcgen05.ld latency is ~110 cycles; a tile-wide load-then-wait exposed it twice per tile).
    constexpr int kGroups = kPts / 8;
    float px[2][8], py[2][8], pz[2][8];
    tmem_ld8(lane_addr + 32 + 0 * kTcTile + col0, px[0]);
    tmem_ld8(lane_addr + 32 + 1 * kTcTile + col0, py[0]);
    tmem_ld8(lane_addr + 32 + 2 * kTcTile + col0, pz[0]);
    tmem_wait_ld8x3(px[0], py[0], pz[0]);
#pragma unroll
    for (int qd = 0; qd < kGroups; ++qd) {
      const int cur = qd & 1, nxt = cur ^ 1;
      if (qd < kGroups - 1) {
        tmem_ld8(lane_addr + 32 + 0 * kTcTile + col0 + 8 * (qd + 1), px[nxt]);
        tmem_ld8(lane_addr + 32 + 1 * kTcTile + col0 + 8 * (qd + 1), py[nxt]);
        tmem_ld8(lane_addr + 32 + 2 * kTcTile + col0 + 8 * (qd + 1), pz[nxt]);
      }
      // Two points per instruction (packed f32x2): TMEM columns j, j+1 land in adjacent registers.
      // 1 / pz = rsqrt(pz)^2: one MUFU, NaN for pz < 0 and +inf for pz == 0, so points behind the
      // camera fail (e < thr) without a clamp; 2 ulp of rsqrt squared stay inside the f32 class.
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const ulonglong2 u4 = *reinterpret_cast<const ulonglong2*>(nu + 8 * qd + 4 * q);
        const ulonglong2 v4 = *reinterpret_cast<const ulonglong2*>(nv + 8 * qd + 4 * q);
#pragma unroll
        for (int pr = 0; pr < 2; ++pr) {
          const int j = 4 * q + 2 * pr;
          const u64 rs = pack2(rsq_approx(pz[cur][j]), rsq_approx(pz[cur][j + 1]));
          const u64 inv = fmul2(rs, rs);
          const u64 du = ffma2(pack2(px[cur][j], px[cur][j + 1]), inv, pr ? u4.y : u4.x);
          const u64 dv = ffma2(pack2(py[cur][j], py[cur][j + 1]), inv, pr ? v4.y : v4.x);
          const u64 e2 = ffma2(du, du, fmul2(dv, dv));
          float ea, eb;
          unpack2(e2, ea, eb);
          if (COUNT) {
            msac_add(ea, thr, s, c);
            msac_add(eb, thr, s, c);
          } else {
            s2 = fadd2(s2, pack2(fminf(ea, thr), fminf(eb, thr)));  // NaN / inf -> thr
          }
        }
      }
      if (qd < kGroups - 1) tmem_wait_ld8x3(px[nxt], py[nxt], pz[nxt]);
    }
    live_in_chunk += max(0, min(kPts, n_live - ((t_begin + it) * kTcTile + col0)));
    if (((it + 1) % kTcFlush) == 0 || it + 1 == ntiles) {
      if (COUNT) {
        dacc += w * (static_cast<double>(s) + thr64 * static_cast<double>(live_in_chunk - c));
        dcnt += c;
      } else {
        // every processed point added min(e, thr); padding (NaN residual) added thr and is taken out again
        float sa, sb;
        unpack2(s2, sa, sb);
        const int processed = (it % kTcFlush + 1) * kPts;
        dacc += w * (static_cast<double>(sa) + static_cast<double>(sb) -
                     static_cast<double>(thr) * static_cast<double>(processed - live_in_chunk));
        s2 = 0ull;
      }
      s = 0.0f;
      c = 0;
      live_in_chunk = 0;
    }
    // everyone is done with TMEM and with this ring slot
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0 && it + kTcRing < ntiles) {
      if (elect_one()) issue_tma(it + kTcRing);
      __syncwarp();
    }
  }

  if (HALVES == 2) {
    // join the two halves of every hypothesis (fixed order: first half + second half)
    TcJoin& J = *reinterpret_cast<TcJoin*>(S.ring);  // the ring is idle now
    if (half == 1) {
      J.s[hyp] = dacc;
      J.c[hyp] = dcnt;
    }
    __syncthreads();
    if (half == 0) {
      dacc += J.s[hyp];
      dcnt += J.c[hyp];
    }
  }
  if (half == 0 && k < a.k) {
    a.part_score[static_cast<size_t>(split) * a.k + k] = dacc;
    if (COUNT) a.part_count[static_cast<size_t>(split) * a.k + k] = dcnt;
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem),
                 "r"(static_cast<uint32_t>(kTcCols))
                 : "memory");
  }
}

// ------------------------------------------------ the same sweep, MMA and epilogue decoupled
// sweep_pose_tc_kernel alternates per CTA between "one lane issues six MMAs, everyone waits for the commit" and
// "everyone reduces the tile, __syncthreads": its four warps idle through every MMA round trip and the other three
// CTAs of the SM have to fill the gap (ncu: ~65 % of the issue slots). Here a CTA is two groups of four epilogue
// warps + 1 producer warp and owns 256 TMEM columns with TWO accumulator sets:
//     producer warp : TMA ring refills; for tile t waits until set t & 1 has been read out, issues the six MMAs,
//                     tcgen05.commit -> acc_full[t & 1]
//     group g       : tiles g, g + 2, ... from set g. The TMEM loads run two 8-point groups ahead of the arithmetic,
//                     and the set is handed back (acc_empty[g]) when its last columns are in registers - half a
//                     tile of arithmetic before the group needs the next one, which is about the MMA round trip.
// Both groups serve the same 128 hypotheses (a warp can only read the TMEM lanes of warp % 4): every hypothesis'
// sums live in two threads and are joined once at the end. No __syncthreads in the loop. Two CTAs per SM (TMEM),
// 16 epilogue warps per SM as before.
//
// Measured on C5 (profiles/r02_tc_halves_ab.txt): 27.27 ms against 27.08 ms for sweep_pose_tc_kernel (27.49 against
// 25.9 ms without counts) - the MMA round trip is NOT what bounds the sweep. The same file holds the probes that say
// what does: without the MUFU 25.5 ms (the XU pipe is not it either), with eight MOVs in place of every
// tcgen05.ld.x8 30.8 ms (more instructions, slower: issue), 331 warp instructions per 32-point tile at 72 % issue
// utilisation with every pipe below 60 %. A variant that split each tile's 32 points over two warps (16 each, both
// accumulator sets per tile) ran at 32.1 ms: the per-tile fixed instructions then weigh twice. Kept as an A/B
// build (-DRLB_TC_PIPE=1); the default is sweep_pose_tc_kernel.
constexpr int kTcpThreads = 32 * 9;
constexpr int kTcpCols = 256;  // A operand [0,24); accumulator set g: [32 + 96 g, 128 + 96 g)

struct TcpSmem {
  float ring[kTcRing][kTcTileFloats];
  uint64_t full[kTcRing];   // TMA -> producer
  uint64_t acc_full[2];     // MMA commit -> group
  uint64_t acc_empty[2];    // group (4 warps) -> producer
  uint32_t tmem_base;
};

// tcgen05.wait::ld, then hand the accumulator set back: one statement with the loaded registers as in/out
// operands, so that the arrive stays where it is written (before the arithmetic on these registers).
__device__ __forceinline__ void tmem_wait_ld8x3_release(float* a, float* b, float* c, uint64_t* bar, uint32_t lane0) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %25, 0;\n\t"
      "tcgen05.wait::ld.sync.aligned;\n\t"
      "tcgen05.fence::before_thread_sync;\n\t"
      "@p mbarrier.arrive.shared::cta.b64 _, [%24];\n\t}"
      : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7]),
        "+f"(b[0]), "+f"(b[1]), "+f"(b[2]), "+f"(b[3]), "+f"(b[4]), "+f"(b[5]), "+f"(b[6]), "+f"(b[7]),
        "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]), "+f"(c[4]), "+f"(c[5]), "+f"(c[6]), "+f"(c[7])
      : "r"(smem_u32(bar)), "r"(lane0)
      : "memory");
}

template <bool COUNT>
__global__ void __launch_bounds__(kTcpThreads, 2)
    sweep_pose_tcp_kernel(const DeviceData d, const SweepArgs a, const float* __restrict__ tiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcpSmem& S = *reinterpret_cast<TcpSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const bool producer = warp == 8;
  const int quarter = warp & 3;
  const int grp = (warp >> 2) & 1;
  const int hyp = quarter * 32 + (tid & 31);
  const int k = blockIdx.x * 128 + hyp;
  const int split = blockIdx.y;
  const int n_live = d.n[0];

  const int total_tiles = d.npad / kTcTile;
  const int t_begin = split * a.stages_per_split;
  const int t_end = min(total_tiles, t_begin + a.stages_per_split);
  const int ntiles = max(0, t_end - t_begin);

  if (producer) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)),
                 "r"(static_cast<uint32_t>(kTcpCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kTcRing; ++s) mbar_init(&S.full[s], 1);
    mbar_init(&S.acc_full[0], 1);
    mbar_init(&S.acc_full[1], 1);
    mbar_init(&S.acc_empty[0], 4);
    mbar_init(&S.acc_empty[1], 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem_base;
  const uint32_t lane_addr = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
  if (warp < 4) {
    // the hypotheses' TMEM lanes of the A operand, as in sweep_pose_tc_kernel
    float r[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float t[3] = {0.f, 0.f, 1.f};
    if (k < a.k) {
      const double* m = a.models + static_cast<size_t>(k) * 12;
      const double dc[3] = {d.centre[0] - m[9], d.centre[1] - m[10], d.centre[2] - m[11]};
      const double sc[3] = {d.fx, d.fy, 1.0};
#pragma unroll
      for (int row = 0; row < 3; ++row) {
        const double r0 = m[row], r1 = m[row + 3], r2 = m[row + 6];
        r[3 * row + 0] = static_cast<float>(r0 * sc[row]);
        r[3 * row + 1] = static_cast<float>(r1 * sc[row]);
        r[3 * row + 2] = static_cast<float>(r2 * sc[row]);
        t[row] = static_cast<float>((r0 * dc[0] + r1 * dc[1] + r2 * dc[2]) * sc[row]);
      }
    }
#pragma unroll
    for (int mat = 0; mat < 3; ++mat) {
      const float v[4] = {r[3 * mat], r[3 * mat + 1], r[3 * mat + 2], t[mat]};
      float hl[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        hl[j] = tf32_hi(v[j]);
        hl[4 + j] = v[j] - hl[j];
      }
      tmem_st8(lane_addr + 8 * mat, hl);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  double dacc = 0.0;
  int dcnt = 0;

  if (producer) {
    auto issue_tma = [&](int it) {
      const int slot = it % kTcRing;
      mbar_expect_tx(&S.full[slot], kTcTileBytes);
      tma_load_1d(S.ring[slot], tiles + static_cast<size_t>(t_begin + it) * kTcTileFloats, kTcTileBytes,
                  &S.full[slot]);
    };
    if (elect_one()) {
      for (int it = 0; it < kTcRing && it < ntiles; ++it) issue_tma(it);
    }
    __syncwarp();
    for (int it = 0; it < ntiles; ++it) {
      const int slot = it % kTcRing;
      const int buf = it & 1;
      if (it >= 2) {
        // set `buf` has been read out: its group is in the second half of tile it - 2 and has finished tile
        // it - 4, -u / -v included, so THAT tile's ring slot is free
        mbar_wait(&S.acc_empty[buf], ((it >> 1) - 1) & 1);
        if (it >= 4 && it + 4 < ntiles) {
          if (elect_one()) issue_tma(it + 4);
          __syncwarp();
        }
      }
      mbar_wait(&S.full[slot], (it / kTcRing) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (elect_one()) {
        const uint32_t b_addr = smem_u32(S.ring[slot]);
        const uint32_t acc = tmem + 32 + 96 * buf;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint64_t db = umma_desc(b_addr + kk * 2 * 512, 512, 128);
#pragma unroll
          for (int mat = 0; mat < 3; ++mat) umma_tf32_ts(acc + mat * kTcTile, tmem + 8 * mat, db, kk);
        }
        umma_commit(&S.acc_full[buf]);
      }
      __syncwarp();
    }
  } else {
    const double thr64 = a.thr[0];
    const float thr = static_cast<float>(thr64);
    const double w = a.w[0];
    const uint32_t lane0 = (tid & 31) == 0 ? 1u : 0u;
    const uint32_t acc = lane_addr + 32 + 96 * grp;
    float s = 0.0f;
    int c = 0;
    u64 s2 = 0ull;
    int live_in_chunk = 0;
    int visit = 0;
    for (int it = grp; it < ntiles; it += 2, ++visit) {
      const int slot = it % kTcRing;
      mbar_wait(&S.acc_full[grp], visit & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const float* nu = S.ring[slot] + 512;
      const float* nv = nu + 32;
      float px[3][8], py[3][8], pz[3][8];
      tmem_ld8(acc + 0 * kTcTile, px[0]);
      tmem_ld8(acc + 1 * kTcTile, py[0]);
      tmem_ld8(acc + 2 * kTcTile, pz[0]);
      tmem_ld8(acc + 0 * kTcTile + 8, px[1]);
      tmem_ld8(acc + 1 * kTcTile + 8, py[1]);
      tmem_ld8(acc + 2 * kTcTile + 8, pz[1]);
      tmem_wait_ld8x3(px[0], py[0], pz[0]);
      tmem_wait_ld8x3(px[1], py[1], pz[1]);
#pragma unroll
      for (int qd = 0; qd < 4; ++qd) {
        const int cur = qd % 3, nxt = (qd + 2) % 3;
        if (qd < 2) {
          tmem_ld8(acc + 0 * kTcTile + 8 * (qd + 2), px[nxt]);
          tmem_ld8(acc + 1 * kTcTile + 8 * (qd + 2), py[nxt]);
          tmem_ld8(acc + 2 * kTcTile + 8 * (qd + 2), pz[nxt]);
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const ulonglong2 u4 = *reinterpret_cast<const ulonglong2*>(nu + 8 * qd + 4 * q);
          const ulonglong2 v4 = *reinterpret_cast<const ulonglong2*>(nv + 8 * qd + 4 * q);
#pragma unroll
          for (int pr = 0; pr < 2; ++pr) {
            const int j = 4 * q + 2 * pr;
            const u64 rs = pack2(rsq_approx(pz[cur][j]), rsq_approx(pz[cur][j + 1]));
            const u64 inv = fmul2(rs, rs);
            const u64 du = ffma2(pack2(px[cur][j], px[cur][j + 1]), inv, pr ? u4.y : u4.x);
            const u64 dv = ffma2(pack2(py[cur][j], py[cur][j + 1]), inv, pr ? v4.y : v4.x);
            const u64 e2 = ffma2(du, du, fmul2(dv, dv));
            float ea, eb;
            unpack2(e2, ea, eb);
            if (COUNT) {
              msac_add(ea, thr, s, c);
              msac_add(eb, thr, s, c);
            } else {
              s2 = fadd2(s2, pack2(fminf(ea, thr), fminf(eb, thr)));
            }
          }
        }
        if (qd == 0) tmem_wait_ld8x3(px[nxt], py[nxt], pz[nxt]);
        // the last columns are in registers: the set goes back to the producer, two groups of arithmetic to go
        if (qd == 1) tmem_wait_ld8x3_release(px[nxt], py[nxt], pz[nxt], &S.acc_empty[grp], lane0);
      }
      live_in_chunk += max(0, min(kTcTile, n_live - (t_begin + it) * kTcTile));
      if (((visit + 1) % kTcFlush) == 0 || it + 2 >= ntiles) {
        if (COUNT) {
          dacc += w * (static_cast<double>(s) + thr64 * static_cast<double>(live_in_chunk - c));
          dcnt += c;
        } else {
          float sa, sb;
          unpack2(s2, sa, sb);
          const int processed = (visit % kTcFlush + 1) * kTcTile;
          dacc += w * (static_cast<double>(sa) + static_cast<double>(sb) -
                       static_cast<double>(thr) * static_cast<double>(processed - live_in_chunk));
          s2 = 0ull;
        }
        s = 0.0f;
        c = 0;
        live_in_chunk = 0;
      }
    }
  }

  // join the two groups' sums of every hypothesis (fixed order: even tiles + odd tiles)
  __syncthreads();  // every tile has landed and has been consumed: the ring is idle
  TcJoin& J = *reinterpret_cast<TcJoin*>(S.ring);
  if (!producer && grp == 1) {
    J.s[hyp] = dacc;
    J.c[hyp] = dcnt;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp < 4 && k < a.k) {
    a.part_score[static_cast<size_t>(split) * a.k + k] = dacc + J.s[hyp];
    if (COUNT) a.part_count[static_cast<size_t>(split) * a.k + k] = dcnt + J.c[hyp];
  }
  if (producer) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem),
                 "r"(static_cast<uint32_t>(kTcpCols))
                 : "memory");
  }
}

// ------------------------------------------------ filtered f64 sweep ("f64 scores at f32 speed")
// The reference's ScoreModel (ransac.h:295-312) adds min(e, thr) per point: a point whose squared error is
// clearly above the threshold contributes exactly thr, whatever its error is. This kernel classifies every
// (hypothesis, point) pair with the tensor-core f32 front end above and evaluates in f64 only the pairs that
// could be inliers:
//     near  <=>  e32 < thr_hi(k)   or   |pz32| < pzmin(k)
// with a per-hypothesis band derived from an error bound of the f32 evaluation (below), so that every pair
// NOT flagged has e64 >= thr (or lies behind the camera) and adds thr. The flagged pairs go through
// PoseF<double>::eval's arithmetic (score_kernels.cu) on the recentred f64 copy: inlier decisions - hence
// counts - are those of the plain f64 sweep, bit for bit.
//
// Work distribution: the front end leaves one 32-bit mask per thread (hypothesis) and tile (32 points). A good
// hypothesis has half of its points flagged, a bad one none, so lane-local processing would serialise; the
// warp transposes its 32 x 32 bit matrix (5 shuffle steps) and lane j takes POINT j of the tile against the
// flagged hypotheses of the warp, whose f64 coefficients sit in shared memory. Sums are accumulated as
// fixed-point integers (e * 2^45 / thr, split in two 32-bit halves) with shared-memory reductions, so they do
// not depend on the order of arrival: results are bit-reproducible like everything else here, and differ from
// a sequential f64 sum by <= 2^-46 thr per inlier (relative 1e-14 on C5; the promise is 1e-9).
//
// Error bound of the front end, per hypothesis k: with Rmax = max |X'|, B_i = |row_i| Rmax + |t_i| bounds every
// partial product sum of row i, and the TF32x2 split + f32 accumulation keep |p32_i - p64_i| <= eta B_i with
// eta = 2^-19 (8x the 2^-22 per-term figure above). For pz >= pzmin = B_z / 64 the projected coordinate is off by
// <= 2^-13 (B_xy / B_z + |proj|), and |proj| <= U = umax + sqrt(thr) + 1 for any point that could be an inlier
// (points projected farther out have e64 > thr as long as the bound stays below 1 px, which is checked):
//     delta = 2^-13 (rho + U) + 2^-19 U      (the second term: rsqrt^2, the stored -u / -v, the final FMAs)
//     thr_hi = (sqrt(thr) + 1.5 delta)^2.
// A hypothesis whose bound is useless (delta >= 0.5 px) has thr_hi = +inf: everything in front of the camera is
// evaluated in f64 (correct, slow). Invalid hypotheses (valid[k] == 0) and non-finite models are never flagged:
// finalize_kernel overrides their score (DBL_MAX / NaN) as for every other sweep.
constexpr int kTcxCopies = 4;        // replicas of every accumulator (lane & 3): same-address reductions 4x rarer
constexpr int kTcxFlush = 16;        // tiles between accumulator drains: 16 * 8 entries * 2^23 < 2^32 per replica
constexpr int kTcxHalfBits = 23;     // q = e * 2^45 / thr < 2^45: low 23 bits + high 22 bits

struct TcxSmem {
  float ring[kTcRing][kTcTileFloats];           // f32 operand tiles
  double pts[kTcRing][5][kTcTile];              // the same points, recentred f64 channels X Y Z u v
  double coef[12][kTcThreads];                  // PoseF<double>::Model of every hypothesis of the CTA
  uint32_t acc[3][kTcThreads][kTcxCopies];      // lo, hi, count
  uint32_t nanf[kTcThreads];
  uint64_t full[kTcRing];
  uint64_t mma_done;
  uint32_t tmem_base;
};

// The reciprocal of the f64 sweep (score_kernels.cu: rcp_f64_sweep) - keep the two in step.
__device__ __forceinline__ double rcp_f64_sweep_tc(double x) {
  double r0;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r0) : "d"(x));
  const double e = fma(-x, r0, 1.0);
  const double r = fma(fma(e, e, e), r0, r0);
  return ((__double2hiint(x) & 0x7ff00000) == 0) ? r0 : r;
}

// bit j of lane h  ->  bit h of lane j
__device__ __forceinline__ uint32_t transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const int j = 16 >> s;
    const uint32_t m = (s == 0) ? 0x0000ffffu : (s == 1) ? 0x00ff00ffu : (s == 2) ? 0x0f0f0f0fu
                     : (s == 3) ? 0x33333333u : 0x55555555u;
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
    x = (lane & j) ? (((y & ~m) >> j) | (x & ~m)) : ((x & m) | ((y & m) << j));
  }
  return x;
}

// mask |= bit when the pair could be an inlier: e < thr_hi or |pz| < pzmin (NaN compares false).
// (bit is a constant after unrolling: the OR becomes a LOP3 with an immediate)
__device__ __forceinline__ void near_test(float e, float pz, float thr_hi, float pzmin, uint32_t bit, uint32_t& mask) {
  asm("{\n\t.reg .pred p;\n\t.reg .f32 a;\n\tabs.f32 a, %2;\n\tsetp.lt.f32 p, %1, %3;\n\t"
      "setp.lt.or.f32 p, a, %4, p;\n\t@p or.b32 %0, %0, %5;\n\t}"
      : "+r"(mask)
      : "f"(e), "f"(pz), "f"(thr_hi), "f"(pzmin), "r"(bit));
}

// PREPASS = true: the front end alone over a strided sample of the tiles; part_count[split][k] = number of
// flagged pairs of hypothesis k in this split's sample (the partition kernel turns that into the dense / sparse
// hypothesis lists). PREPASS = false: the hypotheses a.ids[0 .. *a.k_dev) over all tiles, front end + back end,
// partial sums to part_score / part_count[split][kk] in the compact index space.
template <bool PREPASS>
__global__ void __launch_bounds__(kTcThreads, 4)
    sweep_pose_tcx_kernel(const DeviceData d, const SweepArgs a, const float* __restrict__ tiles,
                          const float* __restrict__ bounds) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcxSmem& S = *reinterpret_cast<TcxSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int k_count = (a.k_dev != nullptr) ? *a.k_dev : a.k;
  if (static_cast<int>(blockIdx.x) * kTcThreads >= k_count) return;  // the list is shorter than its capacity
  const int kk = blockIdx.x * kTcThreads + tid;
  const int split = blockIdx.y;
  const int n_live = d.n[0];

  // tiles of this CTA: t_begin + it * t_step, it < ntiles
  const int total_tiles = d.npad / kTcTile;
  const int t_step = PREPASS ? a.tile_stride : 1;
  const int t_begin = split * a.stages_per_split * t_step;
  const int ntiles = PREPASS ? max(0, min(a.stages_per_split, (total_tiles - t_begin + t_step - 1) / t_step))
                             : max(0, min(total_tiles, t_begin + a.stages_per_split) - t_begin);

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)),
                 "r"(static_cast<uint32_t>(kTcCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 32) {
#pragma unroll
    for (int s = 0; s < kTcRing; ++s) mbar_init(&S.full[s], 1);
    mbar_init(&S.mma_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (!PREPASS) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int cpy = 0; cpy < kTcxCopies; ++cpy) S.acc[i][tid][cpy] = 0u;
    S.nanf[tid] = 0u;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem_base;
  const uint32_t lane_addr = tmem + (static_cast<uint32_t>(warp * 32) << 16);

  const double thr64 = a.thr[0];
  float thr_hi = -1.0f, pzmin = 0.0f;  // never flagged
  {
    float r[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float t[3] = {0.f, 0.f, 1.f};
    double c64[12] = {0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 1.};
    if (kk < k_count) {
      const int k = (a.ids != nullptr) ? a.ids[kk] : kk;
      const double* m = a.models + static_cast<size_t>(k) * 12;
      const double dc[3] = {d.centre[0] - m[9], d.centre[1] - m[10], d.centre[2] - m[11]};
      const double sc[3] = {d.fx, d.fy, 1.0};
      bool finite = true;
      double B[3];
#pragma unroll
      for (int row = 0; row < 3; ++row) {
        const double r0 = m[row], r1 = m[row + 3], r2 = m[row + 6];
        const double tt = r0 * dc[0] + r1 * dc[1] + r2 * dc[2];  // PoseF<double>::load, operation for operation
        c64[3 * row + 0] = r0 * sc[row];
        c64[3 * row + 1] = r1 * sc[row];
        c64[3 * row + 2] = r2 * sc[row];
        c64[9 + row] = tt * sc[row];
        r[3 * row + 0] = static_cast<float>(r0 * sc[row]);
        r[3 * row + 1] = static_cast<float>(r1 * sc[row]);
        r[3 * row + 2] = static_cast<float>(r2 * sc[row]);
        t[row] = static_cast<float>((r0 * dc[0] + r1 * dc[1] + r2 * dc[2]) * sc[row]);
        const double nr = sqrt(c64[3 * row] * c64[3 * row] + c64[3 * row + 1] * c64[3 * row + 1] +
                               c64[3 * row + 2] * c64[3 * row + 2]);
        B[row] = nr * static_cast<double>(bounds[0]) + fabs(c64[9 + row]);
        finite = finite && (B[row] - B[row] == 0.0);
      }
      const bool is_valid = (a.valid == nullptr) || (a.valid[k] != 0);
      if (is_valid && finite && B[2] > 0.0 && thr64 > 0.0) {
        const double rt = sqrt(thr64);
        const double U = static_cast<double>(bounds[1]) + rt + 1.0;
        const double rho = fmax(B[0], B[1]) / B[2];
        const double delta = (rho + U) * (1.0 / 8192.0) + U * (1.0 / 524288.0);
        const double hi = (rt + 1.5 * delta) * (rt + 1.5 * delta);
        thr_hi = (delta < 0.5) ? static_cast<float>(hi) * 1.000001f : __int_as_float(0x7f800000);
        pzmin = static_cast<float>(B[2] * (1.0 / 64.0)) * 1.000001f;
      }
    }
    if (!PREPASS) {
#pragma unroll
      for (int i = 0; i < 12; ++i) S.coef[i][tid] = c64[i];
    }
#pragma unroll
    for (int mat = 0; mat < 3; ++mat) {
      const float v[4] = {r[3 * mat], r[3 * mat + 1], r[3 * mat + 2], t[mat]};
      float hl[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        hl[j] = tf32_hi(v[j]);
        hl[4 + j] = v[j] - hl[j];
      }
      tmem_st8(lane_addr + 8 * mat, hl);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  const double* soa64 = d.soa64;
  const size_t npad = static_cast<size_t>(d.npad);
  auto issue_tma = [&](int it) {
    const int slot = it % kTcRing;
    const size_t tile = static_cast<size_t>(t_begin + it * t_step);
    mbar_expect_tx(&S.full[slot], kTcTileBytes + (PREPASS ? 0 : 5 * kTcTile * 8));
    tma_load_1d(S.ring[slot], tiles + tile * kTcTileFloats, kTcTileBytes, &S.full[slot]);
    if (!PREPASS) {
#pragma unroll
      for (int ch = 0; ch < 5; ++ch)
        tma_load_1d(S.pts[slot][ch], soa64 + ch * npad + tile * kTcTile, kTcTile * 8, &S.full[slot]);
    }
  };
  if (tid == 0) {
    for (int it = 0; it < kTcRing && it < ntiles; ++it) issue_tma(it);
  }
  __syncwarp();

  const double scale = 35184372088832.0 / thr64;  // 2^45 / thr
  unsigned long long qsum = 0ull;
  int cnt = 0;

  // warp 0: the six MMAs of tile `jt` (its operands have landed) and their commit
  auto issue_mma = [&](int jt) {
    const int sl = jt % kTcRing;
    mbar_wait(&S.full[sl], (jt / kTcRing) & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (elect_one()) {
      const uint32_t b_addr = smem_u32(S.ring[sl]);
#pragma unroll
      for (int kx = 0; kx < 2; ++kx) {
        const uint64_t db = umma_desc(b_addr + kx * 2 * 512, 512, 128);
#pragma unroll
        for (int mat = 0; mat < 3; ++mat) umma_tf32_ts(tmem + 32 + mat * kTcTile, tmem + 8 * mat, db, kx);
      }
      umma_commit(&S.mma_done);
    }
    __syncwarp();
  };
  if (warp == 0 && ntiles > 0) issue_mma(0);

  // Per tile: f32 front end (TMEM read-back -> mask), barrier, the NEXT tile's MMAs are issued, and the f64 back
  // end of this tile runs while the tensor core works. The back end needs the tile's f64 points, not TMEM: the
  // ring slot of tile it is refilled one iteration later than in the plain tensor-core sweep.
  for (int it = 0; it < ntiles; ++it) {
    const int slot = it % kTcRing;
    mbar_wait(&S.mma_done, it & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    const float* nu = S.ring[slot] + 512;
    const float* nv = nu + 32;
    uint32_t mask = 0u;
    float px[2][8], py[2][8], pz[2][8];
    tmem_ld8(lane_addr + 32 + 0 * kTcTile, px[0]);
    tmem_ld8(lane_addr + 32 + 1 * kTcTile, py[0]);
    tmem_ld8(lane_addr + 32 + 2 * kTcTile, pz[0]);
    tmem_wait_ld8x3(px[0], py[0], pz[0]);
#pragma unroll
    for (int qd = 0; qd < 4; ++qd) {
      const int cur = qd & 1, nxt = cur ^ 1;
      if (qd < 3) {
        tmem_ld8(lane_addr + 32 + 0 * kTcTile + 8 * (qd + 1), px[nxt]);
        tmem_ld8(lane_addr + 32 + 1 * kTcTile + 8 * (qd + 1), py[nxt]);
        tmem_ld8(lane_addr + 32 + 2 * kTcTile + 8 * (qd + 1), pz[nxt]);
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const ulonglong2 u4 = *reinterpret_cast<const ulonglong2*>(nu + 8 * qd + 4 * q);
        const ulonglong2 v4 = *reinterpret_cast<const ulonglong2*>(nv + 8 * qd + 4 * q);
#pragma unroll
        for (int pr = 0; pr < 2; ++pr) {
          const int j = 4 * q + 2 * pr;
          const u64 rs = pack2(rsq_approx(pz[cur][j]), rsq_approx(pz[cur][j + 1]));
          const u64 inv = fmul2(rs, rs);
          const u64 du = ffma2(pack2(px[cur][j], px[cur][j + 1]), inv, pr ? u4.y : u4.x);
          const u64 dv = ffma2(pack2(py[cur][j], py[cur][j + 1]), inv, pr ? v4.y : v4.x);
          const u64 e2 = ffma2(du, du, fmul2(dv, dv));
          float ea, eb;
          unpack2(e2, ea, eb);
          near_test(ea, pz[cur][j], thr_hi, pzmin, 1u << (8 * qd + j), mask);
          near_test(eb, pz[cur][j + 1], thr_hi, pzmin, 1u << (8 * qd + j + 1), mask);
        }
      }
      if (qd < 3) tmem_wait_ld8x3(px[nxt], py[nxt], pz[nxt]);
    }
    // padding is never an inlier (its f64 channels are zeros)
    const int nl = n_live - (t_begin + it * t_step) * kTcTile;
    if (nl < kTcTile) mask &= (nl <= 0) ? 0u : ((1u << nl) - 1u);

    // everyone has read its accumulators (and finished the back end of the previous tile)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
      if (it + 1 < ntiles) issue_mma(it + 1);
      if (it >= 1 && it - 1 + kTcRing < ntiles) {
        if (elect_one()) issue_tma(it - 1 + kTcRing);
        __syncwarp();
      }
    }
    if (PREPASS) {
      cnt += __popc(mask);
      continue;
    }

    // ---- f64 back end of this tile (the tensor core is working on the next one) ----
    // The hypotheses of this launch have FEW flagged pairs (the dense ones went to the plain f64 sweep), spread
    // over hypotheses and points: the warp transposes its 32 x 32 bit matrix and lane j takes POINT j of the
    // tile against the flagged hypotheses of the warp.
    if (__any_sync(0xffffffffu, mask != 0u)) {
      uint32_t todo = transpose32(mask, lane);  // bit h: hypothesis warp * 32 + h against MY point of the tile
      const double X = S.pts[slot][0][lane], Y = S.pts[slot][1][lane], Z = S.pts[slot][2][lane];
      const double U = S.pts[slot][3][lane], V = S.pts[slot][4][lane];
      const int cpy = lane & (kTcxCopies - 1);
      while (todo) {
        const int hh = warp * 32 + (__ffs(todo) - 1);
        todo &= todo - 1u;
        // PoseF<double>::eval (score_kernels.cu), operation for operation
        const double pz64 = fma(S.coef[6][hh], X, fma(S.coef[7][hh], Y, fma(S.coef[8][hh], Z, S.coef[11][hh])));
        const double px64 = fma(S.coef[0][hh], X, fma(S.coef[1][hh], Y, fma(S.coef[2][hh], Z, S.coef[9][hh])));
        const double py64 = fma(S.coef[3][hh], X, fma(S.coef[4][hh], Y, fma(S.coef[5][hh], Z, S.coef[10][hh])));
        const double inv = rcp_f64_sweep_tc(pz64);
        const double du = fma(px64, inv, -U);
        const double dv = fma(py64, inv, -V);
        double e = fma(du, du, dv * dv);
        e = (pz64 < 0.0) ? DBL_MAX : e;
        if (e < thr64) {
          const unsigned long long q = __double2ull_rn(e * scale);
          atomicAdd(&S.acc[0][hh][cpy], static_cast<uint32_t>(q) & ((1u << kTcxHalfBits) - 1u));
          atomicAdd(&S.acc[1][hh][cpy], static_cast<uint32_t>(q >> kTcxHalfBits));
          atomicAdd(&S.acc[2][hh][cpy], 1u);
        } else if (e != e) {
          S.nanf[hh] = 1u;  // std::min(e, thr) propagates NaN (ransac.h:308)
        }
      }
      __syncwarp();
    }
    if (((it + 1) % kTcxFlush) == 0 || it + 1 == ntiles) {
      // a hypothesis' accumulators are only touched by its own warp
      __syncwarp();
      const uint4 lo = *reinterpret_cast<const uint4*>(S.acc[0][tid]);
      const uint4 hi = *reinterpret_cast<const uint4*>(S.acc[1][tid]);
      const uint4 cc = *reinterpret_cast<const uint4*>(S.acc[2][tid]);
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(S.acc[0][tid]) = z;
      *reinterpret_cast<uint4*>(S.acc[1][tid]) = z;
      *reinterpret_cast<uint4*>(S.acc[2][tid]) = z;
      qsum += static_cast<unsigned long long>(lo.x) + lo.y + lo.z + lo.w;
      qsum += (static_cast<unsigned long long>(hi.x) + hi.y + hi.z + hi.w) << kTcxHalfBits;
      cnt += static_cast<int>(cc.x + cc.y + cc.z + cc.w);
      __syncwarp();
    }
  }

  if (kk < k_count) {
    if (PREPASS) {
      a.part_count[static_cast<size_t>(split) * a.k + kk] = cnt;
    } else {
      const int live = max(0, min(ntiles * kTcTile, n_live - t_begin * kTcTile));
      double sc = thr64 * static_cast<double>(live - cnt) + static_cast<double>(qsum) * (thr64 * (1.0 / 35184372088832.0));
      if (S.nanf[tid]) sc = __longlong_as_double(0x7ff8000000000000LL);
      a.part_score[static_cast<size_t>(split) * a.k + kk] = a.w[0] * sc;
      a.part_count[static_cast<size_t>(split) * a.k + kk] = cnt;
    }
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem),
                 "r"(static_cast<uint32_t>(kTcCols))
                 : "memory");
  }
}

// ---- the dense / sparse split of a batch -------------------------------------------------------------------
// One CTA. flagged[sp][k] (the prepass' part_count) -> per hypothesis: dense when at least 1 / 128 of the sampled
// pairs were flagged (there the plain f64 sweep is cheaper per pair than flag + gather), else sparse. Ordered
// compaction (ids ascending in both lists): everything downstream is reproducible.
//   lists[0 .. k)     ids of the dense hypotheses, lists[k .. 2k) ids of the sparse ones
//   pos[k]            position of hypothesis k in its list; dense ones as ~position (negative)
//   counts[0], [1]    lengths of the two lists
__global__ void __launch_bounds__(1024)
    tcx_partition_kernel(const int32_t* __restrict__ flagged, int n_splits, int k, int sample_points,
                         int32_t* __restrict__ lists, int32_t* __restrict__ pos, int32_t* __restrict__ counts) {
  __shared__ int woff[2][32];
  __shared__ int base[2];
  __shared__ int chunk_total[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 2) base[tid] = 0;
  __syncthreads();
  const int dense_min = max(1, sample_points / 128);
  const uint32_t below = (1u << lane) - 1u;
  for (int k0 = 0; k0 < k; k0 += 1024) {
    const int kk = k0 + tid;
    const bool live = kk < k;
    int nf = 0;
    if (live)
      for (int sp = 0; sp < n_splits; ++sp) nf += flagged[static_cast<size_t>(sp) * k + kk];
    const bool dense = live && nf >= dense_min;
    const uint32_t bd = __ballot_sync(0xffffffffu, dense);
    const uint32_t bs = __ballot_sync(0xffffffffu, live && !dense);
    if (lane == 0) {
      woff[0][warp] = __popc(bd);
      woff[1][warp] = __popc(bs);
    }
    __syncthreads();
    if (warp < 2) {  // exclusive scan of the 32 warp counts of list `warp`
      const int v = woff[warp][lane];
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
      }
      woff[warp][lane] = inc - v;
      if (lane == 31) chunk_total[warp] = inc;
    }
    __syncthreads();
    if (live) {
      const int l = dense ? 0 : 1;
      const int p = base[l] + woff[l][warp] + __popc((dense ? bd : bs) & below);
      lists[static_cast<size_t>(l) * k + p] = kk;
      pos[kk] = dense ? ~p : p;
    }
    __syncthreads();
    if (tid < 2) base[tid] += chunk_total[tid];
    __syncthreads();
  }
  if (tid < 2) counts[tid] = base[tid];
}

// Filtered f64 sweep, the launches of one batch (all on `st`, nothing waits on the host):
//   1. prepass     sweep_pose_tcx_kernel<true> over a strided sample of the tiles -> flagged pairs per hypothesis
//   2. partition   dense / sparse hypothesis lists
//   3. dense       the plain f64 sweep (score_kernels.cu) over the dense list, sums into a.acc_q / a.acc_c
//   4. sparse      sweep_pose_tcx_kernel<false> over the sparse list, partials into part_score / part_count
// finalize_kernel then reads a hypothesis' result where a.pos says it is.
cudaError_t launch_sweep_tcx_prepass(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                                     int* launches) {
  ++*launches;
  static PerDeviceOnce configured;
  const int dev = PerDeviceOnce::device();
  if (!configured.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(sweep_pose_tcx_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(TcxSmem)));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(sweep_pose_tcx_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(sizeof(TcxSmem)));
    if (e != cudaSuccess) return e;
    configured.set(dev);
  }
  dim3 grid((a.k + kTcThreads - 1) / kTcThreads, a.n_splits);
  sweep_pose_tcx_kernel<true><<<grid, kTcThreads, sizeof(TcxSmem), st>>>(
      d, a, tiles, tiles + static_cast<size_t>(d.npad / kTcTile) * kTcTileFloats);
  return cudaGetLastError();
}

cudaError_t launch_tcx_partition(const int32_t* flagged, int n_splits, int k, int sample_points, int32_t* lists,
                                 int32_t* pos, int32_t* counts, cudaStream_t st, int* launches) {
  ++*launches;
  tcx_partition_kernel<<<1, 1024, 0, st>>>(flagged, n_splits, k, sample_points, lists, pos, counts);
  return cudaGetLastError();
}

// (the attributes were set by the prepass launch, which always comes first)
cudaError_t launch_sweep_tcx(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                             int* launches) {
  ++*launches;
  dim3 grid((a.k + kTcThreads - 1) / kTcThreads, a.n_splits);
  sweep_pose_tcx_kernel<false><<<grid, kTcThreads, sizeof(TcxSmem), st>>>(
      d, a, tiles, tiles + static_cast<size_t>(d.npad / kTcTile) * kTcTileFloats);
  return cudaGetLastError();
}

int tcx_resident_ctas() {
  static std::atomic<int> cache[64];
  std::atomic<int>& cached = cache[PerDeviceOnce::device()];
  if (const int v = cached.load(std::memory_order_relaxed)) return v;
  cudaFuncSetAttribute(sweep_pose_tcx_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(TcxSmem)));
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sweep_pose_tcx_kernel<false>, kTcThreads,
                                                    sizeof(TcxSmem)) != cudaSuccess)
    n = 1;
  const int v = max(1, min(n, 512 / kTcCols));
  cached.store(v, std::memory_order_relaxed);
  return v;
}

// ------------------------------------------------ the sorted two-speed sweep (variant 12)
// The probes above say the tensor-core sweep is bound by CUDA-core dispatch: ~15 dispatch slots per evaluation,
// of which the reciprocal, the residual and the MSAC adds (9) only matter for pairs that can be inliers - and for
// random minimal-sample hypotheses almost no pair is. Two changes make the common case cheap:
//  1. -u, -v move into the contraction: a = px - u pz and b = py - v pz are bilinear in (hypothesis, point)
//         a = [rx tx | rz tz] . [X 1 | -uX -u]        b = [rz tz | ry ty] . [-vX -v | X 1]      pz = [rz tz | ..] . [X 1 | 0]
//     (K = 8 each, TF32 hi/lo split: A_hi B_hi + A_lo B_hi + A_hi B_lo, nine MMAs per 32-point tile instead of six),
//     and the inlier test needs no reciprocal:  e < thr  <=>  a^2 + b^2 < thr pz^2  (pz > 0).
//  2. the test is a FILTER: q = a^2 + b^2 < thr (1 + 2^-12) pz^2 flags a superset of the inliers (rounding of either
//     side is covered by the margin; the sign of pz is not looked at). A warp whose 32 hypotheses flag none of the
//     eight points of a column group skips them (they add thr each, which the count-based score formula already
//     does); otherwise the group goes through e = q / pz^2 (rsqrt, as before), the MSAC test and the sums.
//     What is added to a hypothesis' sums is decided by the second, exact step alone, so a score does not depend
//     on which other hypotheses share the warp.
// For the skip to happen the hypotheses of a warp must be alike: a prepass over every 64th tile counts each
// hypothesis' flags, tc2_order_kernel buckets the hypotheses by that count (log2 buckets, best first) and the sweep
// walks them in that order (a.ids); finalize_kernel finds a hypothesis' partials through a.pos.
// Built on the producer / epilogue-group skeleton of sweep_pose_tcp_kernel (the nine MMAs of tile t + 1 run under
// the reduction of tile t). Point image: six 512-byte K-chunks per tile, [c2h c0h c1h | c2l c0l c1l] with
// c0 = (X, 1), c1 = (-uX, -u), c2 = (-vX, -v); the zero chunk of pz sits once behind the ring.
constexpr int kTc2TileFloats = 6 * 128;
constexpr int kTc2TileBytes = kTc2TileFloats * 4;

__global__ void tc2_pack_kernel(const DeviceData d, float* __restrict__ tiles) {
  const int total = d.npad;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    float c[3][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    float l[3][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    if (g < d.n[0]) {  // padding stays all zero: a = b = pz = 0 is never flagged and never counted
      const double x[4] = {static_cast<double>(d.soa32[0 * static_cast<size_t>(total) + g]),
                           static_cast<double>(d.soa32[1 * static_cast<size_t>(total) + g]),
                           static_cast<double>(d.soa32[2 * static_cast<size_t>(total) + g]), 1.0};
      const double f[3] = {1.0, static_cast<double>(d.soa32[3 * static_cast<size_t>(total) + g]),   // -u
                           static_cast<double>(d.soa32[4 * static_cast<size_t>(total) + g])};       // -v
#pragma unroll
      for (int ch = 0; ch < 3; ++ch)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double pr = f[ch] * x[j];  // exact in f64 (24 x 24 bits)
          c[ch][j] = tf32_hi(static_cast<float>(pr));
          l[ch][j] = static_cast<float>(pr - static_cast<double>(c[ch][j]));
        }
    }
    float* t = tiles + static_cast<size_t>(g / kTcTile) * kTc2TileFloats;
    const int n = g % kTcTile;
    const int at = ((n >> 3) * 8 + (n & 7)) * 4;  // core matrix n / 8, row n % 8
    const int order[3] = {1, 2, 0};               // chunk slot of c0, c1, c2: [c2 c0 c1]
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      *reinterpret_cast<float4*>(t + order[ch] * 128 + at) = make_float4(c[ch][0], c[ch][1], c[ch][2], c[ch][3]);
      *reinterpret_cast<float4*>(t + (3 + order[ch]) * 128 + at) = make_float4(l[ch][0], l[ch][1], l[ch][2], l[ch][3]);
    }
  }
}

struct Tc2Smem {
  float ring[kTcRing][kTc2TileFloats];
  float zero[128];          // the second K-chunk of pz's B operand
  uint64_t full[kTcRing];
  uint64_t acc_full[2];
  uint64_t acc_empty[2];
  uint32_t tmem_base;
};

// PREPASS: tiles t_begin, t_begin + stride, ...; part_count[split][k] = flags of hypothesis k (identity order).
template <bool PREPASS>
__global__ void __launch_bounds__(kTcpThreads, 2)
    sweep_pose_tc2_kernel(const DeviceData d, const SweepArgs a, const float* __restrict__ tiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  Tc2Smem& S = *reinterpret_cast<Tc2Smem*>(smem_raw);
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const bool producer = warp == 8;
  const int quarter = warp & 3;
  const int grp = (warp >> 2) & 1;
  const int hyp = quarter * 32 + (tid & 31);
  const int kk = blockIdx.x * 128 + hyp;  // position in the sweep order
  const int split = blockIdx.y;
  const int n_live = d.n[0];
  const int stride = PREPASS ? a.tile_stride : 1;

  const int total_tiles = d.npad / kTcTile;
  const int sweep_tiles = PREPASS ? (total_tiles + stride - 1) / stride : total_tiles;
  const int t_begin = split * a.stages_per_split;
  const int t_end = min(sweep_tiles, t_begin + a.stages_per_split);
  const int ntiles = max(0, t_end - t_begin);

  if (producer) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)),
                 "r"(static_cast<uint32_t>(kTcpCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kTcRing; ++s) mbar_init(&S.full[s], 1);
    mbar_init(&S.acc_full[0], 1);
    mbar_init(&S.acc_full[1], 1);
    mbar_init(&S.acc_empty[0], 4);
    mbar_init(&S.acc_empty[1], 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid < 128) S.zero[tid] = 0.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // S.zero is read by the tensor core (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem_base;
  const uint32_t lane_addr = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
  if (warp < 4) {
    // A operand, TMEM columns: [0,8) (rx tx | rz tz) hi, [8,16) (rz tz | ry ty) hi, [16,24) and [24,32) the lo parts
    float r[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float t[3] = {0.f, 0.f, 0.f};
    const int k = (kk < a.k) ? (a.ids ? a.ids[kk] : kk) : -1;
    if (k >= 0) {
      const double* m = a.models + static_cast<size_t>(k) * 12;
      const double dc[3] = {d.centre[0] - m[9], d.centre[1] - m[10], d.centre[2] - m[11]};
      const double sc[3] = {d.fx, d.fy, 1.0};
#pragma unroll
      for (int row = 0; row < 3; ++row) {
        const double r0 = m[row], r1 = m[row + 3], r2 = m[row + 6];
        r[3 * row + 0] = static_cast<float>(r0 * sc[row]);
        r[3 * row + 1] = static_cast<float>(r1 * sc[row]);
        r[3 * row + 2] = static_cast<float>(r2 * sc[row]);
        t[row] = static_cast<float>((r0 * dc[0] + r1 * dc[1] + r2 * dc[2]) * sc[row]);
      }
    }
    const int rows[2][2] = {{0, 2}, {2, 1}};  // (x | z), (z | y)
#pragma unroll
    for (int blk = 0; blk < 2; ++blk) {
      float hi[8], lo[8];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int row = rows[blk][h];
        const float v[4] = {r[3 * row], r[3 * row + 1], r[3 * row + 2], t[row]};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          hi[4 * h + j] = tf32_hi(v[j]);
          lo[4 * h + j] = v[j] - hi[4 * h + j];
        }
      }
      tmem_st8(lane_addr + 8 * blk, hi);
      tmem_st8(lane_addr + 16 + 8 * blk, lo);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  double dacc = 0.0;
  int dcnt = 0;

  if (producer) {
    auto issue_tma = [&](int it) {
      const int slot = it % kTcRing;
      mbar_expect_tx(&S.full[slot], kTc2TileBytes);
      tma_load_1d(S.ring[slot], tiles + static_cast<size_t>(t_begin + it) * stride * kTc2TileFloats, kTc2TileBytes,
                  &S.full[slot]);
    };
    if (elect_one()) {
      for (int it = 0; it < kTcRing && it < ntiles; ++it) issue_tma(it);
    }
    __syncwarp();
    const uint32_t zero_addr = smem_u32(S.zero);
#ifdef RLB_TC2_STAMPS
    long long st_p[4] = {0, 0, 0, 0};
#define TC2_STAMP(arr, i, t0) { const long long t1_ = clock64(); arr[i] += t1_ - t0; t0 = t1_; }
    long long tp = clock64();
#else
#define TC2_STAMP(arr, i, t0)
#endif
    for (int it = 0; it < ntiles; ++it) {
      const int slot = it % kTcRing;
      const int buf = it & 1;
      TC2_STAMP(st_p, 3, tp);
      if (it >= 2) {
        mbar_wait(&S.acc_empty[buf], ((it >> 1) - 1) & 1);
        // the group that handed set `buf` back has finished tile it - 2: the operand tiles are only read by the
        // tensor core, whose MMAs on tile it - 2 completed before that group could start
        if (it - 2 + kTcRing < ntiles) {
          if (elect_one()) issue_tma(it - 2 + kTcRing);
          __syncwarp();
        }
      }
      TC2_STAMP(st_p, 0, tp);
      mbar_wait(&S.full[slot], (it / kTcRing) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      TC2_STAMP(st_p, 1, tp);
      if (elect_one()) {
        const uint32_t b_addr = smem_u32(S.ring[slot]);
        const uint32_t acc = tmem + 32 + 96 * buf;
#pragma unroll
        for (int part = 0; part < 3; ++part) {
          // part 0: A hi x B hi, 1: A lo x B hi, 2: A hi x B lo
          const uint32_t a_off = (part == 1) ? 16u : 0u;
          const uint32_t b_off = (part == 2) ? 3u * 512u : 0u;
          const uint32_t c0 = b_addr + b_off + 512u;  // chunk order [c2 c0 c1]
          umma_tf32_ts(acc + 0 * kTcTile, tmem + a_off + 0, umma_desc(c0, 512, 128), part);                // a
          umma_tf32_ts(acc + 1 * kTcTile, tmem + a_off + 8, umma_desc(b_addr + b_off, 512, 128), part);    // b
#ifdef RLB_TC2_PROBE_PZ  // timing probe (wrong results): an ordinary descriptor for pz
          umma_tf32_ts(acc + 2 * kTcTile, tmem + a_off + 8, umma_desc(c0, 512, 128), part);
#else
          umma_tf32_ts(acc + 2 * kTcTile, tmem + a_off + 8, umma_desc(c0, zero_addr - c0, 128), part);     // pz
#endif
        }
        umma_commit(&S.acc_full[buf]);
      }
      __syncwarp();
      TC2_STAMP(st_p, 2, tp);
    }
#ifdef RLB_TC2_STAMPS
    if (!PREPASS && blockIdx.x == 300 && blockIdx.y == 1 && (tid & 31) == 0)
      printf("producer: tiles %d  wait_empty %lld  wait_full %lld  mma_issue %lld  other %lld (cycles per tile)\n", ntiles,
             st_p[0] / ntiles, st_p[1] / ntiles, st_p[2] / ntiles, st_p[3] / ntiles);
#endif
  } else {
    const double thr64 = a.thr[0];
    const float thr = static_cast<float>(thr64);
    const float thrw = thr * (1.0f + 1.0f / 4096.0f);
    const u64 thrw2 = pack2(thrw, thrw);
    const double w = a.w[0];
    const uint32_t lane0 = (tid & 31) == 0 ? 1u : 0u;
    const uint32_t acc = lane_addr + 32 + 96 * grp;
    float s = 0.0f;
    int c = 0;
    int live_in_chunk = 0;
    int visit = 0;
#ifdef RLB_TC2_STAMPS
    long long st_e[4] = {0, 0, 0, 0};
    long long te = clock64();
#endif
    for (int it = grp; it < ntiles; it += 2, ++visit) {
      TC2_STAMP(st_e, 3, te);
      mbar_wait(&S.acc_full[grp], visit & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      TC2_STAMP(st_e, 0, te);
      float pa[3][8], pb[3][8], pz[3][8];
      tmem_ld8(acc + 0 * kTcTile, pa[0]);
      tmem_ld8(acc + 1 * kTcTile, pb[0]);
      tmem_ld8(acc + 2 * kTcTile, pz[0]);
      tmem_ld8(acc + 0 * kTcTile + 8, pa[1]);
      tmem_ld8(acc + 1 * kTcTile + 8, pb[1]);
      tmem_ld8(acc + 2 * kTcTile + 8, pz[1]);
      tmem_wait_ld8x3(pa[0], pb[0], pz[0]);
      tmem_wait_ld8x3(pa[1], pb[1], pz[1]);
      TC2_STAMP(st_e, 1, te);
#pragma unroll
      for (int qd = 0; qd < 4; ++qd) {
        const int cur = qd % 3, nxt = (qd + 2) % 3;
        if (qd < 2) {
          tmem_ld8(acc + 0 * kTcTile + 8 * (qd + 2), pa[nxt]);
          tmem_ld8(acc + 1 * kTcTile + 8 * (qd + 2), pb[nxt]);
          tmem_ld8(acc + 2 * kTcTile + 8 * (qd + 2), pz[nxt]);
        }
        u64 q[4];
        bool flag = false;
#ifdef RLB_TC2_PROBE_NOEPI
#pragma unroll
        for (int pr = 0; pr < 4; ++pr) q[pr] = pack2(pa[cur][2 * pr], pb[cur][2 * pr + 1]);
#else
#pragma unroll
        for (int pr = 0; pr < 4; ++pr) {
          const int j = 2 * pr;
          const u64 a2 = pack2(pa[cur][j], pa[cur][j + 1]);
          const u64 b2 = pack2(pb[cur][j], pb[cur][j + 1]);
          const u64 z2 = pack2(pz[cur][j], pz[cur][j + 1]);
          q[pr] = ffma2(a2, a2, fmul2(b2, b2));
          const u64 t2 = fmul2(fmul2(thrw2, z2), z2);
          float qa, qb, ta, tb;
          unpack2(q[pr], qa, qb);
          unpack2(t2, ta, tb);
          if (PREPASS) {
            c += (qa < ta) ? 1 : 0;
            c += (qb < tb) ? 1 : 0;
          } else {
            flag = flag || (qa < ta) || (qb < tb);
          }
        }
#endif
#ifdef RLB_TC2_PROBE_NOSLOW  // timing probe (wrong results): the filter alone
        if (false) {
#elif defined(RLB_TC2_PROBE_NOEPI)  // timing probe (wrong results): TMEM loads only
        if (pa[cur][0] == 1234.5f) {
#else
        if (!PREPASS && __any_sync(0xffffffffu, flag)) {
#endif
#pragma unroll
          for (int pr = 0; pr < 4; ++pr) {
            const int j = 2 * pr;
            const u64 rs = pack2(rsq_approx(pz[cur][j]), rsq_approx(pz[cur][j + 1]));
            const u64 r2 = fmul2(rs, rs);
            const u64 e2 = fmul2(q[pr], fmul2(r2, r2));  // pz < 0: NaN, pz == 0: inf or NaN - never below thr
            float ea, eb;
            unpack2(e2, ea, eb);
            msac_add(ea, thr, s, c);
            msac_add(eb, thr, s, c);
          }
        }
        if (qd == 0) tmem_wait_ld8x3(pa[nxt], pb[nxt], pz[nxt]);
        if (qd == 1) tmem_wait_ld8x3_release(pa[nxt], pb[nxt], pz[nxt], &S.acc_empty[grp], lane0);
        if (qd == 1) TC2_STAMP(st_e, 2, te);
      }
      if (!PREPASS) {
        live_in_chunk += max(0, min(kTcTile, n_live - (t_begin + it) * kTcTile));
        if (((visit + 1) % kTcFlush) == 0 || it + 2 >= ntiles) {
          dacc += w * (static_cast<double>(s) + thr64 * static_cast<double>(live_in_chunk - c));
          dcnt += c;
          s = 0.0f;
          c = 0;
          live_in_chunk = 0;
        }
      }
    }
    if (PREPASS) dcnt = c;
#ifdef RLB_TC2_STAMPS
    if (!PREPASS && blockIdx.x == 300 && blockIdx.y == 1 && (tid & 31) == 0 && (warp & 3) == 0)
      printf("group %d: visits %d  wait_full %lld  first_loads %lld  until_release %lld  rest %lld (cycles per tile)\n", grp,
             visit, st_e[0] / max(visit, 1), st_e[1] / max(visit, 1), st_e[2] / max(visit, 1), st_e[3] / max(visit, 1));
#endif
  }

  __syncthreads();  // every tile has landed and has been consumed: the ring is idle
  TcJoin& J = *reinterpret_cast<TcJoin*>(S.ring);
  if (!producer && grp == 1) {
    J.s[hyp] = dacc;
    J.c[hyp] = dcnt;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp < 4 && kk < a.k) {
    if (!PREPASS) a.part_score[static_cast<size_t>(split) * a.k + kk] = dacc + J.s[hyp];
    if (PREPASS || a.want_counts) a.part_count[static_cast<size_t>(split) * a.k + kk] = dcnt + J.c[hyp];
  }
  if (producer) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem),
                 "r"(static_cast<uint32_t>(kTcpCols))
                 : "memory");
  }
}

// Sweep order from the prepass counts: bucket 0 = no flag in the sample, bucket b = 1 + floor(log2(count));
// the order is buckets descending (the slow warps first), any order inside a bucket (shared-memory cursors: a
// hypothesis' sums do not depend on its place, so the order need not be reproducible).
//   order[0 .. k) hypothesis ids in sweep order, pos[k] = place of hypothesis k
__global__ void __launch_bounds__(1024)
    tc2_order_kernel(const int32_t* __restrict__ flagged, int n_splits, int k, int32_t* __restrict__ order,
                     int32_t* __restrict__ pos) {
  __shared__ int hist[32];
  __shared__ int cursor[32];
  const int tid = threadIdx.x;
  if (tid < 32) hist[tid] = 0;
  __syncthreads();
  auto bucket_of = [&](int kk) {
    int nf = 0;
    for (int sp = 0; sp < n_splits; ++sp) nf += flagged[static_cast<size_t>(sp) * k + kk];
    return nf <= 0 ? 0 : 32 - __clz(nf);
  };
  for (int kk = tid; kk < k; kk += 1024) atomicAdd(&hist[bucket_of(kk)], 1);
  __syncthreads();
  if (tid == 0) {
    int at = 0;
    for (int b = 31; b >= 0; --b) {
      cursor[b] = at;
      at += hist[b];
    }
  }
  __syncthreads();
  for (int kk = tid; kk < k; kk += 1024) {
    const int p = atomicAdd(&cursor[bucket_of(kk)], 1);
    order[p] = kk;
    pos[kk] = p;
  }
}

size_t tc2_tiles_bytes(int npad) { return static_cast<size_t>(npad / kTcTile) * kTc2TileBytes; }

cudaError_t launch_tc2_pack(const DeviceData& d, float* tiles, cudaStream_t st, int* launches) {
  ++*launches;
  const int blocks = max(1, min(2048, (d.npad + 255) / 256));
  tc2_pack_kernel<<<blocks, 256, 0, st>>>(d, tiles);
  return cudaGetLastError();
}

static cudaError_t tc2_configure() {
  static PerDeviceOnce configured;
  const int dev = PerDeviceOnce::device();
  if (configured.done(dev)) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(sweep_pose_tc2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(sizeof(Tc2Smem)));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(sweep_pose_tc2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             static_cast<int>(sizeof(Tc2Smem)));
  if (e == cudaSuccess) configured.set(dev);
  return e;
}

int tc2_resident_ctas() {
  static std::atomic<int> cache[64];
  std::atomic<int>& cached = cache[PerDeviceOnce::device()];
  if (const int v = cached.load(std::memory_order_relaxed)) return v;
  tc2_configure();
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sweep_pose_tc2_kernel<false>, kTcpThreads,
                                                    sizeof(Tc2Smem)) != cudaSuccess)
    n = 1;
  const int v = max(1, min(n, 512 / kTcpCols));
  cached.store(v, std::memory_order_relaxed);
  return v;
}

// prepass = true: a.tile_stride, a.n_splits / stages_per_split over the sampled tiles, counts to a.part_count
cudaError_t launch_sweep_tc2(const DeviceData& d, const SweepArgs& a, const float* tiles, bool prepass,
                             cudaStream_t st, int* launches) {
  ++*launches;
  const cudaError_t e = tc2_configure();
  if (e != cudaSuccess) return e;
  dim3 grid((a.k + 127) / 128, a.n_splits);
  if (prepass)
    sweep_pose_tc2_kernel<true><<<grid, kTcpThreads, sizeof(Tc2Smem), st>>>(d, a, tiles);
  else
    sweep_pose_tc2_kernel<false><<<grid, kTcpThreads, sizeof(Tc2Smem), st>>>(d, a, tiles);
  return cudaGetLastError();
}

cudaError_t launch_tc2_order(const int32_t* flagged, int n_splits, int k, int32_t* order, int32_t* pos,
                             cudaStream_t st, int* launches) {
  ++*launches;
  tc2_order_kernel<<<1, 1024, 0, st>>>(flagged, n_splits, k, order, pos);
  return cudaGetLastError();
}

int tc_tile_points() { return kTcTile; }
size_t tc_tiles_bytes(int npad) { return static_cast<size_t>(npad / kTcTile) * kTcTileBytes + 256; }  // + bounds
int tc_hyps_per_cta() { return kTcThreads; }

#ifndef RLB_TC_HALVES
#define RLB_TC_HALVES 1  // 1 (default): 128-thread CTAs, 4 per SM; 2: 256-thread CTAs, 3 per SM (A/B: slower, see above)
#endif
constexpr int kTcHalves = RLB_TC_HALVES;

#ifndef RLB_TC_PIPE
#define RLB_TC_PIPE 0  // 0 (default): sweep_pose_tc_kernel; 1: sweep_pose_tcp_kernel (A/B: no faster, see its header)
#endif

int tc_resident_ctas() {
  static std::atomic<int> cache[64];  // per device; 0 = not asked yet
  std::atomic<int>& cached = cache[PerDeviceOnce::device()];
  if (const int v = cached.load(std::memory_order_relaxed)) return v;
  int n = 0;
#if RLB_TC_PIPE
  cudaFuncSetAttribute(sweep_pose_tcp_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(TcpSmem)));
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sweep_pose_tcp_kernel<true>, kTcpThreads,
                                                    sizeof(TcpSmem)) != cudaSuccess)
    n = 1;
  const int v = max(1, min(n, 512 / kTcpCols));  // TMEM: 512 columns per SM
#else
  cudaFuncSetAttribute(sweep_pose_tc_kernel<true, kTcHalves>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(TcSmem)));
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sweep_pose_tc_kernel<true, kTcHalves>, 128 * kTcHalves,
                                                    sizeof(TcSmem)) != cudaSuccess)
    n = 1;
  const int v = max(1, min(n, 512 / kTcCols));  // TMEM: 512 columns per SM
#endif
  cached.store(v, std::memory_order_relaxed);
  return v;
}

cudaError_t launch_sweep_tc(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                            int* launches) {
  ++*launches;
  static PerDeviceOnce configured;
  const int dev = PerDeviceOnce::device();
  dim3 grid((a.k + 127) / 128, a.n_splits);
#if RLB_TC_PIPE
  if (!configured.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(sweep_pose_tcp_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(TcpSmem)));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(sweep_pose_tcp_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(sizeof(TcpSmem)));
    if (e != cudaSuccess) return e;
    configured.set(dev);
  }
  if (a.want_counts)
    sweep_pose_tcp_kernel<true><<<grid, kTcpThreads, sizeof(TcpSmem), st>>>(d, a, tiles);
  else
    sweep_pose_tcp_kernel<false><<<grid, kTcpThreads, sizeof(TcpSmem), st>>>(d, a, tiles);
#else
  if (!configured.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(sweep_pose_tc_kernel<true, kTcHalves>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(TcSmem)));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(sweep_pose_tc_kernel<false, kTcHalves>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(sizeof(TcSmem)));
    if (e != cudaSuccess) return e;
    configured.set(dev);
  }
  if (a.want_counts)
    sweep_pose_tc_kernel<true, kTcHalves><<<grid, 128 * kTcHalves, sizeof(TcSmem), st>>>(d, a, tiles);
  else
    sweep_pose_tc_kernel<false, kTcHalves><<<grid, 128 * kTcHalves, sizeof(TcSmem), st>>>(d, a, tiles);
#endif
  return cudaGetLastError();
}

}  // namespace rlb
