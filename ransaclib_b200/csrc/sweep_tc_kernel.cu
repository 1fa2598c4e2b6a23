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
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
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
__global__ void tc_pack_kernel(const DeviceData d, float* __restrict__ tiles) {
  const int total = d.npad;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    const float x = d.soa32[0 * static_cast<size_t>(total) + g];
    const float y = d.soa32[1 * static_cast<size_t>(total) + g];
    const float z = d.soa32[2 * static_cast<size_t>(total) + g];
    const float nu = d.soa32[3 * static_cast<size_t>(total) + g];
    const float nv = d.soa32[4 * static_cast<size_t>(total) + g];
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
}

cudaError_t launch_tc_pack(const DeviceData& d, float* tiles, cudaStream_t st, int* launches) {
  ++*launches;
  const int blocks = max(1, min(2048, (d.npad + 255) / 256));
  tc_pack_kernel<<<blocks, 256, 0, st>>>(d, tiles);
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
template <bool COUNT>
__global__ void __launch_bounds__(kTcThreads, 4)
    sweep_pose_tc_kernel(const DeviceData d, const SweepArgs a, const float* __restrict__ tiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcSmem& S = *reinterpret_cast<TcSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int k = blockIdx.x * kTcThreads + tid;
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
  const uint32_t lane_addr = tmem + (static_cast<uint32_t>(warp * 32) << 16);
  {
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

    const float* nu = S.ring[slot] + 512;
    const float* nv = nu + 32;
    // Quarter-tiles of 8 points: the TMEM loads of the next quarter are in flight while this one is
    // reduced (tcgen05.ld latency is ~110 cycles; a tile-wide load-then-wait exposed it twice per tile).
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
      if (qd < 3) tmem_wait_ld8x3(px[nxt], py[nxt], pz[nxt]);
    }
    live_in_chunk += max(0, min(kTcTile, n_live - (t_begin + it) * kTcTile));
    if (((it + 1) % kTcFlush) == 0 || it + 1 == ntiles) {
      if (COUNT) {
        dacc += w * (static_cast<double>(s) + thr64 * static_cast<double>(live_in_chunk - c));
        dcnt += c;
      } else {
        // every processed point added min(e, thr); padding (NaN residual) added thr and is taken out again
        float sa, sb;
        unpack2(s2, sa, sb);
        const int processed = (it % kTcFlush + 1) * kTcTile;
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

  if (k < a.k) {
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

int tc_tile_points() { return kTcTile; }
size_t tc_tiles_bytes(int npad) { return static_cast<size_t>(npad / kTcTile) * kTcTileBytes; }
int tc_hyps_per_cta() { return kTcThreads; }

int tc_resident_ctas() {
  static std::atomic<int> cache[64];  // per device; 0 = not asked yet
  std::atomic<int>& cached = cache[PerDeviceOnce::device()];
  if (const int v = cached.load(std::memory_order_relaxed)) return v;
  cudaFuncSetAttribute(sweep_pose_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(sizeof(TcSmem)));
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sweep_pose_tc_kernel<true>, kTcThreads, sizeof(TcSmem)) !=
      cudaSuccess)
    n = 1;
  const int v = max(1, min(n, 512 / kTcCols));  // TMEM: 512 columns per SM
  cached.store(v, std::memory_order_relaxed);
  return v;
}

cudaError_t launch_sweep_tc(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                            int* launches) {
  ++*launches;
  static PerDeviceOnce configured;
  const int dev = PerDeviceOnce::device();
  if (!configured.done(dev)) {
    cudaError_t e = cudaFuncSetAttribute(sweep_pose_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(sizeof(TcSmem)));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(sweep_pose_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(sizeof(TcSmem)));
    if (e != cudaSuccess) return e;
    configured.set(dev);
  }
  dim3 grid((a.k + kTcThreads - 1) / kTcThreads, a.n_splits);
  if (a.want_counts)
    sweep_pose_tc_kernel<true><<<grid, kTcThreads, sizeof(TcSmem), st>>>(d, a, tiles);
  else
    sweep_pose_tc_kernel<false><<<grid, kTcThreads, sizeof(TcSmem), st>>>(d, a, tiles);
  return cudaGetLastError();
}

}  // namespace rlb
