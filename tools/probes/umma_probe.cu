// Micro-probe: latency / throughput of small tcgen05.mma kind::tf32 (M = 128, K = 8) on sm_100a, as used by
// the tensor-core sweep. Build: nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
// Prints cycles for: n back-to-back MMAs + commit + wait (one issuing thread), TS (A in TMEM) and SS (A in
// shared memory), N = 16 / 32 / 64, independent accumulators vs one accumulator; tcgen05.ld latency.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffffu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3fffu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  return d;
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{.reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;}" ::"r"(d),
               "r"(a), "l"(b), "r"(idesc), "r"(acc)
               : "memory");
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{.reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;}" ::"r"(d),
               "l"(a), "l"(b), "r"(idesc), "r"(acc)
               : "memory");
}
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
}

struct Smem {
  float a[4096];   // 16 KB: A operand for SS
  float b[8192];   // 32 KB: B operand (up to N = 256 x K = 8 ... )
  uint64_t bar;
  uint32_t tmem;
};

// mode bit 0: SS instead of TS; bit 1: all MMAs accumulate into one D instead of rotating over 3
template <int N>
__global__ void probe(long long* out, int n_mma, int mode, int reps) {
  extern __shared__ __align__(128) unsigned char raw[];
  Smem& S = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x;
  for (int i = tid; i < 4096; i += blockDim.x) S.a[i] = 1.0f;
  for (int i = tid; i < 8192; i += blockDim.x) S.b[i] = 1.0f;
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&S.bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
  const uint64_t da = umma_desc(smem_u32(S.a), 2048, 128);
  const uint64_t db = umma_desc(smem_u32(S.b), (N / 8) * 128, 128);
  long long best = 1ll << 60;
  uint32_t parity = 0;
  if (tid == 0) {
    for (int r = 0; r < reps; ++r) {
      const long long t0 = clock64();
      for (int i = 0; i < n_mma; ++i) {
        const uint32_t d = tmem + 64 + ((mode & 2) ? 0 : (i % 3) * N);
        if (mode & 1)
          mma_ss(d, da, db, idesc, i >= 3);
        else
          mma_ts(d, tmem + 8 * (i % 3), db, idesc, i >= 3);
      }
      commit(&S.bar);
      wait(&S.bar, parity);
      parity ^= 1;
      const long long t1 = clock64();
      if (t1 - t0 < best) best = t1 - t0;
    }
    out[blockIdx.x] = best;
  }
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// tcgen05.ld latency: x16 loads back to back, then wait::ld
__global__ void probe_ld(long long* out, int n_ld, int reps) {
  __shared__ uint32_t tm;
  const int tid = threadIdx.x;
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tm)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tm + ((uint32_t)((tid >> 5) * 32) << 16);
  long long best = 1ll << 60;
  uint32_t acc = 0;
  for (int r = 0; r < reps; ++r) {
    const long long t0 = clock64();
    for (int i = 0; i < n_ld; ++i) {
      uint32_t v[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(base + 16 * (i & 15)));
      acc += v[0] ^ v[15];
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const long long t1 = clock64();
    if (t1 - t0 < best) best = t1 - t0;
  }
  if (tid % 32 == 0) out[tid / 32] = best + (acc == 0x12345 ? 1 : 0);
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512u) : "memory");
}


__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
}
template <int KIND>
__device__ __forceinline__ void mma_ss_k(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{.reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{.reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
struct Smem2 {
  float a[4096];
  float b[8192];
  uint64_t bar[4];
  uint32_t tmem;
};
// `issuers` warps each issue n_mma MMAs (straight-line groups of 6, elected lane), own barrier, own D columns.
template <int N, int M, int KIND, int TS>
__global__ void probe2(long long* out, int n_groups, int issuers, int reps) {
  extern __shared__ __align__(128) unsigned char raw[];
  Smem2& S = *reinterpret_cast<Smem2*>(raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 4096; i += blockDim.x) S.a[i] = 0.0f;
  for (int i = tid; i < 8192; i += blockDim.x) S.b[i] = 0.0f;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&S.bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = S.tmem;
  // KIND 0: tf32 (a_format = b_format = 2), KIND 1: bf16 (1)
  const uint32_t fmt = KIND == 0 ? 2u : 1u;
  const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
  const uint64_t da = umma_desc(smem_u32(S.a), 2048, 128);
  const uint64_t db = umma_desc(smem_u32(S.b), (N / 8) * 128, 128);
  long long best = 1ll << 60, best_issue = 1ll << 60;
  uint32_t parity = 0;
  if (warp < issuers) {
    const uint32_t d0 = tmem + warp * 128;
    for (int r = 0; r < reps; ++r) {
      __syncwarp();
      const long long t0 = clock64();
      if (elect_one()) {
        for (int g = 0; g < n_groups; ++g) {
          if (TS) {
            mma_ts(d0, tmem + 480, db, idesc, 0);
            mma_ts(d0 + N % 128, tmem + 488, db, idesc, 0);
            mma_ts(d0, tmem + 496, db, idesc, 1);
            mma_ts(d0 + N % 128, tmem + 480, db, idesc, 1);
            mma_ts(d0, tmem + 488, db, idesc, 1);
            mma_ts(d0 + N % 128, tmem + 496, db, idesc, 1);
          } else {
            mma_ss_k<KIND>(d0, da, db, idesc, 0);
            mma_ss_k<KIND>(d0 + N % 128, da, db, idesc, 0);
            mma_ss_k<KIND>(d0, da, db, idesc, 1);
            mma_ss_k<KIND>(d0 + N % 128, da, db, idesc, 1);
            mma_ss_k<KIND>(d0, da, db, idesc, 1);
            mma_ss_k<KIND>(d0 + N % 128, da, db, idesc, 1);
          }
        }
        commit(&S.bar[warp]);
      }
      __syncwarp();
      const long long t1 = clock64();
      wait(&S.bar[warp], parity);
      parity ^= 1;
      const long long t2 = clock64();
      if (t2 - t0 < best) best = t2 - t0;
      if (t1 - t0 < best_issue) best_issue = t1 - t0;
    }
    if ((tid & 31) == 0) {
      out[2 * warp] = best;
      out[2 * warp + 1] = best_issue;
    }
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}
template <int N, int M, int KIND, int TS>
void run2(long long* d_out) {
  cudaFuncSetAttribute(probe2<N, M, KIND, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem2));
  for (int issuers = 1; issuers <= 2; ++issuers) {
    printf("probe2 %s %s M=%d N=%d issuers=%d (total/issue cycles):", KIND ? "bf16" : "tf32", TS ? "TS" : "SS", M, N, issuers);
    for (int g : {1, 2, 4, 8}) {
      probe2<N, M, KIND, TS><<<1, 128, sizeof(Smem2)>>>(d_out, g, issuers, 20);
      long long h[4];
      cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
      printf("  n%d=%lld/%lld", 6 * g, h[0], h[1]);
      if (issuers == 2) printf("|%lld/%lld", h[2], h[3]);
    }
    printf("  [%s]\n", cudaGetErrorString(cudaDeviceSynchronize()));
  }
}

template <int N>
void run(long long* d_out, const char* name) {
  cudaFuncSetAttribute(probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
  for (int mode = 0; mode < 4; ++mode) {
    printf("%s N=%d %s %s:", name, N, (mode & 1) ? "SS" : "TS", (mode & 2) ? "oneD" : "3D");
    for (int n : {0, 1, 2, 3, 6, 12, 24, 48}) {
      probe<N><<<1, 128, sizeof(Smem)>>>(d_out, n, mode, 20);
      long long h = 0;
      cudaMemcpy(&h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
      printf(" n%d=%lld", n, h);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("  [%s]\n", cudaGetErrorString(e));
  }
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 64 * sizeof(long long));
  run2<16, 128, 0, 0>(d_out);
  run2<16, 128, 0, 1>(d_out);
  run2<32, 128, 0, 0>(d_out);
  run2<32, 128, 0, 1>(d_out);
  run2<64, 128, 0, 0>(d_out);
  run2<64, 128, 0, 1>(d_out);
  run2<128, 128, 0, 0>(d_out);
  run2<128, 128, 0, 1>(d_out);
  for (int n : {1, 2, 3, 6, 12}) {
    probe_ld<<<1, 128>>>(d_out, n, 20);
    long long h[4];
    cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("ld x16: n=%d cycles=%lld %lld %lld %lld\n", n, h[0], h[1], h[2], h[3]);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
