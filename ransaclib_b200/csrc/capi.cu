// C ABI of ransaclib_b200 (see include/ransaclib_b200.h). Host-side plumbing only:
// device memory, the stream, scratch buffers and launch sequencing. There is no CPU
// implementation of any entry point: if CUDA is unavailable every call fails.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <vector>

#include "rlb_internal.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define RLB_CUDA(call)                                                                      \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess)                                                                 \
      return fail(RLB_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),      \
                  __FILE__, __LINE__);                                                      \
  } while (0)

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max<size_t>(bytes, 256);
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T>
  T* as() const {
    return static_cast<T*>(p);
  }
};

}  // namespace

struct rlb_ctx {
  int device = 0;
  // Scorers keep pointers into their context (stream, pool). A context destroyed while some are alive - a
  // garbage collector finalising objects in no particular order does that - only marks itself; the last
  // rlb_scorer_destroy then completes the destruction.
  int live_scorers = 0;
  bool destroy_pending = false;
  int sms = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  int launches_i = 0;  // per-call counter target for the launch_* helpers
  int last_variant = -1, last_splits = 0;  // plan of the last sweep (rlb_ctx_last_sweep_plan)
  std::map<int, int> occ_cache;            // sweep planner: resident CTAs per SM of each kernel variant on THIS device
  int64_t launches = 0;
  unsigned int epoch = 0;
  // sweep scratch
  DevBuf models, valid, samples, scores, counts, part_score, part_count, block_best, small;
  // filtered f64 sweep: {dense ids [k], sparse ids [k], pos [k], counts [4]} and the dense list's accumulators
  DevBuf tcx_lists, tcx_acc;
  // inlier scratch
  DevBuf one_model, idx, tile_score, status, header;
  // pinned host staging for tiny results
  void* h_small = nullptr;
  // pinned staging of one hypothesis batch (samples up, models / valid / scores / counts back): the callers'
  // buffers are pageable, and every pageable cudaMemcpyAsync is a synchronous staged copy of its own
  void* h_stage = nullptr;
  size_t stage_cap = 0;
  // pinned copy of an inlier list too long for the zero-copy path (rlb_get_inliers_view)
  void* h_list = nullptr;
  size_t list_cap = 0;
  // optional per-kernel timing (bench.py's roofline numbers)
  bool timing = false;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // sweep begin/end, inlier begin/end
  bool ev_valid[2] = {false, false};
  // Data-set memory pool: scorers are created and destroyed per query (batched localization, the
  // e2e path), and cudaMalloc/cudaFree cost milliseconds and synchronise the device. Freed blocks
  // are kept (up to kPoolLimit bytes) and reused by the next scorer of a similar size.
  std::multimap<size_t, void*> pool_free_blocks;
  std::map<void*, size_t> pool_sizes;
  size_t pool_idle_bytes = 0;
  // zero-copy result buffer of the K=1 pass for small data sets: {InlierHeader; int32 idx[]} in
  // mapped pinned host memory, written by the kernel itself (one launch + one sync per pass)
  void* h_result = nullptr;
  void* d_result = nullptr;  // device alias of h_result
  size_t result_cap = 0;
  // device-resident batched localization: per-slot scratch and the ticket words of its launches
  DevBuf loc_soa, loc_lists, loc_words;
  // device-resident line LO-MSAC: per-CTA lists, {ticket, line, result} words + inlier list of the
  // single-problem entry point
  DevBuf ll_lists, ll_io;
  // host-pointer batched localization: one copy stream, and per uploaded chunk one launch stream
  // + "chunk uploaded" / "chunk done" events (the kernels are ordered behind their copies by
  // events; they never wait on the host)
  cudaStream_t copy_stream = nullptr;
  cudaStream_t loc_streams[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t loc_ev[9] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // set by rlb_sharded_round_dev for the next score_dev call
  double* pending_record = nullptr;
  long long pending_offset = 0;
};

namespace rlb {
// for the other translation units of the library (comm.cu)
int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
void* ctx_stream(const rlb_ctx* c) { return c->stream; }
int ctx_device(const rlb_ctx* c) { return c->device; }
}  // namespace rlb

struct rlb_scorer {
  rlb_ctx* ctx = nullptr;
  rlb::DeviceData d;
  void* raw0 = nullptr;
  void* raw1 = nullptr;
  void* rays = nullptr;
  void* soa64 = nullptr;
  void* soa32 = nullptr;
  void* tc_tiles = nullptr;  // tensor-core operand tiles of the points, packed on first use
  void* tc2_tiles = nullptr; // the operand image of the sorted two-speed sweep (variant 12), packed on first use
  rlb::CustomModule* custom = nullptr;  // RLB_KIND_CUSTOM: the NVRTC-compiled functor kernels
  void* custom_err = nullptr;           // RLB_KIND_CUSTOM: [npad] per-point errors of the K = 1 passes
};

namespace {

using rlb::BestRec;
using rlb::DeviceData;
using rlb::InlierArgs;
using rlb::InlierHeader;
using rlb::SweepArgs;

// layout of ctx->small: [0] BestRec best, [64] uint ticket, [128] uint control[2], [192] best model (12 d)
constexpr size_t kSmallBest = 0, kSmallTicket = 64, kSmallControl = 128, kSmallBestModel = 192,
                 kSmallBytes = 512;

inline int pad_to(int n, int a) { return (n + a - 1) / a * a; }

void account(rlb_ctx* c) {
  c->launches += c->launches_i;
  c->launches_i = 0;
}

int ctx_use(rlb_ctx* c) {
  RLB_CUDA(cudaSetDevice(c->device));
  return RLB_OK;
}

// Sub-sampled mean as the recentring point (any point inside the cloud works; it only
// has to make |X - centre| small so f32 keeps its 24 bits for the part that matters).
void subsampled_mean(const double* p, int n, int stride, int dims, double* out) {
  for (int c = 0; c < 3; ++c) out[c] = 0.0;
  if (n <= 0) return;
  const int step = std::max(1, n / 4096);
  int m = 0;
  for (int i = 0; i < n; i += step, ++m)
    for (int c = 0; c < dims; ++c) out[c] += p[static_cast<size_t>(i) * stride + c];
  for (int c = 0; c < dims; ++c) {
    out[c] /= static_cast<double>(m);
    if (!(out[c] - out[c] == 0.0)) out[c] = 0.0;
  }
}

constexpr size_t kPoolLimit = size_t(2) << 30;

int pool_alloc(rlb_ctx* c, size_t bytes, void** out) {
  bytes = std::max<size_t>((bytes + 511) / 512 * 512, 512);
  auto it = c->pool_free_blocks.lower_bound(bytes);
  if (it != c->pool_free_blocks.end() && it->first <= bytes * 2) {
    *out = it->second;
    c->pool_idle_bytes -= it->first;
    c->pool_free_blocks.erase(it);
    return RLB_OK;
  }
  RLB_CUDA(cudaMalloc(out, bytes));
  c->pool_sizes[*out] = bytes;
  return RLB_OK;
}

void pool_release(rlb_ctx* c, void* p) {
  if (!p) return;
  auto it = c->pool_sizes.find(p);
  if (it == c->pool_sizes.end()) {
    cudaFree(p);
    return;
  }
  if (c->pool_idle_bytes + it->second > kPoolLimit) {
    cudaFree(p);
    c->pool_sizes.erase(it);
    return;
  }
  c->pool_free_blocks.emplace(it->second, p);
  c->pool_idle_bytes += it->second;
}

// Pinned staging block of the context, at least `bytes` long.
int ensure_stage(rlb_ctx* c, size_t bytes) {
  if (bytes <= c->stage_cap) return RLB_OK;
  if (c->h_stage) cudaFreeHost(c->h_stage);
  c->h_stage = nullptr;
  c->stage_cap = 0;
  const size_t cap = std::max<size_t>(bytes, 256 * 1024);
  RLB_CUDA(cudaMallocHost(&c->h_stage, cap));
  c->stage_cap = cap;
  return RLB_OK;
}

void pool_drain(rlb_ctx* c) {
  for (auto& kv : c->pool_free_blocks) cudaFree(kv.second);
  c->pool_free_blocks.clear();
  c->pool_sizes.clear();
  c->pool_idle_bytes = 0;
}

int finish_scorer(rlb_scorer* s) {
  rlb_ctx* c = s->ctx;
  DeviceData& d = s->d;
  const size_t ch = static_cast<size_t>(d.channels) * d.npad;
  {
    int rc = pool_alloc(c, ch * sizeof(double), &s->soa64);
    if (!rc) rc = pool_alloc(c, ch * sizeof(float), &s->soa32);
    if (rc) return rc;
  }
  d.soa64 = static_cast<const double*>(s->soa64);
  d.soa32 = static_cast<const float*>(s->soa32);
  RLB_CUDA(rlb::launch_repack(d, static_cast<double*>(s->soa64), static_cast<float*>(s->soa32),
                              c->stream, &c->launches_i));
  if (d.kind == RLB_KIND_POSE) {
    {
      const int rc = pool_alloc(c, static_cast<size_t>(std::max(d.n[0], 1)) * 3 * sizeof(double), &s->rays);
      if (rc) return rc;
    }
    d.rays = static_cast<const double*>(s->rays);
    RLB_CUDA(rlb::launch_rays(d, static_cast<double*>(s->rays), c->stream, &c->launches_i));
  }
  account(c);
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  return RLB_OK;
}

int upload(rlb_ctx* c, void** dst, const double* src, size_t count, cudaStream_t st) {
  {
    const int rc = pool_alloc(c, std::max<size_t>(count, 1) * sizeof(double), dst);
    if (rc) return rc;
  }
  if (count) RLB_CUDA(cudaMemcpyAsync(*dst, src, count * sizeof(double), cudaMemcpyHostToDevice, st));
  return RLB_OK;
}

int check_scorer(const rlb_scorer* s) {
  if (!s || !s->ctx) return fail(RLB_E_ARG, "null scorer");
  return RLB_OK;
}

void fill_thr(const DeviceData& d, const double* thr, const double* w, double* thr_out,
              double* w_out) {
  for (int t = 0; t < rlb::kMaxTypes; ++t) {
    thr_out[t] = thr ? thr[std::min(t, d.num_types - 1)] : 0.0;
    w_out[t] = w ? w[std::min(t, d.num_types - 1)] : 1.0;
  }
}

int ensure_first_use(rlb_ctx* c) {
  if (c->small.cap == 0) {
    RLB_CUDA(c->small.ensure(kSmallBytes));
    RLB_CUDA(cudaMemsetAsync(c->small.p, 0, kSmallBytes, c->stream));
  }
  return RLB_OK;
}

// Enqueue sweep + finalize for K models that are already on the device.
int score_dev(rlb_scorer* s, int precision, int k, const double* d_models, const uint8_t* d_valid,
              const double* thr, const double* w, double* d_scores, int32_t* d_counts,
              BestRec* d_best, double* d_best_model) {
  rlb_ctx* c = s->ctx;
  const DeviceData& d = s->d;
  if (precision != RLB_F64 && precision != RLB_F32 && precision != RLB_F64_PLAIN && precision != RLB_F64_FILTERED &&
      precision != RLB_F32_SORTED && precision != RLB_F32_UNSORTED)
    return fail(RLB_E_ARG, "bad precision %d", precision);
  const bool plain64 = (precision == RLB_F64_PLAIN);        // f64 without the f32 pre-classification
  const bool filtered64 = (precision == RLB_F64_FILTERED);  // ... with it, whatever the size of the batch
  if (plain64 || filtered64) precision = RLB_F64;
  const bool sorted32 = (precision == RLB_F32_SORTED);      // the sorted two-speed f32 sweep whatever the size
  const bool unsorted32 = (precision == RLB_F32_UNSORTED);  // f32 without it
  if (sorted32 || unsorted32) precision = RLB_F32;
  if (k <= 0) return fail(RLB_E_ARG, "k must be positive (got %d)", k);
  if (!thr) return fail(RLB_E_ARG, "sq_thr is required");
  {
    const int rc0 = ensure_first_use(c);
    if (rc0) return rc0;
  }

  SweepArgs a = {};
  a.k = k;
  a.models = d_models;
  a.valid = d_valid;
  fill_thr(d, thr, w, a.thr, a.w);
  // Grid plan. Every CTA of a sweep does the same amount of work, so CTAs run in waves of
  // (resident CTAs per SM) x (SMs); the plan picks the register tile H (hypotheses per thread) and
  // the number of point-axis splits that fill whole waves: score = fill of the last wave x balance
  // of the stages over the splits x relative speed of the tile. (K = 65,536 on one GPU: 62 instead
  // of 64 splits; K = 8192 per GPU at 8 GPUs: 55 splits = 440 CTAs for 444 slots instead of 512.)
  static const int forced_variant = [] {
    const char* v = getenv("RLB_SWEEP_VARIANT");
    return v ? atoi(v) : -1;
  }();
  static const int forced_splits = [] {
    const char* v = getenv("RLB_SWEEP_SPLITS");
    return v ? atoi(v) : 0;
  }();
  a.record = c->pending_record;
  a.id_offset = c->pending_offset;
  a.want_counts = (d_counts != nullptr) ? 1 : 0;
  c->pending_record = nullptr;
  c->pending_offset = 0;
  const int total_stages = d.npad / rlb::kStage;
  struct Cand {
    int variant;
    double speed;
  };
  Cand cands[4];
  int ncand = 0;
  if (forced_variant >= 0) {
    cands[ncand++] = {forced_variant, 1.0};
  } else if (d.kind == RLB_KIND_POSE && precision == RLB_F32) {
    cands[ncand++] = {3, 1.00};  // H = 8: 1.90e12 evals/s on C5
    cands[ncand++] = {0, 0.92};  // H = 4
    cands[ncand++] = {2, 0.80};  // H = 2
  } else if (d.kind == RLB_KIND_POSE) {
    cands[ncand++] = {7, 1.00};  // f64, H = 4
    cands[ncand++] = {-1, 0.96};  // f64, H = 2
    cands[ncand++] = {8, 0.85};  // f64, H = 1: batches of up to ~100 (the LO's deferred scores, small rounds)
  } else {
    cands[ncand++] = {-1, 1.0};  // H = 4
    // H = 1: a batch of up to ~400 hypotheses (the drivers' rounds) leaves 3 of 4 register tiles of the H = 4
    // kernel evaluating nothing; tile_use below picks this one for those.
    if (d.kind != RLB_KIND_CUSTOM) cands[ncand++] = {10, 0.8};
  }
  double best_score = -1.0;
  int splits = 1, sps = total_stages;
  a.variant = cands[0].variant;
  for (int ci = 0; ci < ncand; ++ci) {
    const int H = rlb::sweep_hyps_per_thread(d.kind, precision, cands[ci].variant);
    // (kind, precision, variant) -> resident CTAs per SM; cached per context, i.e. per device and per host thread
    const int occ_key = (d.kind * 4 + precision) * 64 + (cands[ci].variant + 1);
    auto occ_it = c->occ_cache.find(occ_key);
    if (occ_it == c->occ_cache.end())
      occ_it = c->occ_cache.emplace(occ_key, rlb::sweep_resident_ctas(d.kind, precision, cands[ci].variant)).first;
    const int slots = c->sms * occ_it->second;
    const int k_tiles = (k + rlb::kSweepThreads * H - 1) / (rlb::kSweepThreads * H);
    // partials are [splits][k]. 64 splits fill the device once a batch has a few hypothesis tiles; a small
    // batch over many points (the hybrid drivers' rounds: ~100 hypotheses x 200k points) would leave it on 64
    // CTAs whose threads each walk 3000 points - there the point axis is cut finer (220 -> 75 us per round).
    static const int fine_max = [] {
      const char* v = getenv("RLB_SWEEP_FINE_MAX");  // A/B timing
      return v ? atoi(v) : 256;  // 512 (one stage per CTA) measured 3 % slower on configs[2]
    }();
    const int fine = std::min(fine_max, slots / std::max(k_tiles, 1));
    const int max_s = std::max(1, std::min(total_stages, std::max(64, fine)));
    for (int s0 = 1; s0 <= max_s; ++s0) {
      if (forced_splits > 0 && s0 != std::min(forced_splits, max_s)) continue;
      const int sp = (total_stages + s0 - 1) / s0;
      const int s_eff = (total_stages + sp - 1) / sp;
      const long long ctas = static_cast<long long>(k_tiles) * s_eff;
      const long long waves = (ctas + slots - 1) / slots;
      const double fill = static_cast<double>(ctas) / static_cast<double>(waves * slots);
      const double balance = static_cast<double>(total_stages) / (static_cast<double>(s_eff) * sp);
      // lanes of the last hypothesis tile that carry no hypothesis
      const double tile_use = static_cast<double>(k) / (static_cast<double>(k_tiles) * rlb::kSweepThreads * H);
      // a few more, shorter CTAs are preferred on ties (shorter tail)
      const double sc = fill * balance * tile_use * cands[ci].speed * (1.0 + 1e-4 * s_eff);
      if (sc > best_score) {
        best_score = sc;
        a.variant = cands[ci].variant;
        splits = s_eff;
        sps = sp;
      }
    }
  }
  a.n_splits = splits;
  a.stages_per_split = sps;
  // Tensor-core sweep (pose, f32 arithmetic class): the projection runs as tcgen05.mma on TF32x2-split
  // operands. Its unit along the point axis is a 32-point tile. Measured 2.40e12 evals/s with counts and
  // 2.51e12 without on C5 against 1.90e12 for the FFMA2 sweep (profiles/r01r_*), so it is planned with the
  // same fill / balance score and wins unless the batch is too small to pay for packing the tiles.
  //
  // Filtered f64 sweep (pose, f64 results; variant 11): the same tensor-core front end classifies every pair in
  // f32 and only the pairs that could be inliers are evaluated in f64 (sweep_tc_kernel.cu). Scores within 1e-12
  // of the plain f64 sweep, counts identical; 6.4e11 -> 1.5e12 evals/s on C5. RLB_F64_FILTER=0 keeps the plain
  // f64 sweep (A/B, and batches of nothing but good hypotheses, where every second pair needs f64 anyway).
  static const bool f64_filter = [] {
    const char* v = getenv("RLB_F64_FILTER");
    return !(v && atoi(v) == 0);
  }();
  bool use_tc = false;
  const bool tc_f32 = precision == RLB_F32 && (forced_variant == 9 || (forced_variant < 0 && k >= 512));
  // (four launches instead of one: only where the plain sweep would take a few hundred microseconds or more -
  // configs[1]'s batches of <= 8192 hypotheses x 10k points stay with it)
  const bool tc_f64 = precision == RLB_F64 && !plain64 && a.thr[0] > 0.0 &&
                      (forced_variant == 11 || (filtered64 && d.kind == RLB_KIND_POSE) ||
                       (forced_variant < 0 && f64_filter && k >= 512 &&
                        static_cast<long long>(k) * d.n[0] >= (1LL << 28)));
  if (d.kind == RLB_KIND_POSE && (tc_f32 || tc_f64)) {
    const int total_tiles = d.npad / rlb::tc_tile_points();
    const int slots = c->sms * (tc_f64 ? rlb::tcx_resident_ctas() : rlb::tc_resident_ctas());
    const int hpc = rlb::tc_hyps_per_cta();
    const int k_tiles = (k + hpc - 1) / hpc;
    const double tile_use = static_cast<double>(k) / (static_cast<double>(k_tiles) * hpc);
    const double speed = tc_f64 ? 2.3 : ((d_counts != nullptr) ? 1.26 : 1.32);
    double best = -1.0;
    int tc_splits = 1, tc_tps = total_tiles;
    // Up to 64 splits, more (up to 256) while the grid would otherwise stay below ~8 waves of resident CTAs: a
    // hypothesis shard of one of eight GPUs (K / 8 = 8192 = 64 tiles of 128) ran on 37 splits = 4 waves, whose
    // last wave's imbalance cost 4 % of the sweep rate (SCALE_r01: 2.31e12 against 2.41e12 evals/s per GPU).
    const int max_tc_splits = std::min(total_tiles, std::max(64, std::min(256, (8 * slots + k_tiles - 1) / k_tiles)));
    for (int s0 = 1; s0 <= max_tc_splits; ++s0) {
      if (forced_splits > 0 && s0 != std::min(forced_splits, max_tc_splits)) continue;
      const int tp = (total_tiles + s0 - 1) / s0;
      const int s_eff = (total_tiles + tp - 1) / tp;
      const long long ctas = static_cast<long long>(k_tiles) * s_eff;
      const long long waves = (ctas + slots - 1) / slots;
      const double fill = static_cast<double>(ctas) / static_cast<double>(waves * slots);
      const double balance = static_cast<double>(total_tiles) / (static_cast<double>(s_eff) * tp);
      const double sc = fill * balance * tile_use * speed * (1.0 + 1e-4 * s_eff);
      if (sc > best) {
        best = sc;
        tc_splits = s_eff;
        tc_tps = tp;
      }
    }
    if (forced_variant == 9 || forced_variant == 11 || filtered64 || best > best_score) {
      use_tc = true;
      a.variant = tc_f64 ? 11 : 9;
      a.n_splits = splits = tc_splits;
      a.stages_per_split = tc_tps;
    }
  }
  // The sorted two-speed form (prepass + order + sweep). Opt-in: RLB_TC_SORTED=1 picks it for large f32 sweeps
  // (K >= 2048, K x N >= 2^30), the precision code RLB_F32_SORTED for any. On C5 it is SLOWER than variant 9 (30.0
  // against 27.1 ms, profiles/r02_tc_sorted_ab.txt): nine dependent-latency MMAs per tile instead of six cap the
  // pipeline at 24.9 ms even with no arithmetic at all, and at a 12 px threshold a third of the 8-point groups of
  // a warp of random hypotheses still holds a flagged pair.
  static const bool tc_sorted = [] {
    const char* v = getenv("RLB_TC_SORTED");
    return v && atoi(v) == 1;
  }();
  if (d.kind == RLB_KIND_POSE && precision == RLB_F32 && a.thr[0] > 0.0 &&
      !unsorted32 &&
      (forced_variant == 12 || sorted32 ||
       (forced_variant < 0 && use_tc && a.variant == 9 && tc_sorted && k >= 2048 &&
        static_cast<long long>(k) * d.n[0] >= (1LL << 30)))) {
    const int total_tiles = d.npad / rlb::tc_tile_points();
    const int slots = c->sms * rlb::tc2_resident_ctas();
    const int k_tiles = (k + 127) / 128;
    // as many splits as give ~8 waves of resident CTAs (the CTAs of the flagged hypotheses run ~1.7x longer)
    const int want = std::max(1, std::min(std::min(256, total_tiles), (8 * slots + k_tiles - 1) / k_tiles));
    const int tp = (total_tiles + want - 1) / want;
    use_tc = true;
    a.variant = 12;
    a.n_splits = splits = (total_tiles + tp - 1) / tp;
    a.stages_per_split = tp;
  }
  if (use_tc && a.variant == 12) {
    if (!s->tc2_tiles) {
      const int rc = pool_alloc(c, rlb::tc2_tiles_bytes(d.npad), &s->tc2_tiles);
      if (rc) return rc;
      RLB_CUDA(rlb::launch_tc2_pack(d, static_cast<float*>(s->tc2_tiles), c->stream, &c->launches_i));
    }
  } else if (use_tc && !s->tc_tiles) {
    const int rc = pool_alloc(c, rlb::tc_tiles_bytes(d.npad), &s->tc_tiles);
    if (rc) return rc;
    RLB_CUDA(rlb::launch_tc_pack(d, static_cast<float*>(s->tc_tiles), c->stream, &c->launches_i));
  }
  int custom_pps = 128;
  if (d.kind == RLB_KIND_CUSTOM) {
    // user functor (custom_functor.cu): 128 hypotheses per CTA, whole 128-point chunks per split, ~8 CTAs per SM
    const int k_tiles = (k + 127) / 128;
    // extent of the padded point axis that holds live points (two types: the second segment starts at off[1])
    const int extent = (d.num_types > 1 && d.n[1] > 0) ? d.off[1] + d.n[1] : d.n[0];
    const int chunks = (extent + 127) / 128;
    if (chunks > 0) {
      // (up to 256 splits: the drivers' rounds of a few hundred hypotheses over 10^4 - 10^5 points otherwise run on
      // two or three warps per SM, where dependent FP64 latency is all there is)
      const int want = std::max(1, std::min(std::min(256, chunks), (c->sms * 8 + k_tiles - 1) / k_tiles));
      custom_pps = ((chunks + want - 1) / want) * 128;
      splits = (extent + custom_pps - 1) / custom_pps;
    } else {
      splits = 1;
    }
    a.n_splits = splits;
    a.stages_per_split = custom_pps;
    a.variant = -2;
  }
  c->last_variant = a.variant;
  c->last_splits = splits;
  RLB_CUDA(c->part_score.ensure(static_cast<size_t>(splits) * k * sizeof(double)));
  RLB_CUDA(c->part_count.ensure(static_cast<size_t>(splits) * k * sizeof(int32_t)));
  a.part_score = c->part_score.as<double>();
  a.part_count = c->part_count.as<int32_t>();
  const int fblocks = (k + 255) / 256;
  RLB_CUDA(c->block_best.ensure(static_cast<size_t>(fblocks) * sizeof(BestRec)));

  if (c->timing) RLB_CUDA(cudaEventRecord(c->ev[0], c->stream));
  if (d.kind == RLB_KIND_CUSTOM) {
    std::string err;
    if (!rlb::launch_custom_sweep(s->custom, d, a, custom_pps, c->stream, &c->launches_i, &err))
      return fail(RLB_E_CUDA, "custom sweep: %s", err.c_str());
  } else if (use_tc && a.variant == 11) {
    // Filtered f64 sweep: prepass over a strided sample of the tiles -> dense / sparse hypothesis lists -> the
    // plain f64 sweep over the dense list, the tensor-core front end + f64 back end over the sparse one. All on
    // this stream; the list lengths stay on the device (grids are sized for k, CTAs beyond a list return at once).
    const float* tiles = static_cast<const float*>(s->tc_tiles);
    const int tile_pts = rlb::tc_tile_points();
    const int total_tiles = d.npad / tile_pts;
    const int hpc = rlb::tc_hyps_per_cta();
    const int k_tiles = (k + hpc - 1) / hpc;
    const int slots = c->sms * rlb::tcx_resident_ctas();
    RLB_CUDA(c->tcx_lists.ensure((static_cast<size_t>(k) * 3 + 4) * sizeof(int32_t)));
    RLB_CUDA(c->tcx_acc.ensure(static_cast<size_t>(k) * 12));
    int32_t* lists = c->tcx_lists.as<int32_t>();
    int32_t* pos = lists + 2 * static_cast<size_t>(k);
    int32_t* list_counts = pos + k;
    unsigned long long* acc_q = c->tcx_acc.as<unsigned long long>();
    unsigned int* acc_c = reinterpret_cast<unsigned int*>(acc_q + k);
    // 1. prepass: every 64th tile (at least 8), cut into as many splits as fill the device
    const int n_sample = std::min(total_tiles, std::max(8, total_tiles / 64));
    const int stride = std::max(1, total_tiles / n_sample);
    SweepArgs pa = a;
    pa.tile_stride = stride;
    pa.n_splits = std::max(1, std::min(n_sample, (slots + k_tiles - 1) / k_tiles));
    pa.stages_per_split = (n_sample + pa.n_splits - 1) / pa.n_splits;
    pa.n_splits = (n_sample + pa.stages_per_split - 1) / pa.stages_per_split;
    RLB_CUDA(c->part_count.ensure(static_cast<size_t>(std::max(splits, pa.n_splits)) * k * sizeof(int32_t)));
    pa.part_count = c->part_count.as<int32_t>();
    a.part_count = pa.part_count;
    RLB_CUDA(rlb::launch_sweep_tcx_prepass(d, pa, tiles, c->stream, &c->launches_i));
    // 2. partition
    const long long sampled = std::min<long long>(d.n[0], static_cast<long long>(n_sample) * tile_pts);
    RLB_CUDA(rlb::launch_tcx_partition(pa.part_count, pa.n_splits, k, static_cast<int>(sampled), lists, pos,
                                       list_counts, c->stream, &c->launches_i));
    RLB_CUDA(cudaMemsetAsync(acc_q, 0, static_cast<size_t>(k) * 12, c->stream));
    // 3. dense list: plain f64 sweep (H = 4), up to 256 splits so that a short list still fills the device
    {
      SweepArgs da = a;
      da.ids = lists;
      da.k_dev = list_counts;
      da.acc_q = acc_q;
      da.acc_c = acc_c;
      da.variant = 7;
      da.n_splits = std::max(1, std::min(total_stages, 256));
      da.stages_per_split = (total_stages + da.n_splits - 1) / da.n_splits;
      da.n_splits = (total_stages + da.stages_per_split - 1) / da.stages_per_split;
      // scores are below thr * w * n: the fixed-point sum of all splits stays below 2^62
      int e2 = 0;
      frexp(a.thr[0] * std::max(a.w[0], 1.0) * static_cast<double>(std::max(d.n[0], 1)), &e2);
      da.acc_shift = 61 - e2;
      a.acc_shift = da.acc_shift;
      RLB_CUDA(rlb::launch_sweep(d, da, RLB_F64, c->stream, &c->launches_i));
    }
    // 4. sparse list
    a.ids = lists + k;
    a.k_dev = list_counts + 1;
    RLB_CUDA(rlb::launch_sweep_tcx(d, a, tiles, c->stream, &c->launches_i));
    a.ids = nullptr;  // finalize walks all k hypotheses
    a.k_dev = nullptr;
    a.pos = pos;
    a.acc_q = acc_q;
    a.acc_c = acc_c;
  } else if (use_tc && a.variant == 12) {
    // Sorted two-speed sweep: prepass over every 64th tile -> sweep order by flag count -> the sweep in that order.
    const float* tiles = static_cast<const float*>(s->tc2_tiles);
    const int total_tiles = d.npad / rlb::tc_tile_points();
    const int k_tiles = (k + 127) / 128;
    const int slots = c->sms * rlb::tc2_resident_ctas();
    RLB_CUDA(c->tcx_lists.ensure(static_cast<size_t>(k) * 2 * sizeof(int32_t)));
    int32_t* order = c->tcx_lists.as<int32_t>();
    int32_t* pos = order + k;
    const int stride = 64;
    const int n_sample = (total_tiles + stride - 1) / stride;
    SweepArgs pa = a;
    pa.tile_stride = stride;
    pa.n_splits = std::max(1, std::min(n_sample, (slots + k_tiles - 1) / k_tiles));
    pa.stages_per_split = (n_sample + pa.n_splits - 1) / pa.n_splits;
    pa.n_splits = (n_sample + pa.stages_per_split - 1) / pa.stages_per_split;
    RLB_CUDA(c->part_count.ensure(static_cast<size_t>(std::max(splits, pa.n_splits)) * k * sizeof(int32_t)));
    pa.part_count = c->part_count.as<int32_t>();
    a.part_count = pa.part_count;
    RLB_CUDA(rlb::launch_sweep_tc2(d, pa, tiles, true, c->stream, &c->launches_i));
    RLB_CUDA(rlb::launch_tc2_order(pa.part_count, pa.n_splits, k, order, pos, c->stream, &c->launches_i));
    a.ids = order;
    RLB_CUDA(rlb::launch_sweep_tc2(d, a, tiles, false, c->stream, &c->launches_i));
    a.ids = nullptr;  // finalize walks all k hypotheses and finds their partials through pos
    a.pos = pos;
  } else if (use_tc) {
    RLB_CUDA(rlb::launch_sweep_tc(d, a, static_cast<const float*>(s->tc_tiles), c->stream, &c->launches_i));
  } else {
    RLB_CUDA(rlb::launch_sweep(d, a, precision, c->stream, &c->launches_i));
  }
  if (c->timing) {
    RLB_CUDA(cudaEventRecord(c->ev[1], c->stream));
    c->ev_valid[0] = true;
  }
  RLB_CUDA(rlb::launch_finalize(
      d, a, d_scores, d_counts, c->block_best.as<BestRec>(),
      reinterpret_cast<unsigned int*>(c->small.as<char>() + kSmallTicket), d_best, d_best_model,
      c->stream, &c->launches_i));
  account(c);
  return RLB_OK;
}

int inliers_dev(rlb_scorer* s, const double* d_model, const double* thr_inl, const double* thr_score,
                const double* w, int32_t* d_idx, void* d_header, const double* host_model = nullptr) {
  rlb_ctx* c = s->ctx;
  const DeviceData& d = s->d;
  if (!thr_inl) return fail(RLB_E_ARG, "sq_thr_inl is required");
  const int tiles = d.npad / rlb::kInlTile;
  int rc = ensure_first_use(c);
  if (rc) return rc;
  // status words: one per tile + one per group of 256 tiles
  const size_t status_words = static_cast<size_t>(tiles) + (tiles + 255) / 256;
  if (status_words * 8 > c->status.cap) {
    RLB_CUDA(c->status.ensure(status_words * 8));
    RLB_CUDA(cudaMemsetAsync(c->status.p, 0, c->status.cap, c->stream));
  }
  RLB_CUDA(c->tile_score.ensure(static_cast<size_t>(tiles) * sizeof(double)));
  InlierArgs a;
  a.model = d_model;
  if (host_model) {
    a.model = nullptr;
    for (int i = 0; i < rlb::model_size_of(d); ++i) a.model_v[i] = host_model[i];
  }
  fill_thr(d, thr_inl, w, a.thr_inl, a.w);
  double dummy[rlb::kMaxTypes];
  fill_thr(d, thr_score ? thr_score : thr_inl, w, a.thr_score, dummy);
  a.want_score = thr_score ? 1 : 0;
  a.idx = d_idx;
  a.tile_score = c->tile_score.as<double>();
  a.status = c->status.as<unsigned long long>();
  a.control = reinterpret_cast<unsigned int*>(c->small.as<char>() + kSmallControl);
  c->epoch = (c->epoch % 0x3fffffffu) + 1u;
  a.epoch = c->epoch;
  a.header = d_header;
  if (c->timing) RLB_CUDA(cudaEventRecord(c->ev[2], c->stream));
  if (d.kind == RLB_KIND_CUSTOM) {
    // the user's functor fills the error array, the shared compaction kernel reads it as a 1-channel data set
    const double* dm = d_model;
    if (host_model) {
      RLB_CUDA(c->one_model.ensure(12 * sizeof(double)));
      RLB_CUDA(cudaMemcpyAsync(c->one_model.p, host_model, d.model_doubles * sizeof(double), cudaMemcpyHostToDevice,
                               c->stream));
      dm = c->one_model.as<double>();
    }
    std::string err;
    if (!rlb::launch_custom_errors(s->custom, d, dm, static_cast<double*>(s->custom_err), c->stream,
                                   &c->launches_i, &err))
      return fail(RLB_E_CUDA, "custom errors: %s", err.c_str());
    DeviceData de = d;
    de.soa64 = static_cast<const double*>(s->custom_err);
    de.channels = 1;
    RLB_CUDA(rlb::launch_inliers(de, a, c->stream, &c->launches_i));
  } else {
    RLB_CUDA(rlb::launch_inliers(d, a, c->stream, &c->launches_i));
  }
  if (c->timing) {
    RLB_CUDA(cudaEventRecord(c->ev[3], c->stream));
    c->ev_valid[1] = true;
  }
  account(c);
  return RLB_OK;
}

int solver_sample_size(const DeviceData& d, int solver) {
  if (d.kind == RLB_KIND_CUSTOM) return -1;  // hypotheses of a user-defined solver come from its host MinimalSolver
  if (d.kind == RLB_KIND_POSE) return solver == 0 ? 4 : -1;
  if (d.kind == RLB_KIND_LINE) return solver == 0 ? 2 : -1;
  return solver == 0 ? 2 : (solver == 1 ? 1 : -1);
}

int solve_dev(rlb_scorer* s, int solver, int k, const int32_t* d_samples, double sq_thr,
              double* d_models, uint8_t* d_valid) {
  rlb_ctx* c = s->ctx;
  const DeviceData& d = s->d;
  if (k <= 0) return fail(RLB_E_ARG, "k must be positive (got %d)", k);
  if (solver_sample_size(d, solver) < 0) return fail(RLB_E_KIND, "solver %d undefined for kind %d", solver, d.kind);
  if (d.kind == RLB_KIND_POSE) {
    RLB_CUDA(rlb::launch_solve_p3p(d, k, d_samples, sq_thr, d_models, d_valid, c->stream, &c->launches_i));
  } else if (solver == 0) {
    RLB_CUDA(rlb::launch_solve_line2pt(d, k, d_samples, d_models, d_valid, c->stream, &c->launches_i));
  } else {
    RLB_CUDA(rlb::launch_solve_point_normal(d, k, d_samples, d_models, d_valid, c->stream, &c->launches_i));
  }
  account(c);
  return RLB_OK;
}

int check_samples(const DeviceData& d, int solver, int k, const int32_t* samples) {
  const int ss = solver_sample_size(d, solver);
  if (ss < 0) return fail(RLB_E_KIND, "solver %d undefined for kind %d", solver, d.kind);
  const int type = (d.kind == RLB_KIND_HYBRID_LINE && solver == 1) ? 1 : 0;
  const int n = d.n[type];
  for (size_t i = 0; i < static_cast<size_t>(k) * ss; ++i)
    if (samples[i] < 0 || samples[i] >= n)
      return fail(RLB_E_ARG, "sample index %d out of range [0,%d)", samples[i], n);
  return RLB_OK;
}

}  // namespace

extern "C" {

const char* rlb_last_error(void) { return g_err; }
int rlb_version(void) { return 100; }

int rlb_ctx_create(int device, rlb_ctx** out) {
  if (!out) return fail(RLB_E_ARG, "null out pointer");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(RLB_E_CUDA, "no CUDA device available (%s); ransaclib_b200 has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= count) return fail(RLB_E_ARG, "device %d out of range [0,%d)", device, count);
  RLB_CUDA(cudaSetDevice(device));
  rlb_ctx* c = new (std::nothrow) rlb_ctx();
  if (!c) return fail(RLB_E_NOMEM, "out of host memory");
  c->device = device;
  cudaDeviceProp prop;
  RLB_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sms = prop.multiProcessorCount;
  RLB_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  RLB_CUDA(cudaMallocHost(&c->h_small, 4096));
  for (int i = 0; i < 4; ++i) RLB_CUDA(cudaEventCreate(&c->ev[i]));
  *out = c;
  return RLB_OK;
}

int rlb_ctx_destroy(rlb_ctx* c) {
  if (!c) return RLB_OK;
  if (c->live_scorers > 0) {
    c->destroy_pending = true;  // completed by the last rlb_scorer_destroy
    return RLB_OK;
  }
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  DevBuf* bufs[] = {&c->models, &c->valid, &c->samples, &c->scores, &c->counts, &c->part_score,
                    &c->part_count, &c->block_best, &c->small, &c->one_model, &c->idx,
                    &c->tile_score, &c->status, &c->header, &c->tcx_lists, &c->tcx_acc};
  for (DevBuf* b : bufs) b->release();
  pool_drain(c);
  if (c->h_small) cudaFreeHost(c->h_small);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->h_list) cudaFreeHost(c->h_list);
  if (c->h_result) cudaFreeHost(c->h_result);
  c->loc_soa.release();
  c->loc_lists.release();
  c->ll_lists.release();
  c->ll_io.release();
  c->loc_words.release();
  for (int i = 0; i < 9; ++i)
    if (c->loc_ev[i]) cudaEventDestroy(c->loc_ev[i]);
  for (int i = 0; i < 4; ++i)
    if (c->loc_streams[i]) cudaStreamDestroy(c->loc_streams[i]);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (int i = 0; i < 4; ++i)
    if (c->ev[i]) cudaEventDestroy(c->ev[i]);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
  return RLB_OK;
}

int rlb_ctx_set_stream(rlb_ctx* c, void* stream) {
  if (!c) return fail(RLB_E_ARG, "null ctx");
  c->stream = static_cast<cudaStream_t>(stream);
  return RLB_OK;
}

int rlb_ctx_use_own_stream(rlb_ctx* c) {
  if (!c) return fail(RLB_E_ARG, "null ctx");
  c->stream = c->own_stream;
  return RLB_OK;
}

int rlb_ctx_synchronize(rlb_ctx* c) {
  if (!c) return fail(RLB_E_ARG, "null ctx");
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  return RLB_OK;
}

int64_t rlb_ctx_launch_count(const rlb_ctx* c) { return c ? c->launches : 0; }
int rlb_ctx_sm_count(const rlb_ctx* c) { return c ? c->sms : 0; }

int rlb_scorer_create_line(rlb_ctx* c, int n, const double* xy, rlb_scorer** out) {
  if (!c || !out || n < 0 || (n > 0 && !xy)) return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  rlb_scorer* s = new (std::nothrow) rlb_scorer();
  if (!s) return fail(RLB_E_NOMEM, "out of host memory");
  s->ctx = c;
  ++c->live_scorers;
  DeviceData& d = s->d;
  memset(&d, 0, sizeof(d));
  d.kind = RLB_KIND_LINE;
  d.num_types = 1;
  d.n[0] = n;
  d.off[0] = 0;
  d.off[1] = d.npad = std::max(pad_to(n, rlb::kSegAlign), rlb::kSegAlign);
  d.channels = 2;
  subsampled_mean(xy, n, 2, 2, d.centre);
  rc = upload(c, &s->raw0, xy, static_cast<size_t>(n) * 2, c->stream);
  if (!rc) {
    d.raw0 = static_cast<const double*>(s->raw0);
    rc = finish_scorer(s);
  }
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  *out = s;
  return RLB_OK;
}

int rlb_scorer_create_hybrid_line(rlb_ctx* c, int n0, const double* xy0, int n1,
                                  const double* xynn1, rlb_scorer** out) {
  if (!c || !out || n0 < 0 || n1 < 0 || (n0 > 0 && !xy0) || (n1 > 0 && !xynn1))
    return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  rlb_scorer* s = new (std::nothrow) rlb_scorer();
  if (!s) return fail(RLB_E_NOMEM, "out of host memory");
  s->ctx = c;
  ++c->live_scorers;
  DeviceData& d = s->d;
  memset(&d, 0, sizeof(d));
  d.kind = RLB_KIND_HYBRID_LINE;
  d.num_types = 2;
  d.n[0] = n0;
  d.n[1] = n1;
  d.off[0] = 0;
  d.off[1] = pad_to(n0, rlb::kSegAlign);
  d.npad = std::max(d.off[1] + pad_to(n1, rlb::kSegAlign), rlb::kSegAlign);
  d.channels = 2;
  if (n0 > 0)
    subsampled_mean(xy0, n0, 2, 2, d.centre);
  else
    subsampled_mean(xynn1, n1, 4, 2, d.centre);
  rc = upload(c, &s->raw0, xy0, static_cast<size_t>(n0) * 2, c->stream);
  if (!rc) rc = upload(c, &s->raw1, xynn1, static_cast<size_t>(n1) * 4, c->stream);
  if (!rc) {
    d.raw0 = static_cast<const double*>(s->raw0);
    d.raw1 = static_cast<const double*>(s->raw1);
    rc = finish_scorer(s);
  }
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  *out = s;
  return RLB_OK;
}

int rlb_scorer_create_pose(rlb_ctx* c, int n, const double* p2d, const double* p3d, double fx,
                           double fy, rlb_scorer** out) {
  if (!c || !out || n < 0 || (n > 0 && (!p2d || !p3d))) return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  rlb_scorer* s = new (std::nothrow) rlb_scorer();
  if (!s) return fail(RLB_E_NOMEM, "out of host memory");
  s->ctx = c;
  ++c->live_scorers;
  DeviceData& d = s->d;
  memset(&d, 0, sizeof(d));
  d.kind = RLB_KIND_POSE;
  d.num_types = 1;
  d.n[0] = n;
  d.off[0] = 0;
  d.off[1] = d.npad = std::max(pad_to(n, rlb::kSegAlign), rlb::kSegAlign);
  d.channels = 5;
  d.fx = fx;
  d.fy = fy;
  subsampled_mean(p3d, n, 3, 3, d.centre);
  rc = upload(c, &s->raw0, p2d, static_cast<size_t>(n) * 2, c->stream);
  if (!rc) rc = upload(c, &s->raw1, p3d, static_cast<size_t>(n) * 3, c->stream);
  if (!rc) {
    d.raw0 = static_cast<const double*>(s->raw0);
    d.raw1 = static_cast<const double*>(s->raw1);
    rc = finish_scorer(s);
  }
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  *out = s;
  return RLB_OK;
}

}  // extern "C"

namespace rlb {
// Two-step construction of a pose scorer whose raw arrays are filled ON THE DEVICE by the caller (a device
// copy, or an upload / ncclBroadcast straight into them: comm.cu). begin: allocates; *raw2d (n x 2) and
// *raw3d (n x 3) are the scorer's own arrays. finish: everything written to them must be ordered before it
// on the context stream; derives the recentring point exactly as the host-pointer constructor does (the
// same sub-sampled rows, read back with one strided copy) and builds the SoA copies and the rays.
int scorer_pose_begin(rlb_ctx* c, int n, double fx, double fy, rlb_scorer** out, double** raw2d, double** raw3d) {
  if (!c || !out || n < 0) return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  rlb_scorer* s = new (std::nothrow) rlb_scorer();
  if (!s) return fail(RLB_E_NOMEM, "out of host memory");
  s->ctx = c;
  ++c->live_scorers;
  DeviceData& d = s->d;
  memset(&d, 0, sizeof(d));
  d.kind = RLB_KIND_POSE;
  d.num_types = 1;
  d.n[0] = n;
  d.off[0] = 0;
  d.off[1] = d.npad = std::max(pad_to(n, rlb::kSegAlign), rlb::kSegAlign);
  d.channels = 5;
  d.fx = fx;
  d.fy = fy;
  rc = pool_alloc(c, static_cast<size_t>(std::max(n, 1)) * 2 * sizeof(double), &s->raw0);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(std::max(n, 1)) * 3 * sizeof(double), &s->raw1);
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  d.raw0 = static_cast<const double*>(s->raw0);
  d.raw1 = static_cast<const double*>(s->raw1);
  *raw2d = static_cast<double*>(s->raw0);
  *raw3d = static_cast<double*>(s->raw1);
  *out = s;
  return RLB_OK;
}

// rows 0, step, 2 step, ... of an n x 3 array, contiguous (the sub-sample the recentring point is the mean of)
__global__ void gather_rows3_kernel(const double* __restrict__ src, int step, int m, double* __restrict__ dst) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= m) return;
  const double* p = src + static_cast<size_t>(r) * step * 3;
  dst[3 * r] = p[0];
  dst[3 * r + 1] = p[1];
  dst[3 * r + 2] = p[2];
}

// host_p3d: the caller's copy of the WHOLE points3D array when this process has one (same bytes as on the
// device): the recentring point then needs no device read-back at all.
int scorer_pose_finish(rlb_scorer* s, const double* host_p3d) {
  rlb_ctx* c = s->ctx;
  int rc = ctx_use(c);
  if (rc) return rc;
  DeviceData& d = s->d;
  const int n = d.n[0];
  if (host_p3d != nullptr) {
    subsampled_mean(host_p3d, n, 3, 3, d.centre);
    return finish_scorer(s);
  }
  const int step = std::max(1, n / 4096);
  const int m = (n > 0) ? (n + step - 1) / step : 0;
  std::vector<double> sub(static_cast<size_t>(std::max(m, 1)) * 3, 0.0);
  if (m > 0) {
    // gathered on the device into one contiguous block (a strided cudaMemcpy2D of 4096 24-byte rows costs ~0.5 ms)
    const size_t bytes = static_cast<size_t>(m) * 3 * sizeof(double);
    RLB_CUDA(c->scores.ensure(bytes));
    rc = ensure_stage(c, bytes);
    if (rc) return rc;
    gather_rows3_kernel<<<(m + 255) / 256, 256, 0, c->stream>>>(d.raw1, step, m, c->scores.as<double>());
    ++c->launches_i;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(c->h_stage, c->scores.p, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(RLB_E_CUDA, "reading the sub-sample back failed: %s", cudaGetErrorString(e));
    memcpy(sub.data(), c->h_stage, bytes);
  }
  subsampled_mean(sub.data(), m, 3, 3, d.centre);
  return finish_scorer(s);
}
}  // namespace rlb

extern "C" {

int rlb_scorer_create_pose_dev(rlb_ctx* c, int n, const double* d_p2d, const double* d_p3d, double fx,
                               double fy, rlb_scorer** out) {
  if (!c || !out || n < 0 || (n > 0 && (!d_p2d || !d_p3d))) return fail(RLB_E_ARG, "bad arguments");
  rlb_scorer* s = nullptr;
  double *r2 = nullptr, *r3 = nullptr;
  int rc = rlb::scorer_pose_begin(c, n, fx, fy, &s, &r2, &r3);
  if (rc) return rc;
  if (n > 0) {
    cudaError_t e = cudaMemcpyAsync(r2, d_p2d, static_cast<size_t>(n) * 2 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(r3, d_p3d, static_cast<size_t>(n) * 3 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream);
    if (e != cudaSuccess) rc = fail(RLB_E_CUDA, "device copy of the point set failed: %s", cudaGetErrorString(e));
  }
  if (!rc) rc = rlb::scorer_pose_finish(s);
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  *out = s;
  return RLB_OK;
}

int rlb_custom_compile_check(const char* eval_source, int dims, int model_doubles, size_t* cubin_bytes) {
  if (!eval_source) return fail(RLB_E_ARG, "null source");
  if (dims < 1 || dims > 8) return fail(RLB_E_ARG, "dims must be 1..8 (got %d)", dims);
  if (model_doubles < 1 || model_doubles > 12) return fail(RLB_E_ARG, "model_doubles must be 1..12 (got %d)", model_doubles);
  std::string cubin, err;
  // a source that defines rlb_eval_t (the HybridSolver form) is compiled as such
  const bool hybrid = strstr(eval_source, "rlb_eval_t") != nullptr;
  if (!rlb::custom_compile(eval_source, dims, model_doubles, hybrid, &cubin, &err)) return fail(RLB_E_ARG, "%s", err.c_str());
  if (cubin_bytes) *cubin_bytes = cubin.size();
  return RLB_OK;
}

namespace {

// One (num_types = 1) or two data types of a user-defined solver: rows of dims_t doubles per type, stored
// channel-major with max(dims) channels, type 1 from the next multiple of kSegAlign.
int create_custom(rlb_ctx* c, int num_types, const int* n, const int* dims_t, const double* const* data,
                  int model_doubles, const char* eval_source, rlb_scorer** out) {
  if (!c || !out || !eval_source) return fail(RLB_E_ARG, "bad arguments");
  int dims = 0;
  for (int t = 0; t < num_types; ++t) {
    if (n[t] < 0 || (n[t] > 0 && !data[t])) return fail(RLB_E_ARG, "bad arguments (data type %d)", t);
    if (dims_t[t] < 1 || dims_t[t] > 8) return fail(RLB_E_ARG, "dims must be 1..8 (got %d)", dims_t[t]);
    dims = std::max(dims, dims_t[t]);
  }
  if (model_doubles < 1 || model_doubles > 12) return fail(RLB_E_ARG, "model_doubles must be 1..12 (got %d)", model_doubles);
  int rc = ctx_use(c);
  if (rc) return rc;
  std::string cubin, err;
  if (!rlb::custom_compile(eval_source, dims, model_doubles, num_types > 1, &cubin, &err))
    return fail(RLB_E_ARG, "%s", err.c_str());
  rc = ensure_first_use(c);  // a runtime call: the primary context the driver-API module is loaded into exists
  if (rc) return rc;
  rlb_scorer* s = new (std::nothrow) rlb_scorer();
  if (!s) return fail(RLB_E_NOMEM, "out of host memory");
  s->ctx = c;
  ++c->live_scorers;
  DeviceData& d = s->d;
  memset(&d, 0, sizeof(d));
  d.kind = RLB_KIND_CUSTOM;
  d.num_types = num_types;
  d.n[0] = n[0];
  d.off[0] = 0;
  const int span0 = std::max(pad_to(n[0], rlb::kSegAlign), rlb::kSegAlign);
  int span1 = 0;
  if (num_types > 1) {
    d.n[1] = n[1];
    span1 = std::max(pad_to(n[1], rlb::kSegAlign), rlb::kSegAlign);
  }
  d.off[1] = span0;
  d.npad = span0 + span1;
  d.channels = dims;
  d.model_doubles = model_doubles;
  s->custom = rlb::custom_load(cubin, dims, model_doubles, &err);
  if (!s->custom) {
    rlb_scorer_destroy(s);
    return fail(RLB_E_CUDA, "loading the compiled functor: %s", err.c_str());
  }
  rc = upload(c, &s->raw0, data[0], static_cast<size_t>(n[0]) * dims_t[0], c->stream);
  if (!rc && num_types > 1) rc = upload(c, &s->raw1, data[1], static_cast<size_t>(n[1]) * dims_t[1], c->stream);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(dims) * d.npad * sizeof(double), &s->soa64);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(d.npad) * sizeof(double), &s->custom_err);
  if (rc) {
    rlb_scorer_destroy(s);
    return rc;
  }
  d.raw0 = static_cast<const double*>(s->raw0);
  d.raw1 = static_cast<const double*>(s->raw1);
  d.soa64 = static_cast<const double*>(s->soa64);
  cudaError_t ce = rlb::launch_custom_repack(d.raw0, n[0], dims_t[0], dims, d.npad, 0, span0,
                                             static_cast<double*>(s->soa64), c->stream, &c->launches_i);
  if (ce == cudaSuccess && num_types > 1)
    ce = rlb::launch_custom_repack(d.raw1, n[1], dims_t[1], dims, d.npad, span0, span1,
                                   static_cast<double*>(s->soa64), c->stream, &c->launches_i);
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(c->stream);
  account(c);
  if (ce != cudaSuccess) {
    rlb_scorer_destroy(s);
    return fail(RLB_E_CUDA, "custom repack: %s", cudaGetErrorString(ce));
  }
  *out = s;
  return RLB_OK;
}

}  // namespace

int rlb_scorer_create_custom(rlb_ctx* c, int n, int dims, const double* data, int model_doubles,
                             const char* eval_source, rlb_scorer** out) {
  return create_custom(c, 1, &n, &dims, &data, model_doubles, eval_source, out);
}

int rlb_scorer_create_custom_hybrid(rlb_ctx* c, int n0, int dims0, const double* data0, int n1, int dims1,
                                    const double* data1, int model_doubles, const char* eval_source,
                                    rlb_scorer** out) {
  const int n[2] = {n0, n1}, dims[2] = {dims0, dims1};
  const double* data[2] = {data0, data1};
  return create_custom(c, 2, n, dims, data, model_doubles, eval_source, out);
}

int rlb_scorer_destroy(rlb_scorer* s) {
  if (!s) return RLB_OK;
  if (s->ctx) {
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
  }
  rlb::custom_unload(s->custom);
  void* ptrs[] = {s->raw0, s->raw1, s->rays, s->soa64, s->soa32, s->tc_tiles, s->tc2_tiles, s->custom_err};
  for (void* p : ptrs) {
    if (!p) continue;
    if (s->ctx) pool_release(s->ctx, p);
    else cudaFree(p);
  }
  rlb_ctx* c = s->ctx;
  delete s;
  if (c && --c->live_scorers == 0 && c->destroy_pending) return rlb_ctx_destroy(c);
  return RLB_OK;
}

int rlb_scorer_kind(const rlb_scorer* s) { return s ? s->d.kind : RLB_E_ARG; }
int rlb_scorer_num_types(const rlb_scorer* s) { return s ? s->d.num_types : RLB_E_ARG; }
int rlb_scorer_num_data(const rlb_scorer* s, int type) {
  if (!s || type < 0 || type >= s->d.num_types) return RLB_E_ARG;
  return s->d.n[type];
}
int rlb_scorer_model_size(const rlb_scorer* s) { return s ? rlb::model_size_of(s->d) : RLB_E_ARG; }

int rlb_scorer_get_rays(const rlb_scorer* s, double* rays) {
  int rc = check_scorer(s);
  if (rc) return rc;
  if (s->d.kind != RLB_KIND_POSE) return fail(RLB_E_KIND, "rays exist for pose scorers only");
  if (!rays) return fail(RLB_E_ARG, "null rays");
  RLB_CUDA(cudaSetDevice(s->ctx->device));
  RLB_CUDA(cudaMemcpyAsync(rays, s->rays, static_cast<size_t>(s->d.n[0]) * 3 * sizeof(double),
                           cudaMemcpyDeviceToHost, s->ctx->stream));
  RLB_CUDA(cudaStreamSynchronize(s->ctx->stream));
  return RLB_OK;
}

int rlb_score_models_dev(rlb_scorer* s, int precision, int k, const double* d_models,
                         const uint8_t* d_valid, const double* sq_thr, const double* weights,
                         double* d_scores, int32_t* d_counts, void* d_best) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rc = ctx_use(s->ctx);
  if (rc) return rc;
  if (!d_models || !d_best) return fail(RLB_E_ARG, "d_models and d_best are required");
  rc = ensure_first_use(s->ctx);
  if (rc) return rc;
  return score_dev(s, precision, k, d_models, d_valid, sq_thr, weights, d_scores, d_counts,
                   static_cast<BestRec*>(d_best), nullptr);
}

int rlb_score_models(rlb_scorer* s, int precision, int k, const double* models,
                     const uint8_t* valid, const double* sq_thr, const double* weights,
                     double* scores, int32_t* counts, int32_t* best_id, double* best_score) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  if (k <= 0) return fail(RLB_E_ARG, "k must be positive (got %d)", k);
  if (!models) return fail(RLB_E_ARG, "null models");
  rc = ensure_first_use(c);
  if (rc) return rc;
  const int md = rlb::model_size_of(s->d);
  const size_t b_models = static_cast<size_t>(k) * md * sizeof(double), b_valid = valid ? static_cast<size_t>(k) : 0;
  const size_t b_scores = static_cast<size_t>(k) * sizeof(double), b_counts = static_cast<size_t>(k) * sizeof(int32_t);
  RLB_CUDA(c->models.ensure(b_models));
  RLB_CUDA(c->scores.ensure(b_scores));
  RLB_CUDA(c->counts.ensure(b_counts));
  // Small batches (the drivers' rounds) through pinned staging, as in rlb_hypothesize_and_score.
  const size_t o_valid = (b_models + 63) & ~size_t(63), o_scores = (o_valid + b_valid + 63) & ~size_t(63);
  const size_t o_counts = o_scores + b_scores, stage_bytes = o_counts + b_counts;
  char* stage = nullptr;
  if (stage_bytes <= (size_t(1) << 20)) {
    rc = ensure_stage(c, stage_bytes);
    if (rc) return rc;
    stage = static_cast<char*>(c->h_stage);
    memcpy(stage, models, b_models);
    if (valid) memcpy(stage + o_valid, valid, b_valid);
  }
  RLB_CUDA(cudaMemcpyAsync(c->models.p, stage ? static_cast<const void*>(stage) : models, b_models,
                           cudaMemcpyHostToDevice, c->stream));
  const uint8_t* d_valid = nullptr;
  if (valid) {
    RLB_CUDA(c->valid.ensure(b_valid));
    RLB_CUDA(cudaMemcpyAsync(c->valid.p, stage ? static_cast<const void*>(stage + o_valid) : valid, b_valid,
                             cudaMemcpyHostToDevice, c->stream));
    d_valid = c->valid.as<uint8_t>();
  }
  BestRec* d_best = reinterpret_cast<BestRec*>(c->small.as<char>() + kSmallBest);
  rc = score_dev(s, precision, k, c->models.as<double>(), d_valid, sq_thr, weights,
                 c->scores.as<double>(), counts ? c->counts.as<int32_t>() : nullptr, d_best, nullptr);
  if (rc) return rc;
  if (scores)
    RLB_CUDA(cudaMemcpyAsync(stage ? static_cast<void*>(stage + o_scores) : scores, c->scores.p, b_scores,
                             cudaMemcpyDeviceToHost, c->stream));
  if (counts)
    RLB_CUDA(cudaMemcpyAsync(stage ? static_cast<void*>(stage + o_counts) : counts, c->counts.p, b_counts,
                             cudaMemcpyDeviceToHost, c->stream));
  BestRec* h_best = static_cast<BestRec*>(c->h_small);
  RLB_CUDA(cudaMemcpyAsync(h_best, d_best, sizeof(BestRec), cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  if (stage) {
    if (scores) memcpy(scores, stage + o_scores, b_scores);
    if (counts) memcpy(counts, stage + o_counts, b_counts);
  }
  if (best_id) *best_id = static_cast<int32_t>(h_best->id);
  if (best_score) *best_score = h_best->score;
  return RLB_OK;
}

int rlb_get_inliers_dev(rlb_scorer* s, const double* d_model, const double* sq_thr_inl,
                        const double* sq_thr_score, const double* weights, int32_t* d_idx,
                        void* d_header) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rc = ctx_use(s->ctx);
  if (rc) return rc;
  if (!d_model || !d_header) return fail(RLB_E_ARG, "d_model and d_header are required");
  return inliers_dev(s, d_model, sq_thr_inl, sq_thr_score, weights, d_idx, d_header);
}

// The same pass with its scratch in a fixed file-scope device array (score_kernels.cu) instead of the context's
// growable buffers: no cudaMalloc / cudaFree / allocating routine on this path, ever. No counterpart in the
// reference. Calls on one device share the array: do not overlap them.
int rlb_get_inliers_noalloc_dev(rlb_scorer* s, const double* d_model, const double* sq_thr_inl,
                                const double* sq_thr_score, const double* weights, int32_t* d_idx,
                                void* d_header) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  if (!d_model || !d_header) return fail(RLB_E_ARG, "d_model and d_header are required");
  if (!sq_thr_inl) return fail(RLB_E_ARG, "sq_thr_inl is required");
  const DeviceData& d = s->d;
  if (d.kind == RLB_KIND_CUSTOM)
    return fail(RLB_E_KIND, "rlb_get_inliers_noalloc_dev serves the built-in kinds (line, hybrid line, pose)");
  if (d.npad > rlb::inliers_noalloc_max_points())
    return fail(RLB_E_ARG,
                "data set of %d (padded) points exceeds the fixed scratch of rlb_get_inliers_noalloc_dev: at most "
                "RLB_NOALLOC_MAX_POINTS = %d; use rlb_get_inliers_dev",
                d.npad, rlb::inliers_noalloc_max_points());
  InlierArgs a;
  a.model = d_model;
  fill_thr(d, sq_thr_inl, weights, a.thr_inl, a.w);
  double dummy[rlb::kMaxTypes];
  fill_thr(d, sq_thr_score ? sq_thr_score : sq_thr_inl, weights, a.thr_score, dummy);
  a.want_score = sq_thr_score ? 1 : 0;
  a.idx = d_idx;
  a.tile_score = nullptr;  // filled by the launcher from the device's static scratch
  a.status = nullptr;
  a.control = nullptr;
  a.epoch = 0u;
  a.header = d_header;
  RLB_CUDA(rlb::launch_inliers_noalloc(d, a, c->stream, &c->launches_i));
  account(c);
  return RLB_OK;
}

namespace {
// idx: copy the list there; view: hand out a pointer to the context's own pinned copy instead (valid until the
// next call on the context). Either, both or none.
int get_inliers_impl(rlb_scorer* s, const double* model, const double* sq_thr_inl,
                     const double* sq_thr_score, const double* weights, int32_t* idx, const int32_t** view,
                     int32_t* counts, double* score) {
  const bool want_list = idx != nullptr || view != nullptr;
  if (view) *view = nullptr;
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  if (!model) return fail(RLB_E_ARG, "null model");
  const DeviceData& d = s->d;
  const int md = rlb::model_size_of(d);
  const int ntot = d.n[0] + (d.num_types > 1 ? d.n[1] : 0);
  // points; above this the list is fetched by DMA (RLB_ZERO_COPY_MAX overrides: A/B timing)
  static const int kZeroCopyMax = [] {
    const char* e = std::getenv("RLB_ZERO_COPY_MAX");
    return e ? std::atoi(e) : 262144;  // configs[2] (2 x 100k points): 10.7 -> 8.6 ms against 65536
  }();
  InlierHeader* h = nullptr;
  if (ntot <= kZeroCopyMax) {
    // Small data sets (the latency-critical LO chains): the kernel writes header + list straight
    // into mapped pinned memory; the model travels in the launch parameters. 1 launch + 1 sync.
    const size_t need = sizeof(InlierHeader) + static_cast<size_t>(std::max(ntot, 1)) * sizeof(int32_t);
    if (need > c->result_cap) {
      if (c->h_result) cudaFreeHost(c->h_result);
      c->h_result = nullptr;
      c->result_cap = 0;
      const size_t cap = std::max<size_t>(need, 64 * 1024);
      RLB_CUDA(cudaHostAlloc(&c->h_result, cap, cudaHostAllocMapped));
      RLB_CUDA(cudaHostGetDevicePointer(&c->d_result, c->h_result, 0));
      c->result_cap = cap;
    }
    char* dres = static_cast<char*>(c->d_result);
    rc = inliers_dev(s, nullptr, sq_thr_inl, sq_thr_score, weights,
                     want_list ? reinterpret_cast<int32_t*>(dres + sizeof(InlierHeader)) : nullptr, dres, model);
    if (rc) return rc;
    RLB_CUDA(cudaStreamSynchronize(c->stream));
    h = static_cast<InlierHeader*>(c->h_result);
    const int32_t* list = reinterpret_cast<const int32_t*>(static_cast<char*>(c->h_result) + sizeof(InlierHeader));
    if (idx && h->count[1] > 0) memcpy(idx, list, static_cast<size_t>(h->count[1]) * sizeof(int32_t));
    if (view) *view = list;
  } else {
    RLB_CUDA(c->header.ensure(sizeof(InlierHeader)));
    if (want_list) RLB_CUDA(c->idx.ensure(static_cast<size_t>(std::max(ntot, 1)) * sizeof(int32_t)));
    rc = inliers_dev(s, nullptr, sq_thr_inl, sq_thr_score, weights, want_list ? c->idx.as<int32_t>() : nullptr,
                     c->header.p, model);
    if (rc) return rc;
    h = reinterpret_cast<InlierHeader*>(static_cast<char*>(c->h_small) + 64);
    RLB_CUDA(cudaMemcpyAsync(h, c->header.p, sizeof(InlierHeader), cudaMemcpyDeviceToHost, c->stream));
    const bool one_shot = ntot <= (1 << 22);  // beyond that the unused tail of the range costs more than a second sync
    if (want_list) {
      // The list's length is in the header, which is not here yet: the whole candidate range goes to the context's
      // pinned block behind the header (4 B x points at PCIe rate: 80 us for 1M points) and ONE synchronisation
      // covers both, instead of header -> sync -> list -> sync with a pageable destination.
      const size_t cap = static_cast<size_t>(ntot) * sizeof(int32_t);
      if (cap > c->list_cap) {
        if (c->h_list) cudaFreeHost(c->h_list);
        c->h_list = nullptr;
        c->list_cap = 0;
        RLB_CUDA(cudaMallocHost(&c->h_list, cap));
        c->list_cap = cap;
      }
      if (one_shot) RLB_CUDA(cudaMemcpyAsync(c->h_list, c->idx.p, cap, cudaMemcpyDeviceToHost, c->stream));
    }
    RLB_CUDA(cudaStreamSynchronize(c->stream));
    const size_t bytes = static_cast<size_t>(std::max(h->count[1], 0)) * sizeof(int32_t);
    if (want_list && !one_shot && bytes > 0) {
      RLB_CUDA(cudaMemcpyAsync(c->h_list, c->idx.p, bytes, cudaMemcpyDeviceToHost, c->stream));
      RLB_CUDA(cudaStreamSynchronize(c->stream));
    }
    if (view && bytes > 0) *view = static_cast<const int32_t*>(c->h_list);
    if (idx && bytes > 0) memcpy(idx, c->h_list, bytes);
  }
  const int total = h->count[1];
  if (counts) {
    counts[0] = h->count[0];
    if (d.num_types > 1) counts[1] = total - h->count[0];
  }
  if (score) *score = h->score;
  return RLB_OK;
}
}  // namespace

int rlb_get_inliers(rlb_scorer* s, const double* model, const double* sq_thr_inl,
                    const double* sq_thr_score, const double* weights, int32_t* idx,
                    int32_t* counts, double* score) {
  return get_inliers_impl(s, model, sq_thr_inl, sq_thr_score, weights, idx, nullptr, counts, score);
}

int rlb_get_inliers_view(rlb_scorer* s, const double* model, const double* sq_thr_inl,
                         const double* sq_thr_score, const double* weights, const int32_t** idx_view,
                         int32_t* counts, double* score) {
  if (!idx_view) return fail(RLB_E_ARG, "null idx_view");
  return get_inliers_impl(s, model, sq_thr_inl, sq_thr_score, weights, nullptr, idx_view, counts, score);
}

int rlb_eval_errors(rlb_scorer* s, int precision, const double* model, int type, double* err) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  const DeviceData& d = s->d;
  if (!model || !err) return fail(RLB_E_ARG, "null pointer");
  if (type < 0 || type >= d.num_types) return fail(RLB_E_ARG, "bad data type %d", type);
  const int md = rlb::model_size_of(d);
  const int n = d.n[type];
  if (n == 0) return RLB_OK;
  RLB_CUDA(c->one_model.ensure(12 * sizeof(double)));
  RLB_CUDA(c->scores.ensure(static_cast<size_t>(n) * sizeof(double)));
  RLB_CUDA(cudaMemcpyAsync(c->one_model.p, model, md * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (d.kind == RLB_KIND_CUSTOM) {
    std::string cerr;
    if (!rlb::launch_custom_errors(s->custom, d, c->one_model.as<double>(), static_cast<double*>(s->custom_err),
                                   c->stream, &c->launches_i, &cerr))
      return fail(RLB_E_CUDA, "custom errors: %s", cerr.c_str());
    RLB_CUDA(cudaMemcpyAsync(c->scores.p, static_cast<const double*>(s->custom_err) + d.off[type],
                             static_cast<size_t>(n) * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  } else {
    RLB_CUDA(rlb::launch_eval_errors(d, precision, c->one_model.as<double>(), type,
                                     c->scores.as<double>(), c->stream, &c->launches_i));
  }
  account(c);
  RLB_CUDA(cudaMemcpyAsync(err, c->scores.p, static_cast<size_t>(n) * sizeof(double),
                           cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  return RLB_OK;
}

static int solve_host(rlb_scorer* s, int solver, int k, const int32_t* samples, double sq_thr,
                      double* models, uint8_t* valid) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  if (k <= 0) return fail(RLB_E_ARG, "k must be positive (got %d)", k);
  if (!samples || !models || !valid) return fail(RLB_E_ARG, "null pointer");
  rc = check_samples(s->d, solver, k, samples);
  if (rc) return rc;
  const int ss = solver_sample_size(s->d, solver);
  const int md = rlb::model_size_of(s->d);
  RLB_CUDA(c->samples.ensure(static_cast<size_t>(k) * ss * sizeof(int32_t)));
  RLB_CUDA(c->models.ensure(static_cast<size_t>(k) * md * sizeof(double)));
  RLB_CUDA(c->valid.ensure(static_cast<size_t>(k)));
  RLB_CUDA(cudaMemcpyAsync(c->samples.p, samples, static_cast<size_t>(k) * ss * sizeof(int32_t),
                           cudaMemcpyHostToDevice, c->stream));
  rc = solve_dev(s, solver, k, c->samples.as<int32_t>(), sq_thr, c->models.as<double>(),
                 c->valid.as<uint8_t>());
  if (rc) return rc;
  RLB_CUDA(cudaMemcpyAsync(models, c->models.p, static_cast<size_t>(k) * md * sizeof(double),
                           cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaMemcpyAsync(valid, c->valid.p, static_cast<size_t>(k), cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  return RLB_OK;
}

int rlb_solve_line2pt_batch(rlb_scorer* s, int k, const int32_t* samples, double* models,
                            uint8_t* valid) {
  if (s && s->d.kind == RLB_KIND_POSE) return fail(RLB_E_KIND, "2-point line solver on a pose scorer");
  return solve_host(s, 0, k, samples, 0.0, models, valid);
}
int rlb_solve_point_normal_batch(rlb_scorer* s, int k, const int32_t* samples, double* models,
                                 uint8_t* valid) {
  if (s && s->d.kind != RLB_KIND_HYBRID_LINE)
    return fail(RLB_E_KIND, "point+normal solver needs a hybrid line scorer");
  return solve_host(s, 1, k, samples, 0.0, models, valid);
}
int rlb_solve_p3p_batch(rlb_scorer* s, int k, const int32_t* samples, double sq_thr,
                        double* models, uint8_t* valid) {
  if (s && s->d.kind != RLB_KIND_POSE) return fail(RLB_E_KIND, "P3P needs a pose scorer");
  return solve_host(s, 0, k, samples, sq_thr, models, valid);
}

int rlb_hypothesize_and_score_dev(rlb_scorer* s, int precision, int solver, int k,
                                  const int32_t* d_samples, const double* sq_thr,
                                  const double* weights, double* d_models, uint8_t* d_valid,
                                  double* d_scores, int32_t* d_counts, void* d_best) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rc = ctx_use(s->ctx);
  if (rc) return rc;
  if (!d_samples || !d_models || !d_valid || !d_best || !sq_thr)
    return fail(RLB_E_ARG, "d_samples, d_models, d_valid, d_best, sq_thr are required");
  rc = ensure_first_use(s->ctx);
  if (rc) return rc;
  rc = solve_dev(s, solver, k, d_samples, sq_thr[0], d_models, d_valid);
  if (rc) return rc;
  return score_dev(s, precision, k, d_models, d_valid, sq_thr, weights, d_scores, d_counts,
                   static_cast<BestRec*>(d_best), nullptr);
}

int rlb_sharded_round_dev(rlb_scorer* s, int precision, int solver, int k, const int32_t* d_samples,
                          const double* sq_thr, const double* weights, int64_t id_offset,
                          double* d_models, uint8_t* d_valid, double* d_scores, int32_t* d_counts,
                          double* d_record) {
  int rc = check_scorer(s);
  if (rc) return rc;
  if (!d_record) return fail(RLB_E_ARG, "d_record is required");
  rc = ensure_first_use(s->ctx);
  if (rc) return rc;
  s->ctx->pending_record = d_record;
  s->ctx->pending_offset = id_offset;
  void* d_best = s->ctx->small.as<char>() + kSmallBest;
  rc = rlb_hypothesize_and_score_dev(s, precision, solver, k, d_samples, sq_thr, weights, d_models, d_valid,
                                     d_scores, d_counts, d_best);
  s->ctx->pending_record = nullptr;
  return rc;
}

int rlb_select_winner_dev(rlb_ctx* c, const double* d_records, int world, int record_doubles,
                          double* d_winner) {
  if (!c || !d_records || !d_winner || world <= 0 || record_doubles < 3)
    return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  RLB_CUDA(rlb::launch_select_winner(d_records, world, record_doubles, d_winner, c->stream, &c->launches_i));
  account(c);
  return RLB_OK;
}

int rlb_hypothesize_and_score(rlb_scorer* s, int precision, int solver, int k,
                              const int32_t* samples, const double* sq_thr, const double* weights,
                              double* models, uint8_t* valid, double* scores, int32_t* counts,
                              int32_t* best_id, double* best_score, double* best_model) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  rc = ctx_use(c);
  if (rc) return rc;
  if (k <= 0) return fail(RLB_E_ARG, "k must be positive (got %d)", k);
  if (!samples || !sq_thr) return fail(RLB_E_ARG, "samples and sq_thr are required");
  rc = check_samples(s->d, solver, k, samples);
  if (rc) return rc;
  rc = ensure_first_use(c);
  if (rc) return rc;
  const int ss = solver_sample_size(s->d, solver);
  const int md = rlb::model_size_of(s->d);
  RLB_CUDA(c->samples.ensure(static_cast<size_t>(k) * ss * sizeof(int32_t)));
  RLB_CUDA(c->models.ensure(static_cast<size_t>(k) * md * sizeof(double)));
  RLB_CUDA(c->valid.ensure(static_cast<size_t>(k)));
  RLB_CUDA(c->scores.ensure(static_cast<size_t>(k) * sizeof(double)));
  RLB_CUDA(c->counts.ensure(static_cast<size_t>(k) * sizeof(int32_t)));
  // Small batches (the drivers' speculative rounds) go through pinned staging: truly asynchronous copies,
  // one synchronisation, then plain memcpy into the callers' pageable buffers.
  const size_t b_samples = static_cast<size_t>(k) * ss * sizeof(int32_t);
  const size_t b_models = static_cast<size_t>(k) * md * sizeof(double), b_valid = static_cast<size_t>(k);
  const size_t b_scores = static_cast<size_t>(k) * sizeof(double), b_counts = static_cast<size_t>(k) * sizeof(int32_t);
  const size_t o_models = (b_samples + 63) & ~size_t(63), o_valid = o_models + b_models;
  const size_t o_scores = (o_valid + b_valid + 63) & ~size_t(63), o_counts = o_scores + b_scores;
  const size_t stage_bytes = o_counts + b_counts;
  const bool staged = stage_bytes <= (size_t(1) << 20);
  char* stage = nullptr;
  if (staged) {
    rc = ensure_stage(c, stage_bytes);
    if (rc) return rc;
    stage = static_cast<char*>(c->h_stage);
    memcpy(stage, samples, b_samples);
  }
  RLB_CUDA(cudaMemcpyAsync(c->samples.p, staged ? static_cast<const void*>(stage) : samples, b_samples,
                           cudaMemcpyHostToDevice, c->stream));
  rc = solve_dev(s, solver, k, c->samples.as<int32_t>(), sq_thr[0], c->models.as<double>(),
                 c->valid.as<uint8_t>());
  if (rc) return rc;
  BestRec* d_best = reinterpret_cast<BestRec*>(c->small.as<char>() + kSmallBest);
  double* d_best_model = reinterpret_cast<double*>(c->small.as<char>() + kSmallBestModel);
  rc = score_dev(s, precision, k, c->models.as<double>(), c->valid.as<uint8_t>(), sq_thr, weights,
                 c->scores.as<double>(), counts ? c->counts.as<int32_t>() : nullptr, d_best, d_best_model);
  if (rc) return rc;
  if (models)
    RLB_CUDA(cudaMemcpyAsync(staged ? static_cast<void*>(stage + o_models) : models, c->models.p, b_models,
                             cudaMemcpyDeviceToHost, c->stream));
  if (valid)
    RLB_CUDA(cudaMemcpyAsync(staged ? static_cast<void*>(stage + o_valid) : valid, c->valid.p, b_valid,
                             cudaMemcpyDeviceToHost, c->stream));
  if (scores)
    RLB_CUDA(cudaMemcpyAsync(staged ? static_cast<void*>(stage + o_scores) : scores, c->scores.p, b_scores,
                             cudaMemcpyDeviceToHost, c->stream));
  if (counts)
    RLB_CUDA(cudaMemcpyAsync(staged ? static_cast<void*>(stage + o_counts) : counts, c->counts.p, b_counts,
                             cudaMemcpyDeviceToHost, c->stream));
  char* hs = static_cast<char*>(c->h_small);
  RLB_CUDA(cudaMemcpyAsync(hs, c->small.as<char>() + kSmallBest, sizeof(BestRec), cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaMemcpyAsync(hs + 128, d_best_model, 12 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  if (staged) {
    if (models) memcpy(models, stage + o_models, b_models);
    if (valid) memcpy(valid, stage + o_valid, b_valid);
    if (scores) memcpy(scores, stage + o_scores, b_scores);
    if (counts) memcpy(counts, stage + o_counts, b_counts);
  }
  const BestRec* hb = reinterpret_cast<const BestRec*>(hs);
  if (best_id) *best_id = static_cast<int32_t>(hb->id);
  if (best_score) *best_score = hb->score;
  if (best_model && hb->id >= 0) memcpy(best_model, hs + 128, md * sizeof(double));
  return RLB_OK;
}

// ---- device-resident batched localization -------------------------------------------------
namespace {

constexpr int kMaxLocLaunches = 4;

int check_localize_options(const rlb_localize_options* o) {
  if (!o) return fail(RLB_E_ARG, "null options");
  if (o->min_sample_mult < 1 || o->min_sample_mult * 4 > 32)
    return fail(RLB_E_ARG, "min_sample_mult * 4 must be in [4, 32] (got %d)", o->min_sample_mult);
  if (o->non_min_sample_mult < 0 || o->non_min_sample_mult * 4 > 32)
    return fail(RLB_E_ARG, "non_min_sample_mult * 4 must be <= 32 (got %d)", o->non_min_sample_mult);
  if (!(o->sq_thr > 0.0)) return fail(RLB_E_ARG, "sq_thr must be positive");
  if (o->lo_steps < 0 || o->lsq_iters < 0) return fail(RLB_E_ARG, "negative lo_steps / lsq_iters");
  return RLB_OK;
}

// 1 = one query per CTA (default), 0 = one query per warp. RLB_LOC_MODE overrides (A/B timing, tests).
int localize_mode(int /*nq*/) {
  if (const char* e = std::getenv("RLB_LOC_MODE")) return (e[0] == '0') ? 0 : 1;
  return 1;
}

// Scratch geometry of one launch over `nq` queries.
struct LocPlan {
  int mode, ctas, nmax, npad;
  size_t slots, soa_floats, list_ints;  // totals for this launch
};

LocPlan localize_plan(const rlb_ctx* c, int mode, int nq, int max_matches) {
  LocPlan p;
  p.mode = mode;
  const int wpc = rlb::localize_warps_per_cta();
  const int per_cta = (mode == 0) ? wpc : 1;
  p.nmax = std::max(max_matches, 4);
  p.npad = pad_to(p.nmax, 4);
  const bool need_soa = (mode == 0) || p.nmax > rlb::localize_smem_points();
  const size_t per_slot = (need_soa ? static_cast<size_t>(5) * p.npad * sizeof(float) : 0) +
                          static_cast<size_t>(3) * p.nmax * 4;
  p.ctas = std::max(1, std::min((nq + per_cta - 1) / per_cta, c->sms * rlb::localize_ctas_per_sm(mode)));
  const size_t budget = size_t(8) << 30;  // scratch cap; fewer resident queries for huge ones
  while (p.ctas > 1 && static_cast<size_t>(p.ctas) * per_cta * per_slot > budget) p.ctas /= 2;
  p.slots = static_cast<size_t>(p.ctas) * per_cta;
  p.soa_floats = need_soa ? p.slots * 5 * p.npad : 0;
  p.list_ints = p.slots * 3 * p.nmax;
  return p;
}

// Enqueues one grid over the queries [q_first, q_end) on `st`; everything it reads must be ordered before
// it on `st`. soa / lists / ticket: this launch's own scratch (ticket zeroed by the caller).
int localize_enqueue(rlb_ctx* c, cudaStream_t st, const LocPlan& p, int q_first, int q_end,
                     const int64_t* d_offsets, const double* d_p2d, const double* d_p3d, const double* d_fx,
                     const double* d_fy, const uint32_t* d_seeds, const rlb_localize_options* opt,
                     double* d_poses, rlb_localize_result* d_results, int32_t* d_inliers, float* soa,
                     int32_t* lists, unsigned int* ticket) {
  rlb::LocalizeArgs a;
  a.q_first = q_first;
  a.q_end = q_end;
  a.offsets = reinterpret_cast<const long long*>(d_offsets);
  a.p2d = d_p2d;
  a.p3d = d_p3d;
  a.fx = d_fx;
  a.fy = d_fy;
  a.seeds = d_seeds;
  a.opt = *opt;
  a.poses = d_poses;
  a.results = d_results;
  a.inliers = d_inliers;
  a.soa32 = soa;
  a.lists = lists;
  a.npad_max = p.npad;
  a.nmax = p.nmax;
  a.ticket = ticket;
  a.smem_points = rlb::localize_smem_points();
  a.wave = std::max(c->sms, 1);
  RLB_CUDA(rlb::launch_localize(a, p.mode, p.ctas, st, &c->launches_i));
  account(c);
  return RLB_OK;
}

}  // namespace

void rlb_localize_default_options(rlb_localize_options* o, double inlier_threshold_px,
                                  int number_lo_steps, uint32_t min_num_iterations,
                                  uint32_t max_num_iterations) {
  // examples/ransaclib_localization.cc:100-107
  o->min_iters = min_num_iterations;
  o->max_iters = max_num_iterations;
  o->success_prob = 0.9999;
  o->sq_thr = inlier_threshold_px * inlier_threshold_px;
  o->lo_steps = number_lo_steps;
  o->thr_mult = std::sqrt(2.0);
  o->lsq_iters = 4;
  o->min_sample_mult = 7;
  o->non_min_sample_mult = 3;
  o->lo_start = 60u;
  o->final_lsq = 1;
  o->debug_flags = 0;
  o->handback_after = 0u;
  o->handback_min_remaining = 0u;
}

int rlb_localize_batch_dev(rlb_ctx* c, int nq, int max_matches, const int64_t* d_offsets,
                           const double* d_p2d, const double* d_p3d, const double* d_fx,
                           const double* d_fy, const uint32_t* d_seeds, const rlb_localize_options* opt,
                           double* d_poses, rlb_localize_result* d_results, int32_t* d_inliers) {
  if (!c) return fail(RLB_E_ARG, "null context");
  if (nq <= 0) return fail(RLB_E_ARG, "num_queries must be positive (got %d)", nq);
  if (!d_offsets || !d_p2d || !d_p3d || !d_fx || !d_fy || !d_seeds || !d_poses || !d_results)
    return fail(RLB_E_ARG, "null device pointer");
  int rc = check_localize_options(opt);
  if (rc) return rc;
  rc = ctx_use(c);
  if (rc) return rc;
  const LocPlan p = localize_plan(c, localize_mode(nq), nq, max_matches);
  RLB_CUDA(c->loc_soa.ensure(std::max<size_t>(p.soa_floats, 1) * sizeof(float)));
  RLB_CUDA(c->loc_lists.ensure(p.list_ints * sizeof(int32_t)));
  RLB_CUDA(c->loc_words.ensure(64));
  unsigned int* words = c->loc_words.as<unsigned int>();
  RLB_CUDA(cudaMemsetAsync(words, 0, 64, c->stream));
  return localize_enqueue(c, c->stream, p, 0, nq, d_offsets, d_p2d, d_p3d, d_fx, d_fy, d_seeds, opt, d_poses,
                          d_results, d_inliers, p.soa_floats ? c->loc_soa.as<float>() : nullptr,
                          c->loc_lists.as<int32_t>(), words);
}

int rlb_localize_batch(rlb_ctx* c, int nq, const int64_t* offsets, const double* p2d, const double* p3d,
                       const double* fx, const double* fy, const uint32_t* seeds,
                       const rlb_localize_options* opt, double* poses, rlb_localize_result* results,
                       int32_t* inliers) {
  if (!c) return fail(RLB_E_ARG, "null context");
  if (nq <= 0) return fail(RLB_E_ARG, "num_queries must be positive (got %d)", nq);
  if (!offsets || !p2d || !p3d || !fx || !fy || !seeds || !poses || !results)
    return fail(RLB_E_ARG, "null pointer");
  int rc = check_localize_options(opt);
  if (rc) return rc;
  rc = ctx_use(c);
  if (rc) return rc;
  if (offsets[0] != 0) return fail(RLB_E_ARG, "offsets[0] must be 0");
  int64_t max_matches = 0;
  for (int q = 0; q < nq; ++q) {
    const int64_t n = offsets[q + 1] - offsets[q];
    if (n < 0 || n > 0x7fffffff) return fail(RLB_E_ARG, "query %d has %lld matches", q, (long long)n);
    max_matches = std::max(max_matches, n);
  }
  const int64_t total = offsets[nq];
  if (!c->copy_stream) {
    RLB_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < kMaxLocLaunches; ++i)
      RLB_CUDA(cudaStreamCreateWithFlags(&c->loc_streams[i], cudaStreamNonBlocking));
    for (int i = 0; i < 1 + 2 * kMaxLocLaunches; ++i)
      RLB_CUDA(cudaEventCreateWithFlags(&c->loc_ev[i], cudaEventDisableTiming));
  }
  // Chunks of queries: the first one small so that the first grid starts early, few of them so that
  // the grids' tails stay few; a chunk is at least ~8 MB of matches. Chunk k is uploaded on the copy
  // stream, its grid runs on its own launch stream behind the "uploaded" event, and the grids of
  // different chunks overlap (a later grid's CTAs fill the SMs as an earlier grid's CTAs retire).
  int bounds[kMaxLocLaunches + 1];
  int nl = 0;
  bounds[0] = 0;
  {
    const double cuts[kMaxLocLaunches] = {0.125, 0.5, 1.0, 1.0};
    const int64_t min_rows = 200000;
    int q = 0;
    for (int k = 0; k < kMaxLocLaunches && q < nq; ++k) {
      const int64_t want = std::max<int64_t>(static_cast<int64_t>(cuts[k] * static_cast<double>(total)),
                                             offsets[q] + min_rows);
      int q1 = q + 1;
      while (q1 < nq && offsets[q1 + 1] <= want) ++q1;
      if (k == kMaxLocLaunches - 1 || total - offsets[q1] < min_rows) q1 = nq;
      bounds[++nl] = q1;
      q = q1;
    }
  }
  const int mode = localize_mode(nq);
  LocPlan plans[kMaxLocLaunches];
  size_t soa_total = 0, list_total = 0;
  for (int k = 0; k < nl; ++k) {
    plans[k] = localize_plan(c, mode, bounds[k + 1] - bounds[k], static_cast<int>(max_matches));
    soa_total += plans[k].soa_floats;
    list_total += plans[k].list_ints;
  }
  void *d_off = nullptr, *d_p2 = nullptr, *d_p3 = nullptr, *d_f = nullptr, *d_seed = nullptr,
       *d_pose = nullptr, *d_res = nullptr, *d_inl = nullptr;
  auto release = [&]() {
    for (void* p : {d_off, d_p2, d_p3, d_f, d_seed, d_pose, d_res, d_inl}) pool_release(c, p);
  };
  const size_t tot = static_cast<size_t>(std::max<int64_t>(total, 1));
  rc = pool_alloc(c, (nq + 1) * sizeof(int64_t), &d_off);
  if (!rc) rc = pool_alloc(c, tot * 2 * sizeof(double), &d_p2);
  if (!rc) rc = pool_alloc(c, tot * 3 * sizeof(double), &d_p3);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(nq) * 2 * sizeof(double), &d_f);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(nq) * sizeof(uint32_t), &d_seed);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(nq) * 12 * sizeof(double), &d_pose);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(nq) * sizeof(rlb_localize_result), &d_res);
  if (!rc && inliers) rc = pool_alloc(c, tot * sizeof(int32_t), &d_inl);
  if (rc) {
    release();
    return rc;
  }
  cudaStream_t cs = c->copy_stream;
  double* d_fx = static_cast<double*>(d_f);
  double* d_fy = d_fx + nq;
  // From here on every exit goes through `finish`: all streams drained before the blocks go back to the pool.
  auto finish = [&](int code) {
    cudaStreamSynchronize(cs);
    for (int k = 0; k < kMaxLocLaunches; ++k) cudaStreamSynchronize(c->loc_streams[k]);
    cudaStreamSynchronize(c->stream);
    release();
    return code;
  };
#define RLB_LOC_TRY(expr)                                                                       \
  do {                                                                                          \
    cudaError_t e__ = (expr);                                                                   \
    if (e__ != cudaSuccess) return finish(fail(RLB_E_CUDA, "%s: %s", #expr, cudaGetErrorString(e__))); \
  } while (0)
  RLB_LOC_TRY(c->loc_soa.ensure(std::max<size_t>(soa_total, 1) * sizeof(float)));
  RLB_LOC_TRY(c->loc_lists.ensure(list_total * sizeof(int32_t)));
  RLB_LOC_TRY(c->loc_words.ensure(64));
  unsigned int* words = c->loc_words.as<unsigned int>();
  // Order: whatever ran before on the context stream -> reset of the ticket words -> copies and grids.
  RLB_LOC_TRY(cudaMemsetAsync(words, 0, 64, c->stream));
  RLB_LOC_TRY(cudaEventRecord(c->loc_ev[0], c->stream));
  RLB_LOC_TRY(cudaStreamWaitEvent(cs, c->loc_ev[0], 0));
  RLB_LOC_TRY(cudaMemcpyAsync(d_off, offsets, (nq + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, cs));
  RLB_LOC_TRY(cudaMemcpyAsync(d_fx, fx, nq * sizeof(double), cudaMemcpyHostToDevice, cs));
  RLB_LOC_TRY(cudaMemcpyAsync(d_fy, fy, nq * sizeof(double), cudaMemcpyHostToDevice, cs));
  RLB_LOC_TRY(cudaMemcpyAsync(d_seed, seeds, nq * sizeof(uint32_t), cudaMemcpyHostToDevice, cs));
  size_t soa_at = 0, list_at = 0;
  for (int k = 0; k < nl; ++k) {
    const int q0 = bounds[k], q1 = bounds[k + 1];
    const int64_t r0 = offsets[q0], r1 = offsets[q1];
    if (r1 > r0) {
      RLB_LOC_TRY(cudaMemcpyAsync(static_cast<double*>(d_p2) + 2 * r0, p2d + 2 * r0, (r1 - r0) * 2 * sizeof(double),
                                  cudaMemcpyHostToDevice, cs));
      RLB_LOC_TRY(cudaMemcpyAsync(static_cast<double*>(d_p3) + 3 * r0, p3d + 3 * r0, (r1 - r0) * 3 * sizeof(double),
                                  cudaMemcpyHostToDevice, cs));
    }
    cudaEvent_t uploaded = c->loc_ev[1 + 2 * k], done = c->loc_ev[2 + 2 * k];
    cudaStream_t ls = c->loc_streams[k];
    RLB_LOC_TRY(cudaEventRecord(uploaded, cs));
    RLB_LOC_TRY(cudaStreamWaitEvent(ls, uploaded, 0));
    rc = localize_enqueue(c, ls, plans[k], q0, q1, static_cast<int64_t*>(d_off), static_cast<double*>(d_p2),
                          static_cast<double*>(d_p3), d_fx, d_fy, static_cast<uint32_t*>(d_seed), opt,
                          static_cast<double*>(d_pose), static_cast<rlb_localize_result*>(d_res),
                          static_cast<int32_t*>(d_inl),
                          plans[k].soa_floats ? c->loc_soa.as<float>() + soa_at : nullptr,
                          c->loc_lists.as<int32_t>() + list_at, words + k);
    if (rc) return finish(rc);
    soa_at += plans[k].soa_floats;
    list_at += plans[k].list_ints;
    RLB_LOC_TRY(cudaEventRecord(done, ls));
    RLB_LOC_TRY(cudaStreamWaitEvent(c->stream, done, 0));
  }
  RLB_LOC_TRY(cudaMemcpyAsync(poses, d_pose, static_cast<size_t>(nq) * 12 * sizeof(double), cudaMemcpyDeviceToHost,
                              c->stream));
  RLB_LOC_TRY(cudaMemcpyAsync(results, d_res, static_cast<size_t>(nq) * sizeof(rlb_localize_result),
                              cudaMemcpyDeviceToHost, c->stream));
  if (inliers && total > 0)
    RLB_LOC_TRY(cudaMemcpyAsync(inliers, d_inl, static_cast<size_t>(total) * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                c->stream));
  RLB_LOC_TRY(cudaStreamSynchronize(c->stream));
#undef RLB_LOC_TRY
  release();
  return RLB_OK;
}

// ---- device-resident line LO-MSAC -----------------------------------------------------------
namespace {

int check_line_lomsac_options(const rlb_lomsac_options* o) {
  if (!o) return fail(RLB_E_ARG, "null options");
  if (o->min_sample_mult < 1 || o->min_sample_mult * 2 > 32)
    return fail(RLB_E_ARG, "min_sample_mult * 2 must be in [2, 32] (got %d)", o->min_sample_mult);
  if (o->non_min_sample_mult < 0 || o->non_min_sample_mult * 2 > 32)
    return fail(RLB_E_ARG, "non_min_sample_mult * 2 must be <= 32 (got %d)", o->non_min_sample_mult);
  if (o->lo_steps < 0 || o->lsq_iters < 0) return fail(RLB_E_ARG, "negative lo_steps / lsq_iters");
  return RLB_OK;
}

// Enqueues the kernel on the context stream. d_offsets == nullptr: ONE problem of n_single points.
int line_lomsac_enqueue(rlb_ctx* c, int np, int max_points, const int64_t* d_offsets, int n_single,
                        const double* d_xy, const uint32_t* d_seeds, uint32_t seed_single,
                        const rlb_lomsac_options* opt, double* d_lines, rlb_lomsac_result* d_results,
                        int32_t* d_inliers, unsigned int* d_ticket) {
  const int nmax = std::max(max_points, 2);
  // Stage the points in shared memory when the largest problem fits (RLB_LL_SMEM=0: never; A/B timing).
  int smem_points = (nmax <= rlb::line_lomsac_smem_points()) ? nmax : 0;
  if (const char* e = std::getenv("RLB_LL_SMEM"))
    if (e[0] == '0') smem_points = 0;
  int ctas = std::max(1, std::min(np, c->sms * rlb::line_lomsac_ctas_per_sm(smem_points)));
  const size_t budget = size_t(4) << 30;  // scratch cap; fewer resident CTAs for huge problems
  while (ctas > 1 && static_cast<size_t>(ctas) * 3 * nmax * sizeof(int32_t) > budget) ctas /= 2;
  RLB_CUDA(c->ll_lists.ensure(static_cast<size_t>(ctas) * 3 * nmax * sizeof(int32_t)));
  RLB_CUDA(cudaMemsetAsync(d_ticket, 0, 4, c->stream));
  rlb::LineLomsacArgs a;
  a.num_problems = np;
  a.offsets = reinterpret_cast<const long long*>(d_offsets);
  a.n_single = n_single;
  a.xy = d_xy;
  a.seeds = d_seeds;
  a.seed_single = seed_single;
  a.opt = *opt;
  a.lines = d_lines;
  a.results = d_results;
  a.inliers = d_inliers;
  a.lists = c->ll_lists.as<int32_t>();
  a.nmax = nmax;
  a.ticket = d_ticket;
  a.smem_points = smem_points;
  RLB_CUDA(rlb::launch_line_lomsac(a, ctas, c->stream, &c->launches_i));
  account(c);
  return RLB_OK;
}

}  // namespace

void rlb_lomsac_default_options(rlb_lomsac_options* o) {
  // RansacOptions / LORansacOptions defaults (ransac.h:47-92)
  o->min_iters = 100u;
  o->max_iters = 10000u;
  o->success_prob = 0.9999;
  o->sq_thr = 1.0;
  o->lo_steps = 10;
  o->thr_mult = std::sqrt(2.0);
  o->lsq_iters = 4;
  o->min_sample_mult = 7;
  o->non_min_sample_mult = 3;
  o->lo_start = 50u;
  o->final_lsq = 0;
  o->debug_flags = 0;
  o->handback_after = 0u;
  o->handback_min_remaining = 0u;
}

int rlb_line_lomsac_batch_dev(rlb_ctx* c, int np, int max_points, const int64_t* d_offsets, const double* d_xy,
                              const uint32_t* d_seeds, const rlb_lomsac_options* opt, double* d_lines,
                              rlb_lomsac_result* d_results, int32_t* d_inliers) {
  if (!c) return fail(RLB_E_ARG, "null context");
  if (np <= 0) return fail(RLB_E_ARG, "num_problems must be positive (got %d)", np);
  if (!d_offsets || !d_xy || !d_seeds || !d_lines || !d_results) return fail(RLB_E_ARG, "null device pointer");
  int rc = check_line_lomsac_options(opt);
  if (rc) return rc;
  rc = ctx_use(c);
  if (rc) return rc;
  RLB_CUDA(c->ll_io.ensure(256));
  return line_lomsac_enqueue(c, np, max_points, d_offsets, 0, d_xy, d_seeds, 0u, opt, d_lines, d_results,
                             d_inliers, c->ll_io.as<unsigned int>());
}

int rlb_line_lomsac_batch(rlb_ctx* c, int np, const int64_t* offsets, const double* xy, const uint32_t* seeds,
                          const rlb_lomsac_options* opt, double* lines, rlb_lomsac_result* results,
                          int32_t* inliers) {
  if (!c) return fail(RLB_E_ARG, "null context");
  if (np <= 0) return fail(RLB_E_ARG, "num_problems must be positive (got %d)", np);
  if (!offsets || !xy || !seeds || !lines || !results) return fail(RLB_E_ARG, "null pointer");
  int rc = check_line_lomsac_options(opt);
  if (rc) return rc;
  rc = ctx_use(c);
  if (rc) return rc;
  if (offsets[0] != 0) return fail(RLB_E_ARG, "offsets[0] must be 0");
  int64_t max_points = 0;
  for (int q = 0; q < np; ++q) {
    const int64_t n = offsets[q + 1] - offsets[q];
    if (n < 0 || n > 0x7fffffff) return fail(RLB_E_ARG, "problem %d has %lld points", q, (long long)n);
    max_points = std::max(max_points, n);
  }
  const int64_t total = offsets[np];
  const size_t tot = static_cast<size_t>(std::max<int64_t>(total, 1));
  void *d_off = nullptr, *d_xy = nullptr, *d_seed = nullptr, *d_line = nullptr, *d_res = nullptr, *d_inl = nullptr;
  auto release = [&]() {
    for (void* p : {d_off, d_xy, d_seed, d_line, d_res, d_inl}) pool_release(c, p);
  };
  rc = pool_alloc(c, (np + 1) * sizeof(int64_t), &d_off);
  if (!rc) rc = pool_alloc(c, tot * 2 * sizeof(double), &d_xy);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(np) * sizeof(uint32_t), &d_seed);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(np) * 3 * sizeof(double), &d_line);
  if (!rc) rc = pool_alloc(c, static_cast<size_t>(np) * sizeof(rlb_lomsac_result), &d_res);
  if (!rc && inliers) rc = pool_alloc(c, tot * sizeof(int32_t), &d_inl);
  if (rc) {
    release();
    return rc;
  }
  cudaStream_t st = c->stream;
  auto body = [&]() -> int {
    RLB_CUDA(c->ll_io.ensure(256));
    RLB_CUDA(cudaMemcpyAsync(d_off, offsets, (np + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    if (total > 0) RLB_CUDA(cudaMemcpyAsync(d_xy, xy, tot * 2 * sizeof(double), cudaMemcpyHostToDevice, st));
    RLB_CUDA(cudaMemcpyAsync(d_seed, seeds, np * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    RLB_CUDA(cudaMemcpyAsync(d_line, lines, static_cast<size_t>(np) * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
    const int rc2 = line_lomsac_enqueue(c, np, static_cast<int>(max_points), static_cast<int64_t*>(d_off), 0,
                                        static_cast<double*>(d_xy), static_cast<uint32_t*>(d_seed), 0u, opt,
                                        static_cast<double*>(d_line), static_cast<rlb_lomsac_result*>(d_res),
                                        static_cast<int32_t*>(d_inl), c->ll_io.as<unsigned int>());
    if (rc2) return rc2;
    RLB_CUDA(cudaMemcpyAsync(lines, d_line, static_cast<size_t>(np) * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    RLB_CUDA(cudaMemcpyAsync(results, d_res, static_cast<size_t>(np) * sizeof(rlb_lomsac_result),
                             cudaMemcpyDeviceToHost, st));
    if (inliers && total > 0)
      RLB_CUDA(cudaMemcpyAsync(inliers, d_inl, static_cast<size_t>(total) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    RLB_CUDA(cudaStreamSynchronize(st));
    return RLB_OK;
  };
  rc = body();
  if (rc) cudaStreamSynchronize(st);
  release();
  return rc;
}

int rlb_scorer_line_lomsac(rlb_scorer* s, const rlb_lomsac_options* opt, uint32_t seed, double* line,
                           rlb_lomsac_result* result, int32_t* inliers) {
  int rc = check_scorer(s);
  if (rc) return rc;
  rlb_ctx* c = s->ctx;
  if (s->d.kind != RLB_KIND_LINE) return fail(RLB_E_ARG, "rlb_scorer_line_lomsac needs a line scorer");
  if (!line || !result) return fail(RLB_E_ARG, "null pointer");
  rc = check_line_lomsac_options(opt);
  if (rc) return rc;
  rc = ctx_use(c);
  if (rc) return rc;
  const int n = s->d.n[0];
  // ll_io: [0] ticket, [64] line (3 doubles), [128] result, [256] inlier list
  constexpr size_t kLine = 64, kRes = 128, kList = 256;
  RLB_CUDA(c->ll_io.ensure(kList + static_cast<size_t>(std::max(n, 1)) * sizeof(int32_t)));
  char* io = static_cast<char*>(c->ll_io.p);
  char* hs = static_cast<char*>(c->h_small);
  memcpy(hs + 256, line, 3 * sizeof(double));
  RLB_CUDA(cudaMemcpyAsync(io + kLine, hs + 256, 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  rc = line_lomsac_enqueue(c, 1, n, nullptr, n, s->d.raw0, nullptr, seed, opt, reinterpret_cast<double*>(io + kLine),
                           reinterpret_cast<rlb_lomsac_result*>(io + kRes),
                           inliers ? reinterpret_cast<int32_t*>(io + kList) : nullptr,
                           reinterpret_cast<unsigned int*>(io));
  if (rc) return rc;
  RLB_CUDA(cudaMemcpyAsync(hs + 256, io + kLine, 128, cudaMemcpyDeviceToHost, c->stream));
  if (inliers && n > 0)
    RLB_CUDA(cudaMemcpyAsync(inliers, io + kList, static_cast<size_t>(n) * sizeof(int32_t), cudaMemcpyDeviceToHost,
                             c->stream));
  RLB_CUDA(cudaStreamSynchronize(c->stream));
  memcpy(line, hs + 256, 3 * sizeof(double));
  memcpy(result, hs + 256 + (kRes - kLine), sizeof(rlb_lomsac_result));
  return RLB_OK;
}

int rlb_ctx_set_timing(rlb_ctx* c, int enabled) {
  if (!c) return fail(RLB_E_ARG, "null ctx");
  c->timing = enabled != 0;
  c->ev_valid[0] = c->ev_valid[1] = false;
  return RLB_OK;
}

int rlb_ctx_kernel_time_ms(rlb_ctx* c, int which, double* ms) {
  if (!c || !ms || which < 0 || which > 1) return fail(RLB_E_ARG, "bad arguments");
  if (!c->ev_valid[which]) return fail(RLB_E_ARG, "no timed kernel of kind %d yet (enable timing first)", which);
  int rc = ctx_use(c);
  if (rc) return rc;
  RLB_CUDA(cudaEventSynchronize(c->ev[2 * which + 1]));
  float t = 0.f;
  RLB_CUDA(cudaEventElapsedTime(&t, c->ev[2 * which], c->ev[2 * which + 1]));
  *ms = t;
  return RLB_OK;
}

int rlb_ctx_last_sweep_plan(rlb_ctx* c, int* variant, int* splits) {
  if (!c) return fail(RLB_E_ARG, "null context");
  if (variant) *variant = c->last_variant;
  if (splits) *splits = c->last_splits;
  return RLB_OK;
}

int rlb_measure_fma_peak(rlb_ctx* c, int dtype, double* tflops) {
  if (!c || !tflops || dtype < 0 || dtype > 2) return fail(RLB_E_ARG, "bad arguments");
  int rc = ctx_use(c);
  if (rc) return rc;
  float ms = 0;
  double flops = 0;
  RLB_CUDA(rlb::launch_fma_peak(dtype, c->sms, &ms, &flops, c->stream));
  *tflops = flops / (static_cast<double>(ms) * 1e-3) / 1e12;
  return RLB_OK;
}

}  // extern "C"
