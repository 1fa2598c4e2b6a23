// Hypothesis-sharded scoring over the GPUs of one box, with the collective INSIDE the library.
//
// hypotheses are independent, so a round of K hypotheses is cut into
// contiguous id ranges, one per GPU (points replicated on every GPU); each GPU sweeps its range and
// contributes ONE record {score, global id, model} to ONE ncclAllGather per round, and every rank
// picks the lexicographic (score, id) minimum - the lowest-id tie-break of RansacLib/ransac.h:195,288
// survives because the ranges ascend with the rank. The point set reaches the GPUs once: the root
// uploads it over PCIe and ncclBroadcast carries it over NVLink to the others.
//
// Two ways to form the communicator:
//   rlb_comm_create_local  - ONE process drives several GPUs (what the header-only C++ layer uses:
//                            ncclCommInitAll, one rlb_ctx per device owned by the communicator);
//   rlb_comm_init_rank     - one process per GPU (torchrun / MPI style): rank 0 makes a unique id,
//                            the caller hands it to the other processes, every process joins.
//
// NCCL is dlopen()ed on first use (libnccl.so.2), like NVRTC in custom_functor.cu: the library keeps
// loading on machines without NCCL, and inside a PyTorch process it binds to the NCCL that is
// already loaded instead of pulling in a second copy.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>
#include <new>
#include <vector>

#include "rlb_internal.cuh"

namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) {
      api.error = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "");
      return;
    }
    auto sym = [&](const char* n) -> void* {
      void* p = dlsym(api.handle, n);
      if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + n;
      return p;
    };
    api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  });
  return &api;
}

constexpr int kRecMax = 2 + 12;  // {score, global id, model[<= 12]}

struct LocalRank {
  rlb_ctx* ctx = nullptr;
  bool own_ctx = false;
  int device = 0;
  int rank = 0;
  ncclComm_t nccl = nullptr;
  cudaStream_t stream() const { return static_cast<cudaStream_t>(rlb::ctx_stream(ctx)); }
  // round scratch
  double* d_rec = nullptr;     // [kRecMax] this rank's record
  double* d_gather = nullptr;  // [nranks][kRecMax]
  double* d_winner = nullptr;  // [kRecMax]
  double* h_winner = nullptr;  // pinned
  int32_t* d_samples = nullptr;
  double* d_models = nullptr;
  uint8_t* d_valid = nullptr;
  double* d_scores = nullptr;
  int32_t* d_counts = nullptr;
  size_t k_cap = 0;
};

}  // namespace

struct rlb_comm {
  int nranks = 0;
  std::vector<LocalRank> local;
  int64_t collectives = 0;  // ncclAllGather / ncclBroadcast calls issued (per local rank)
};

struct rlb_comm_scorer {
  rlb_comm* comm = nullptr;
  std::vector<rlb_scorer*> sc;  // one per local rank
  int n = 0;
};

namespace {

using rlb::set_error;

// RLB_COMM_TRACE=1: host timestamps of the phases of a scorer creation / a round on stderr (first local rank)
bool comm_trace() {
  static const bool on = [] {
    const char* v = getenv("RLB_COMM_TRACE");
    return v && v[0] == '1';
  }();
  return on;
}
double now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

#define COMM_CUDA(call)                                                                              \
  do {                                                                                               \
    cudaError_t e__ = (call);                                                                        \
    if (e__ != cudaSuccess)                                                                          \
      return set_error(RLB_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                       __LINE__);                                                                    \
  } while (0)
#define COMM_NCCL(call)                                                                                   \
  do {                                                                                                    \
    ncclResult_t r__ = (call);                                                                            \
    if (r__ != ncclSuccess)                                                                               \
      return set_error(RLB_E_COMM, "%s failed: %s (%s:%d)", #call, nccl_api()->GetErrorString(r__), __FILE__, \
                       __LINE__);                                                                         \
  } while (0)

int need_nccl() {
  NcclApi* a = nccl_api();
  if (!a->handle || !a->error.empty()) return set_error(RLB_E_COMM, "NCCL unavailable: %s", a->error.c_str());
  return RLB_OK;
}

int alloc_rank_scratch(rlb_comm* cm, LocalRank& L) {
  COMM_CUDA(cudaSetDevice(L.device));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_rec), kRecMax * sizeof(double)));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_gather), static_cast<size_t>(cm->nranks) * kRecMax * sizeof(double)));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_winner), kRecMax * sizeof(double)));
  COMM_CUDA(cudaMallocHost(reinterpret_cast<void**>(&L.h_winner), kRecMax * sizeof(double)));
  return RLB_OK;
}

int ensure_round_buffers(LocalRank& L, size_t k, int sample_ints, int model_doubles) {
  if (k <= L.k_cap) return RLB_OK;
  COMM_CUDA(cudaSetDevice(L.device));
  COMM_CUDA(cudaStreamSynchronize(L.stream()));
  for (void* p : {static_cast<void*>(L.d_samples), static_cast<void*>(L.d_models), static_cast<void*>(L.d_valid),
                  static_cast<void*>(L.d_scores), static_cast<void*>(L.d_counts)})
    if (p) cudaFree(p);
  L.d_samples = nullptr;
  L.d_models = nullptr;
  L.d_valid = nullptr;
  L.d_scores = nullptr;
  L.d_counts = nullptr;
  L.k_cap = 0;
  const size_t cap = k + k / 4;
  (void)sample_ints;
  (void)model_doubles;
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_samples), cap * 4 * sizeof(int32_t)));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_models), cap * 12 * sizeof(double)));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_valid), cap));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_scores), cap * sizeof(double)));
  COMM_CUDA(cudaMalloc(reinterpret_cast<void**>(&L.d_counts), cap * sizeof(int32_t)));
  L.k_cap = cap;
  return RLB_OK;
}

void free_rank(LocalRank& L) {
  cudaSetDevice(L.device);
  if (L.ctx) cudaStreamSynchronize(L.stream());
  if (L.nccl) nccl_api()->CommDestroy(L.nccl);
  for (void* p : {static_cast<void*>(L.d_rec), static_cast<void*>(L.d_gather), static_cast<void*>(L.d_winner),
                  static_cast<void*>(L.d_samples), static_cast<void*>(L.d_models), static_cast<void*>(L.d_valid),
                  static_cast<void*>(L.d_scores), static_cast<void*>(L.d_counts)})
    if (p) cudaFree(p);
  if (L.h_winner) cudaFreeHost(L.h_winner);
  if (L.own_ctx && L.ctx) rlb_ctx_destroy(L.ctx);
}

// Contiguous, ascending id range of rank r (ransaclib_b200/sharding.py: hypothesis_range).
void shard_range(int k_total, int rank, int world, int* lo, int* hi) {
  const int base = k_total / world, rem = k_total % world;
  *lo = rank * base + std::min(rank, rem);
  *hi = *lo + base + (rank < rem ? 1 : 0);
}

// all_gather of the ranks' records + winner selection on every local rank (enqueued, not synchronised).
int gather_and_select(rlb_comm* cm, int rec_doubles) {
  NcclApi* a = nccl_api();
  if (cm->nranks > 1) {
    COMM_NCCL(a->GroupStart());
    for (LocalRank& L : cm->local) {
      ncclResult_t r = a->AllGather(L.d_rec, L.d_gather, static_cast<size_t>(rec_doubles), ncclDouble, L.nccl, L.stream());
      if (r != ncclSuccess) {
        a->GroupEnd();
        return set_error(RLB_E_COMM, "ncclAllGather failed: %s", a->GetErrorString(r));
      }
    }
    COMM_NCCL(a->GroupEnd());
    ++cm->collectives;
  }
  for (LocalRank& L : cm->local) {
    const double* src = (cm->nranks > 1) ? L.d_gather : L.d_rec;
    int rc = rlb_select_winner_dev(L.ctx, src, cm->nranks, rec_doubles, L.d_winner);
    if (rc) return rc;
  }
  return RLB_OK;
}

}  // namespace

extern "C" {

int rlb_comm_get_unique_id(void* id) {
  if (!id) return set_error(RLB_E_ARG, "null id");
  int rc = need_nccl();
  if (rc) return rc;
  static_assert(sizeof(ncclUniqueId) <= RLB_COMM_ID_BYTES, "unique id does not fit RLB_COMM_ID_BYTES");
  ncclUniqueId u;
  COMM_NCCL(nccl_api()->GetUniqueId(&u));
  memset(id, 0, RLB_COMM_ID_BYTES);
  memcpy(id, &u, sizeof(u));
  return RLB_OK;
}

int rlb_comm_init_rank(rlb_ctx* ctx, int nranks, int rank, const void* id, rlb_comm** out) {
  if (!ctx || !out || nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !id))
    return set_error(RLB_E_ARG, "bad arguments");
  rlb_comm* cm = new (std::nothrow) rlb_comm();
  if (!cm) return set_error(RLB_E_NOMEM, "out of host memory");
  cm->nranks = nranks;
  cm->local.resize(1);
  LocalRank& L = cm->local[0];
  L.ctx = ctx;
  L.device = rlb::ctx_device(ctx);
  L.rank = rank;
  int rc = RLB_OK;
  if (nranks > 1) {
    rc = need_nccl();
    if (!rc) {
      ncclUniqueId u;
      memcpy(&u, id, sizeof(u));
      cudaSetDevice(L.device);
      ncclResult_t r = nccl_api()->CommInitRank(&L.nccl, nranks, u, rank);
      if (r != ncclSuccess) rc = set_error(RLB_E_COMM, "ncclCommInitRank failed: %s", nccl_api()->GetErrorString(r));
    }
  }
  if (!rc) rc = alloc_rank_scratch(cm, L);
  if (rc) {
    rlb_comm_destroy(cm);
    return rc;
  }
  *out = cm;
  return RLB_OK;
}

int rlb_comm_create_local(int ndev, const int* devices, rlb_comm** out) {
  if (!out || ndev < 1) return set_error(RLB_E_ARG, "bad arguments");
  rlb_comm* cm = new (std::nothrow) rlb_comm();
  if (!cm) return set_error(RLB_E_NOMEM, "out of host memory");
  cm->nranks = ndev;
  cm->local.resize(ndev);
  int rc = RLB_OK;
  std::vector<int> devs(ndev);
  for (int i = 0; i < ndev && !rc; ++i) {
    devs[i] = devices ? devices[i] : i;
    LocalRank& L = cm->local[i];
    L.device = devs[i];
    L.rank = i;
    L.own_ctx = true;
    rc = rlb_ctx_create(devs[i], &L.ctx);
  }
  if (!rc && ndev > 1) {
    rc = need_nccl();
    if (!rc) {
      std::vector<ncclComm_t> comms(ndev);
      ncclResult_t r = nccl_api()->CommInitAll(comms.data(), ndev, devs.data());
      if (r != ncclSuccess) {
        rc = set_error(RLB_E_COMM, "ncclCommInitAll failed: %s", nccl_api()->GetErrorString(r));
      } else {
        for (int i = 0; i < ndev; ++i) cm->local[i].nccl = comms[i];
      }
    }
  }
  for (int i = 0; i < ndev && !rc; ++i) rc = alloc_rank_scratch(cm, cm->local[i]);
  if (rc) {
    rlb_comm_destroy(cm);
    return rc;
  }
  *out = cm;
  return RLB_OK;
}

int rlb_comm_destroy(rlb_comm* cm) {
  if (!cm) return RLB_OK;
  for (LocalRank& L : cm->local) free_rank(L);
  delete cm;
  return RLB_OK;
}

int rlb_comm_nranks(const rlb_comm* cm) { return cm ? cm->nranks : 0; }
int rlb_comm_num_local(const rlb_comm* cm) { return cm ? static_cast<int>(cm->local.size()) : 0; }
int rlb_comm_rank(const rlb_comm* cm, int local_index) {
  if (!cm || local_index < 0 || local_index >= static_cast<int>(cm->local.size())) return -1;
  return cm->local[local_index].rank;
}
rlb_ctx* rlb_comm_ctx(rlb_comm* cm, int local_index) {
  if (!cm || local_index < 0 || local_index >= static_cast<int>(cm->local.size())) return nullptr;
  return cm->local[local_index].ctx;
}
int64_t rlb_comm_collective_count(const rlb_comm* cm) { return cm ? cm->collectives : 0; }

int rlb_comm_scorer_create_pose(rlb_comm* cm, int n, const double* p2d, const double* p3d, double fx, double fy,
                                int root, rlb_comm_scorer** out) {
  if (!cm || !out || n < 0 || root < -1 || root >= cm->nranks) return set_error(RLB_E_ARG, "bad arguments");
  const bool everyone = (root == RLB_COMM_ROOT_ALL);
  bool have_root = everyone;
  for (LocalRank& L : cm->local) have_root = have_root || (L.rank == root);
  if (have_root && n > 0 && (!p2d || !p3d)) return set_error(RLB_E_ARG, "the root rank needs the host arrays");
  rlb_comm_scorer* cs = new (std::nothrow) rlb_comm_scorer();
  if (!cs) return set_error(RLB_E_NOMEM, "out of host memory");
  cs->comm = cm;
  cs->n = n;
  cs->sc.assign(cm->local.size(), nullptr);
  NcclApi* a = nccl_api();
  const int G = cm->nranks;
  const double t_begin = comm_trace() ? now_ms() : 0.0;
  double t_up = 0.0, t_nccl = 0.0, t_sync = 0.0;
  std::vector<double*> r2(cm->local.size(), nullptr), r3(cm->local.size(), nullptr);
  // every rank allocates its scorer; host data goes up over PCIe: all of it on the root, or - when every
  // process holds the arrays - the slice [n r / G, n (r + 1) / G) on GPU r, each over its own link
  for (size_t i = 0; i < cm->local.size(); ++i) {
    LocalRank& L = cm->local[i];
    cudaSetDevice(L.device);
    int rc = rlb::scorer_pose_begin(L.ctx, n, fx, fy, &cs->sc[i], &r2[i], &r3[i]);
    int lo = 0, hi = n;
    if (everyone && G > 1) shard_range(n, L.rank, G, &lo, &hi);
    if (!rc && (everyone || L.rank == root) && hi > lo) {
      const size_t m = static_cast<size_t>(hi - lo), o = static_cast<size_t>(lo);
      if (cudaMemcpyAsync(r2[i] + 2 * o, p2d + 2 * o, m * 2 * sizeof(double), cudaMemcpyHostToDevice, L.stream()) != cudaSuccess ||
          cudaMemcpyAsync(r3[i] + 3 * o, p3d + 3 * o, m * 3 * sizeof(double), cudaMemcpyHostToDevice, L.stream()) != cudaSuccess)
        rc = set_error(RLB_E_CUDA, "upload of the point set failed");
    }
    if (rc) {
      rlb_comm_scorer_destroy(cs);
      return rc;
    }
  }
  if (comm_trace()) t_up = now_ms();
  // ... and reaches the other GPUs over NVLink, straight into their scorers' arrays: one broadcast per array
  // from the root, or - every rank holding a slice - two all-gathers (equal slices) / one broadcast per rank's slice
  // (uneven ones)
  if (G > 1 && n > 0) {
    ncclResult_t r = a->GroupStart();
    for (size_t i = 0; i < cm->local.size() && r == ncclSuccess; ++i) {
      LocalRank& L = cm->local[i];
      if (!everyone) {
        r = a->Broadcast(r2[i], r2[i], static_cast<size_t>(n) * 2, ncclDouble, root, L.nccl, L.stream());
        if (r == ncclSuccess) r = a->Broadcast(r3[i], r3[i], static_cast<size_t>(n) * 3, ncclDouble, root, L.nccl, L.stream());
        continue;
      }
      if (n % G == 0) {
        // equal slices: two in-place all-gathers (rank r's slice already sits at offset r * n / G of its arrays)
        const size_t per = static_cast<size_t>(n / G);
        r = a->AllGather(r2[i] + 2 * per * L.rank, r2[i], per * 2, ncclDouble, L.nccl, L.stream());
        if (r == ncclSuccess)
          r = a->AllGather(r3[i] + 3 * per * L.rank, r3[i], per * 3, ncclDouble, L.nccl, L.stream());
        continue;
      }
      for (int src = 0; src < G && r == ncclSuccess; ++src) {
        int lo, hi;
        shard_range(n, src, G, &lo, &hi);
        if (hi <= lo) continue;
        double* b2 = r2[i] + 2 * static_cast<size_t>(lo);
        double* b3 = r3[i] + 3 * static_cast<size_t>(lo);
        r = a->Broadcast(b2, b2, static_cast<size_t>(hi - lo) * 2, ncclDouble, src, L.nccl, L.stream());
        if (r == ncclSuccess) r = a->Broadcast(b3, b3, static_cast<size_t>(hi - lo) * 3, ncclDouble, src, L.nccl, L.stream());
      }
    }
    ncclResult_t rg = a->GroupEnd();
    if (r == ncclSuccess) r = rg;
    if (r != ncclSuccess) {
      set_error(RLB_E_COMM, "ncclBroadcast failed: %s", a->GetErrorString(r));
      rlb_comm_scorer_destroy(cs);
      return RLB_E_COMM;
    }
    cm->collectives += (everyone && n % G != 0) ? 2 * G : 2;
  }
  if (comm_trace()) {
    t_nccl = now_ms();
    cudaSetDevice(cm->local[0].device);
    cudaStreamSynchronize(cm->local[0].stream());
    t_sync = now_ms();
  }
  // every rank derives the same recentring point from the same bytes and builds its SoA copies
  for (size_t i = 0; i < cm->local.size(); ++i) {
    cudaSetDevice(cm->local[i].device);
    int rc = rlb::scorer_pose_finish(cs->sc[i], (everyone && n > 0) ? p3d : nullptr);
    if (rc) {
      rlb_comm_scorer_destroy(cs);
      return rc;
    }
  }
  if (comm_trace() && cm->local[0].rank == 0)
    fprintf(stderr, "[rlb_comm trace] create_pose n=%d G=%d: alloc+upload enqueue %.3f ms, nccl enqueue %.3f ms, "
            "copies+exchange done %.3f ms, centre+repack+rays %.3f ms\n", n, G, t_up - t_begin, t_nccl - t_up,
            t_sync - t_nccl, now_ms() - t_sync);
  *out = cs;
  return RLB_OK;
}

int rlb_comm_scorer_destroy(rlb_comm_scorer* cs) {
  if (!cs) return RLB_OK;
  for (rlb_scorer* s : cs->sc)
    if (s) rlb_scorer_destroy(s);
  delete cs;
  return RLB_OK;
}

rlb_scorer* rlb_comm_scorer_local(rlb_comm_scorer* cs, int local_index) {
  if (!cs || local_index < 0 || local_index >= static_cast<int>(cs->sc.size())) return nullptr;
  return cs->sc[local_index];
}

int rlb_comm_round_dev(rlb_comm_scorer* cs, int precision, int solver, int k_local, const int32_t* d_samples,
                       const double* sq_thr, const double* weights, int64_t id_offset, double* d_models,
                       uint8_t* d_valid, double* d_scores, int32_t* d_counts, double* d_winner) {
  if (!cs || !cs->comm) return set_error(RLB_E_ARG, "null communicator scorer");
  rlb_comm* cm = cs->comm;
  if (cm->local.size() != 1)
    return set_error(RLB_E_ARG, "rlb_comm_round_dev is the one-process-per-GPU form (this communicator drives %d GPUs)",
                     static_cast<int>(cm->local.size()));
  if (!d_winner) return set_error(RLB_E_ARG, "d_winner is required");
  LocalRank& L = cm->local[0];
  const int rec = 2 + rlb_scorer_model_size(cs->sc[0]);
  int rc = rlb_sharded_round_dev(cs->sc[0], precision, solver, k_local, d_samples, sq_thr, weights, id_offset, d_models,
                                 d_valid, d_scores, d_counts, L.d_rec);
  if (rc) return rc;
  rc = gather_and_select(cm, rec);
  if (rc) return rc;
  COMM_CUDA(cudaMemcpyAsync(d_winner, L.d_winner, rec * sizeof(double), cudaMemcpyDeviceToDevice, L.stream()));
  return RLB_OK;
}

int rlb_comm_hypothesize_and_score(rlb_comm_scorer* cs, int precision, int solver, int k_total,
                                   const int32_t* samples, const double* sq_thr, const double* weights,
                                   double* scores, int32_t* counts, int64_t* best_id, double* best_score,
                                   double* best_model) {
  return rlb_comm_hypothesize_and_score_models(cs, precision, solver, k_total, samples, sq_thr, weights, nullptr,
                                               nullptr, scores, counts, best_id, best_score, best_model);
}

int rlb_comm_hypothesize_and_score_models(rlb_comm_scorer* cs, int precision, int solver, int k_total,
                                          const int32_t* samples, const double* sq_thr, const double* weights,
                                          double* models, uint8_t* valid, double* scores, int32_t* counts,
                                          int64_t* best_id, double* best_score, double* best_model) {
  if (!cs || !cs->comm) return set_error(RLB_E_ARG, "null communicator scorer");
  if (k_total <= 0 || !samples || !sq_thr) return set_error(RLB_E_ARG, "k_total, samples and sq_thr are required");
  rlb_comm* cm = cs->comm;
  const int kind = rlb_scorer_kind(cs->sc[0]);
  const int ss = (kind == RLB_KIND_POSE) ? 4 : (solver == 1 ? 1 : 2);
  const int msize = rlb_scorer_model_size(cs->sc[0]);
  const int rec = 2 + msize;
  const bool trace = comm_trace() && cm->local[0].rank == 0;
  const double t_begin = trace ? now_ms() : 0.0;
  double t_round = 0.0, t_gather = 0.0;
  // 1. every local rank: its shard of the samples up, solve + sweep + local argmin -> record
  for (size_t i = 0; i < cm->local.size(); ++i) {
    LocalRank& L = cm->local[i];
    int lo, hi;
    shard_range(k_total, L.rank, cm->nranks, &lo, &hi);
    const int k = hi - lo;
    COMM_CUDA(cudaSetDevice(L.device));
    if (k <= 0) {  // more ranks than hypotheses: a record that cannot win
      const double none[kRecMax] = {1.7976931348623157e308, -1.0};
      COMM_CUDA(cudaMemcpyAsync(L.d_rec, none, rec * sizeof(double), cudaMemcpyHostToDevice, L.stream()));
      COMM_CUDA(cudaStreamSynchronize(L.stream()));
      continue;
    }
    int rc = ensure_round_buffers(L, static_cast<size_t>(k), ss, msize);
    if (rc) return rc;
    COMM_CUDA(cudaMemcpyAsync(L.d_samples, samples + static_cast<size_t>(lo) * ss, static_cast<size_t>(k) * ss * sizeof(int32_t),
                              cudaMemcpyHostToDevice, L.stream()));
    rc = rlb_sharded_round_dev(cs->sc[i], precision, solver, k, L.d_samples, sq_thr, weights, lo, L.d_models, L.d_valid,
                               scores ? L.d_scores : nullptr, counts ? L.d_counts : nullptr, L.d_rec);
    if (rc) return rc;
  }
  if (trace) {  // (tracing synchronises between the phases: durations, not the overlapped timeline)
    cudaSetDevice(cm->local[0].device);
    cudaStreamSynchronize(cm->local[0].stream());
    t_round = now_ms();
  }
  // 2. the round's single collective + winner on every rank
  int rc = gather_and_select(cm, rec);
  if (rc) return rc;
  if (trace) {
    cudaSetDevice(cm->local[0].device);
    cudaStreamSynchronize(cm->local[0].stream());
    t_gather = now_ms();
  }
  // 3. results back: the winner from the first local rank, per-hypothesis outputs from each rank's slice
  for (size_t i = 0; i < cm->local.size(); ++i) {
    LocalRank& L = cm->local[i];
    int lo, hi;
    shard_range(k_total, L.rank, cm->nranks, &lo, &hi);
    COMM_CUDA(cudaSetDevice(L.device));
    if (hi > lo && scores)
      COMM_CUDA(cudaMemcpyAsync(scores + lo, L.d_scores, static_cast<size_t>(hi - lo) * sizeof(double), cudaMemcpyDeviceToHost,
                                L.stream()));
    if (hi > lo && counts)
      COMM_CUDA(cudaMemcpyAsync(counts + lo, L.d_counts, static_cast<size_t>(hi - lo) * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                L.stream()));
    if (hi > lo && models)
      COMM_CUDA(cudaMemcpyAsync(models + static_cast<size_t>(lo) * msize, L.d_models,
                                static_cast<size_t>(hi - lo) * msize * sizeof(double), cudaMemcpyDeviceToHost, L.stream()));
    if (hi > lo && valid)
      COMM_CUDA(cudaMemcpyAsync(valid + lo, L.d_valid, static_cast<size_t>(hi - lo), cudaMemcpyDeviceToHost, L.stream()));
    if (i == 0)
      COMM_CUDA(cudaMemcpyAsync(L.h_winner, L.d_winner, rec * sizeof(double), cudaMemcpyDeviceToHost, L.stream()));
  }
  for (LocalRank& L : cm->local) {
    COMM_CUDA(cudaSetDevice(L.device));
    COMM_CUDA(cudaStreamSynchronize(L.stream()));
  }
  if (trace)
    fprintf(stderr, "[rlb_comm trace] round k_total=%d: samples up + solve + sweep + argmin %.3f ms, all-gather + winner "
            "%.3f ms, results back %.3f ms\n", k_total, t_round - t_begin, t_gather - t_round, now_ms() - t_gather);
  const double* w = cm->local[0].h_winner;
  if (best_score) *best_score = w[0];
  if (best_id) *best_id = static_cast<int64_t>(w[1]);
  if (best_model) memcpy(best_model, w + 2, msize * sizeof(double));
  return RLB_OK;
}

int rlb_comm_get_inliers(rlb_comm_scorer* cs, const double* model, const double* sq_thr_inl, const double* sq_thr_score,
                         const double* weights, int32_t* idx, int32_t* counts, double* score) {
  if (!cs || cs->sc.empty()) return set_error(RLB_E_ARG, "null communicator scorer");
  // K = 1 passes do not shard (latency-bound): the first local GPU answers.
  cudaSetDevice(cs->comm->local[0].device);
  return rlb_get_inliers(cs->sc[0], model, sq_thr_inl, sq_thr_score, weights, idx, counts, score);
}

}  // extern "C"
