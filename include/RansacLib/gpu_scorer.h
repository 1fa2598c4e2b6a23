// ransaclib_b200 — C++ convenience layer over the C ABI (include/ransaclib_b200.h) and
// the opt-in protocol by which a Solver hands its data set to the GPU scorer.
//
// A solver opts in by exposing
//     const ransac_lib::gpu::DeviceScorer& device_scorer() const;
//     static void PackModel(const Model&, double* out);     // model_size doubles
//     static void UnpackModel(const double* in, Model* m);
// and, optionally, a device-side batched minimal solver:
//     int device_minimal_solver() const;           // solver id for rlb_hypothesize_and_score
//     double device_minimal_solver_threshold() const;   // P3P's 4th-point threshold
// and, optionally, a device-resident driver for the WHOLE of EstimateModel (one launch instead of
// one launch per pass; used when the Sampler is the default UniformSampling and RLB_HOST_DRIVER=1 is
// not set in the environment):
//     bool device_lomsac(const rlb_lomsac_options&, uint32_t seed, double* packed_model /* in/out */,
//                        rlb_lomsac_result*, std::vector<int>* inliers) const;   // false: not applicable
// Solvers that do not opt in do not compile against LocallyOptimizedMSAC: there is no
// CPU scoring path in this library.
#ifndef RANSACLIB_B200_RANSACLIB_GPU_SCORER_H_
#define RANSACLIB_B200_RANSACLIB_GPU_SCORER_H_

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../ransaclib_b200.h"

namespace ransac_lib {
namespace gpu {

// Failures of the device path are loud: the reference's API has no error channel
// (EstimateModel returns an inlier count), and silently returning 0 would look like
// "no model found".
inline void Check(int rc, const char* what) {
  if (rc != RLB_OK) {
    std::string msg = std::string("ransaclib_b200: ") + what + " failed: " + rlb_last_error();
    throw std::runtime_error(msg);
  }
}

class Context {
 public:
  explicit Context(int device = 0) { Check(rlb_ctx_create(device, &ctx_), "rlb_ctx_create"); }
  ~Context() { rlb_ctx_destroy(ctx_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  rlb_ctx* get() const { return ctx_; }
  // One lazily created context per host thread and device (contexts are not thread-safe).
  static Context& ThreadDefault(int device = 0) {
    thread_local std::vector<std::unique_ptr<Context>> per_device;
    if (static_cast<int>(per_device.size()) <= device) per_device.resize(device + 1);
    if (!per_device[device]) per_device[device].reset(new Context(device));
    return *per_device[device];
  }

 private:
  rlb_ctx* ctx_ = nullptr;
};

// Constructor tag of the example solvers: keep pointers to the caller's point arrays instead of copying them
// (the arrays must outlive the solver). The reference's estimators copy into Eigen matrices; for a 4.8 MB
// data set that copy costs as much as its upload.
struct BorrowData {};

inline int PrecisionFromEnv() {
  const char* p = std::getenv("RLB_PRECISION");
  return (p && (p[0] == 'f' || p[0] == 'F') && p[1] == '3') ? RLB_F32 : RLB_F64;
}

// Several GPUs of this box driven by this process (rlb_comm_create_local: ncclCommInitAll, one context per
// device owned by the communicator). Creating a communicator costs ~100 ms, so there is one per device count
// for the life of the process; its contexts are shared by every multi-GPU scorer, i.e. multi-GPU solvers of one
// process must not run concurrently.
class LocalCommunicator {
 public:
  static rlb_comm* Get(int ndev) {
    static std::vector<rlb_comm*> per_count;  // never destroyed: NCCL teardown at exit is not worth the risk
    if (ndev < 1) throw std::runtime_error("ransaclib_b200: LocalCommunicator needs at least one device");
    if (static_cast<int>(per_count.size()) <= ndev) per_count.resize(ndev + 1, nullptr);
    if (!per_count[ndev]) Check(rlb_comm_create_local(ndev, nullptr, &per_count[ndev]), "rlb_comm_create_local");
    return per_count[ndev];
  }
};

// RLB_NUM_GPUS=G in the environment: solvers that support it (CalibratedAbsolutePoseEstimator) put their data on
// G GPUs and shard large hypothesis batches over them. 0 / unset: one GPU, no communicator.
inline int NumGpusFromEnv() {
  const char* p = std::getenv("RLB_NUM_GPUS");
  return p ? std::atoi(p) : 0;
}

// Owns one rlb_scorer (a data set resident in HBM) - or, after InitPoseMulti, one replica per GPU of a
// communicator, with s_ the (borrowed) replica of the first GPU.
class DeviceScorer {
 public:
  DeviceScorer() = default;
  ~DeviceScorer() { reset(); }
  DeviceScorer(const DeviceScorer&) = delete;
  DeviceScorer& operator=(const DeviceScorer&) = delete;
  DeviceScorer(DeviceScorer&& o) noexcept
      : s_(o.s_), cs_(o.cs_), multi_min_batch_(o.multi_min_batch_), precision_(o.precision_) {
    o.s_ = nullptr;
    o.cs_ = nullptr;
  }
  DeviceScorer& operator=(DeviceScorer&& o) noexcept {
    if (this != &o) {
      reset();
      s_ = o.s_;
      cs_ = o.cs_;
      multi_min_batch_ = o.multi_min_batch_;
      precision_ = o.precision_;
      o.s_ = nullptr;
      o.cs_ = nullptr;
    }
    return *this;
  }
  void reset() {
    if (cs_) {
      rlb_comm_scorer_destroy(cs_);  // destroys the replicas, s_ among them
    } else if (s_) {
      rlb_scorer_destroy(s_);
    }
    s_ = nullptr;
    cs_ = nullptr;
  }

  void InitLine(Context& c, int n, const double* xy) {
    reset();
    Check(rlb_scorer_create_line(c.get(), n, xy, &s_), "rlb_scorer_create_line");
  }
  void InitHybridLine(Context& c, int n0, const double* xy0, int n1, const double* xynn1) {
    reset();
    Check(rlb_scorer_create_hybrid_line(c.get(), n0, xy0, n1, xynn1, &s_),
          "rlb_scorer_create_hybrid_line");
  }
  void InitPose(Context& c, int n, const double* p2d, const double* p3d, double fx, double fy) {
    reset();
    Check(rlb_scorer_create_pose(c.get(), n, p2d, p3d, fx, fy, &s_), "rlb_scorer_create_pose");
  }

  // The same data set on `ndev` GPUs (every GPU uploads 1/ndev of the points over its own PCIe link, the slices
  // are exchanged over NVLink). Hypothesis batches of at least multi_min_batch() samples are cut into contiguous
  // id ranges, one per GPU, swept concurrently, and settled by ONE ncclAllGather (rlb_comm_*, csrc/comm.cu);
  // smaller batches and every K = 1 pass run on the first GPU. The run is the single-GPU run: scores are per
  // hypothesis and the host replays ransac.h:181-240 over them in id order.
  void InitPoseMulti(int ndev, int n, const double* p2d, const double* p3d, double fx, double fy) {
    reset();
    Check(rlb_comm_scorer_create_pose(LocalCommunicator::Get(ndev), n, p2d, p3d, fx, fy, RLB_COMM_ROOT_ALL, &cs_),
          "rlb_comm_scorer_create_pose");
    s_ = rlb_comm_scorer_local(cs_, 0);
  }
  bool multi_gpu() const { return cs_ != nullptr; }
  int multi_min_batch() const { return multi_min_batch_; }
  void set_multi_min_batch(int k) { multi_min_batch_ = k; }

  // Any other solver of the concept: data is n x dims (row i = what EvaluateModelOnPoint(model, i) reads),
  // eval_source defines `__device__ double rlb_eval(const double* model, const double* point)`.
  void InitCustom(Context& c, int n, int dims, const double* data, int model_doubles, const char* eval_source) {
    reset();
    Check(rlb_scorer_create_custom(c.get(), n, dims, data, model_doubles, eval_source, &s_),
          "rlb_scorer_create_custom");
  }

  // The HybridSolver concept with a user-defined functor: two data types (rows of dims0 / dims1 doubles),
  // eval_source defines `__device__ double rlb_eval_t(const double* model, int type, const double* point)`.
  void InitCustomHybrid(Context& c, int n0, int dims0, const double* data0, int n1, int dims1, const double* data1,
                        int model_doubles, const char* eval_source) {
    reset();
    Check(rlb_scorer_create_custom_hybrid(c.get(), n0, dims0, data0, n1, dims1, data1, model_doubles, eval_source, &s_),
          "rlb_scorer_create_custom_hybrid");
  }

  bool valid() const { return s_ != nullptr; }
  rlb_scorer* get() const { return s_; }
  int model_size() const { return rlb_scorer_model_size(s_); }
  int num_types() const { return rlb_scorer_num_types(s_); }
  // Arithmetic of the K x N sweep: RLB_F64 (default) or RLB_F32 (env RLB_PRECISION=f32 or
  // set_precision). With f32 the driver re-scores near-best hypotheses in f64, so the
  // selected models are the same either way.
  int precision() const { return precision_; }
  void set_precision(int p) { precision_ = p; }

  // ScoreModel for K packed models (+ optional validity mask).
  void ScoreModels(int k, const double* models, const uint8_t* valid, const double* sq_thr,
                   const double* weights, double* scores, int precision) const {
    int32_t best_id;
    double best_score;
    Check(rlb_score_models(s_, precision, k, models, valid, sq_thr, weights, scores, nullptr,
                           &best_id, &best_score),
          "rlb_score_models");
  }

  // Samples in, hypotheses + scores out (device-side minimal solver).
  void HypothesizeAndScore(int solver, int k, const int32_t* samples, const double* sq_thr,
                           const double* weights, double* models, uint8_t* valid, double* scores,
                           int precision) const {
    if (cs_ && k >= multi_min_batch_) {
      int64_t gid;
      double gscore;
      Check(rlb_comm_hypothesize_and_score_models(cs_, precision, solver, k, samples, sq_thr, weights, models, valid,
                                                  scores, nullptr, &gid, &gscore, nullptr),
            "rlb_comm_hypothesize_and_score_models");
      return;
    }
    int32_t best_id;
    double best_score;
    Check(rlb_hypothesize_and_score(s_, precision, solver, k, samples, sq_thr, weights, models,
                                    valid, scores, nullptr, &best_id, &best_score, nullptr),
          "rlb_hypothesize_and_score");
  }

  // Whole LO-MSAC run of the 2D line solver over this (line) data set in one launch
  // (csrc/line_lomsac_kernel.cu). line: 3 doubles in/out; inliers: num_data entries.
  void LineLomsac(const rlb_lomsac_options& opt, uint32_t seed, double* line, rlb_lomsac_result* result,
                  int32_t* inliers) const {
    Check(rlb_scorer_line_lomsac(s_, &opt, seed, line, result, inliers), "rlb_scorer_line_lomsac");
  }

  // Fused K=1 pass: inliers at sq_thr_inl (ascending, per type) and, if sq_thr_score is
  // given, the MSAC score at sq_thr_score. idx may be null (count only).
  void InliersAndScore(const double* model, const double* sq_thr_inl, const double* sq_thr_score,
                       const double* weights, int32_t* idx, int32_t* counts, double* score) const {
    Check(rlb_get_inliers(s_, model, sq_thr_inl, sq_thr_score, weights, idx, counts, score),
          "rlb_get_inliers");
  }

  // The same pass; the list stays in the context's pinned result buffer and the pointer to it is returned
  // (valid until the next call on the scorer's context). For callers that read a sample of the list.
  const int32_t* InliersViewAndScore(const double* model, const double* sq_thr_inl, const double* sq_thr_score,
                                     const double* weights, int32_t* counts, double* score) const {
    const int32_t* view = nullptr;
    Check(rlb_get_inliers_view(s_, model, sq_thr_inl, sq_thr_score, weights, &view, counts, score),
          "rlb_get_inliers_view");
    return view;
  }
 private:
  rlb_scorer* s_ = nullptr;
  rlb_comm_scorer* cs_ = nullptr;  // InitPoseMulti: owns the replicas; s_ is the first one
  int multi_min_batch_ = 2048;     // below this a batch is not worth a collective (one GPU sweeps it in < 0.1 ms)
  int precision_ = PrecisionFromEnv();
};

// ---- opt-in detection ----------------------------------------------------------
template <class S, class = void>
struct HasDeviceScorer : std::false_type {};
template <class S>
struct HasDeviceScorer<S, decltype(void(std::declval<const S&>().device_scorer()))>
    : std::true_type {};

template <class S, class = void>
struct HasDeviceMinimalSolver : std::false_type {};
template <class S>
struct HasDeviceMinimalSolver<S, decltype(void(std::declval<const S&>().device_minimal_solver()))>
    : std::true_type {};

template <class S, class = void>
struct HasDeviceLomsac : std::false_type {};
template <class S>
struct HasDeviceLomsac<S, decltype(void(std::declval<const S&>().device_lomsac(
                              std::declval<const rlb_lomsac_options&>(), 0u, static_cast<double*>(nullptr),
                              static_cast<rlb_lomsac_result*>(nullptr), static_cast<std::vector<int>*>(nullptr))))>
    : std::true_type {};

// RLB_HOST_DRIVER=1: keep the host-driven loop (one launch per pass) even where a device-resident
// driver exists; both produce the same run (tests exercise both).
inline bool HostDriverForced() {
  const char* p = std::getenv("RLB_HOST_DRIVER");
  return p && p[0] == '1';
}

}  // namespace gpu
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_RANSACLIB_GPU_SCORER_H_
