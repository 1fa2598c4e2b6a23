// Internal declarations shared by the kernel translation units and the C ABI.
#ifndef RLB_INTERNAL_CUH_
#define RLB_INTERNAL_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>

#include "../../include/ransaclib_b200.h"

namespace rlb {

// cudaFuncSetAttribute and occupancy queries are per DEVICE, and contexts of several devices (and
// several host threads) share this library: "already configured" is remembered per device in an
// atomic bit mask. The guarded calls are idempotent, so two threads racing only repeat them.
struct PerDeviceOnce {
  std::atomic<unsigned long long> mask{0ull};
  static int device() {
    int d = 0;
    cudaGetDevice(&d);
    return d & 63;
  }
  bool done(int dev) const { return (mask.load(std::memory_order_acquire) >> dev) & 1ull; }
  void set(int dev) { mask.fetch_or(1ull << dev, std::memory_order_release); }
};

// capi.cu: the thread-local error message of rlb_last_error() and two context accessors for the
// other translation units (returns `code`).
int set_error(int code, const char* fmt, ...);
void* ctx_stream(const rlb_ctx* c);  // cudaStream_t
int ctx_device(const rlb_ctx* c);
// capi.cu: two-step construction of a pose scorer whose raw arrays the caller fills on the device
int scorer_pose_begin(rlb_ctx* c, int n, double fx, double fy, rlb_scorer** out, double** raw2d, double** raw3d);
int scorer_pose_finish(rlb_scorer* s, const double* host_p3d = nullptr);

// Points of every data type are stored channel-major (SoA) and zero-padded so that
// each type segment starts at a multiple of kSegAlign points. Stages of the K x N
// sweep (kStage points) and tiles of the inlier pass (kInlTile points) therefore
// never straddle two data types.
constexpr int kSegAlign = 1024;
constexpr int kStage = 512;      // points per shared-memory stage of the sweep
constexpr int kNumStages = 3;    // TMA ring depth
constexpr int kSweepThreads = 128;
constexpr int kInlTile = 1024;   // points per CTA of the inlier pass
constexpr int kInlThreads = 256;
constexpr int kMaxTypes = 2;

struct DeviceData {
  int kind;                 // RLB_KIND_*
  int num_types;            // 1 or 2
  int n[kMaxTypes];         // points per type
  int off[kMaxTypes];       // padded start of each type segment (multiple of kSegAlign)
  int npad;                 // total padded points
  int channels;             // 2 (line kinds) or 5 (pose: X,Y,Z,u,v)
  const double* soa64;      // [channels][npad]
  const float* soa32;       // [channels][npad]; pose: channels 3,4 hold -u,-v, NaN in the padding
  double fx, fy;            // pose intrinsics
  double centre[3];         // f64 point subtracted from the coordinates before the SoA repack
  // raw reference-layout copies for the minimal solvers (f64)
  const double* raw0;       // line: xy (n0 x 2); pose: points2D (n x 2)
  const double* raw1;       // hybrid: xynn (n1 x 4); pose: points3D (n x 3)
  const double* rays;       // pose: n x 3 unit bearings
  int model_doubles;        // RLB_KIND_CUSTOM: doubles per model (the built-in kinds: see model_size_of)
};

struct SweepArgs {
  int k;                       // hypotheses
  const double* models;        // K x model_size (device)
  const uint8_t* valid;        // K or nullptr
  double thr[kMaxTypes];
  double w[kMaxTypes];
  int n_splits;                // CTAs along the point axis
  int stages_per_split;        // stages each split covers
  double* part_score;          // [n_splits][k]
  int32_t* part_count;         // [n_splits][k]
  int variant;                 // 0 = default kernels, 1 = scalar-FFMA sweep (A/B timing)
  double* record;              // optional {score, global id (as f64), model...} of the local winner
  long long id_offset;         // first global hypothesis id of this shard
  int want_counts;             // 0: the caller passed no counts array, part_count may be left unwritten
  // Filtered f64 sweep (sweep_tc_kernel.cu): a launch over a device-side LIST of hypotheses
  const int32_t* ids;          // hypothesis kk of the launch is model ids[kk] (nullptr: kk itself)
  const int32_t* k_dev;        // length of the list on the device, <= k (nullptr: k); CTAs beyond it return at once
  int tile_stride;             // prepass: distance between the sampled tiles
  // ... whose results are added to per-hypothesis integers instead of [splits][k] partials (dense list)
  unsigned long long* acc_q;   // score * 2^acc_shift per hypothesis of the list, zero before the launch
  unsigned int* acc_c;         // inlier count
  int acc_shift;
  // finalize: where hypothesis k's result is. pos[k] >= 0: part_score / part_count[.][pos[k]];
  // pos[k] < 0: acc_q / acc_c[~pos[k]]
  const int32_t* pos;
};

struct BestRec {
  double score;
  long long id;
};

struct InlierArgs {
  const double* model;     // device, model_size doubles (nullptr: use model_v)
  double model_v[12];      // model passed by value with the launch (host-pointer API: no H2D copy)
  double thr_inl[kMaxTypes];
  double thr_score[kMaxTypes];
  double w[kMaxTypes];
  int want_score;
  int32_t* idx;            // device out or nullptr
  double* tile_score;      // [tiles] scratch
  unsigned long long* status;  // [tiles] look-back words (epoch-tagged)
  unsigned int* control;   // [0]=ticket, [1]=done
  unsigned int epoch;
  void* header;            // {double score; int32 count[2]; int32 pad}
};

struct InlierHeader {
  double score;
  int32_t count[2];
  int32_t pad[2];
};

// launchers (score_kernels.cu)
cudaError_t launch_sweep(const DeviceData& d, const SweepArgs& a, int precision, cudaStream_t st,
                         int* launches);
int sweep_hyps_per_thread(int kind, int precision, int variant);
int sweep_resident_ctas(int kind, int precision, int variant);
cudaError_t launch_finalize(const DeviceData& d, const SweepArgs& a, double* scores,
                            int32_t* counts, BestRec* block_best, unsigned int* ticket,
                            BestRec* best, double* best_model, cudaStream_t st, int* launches);
cudaError_t launch_inliers(const DeviceData& d, const InlierArgs& a, cudaStream_t st,
                           int* launches);
// K = 1 pass whose scratch is a fixed file-scope device array (score_kernels.cu): allocates nothing
cudaError_t launch_inliers_noalloc(const DeviceData& d, InlierArgs a, cudaStream_t st, int* launches);
int inliers_noalloc_max_points();
cudaError_t launch_eval_errors(const DeviceData& d, int precision, const double* model, int type,
                               double* err, cudaStream_t st, int* launches);
cudaError_t launch_repack(const DeviceData& d, double* soa64, float* soa32, cudaStream_t st,
                          int* launches);
cudaError_t launch_select_winner(const double* records, int world, int rec_doubles, double* winner,
                                 cudaStream_t st, int* launches);
cudaError_t launch_fma_peak(int dtype, int sms, float* ms, double* flops, cudaStream_t st);

// tensor-core sweep (sweep_tc_kernel.cu)
cudaError_t launch_tc_pack(const DeviceData& d, float* tiles, cudaStream_t st, int* launches);
cudaError_t launch_sweep_tc(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                            int* launches);
// filtered f64 sweep: f32 classification on the tensor cores, f64 only for the pairs that could be inliers
cudaError_t launch_sweep_tcx(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                             int* launches);
cudaError_t launch_sweep_tcx_prepass(const DeviceData& d, const SweepArgs& a, const float* tiles, cudaStream_t st,
                                     int* launches);
cudaError_t launch_tcx_partition(const int32_t* flagged, int n_splits, int k, int sample_points, int32_t* lists,
                                 int32_t* pos, int32_t* counts, cudaStream_t st, int* launches);
int tcx_resident_ctas();
// sorted two-speed f32 sweep (variant 12): -u / -v inside the contraction, reciprocal-free filter, quality order
size_t tc2_tiles_bytes(int npad);
cudaError_t launch_tc2_pack(const DeviceData& d, float* tiles, cudaStream_t st, int* launches);
cudaError_t launch_sweep_tc2(const DeviceData& d, const SweepArgs& a, const float* tiles, bool prepass,
                             cudaStream_t st, int* launches);
cudaError_t launch_tc2_order(const int32_t* flagged, int n_splits, int k, int32_t* order, int32_t* pos,
                             cudaStream_t st, int* launches);
int tc2_resident_ctas();
int tc_tile_points();
size_t tc_tiles_bytes(int npad);
int tc_hyps_per_cta();
int tc_resident_ctas();

// launchers (solver_kernels.cu, compiled with -fmad=false)
cudaError_t launch_rays(const DeviceData& d, double* rays, cudaStream_t st, int* launches);
cudaError_t launch_solve_line2pt(const DeviceData& d, int k, const int32_t* samples, double* models,
                                 uint8_t* valid, cudaStream_t st, int* launches);
cudaError_t launch_solve_point_normal(const DeviceData& d, int k, const int32_t* samples,
                                      double* models, uint8_t* valid, cudaStream_t st,
                                      int* launches);
cudaError_t launch_solve_p3p(const DeviceData& d, int k, const int32_t* samples, double sq_thr,
                             double* models, uint8_t* valid, cudaStream_t st, int* launches);

// Device-resident batched localization (localize_kernels.cu, -fmad=false).
// One launch covers the queries [q_first, q_end) of a batch; everything it reads is resident
// before it starts (the host API launches one grid per uploaded chunk, ordered behind that chunk's
// copies by an event: no kernel ever waits on the host).
struct LocalizeArgs {
  int q_first, q_end;
  const long long* offsets;  // [Q + 1]
  const double* p2d;         // [total][2]
  const double* p3d;         // [total][3]
  const double* fx;          // [Q]
  const double* fy;
  const uint32_t* seeds;
  rlb_localize_options opt;
  double* poses;                 // [Q][12]
  rlb_localize_result* results;  // [Q]
  int32_t* inliers;              // [total] or nullptr
  // scratch of THIS launch, one slot per resident query (mode 0: per warp, mode 1: per CTA)
  float* soa32;    // [slots][5][npad_max]; mode 1 uses it only for queries beyond smem_points
  int32_t* lists;  // [slots][3][nmax]
  int npad_max, nmax;
  unsigned int* ticket;  // query counter of this launch, zero before the launch
  int smem_points;       // mode 1: capacity of the CTA's shared-memory copy of the f32 matches
  int wave;              // mode 1: CTAs per wave (= SM count), rotates the master warp over the schedulers
};
// mode 0: one query per warp (throughput form for very large batches);
// mode 1: one query per CTA, the passes over the matches shared by its four warps.
cudaError_t launch_localize(const LocalizeArgs& a, int mode, int ctas, cudaStream_t st, int* launches);
int localize_warps_per_cta();
int localize_ctas_per_sm(int mode);
int localize_smem_points();  // mode 1: largest query whose f32 copy is staged in shared memory

// Device-resident line LO-MSAC (line_lomsac_kernel.cu, -fmad=false): one problem per CTA.
struct LineLomsacArgs {
  int num_problems;
  const long long* offsets;  // [P + 1], or nullptr for ONE problem of n_single points
  int n_single;
  const double* xy;          // [total][2]
  const uint32_t* seeds;     // [P], or nullptr: seed_single
  uint32_t seed_single;
  rlb_localize_options opt;
  double* lines;                 // [P][3], in/out
  rlb_localize_result* results;  // [P]
  int32_t* inliers;              // [total] or nullptr
  int32_t* lists;                // scratch [ctas][3][nmax]
  int nmax;
  unsigned int* ticket;  // problem counter, zero before the launch
  int smem_points;       // problems up to this size are staged in dynamic shared memory (0: never)
};
cudaError_t launch_line_lomsac(const LineLomsacArgs& a, int ctas, cudaStream_t st, int* launches);
int line_lomsac_smem_points();
int line_lomsac_ctas_per_sm(int smem_points);

inline int model_size_of(int kind) { return kind == RLB_KIND_POSE ? 12 : 3; }
inline int model_size_of(const DeviceData& d) {
  return d.kind == RLB_KIND_CUSTOM ? d.model_doubles : model_size_of(d.kind);
}

// user-defined functors (custom_functor.cu)
struct CustomModule;
bool custom_compile(const char* src, int dims, int md, bool hybrid, std::string* cubin, std::string* err);
CustomModule* custom_load(const std::string& cubin, int dims, int md, std::string* err);
void custom_unload(CustomModule* m);
cudaError_t launch_custom_repack(const double* raw, int n, int width, int dims, int npad, int off, int span,
                                 double* soa, cudaStream_t st, int* launches);
bool launch_custom_sweep(const CustomModule* m, const DeviceData& d, const SweepArgs& a, int pts_per_split,
                         cudaStream_t st, int* launches, std::string* err);
bool launch_custom_errors(const CustomModule* m, const DeviceData& d, const double* d_model, double* d_err,
                          cudaStream_t st, int* launches, std::string* err);

}  // namespace rlb

#endif  // RLB_INTERNAL_CUH_
