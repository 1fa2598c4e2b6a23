/*
 * ransaclib_b200 — C ABI of the B200-native MSAC scoring path.
 *
 * This is the drop-in boundary for ONE path of tsattler/RansacLib: hypothesis
 * scoring / inlier collection (RansacLib/ransac.h:278-336, hybrid_ransac.h:386-462)
 * evaluating the example solvers' EvaluateModelOnPoint
 * (examples/calibrated_absolute_pose_estimator.cc:154-167, line_estimator.cc:110-114,
 * hybrid_line_estimator.cc:100-109) plus the batched minimal solvers that produce
 * the hypotheses (calibrated_absolute_pose_estimator.cc:111-137, line_estimator.cc:53-72,
 * hybrid_line_estimator.cc:111-152). The reference has no FFI for this path (it is
 * a header-only C++ template library); the entry points below are what the
 * template layer in include/RansacLib/ and a pybind/ctypes module bind instead of
 * the CPU loops. See INTEGRATION.md for the reference-side binding.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++ or torch types cross this boundary;
 *  - every function returns RLB_OK (0) or a negative RLB_E_* code; the message is
 *    available from rlb_last_error() (thread-local);
 *  - there is NO CPU fallback: without a CUDA device every call fails loudly;
 *  - host layouts are the reference's: points (n,2)/(n,3)/(n,4) row-major doubles
 *    (== column-major Eigen::Matrix2Xd/Matrix4Xd, std::vector<Eigen::Vector3d>),
 *    line model = 3 doubles (a,b,c), pose model = 12 doubles = column-major 3x4
 *    [R | c] (calibrated_absolute_pose_estimator.h:55,68-70);
 *  - functions ending in _dev take DEVICE pointers and only enqueue work on the
 *    context's stream (inputs already resident in HBM); all others take HOST
 *    pointers and return after the results are on the host.
 */
#ifndef RANSACLIB_B200_H_
#define RANSACLIB_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RLB_OK 0
#define RLB_E_CUDA -1     /* CUDA runtime error (no device, launch failure, ...) */
#define RLB_E_ARG -2      /* invalid argument */
#define RLB_E_KIND -3     /* operation not defined for this scorer kind */
#define RLB_E_NOMEM -4
#define RLB_E_COMM -5     /* NCCL error, or NCCL not loadable (multi-GPU entry points only) */

/* arithmetic the K x N sweep runs in. Inlier lists are always collected in f64. */
#define RLB_F64 0 /* reference-exact semantics, scores within 1e-9 rel */
#define RLB_F32 1 /* recentred f32 inner sums, f64 outer sums; within 1e-4 rel */
/* RLB_F64 on pose batches of >= 512 hypotheses classifies every (hypothesis, point) pair in f32 first and
 * (when K x N >= 2^28: below that one launch of the plain sweep is cheaper than the filtered sweep's four)
 * evaluates in f64 only the pairs that could be inliers; a clear outlier adds exactly thr either way
 * (ransac.h:306-309). Counts are those of the plain f64 sweep, scores agree to 1e-12. RLB_F64_PLAIN evaluates
 * every pair in f64 (the round-1 kernel; also what RLB_F64 runs for lines and small batches). */
#define RLB_F64_PLAIN 2
#define RLB_F64_FILTERED 3 /* the filtered sweep whatever the batch size (RLB_F64 picks it from K x N >= 2^28) */
/* Experimental (A/B): the sorted two-speed tensor-core sweep for f32 pose batches - a reciprocal-free filter
 * a^2 + b^2 < thr pz^2 on every pair (-u, -v inside the contraction), the f32 evaluation only for the 8-point groups in
 * which one of a warp's 32 hypotheses is flagged, hypotheses swept in the order of a prepass' flag counts. Same f32
 * class and promises as RLB_F32; measured slower than the default on configs[4] (profiles/
 * r02_tc_sorted_ab.txt), so RLB_F32 never picks it unless RLB_TC_SORTED=1 is set. The codes force / forbid it. */
#define RLB_F32_SORTED 4
#define RLB_F32_UNSORTED 5

#define RLB_KIND_LINE 0        /* examples/line_estimator */
#define RLB_KIND_HYBRID_LINE 1 /* examples/hybrid_line_estimator (2 data types) */
#define RLB_KIND_POSE 2        /* examples/calibrated_absolute_pose_estimator */
#define RLB_KIND_CUSTOM 3      /* user-defined EvaluateModelOnPoint (rlb_scorer_create_custom) */

typedef struct rlb_ctx rlb_ctx;       /* one device + one stream + scratch; not thread-safe */
typedef struct rlb_scorer rlb_scorer; /* one data set resident in HBM; immutable */

const char* rlb_last_error(void);
int rlb_version(void);

/* ---- context ------------------------------------------------------------- */
int rlb_ctx_create(int device, rlb_ctx** ctx);
/* Scorers point into their context. Destroying a context while scorers created on it are alive is allowed (a garbage
 * collector finalises objects in no particular order): the context then lives on until the last of them is destroyed. */
int rlb_ctx_destroy(rlb_ctx* ctx);
/* Run on an existing cudaStream_t (e.g. torch's current stream). The handle is used
 * as given: NULL is CUDA's legacy default stream. rlb_ctx_use_own_stream goes back to
 * the non-blocking stream the context created for itself. */
int rlb_ctx_set_stream(rlb_ctx* ctx, void* cuda_stream);
int rlb_ctx_use_own_stream(rlb_ctx* ctx);
int rlb_ctx_synchronize(rlb_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
int64_t rlb_ctx_launch_count(const rlb_ctx* ctx);
/* Number of SMs of the context's device. */
int rlb_ctx_sm_count(const rlb_ctx* ctx);

/* ---- data sets ("the solver owns the data", README.md:157) --------------- */
/* LineEstimator(data): xy = n x 2 (line_estimator.cc:47-50). */
int rlb_scorer_create_line(rlb_ctx* ctx, int n, const double* xy, rlb_scorer** out);
/* HybridLineEstimator(points, points_with_normals, .): xy0 = n0 x 2, xynn1 = n1 x 4
 * (hybrid_line_estimator.cc:47-56). */
int rlb_scorer_create_hybrid_line(rlb_ctx* ctx, int n0, const double* xy0, int n1,
                                  const double* xynn1, rlb_scorer** out);
/* CalibratedAbsolutePoseEstimator(fx, fy, ., points2D, rays, points3D)
 * (calibrated_absolute_pose_estimator.cc:99-109); rays are derived on the device
 * as PixelsToViewingRays does (.cc:214-227). */
int rlb_scorer_create_pose(rlb_ctx* ctx, int n, const double* points2D,
                           const double* points3D, double fx, double fy, rlb_scorer** out);
/* Same data set from DEVICE pointers (copied; the caller keeps its buffers). Builds the same scorer as the
 * host-pointer constructor, bit for bit. */
int rlb_scorer_create_pose_dev(rlb_ctx* ctx, int n, const double* d_points2D, const double* d_points3D,
                               double fx, double fy, rlb_scorer** out);
/* Any other solver of the README's concept (reference README.md:105-154): its
 *   double EvaluateModelOnPoint(const Model& model, int i) const
 * is handed over as CUDA source defining
 *   __device__ double rlb_eval(const double* model, const double* point);   // squared residual
 * where `model` is the model packed into model_doubles (1..12) doubles and `point` is row i of data
 * (dims doubles, 1..8). The source is compiled once, at creation, with NVRTC for sm_100a (-fmad=false, so the
 * arithmetic is what a strict-IEEE C++ build of the same expression does); a compile error is returned as
 * RLB_E_ARG with the compiler log in rlb_last_error(). The scorer then works with rlb_score_models[_dev],
 * rlb_get_inliers[_dev] and rlb_eval_errors in f64 (RLB_F32 is accepted and runs the same f64 kernel); the
 * device minimal solvers are not available for it (hypotheses come from the host MinimalSolver). */
int rlb_scorer_create_custom(rlb_ctx* ctx, int n, int dims, const double* data, int model_doubles,
                             const char* eval_source, rlb_scorer** out);
/* The HybridSolver concept (README.md:164-208; EvaluateModelOnPoint(model, t, i) as in
 * examples/hybrid_line_estimator.cc:100-109): two data types with rows of dims0 / dims1 doubles, per-type
 * thresholds and weights as for the built-in hybrid line scorer (hybrid_ransac.h:386-495). eval_source defines
 *     __device__ double rlb_eval_t(const double* model, int type, const double* point);
 * `point` has max(dims0, dims1) entries, those beyond the type's own dims are zero. */
int rlb_scorer_create_custom_hybrid(rlb_ctx* ctx, int n0, int dims0, const double* data0, int n1, int dims1,
                                    const double* data1, int model_doubles, const char* eval_source,
                                    rlb_scorer** out);
/* Compiles eval_source exactly as rlb_scorer_create_custom would, without touching a device (usable on a
 * machine with no GPU): RLB_OK and the size of the sm_100a cubin, or the compiler log as above. */
int rlb_custom_compile_check(const char* eval_source, int dims, int model_doubles, size_t* cubin_bytes);
/* Destroy every scorer BEFORE its context: the data-set memory returns to the context's pool. */
int rlb_scorer_destroy(rlb_scorer* s);
int rlb_scorer_kind(const rlb_scorer* s);
int rlb_scorer_num_types(const rlb_scorer* s);
int rlb_scorer_num_data(const rlb_scorer* s, int type);
int rlb_scorer_model_size(const rlb_scorer* s); /* doubles per model: 3, 12 or the custom model_doubles */
/* Copies the device-computed viewing rays (n x 3) back; pose scorers only. */
int rlb_scorer_get_rays(const rlb_scorer* s, double* rays);

/* ---- the hot path --------------------------------------------------------- */
/* ScoreModel + inlier count for K models and the cross-hypothesis argmin
 * (ransac.h:278-303, 311-323; hybrid_ransac.h:386-421).
 *   models  K x model_size doubles
 *   valid   K bytes or NULL; valid[k]==0 means MinimalSolver returned no model
 *           (ransac.h:185): score = DBL_MAX, count = 0, never wins
 *   sq_thr  one squared threshold per data type; weights one per type or NULL (=1)
 *   scores  K doubles (out, may be NULL); counts K int32 (out, may be NULL)
 *   best_id lowest k with the smallest score < DBL_MAX, or -1; best_score its score
 * A model with a non-finite entry scores NaN and never wins. */
int rlb_score_models(rlb_scorer* s, int precision, int k, const double* models,
                     const uint8_t* valid, const double* sq_thr, const double* weights,
                     double* scores, int32_t* counts, int32_t* best_id, double* best_score);

/* Same with device pointers; asynchronous on the context stream. best is a device
 * pointer to {double score; int64 id} (16 bytes). */
int rlb_score_models_dev(rlb_scorer* s, int precision, int k, const double* d_models,
                         const uint8_t* d_valid, const double* sq_thr, const double* weights,
                         double* d_scores, int32_t* d_counts, void* d_best);

/* GetInliers fused with ScoreModel for ONE model (ransac.h:311-336 and the LO
 * pattern ransac.h:386->391, 401->399, 364->369):
 *   sq_thr_inl    per-type squared threshold for the inlier test (strict <)
 *   sq_thr_score  per-type threshold for the MSAC score, or NULL (no score)
 *   idx           out, ascending indices; type-0 list first, then type-1 list; capacity
 *                 sum of num_data; may be NULL to only count
 *   counts        out, inliers per type
 *   score         out or NULL
 * Always f64. */
int rlb_get_inliers(rlb_scorer* s, const double* model, const double* sq_thr_inl,
                    const double* sq_thr_score, const double* weights, int32_t* idx,
                    int32_t* counts, double* score);
/* Same pass, but instead of copying the list out, *idx_view receives a pointer to the context's own pinned
 * host copy of it (per-type lists concatenated, counts[] entries each): valid until the next call on this
 * context. For callers that read a few entries of a long list - a LeastSquaresFit keeps <= 28 of them
 * (ransac.h:409-420). */
int rlb_get_inliers_view(rlb_scorer* s, const double* model, const double* sq_thr_inl,
                         const double* sq_thr_score, const double* weights, const int32_t** idx_view,
                         int32_t* counts, double* score);
/* Device-output variant: d_idx (capacity sum num_data), d_header = {double score;
 * int32 count[2]; int32 pad} on the device; asynchronous. */
int rlb_get_inliers_dev(rlb_scorer* s, const double* d_model, const double* sq_thr_inl,
                        const double* sq_thr_score, const double* weights, int32_t* d_idx,
                        void* d_header);

/* rlb_get_inliers_dev WITHOUT any allocation. (No counterpart in the reference: added
 * for callers that cannot call cudaMalloc - code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it.) Same arguments, same kernel, same results; the pass's scratch (a score and a
 * status word per tile of 1024 points, ticket counters) lives in a fixed file-scope device array of 66 KB that
 * every GPU loading the library holds until exit, instead of the context's growable buffers. Neither this call
 * nor anything it calls uses cudaMalloc, cudaFree or an allocating library routine; the first call on a device
 * also looks up the array's address (cudaGetSymbolAddress), so make it once before capturing a graph.
 * Limits: built-in kinds only (RLB_E_KIND for custom functors); a data set of more than RLB_NOALLOC_MAX_POINTS
 * points (counted with the per-type padding to multiples of 1024) is rejected with RLB_E_ARG and a message.
 * All calls on one device SHARE the array and it keeps the last call's contents (told apart by an epoch tag;
 * the counters are cleared on the stream before each launch): calls that overlap on the GPU - other streams,
 * contexts or host threads - race. Calls on one stream, one after the other, are fine. */
#define RLB_NOALLOC_MAX_POINTS 4194304
int rlb_get_inliers_noalloc_dev(rlb_scorer* s, const double* d_model, const double* sq_thr_inl,
                                const double* sq_thr_score, const double* weights, int32_t* d_idx,
                                void* d_header);

/* Squared residuals of ONE model on every point of one data type (debug / functor
 * parity tests; EvaluateModelOnPoint itself). err: num_data(type) doubles. */
int rlb_eval_errors(rlb_scorer* s, int precision, const double* model, int type, double* err);

/* ---- batched minimal solvers (one sample per thread) ---------------------- */
/* LineEstimator::MinimalSolver / HybridLineEstimator::TwoPointSolver: samples K x 2
 * indices into data type 0 -> K x 3 lines. */
int rlb_solve_line2pt_batch(rlb_scorer* s, int k, const int32_t* samples, double* models,
                            uint8_t* valid);
/* HybridLineEstimator::PointNormalSolver: samples K indices into data type 1. */
int rlb_solve_point_normal_batch(rlb_scorer* s, int k, const int32_t* samples, double* models,
                                 uint8_t* valid);
/* CalibratedAbsolutePoseEstimator::MinimalSolver: samples K x 4; P3P on the first 3,
 * first solution whose error on the 4th is < sq_thr wins. */
int rlb_solve_p3p_batch(rlb_scorer* s, int k, const int32_t* samples, double sq_thr,
                        double* models, uint8_t* valid);

/* Hypothesise + score in one call: samples go up, hypotheses stay on the device.
 *   solver: 0 = the kind's first minimal solver (2-point line / P3P+4th point),
 *           1 = point+normal (hybrid only)
 *   samples: K x sample_size int32 (sample_size 2, 1 or 4)
 *   models/valid/scores/counts: optional host outputs (may be NULL). */
int rlb_hypothesize_and_score(rlb_scorer* s, int precision, int solver, int k,
                              const int32_t* samples, const double* sq_thr,
                              const double* weights, double* models, uint8_t* valid,
                              double* scores, int32_t* counts, int32_t* best_id,
                              double* best_score, double* best_model);
/* Device-pointer variant (samples already in HBM, outputs stay in HBM). */
int rlb_hypothesize_and_score_dev(rlb_scorer* s, int precision, int solver, int k,
                                  const int32_t* d_samples, const double* sq_thr,
                                  const double* weights, double* d_models, uint8_t* d_valid,
                                  double* d_scores, int32_t* d_counts, void* d_best);

/* ---- hypothesis-sharded rounds: the building blocks (rlb_comm_* below puts the collective around them) - */
/* rlb_hypothesize_and_score_dev for this rank's shard; additionally writes the shard's
 * record = {double score; double global_id (= local id + id_offset, -1 if none);
 * double model[model_size]} to d_record. Ranks must own ascending contiguous id ranges so
 * the lowest-id tie-break of ransac.h:195,288 survives the reduction. */
int rlb_sharded_round_dev(rlb_scorer* s, int precision, int solver, int k, const int32_t* d_samples,
                          const double* sq_thr, const double* weights, int64_t id_offset,
                          double* d_models, uint8_t* d_valid, double* d_scores, int32_t* d_counts,
                          double* d_record);
/* Lexicographic (score, global id) minimum over `world` gathered records -> d_winner (one
 * record; id = -1 when no rank had a model). NaN / DBL_MAX scores never win. */
int rlb_select_winner_dev(rlb_ctx* ctx, const double* d_records, int world, int record_doubles,
                          double* d_winner);

/* ---- multi-GPU: hypothesis-sharded scoring with the collective inside the library ----------
 * a round of K hypotheses is cut into contiguous, ascending id ranges, one per GPU (the
 * point set is replicated); every GPU sweeps its range (rlb_sharded_round_dev) and contributes ONE record
 * {score, global id, model} to ONE ncclAllGather per round; every rank then takes the lexicographic
 * (score, id) minimum (rlb_select_winner_dev), i.e. exactly the model the sequential loop of
 * LocallyOptimizedMSAC::EstimateModel (ransac.h:158-240; strict <, lowest id wins ties: :195,288) would
 * hold after those K iterations. The point set goes up ONCE over PCIe (root) and reaches the other GPUs by
 * ncclBroadcast over NVLink. NCCL is dlopen()ed on first use (libnccl.so.2); without it these entry
 * points return RLB_E_COMM. K = 1 passes (GetInliers, LO rescoring) do not shard: replicas only.
 *
 * A communicator drives `num_local` GPUs of this process: all of them (rlb_comm_create_local: one process,
 * ncclCommInitAll, contexts owned by the communicator) or one (rlb_comm_init_rank: one process per GPU,
 * the caller distributes the unique id). */
typedef struct rlb_comm rlb_comm;
typedef struct rlb_comm_scorer rlb_comm_scorer; /* one replicated data set: a scorer per local GPU */
#define RLB_COMM_ID_BYTES 128
int rlb_comm_get_unique_id(void* id /* RLB_COMM_ID_BYTES */);
int rlb_comm_init_rank(rlb_ctx* ctx, int nranks, int rank, const void* id, rlb_comm** out);
int rlb_comm_create_local(int ndev, const int* devices /* NULL: 0..ndev-1 */, rlb_comm** out);
int rlb_comm_destroy(rlb_comm* comm); /* destroy the communicator's scorers first */
int rlb_comm_nranks(const rlb_comm* comm);
int rlb_comm_num_local(const rlb_comm* comm);
int rlb_comm_rank(const rlb_comm* comm, int local_index);
rlb_ctx* rlb_comm_ctx(rlb_comm* comm, int local_index);
int64_t rlb_comm_collective_count(const rlb_comm* comm); /* NCCL collectives issued so far */

/* CalibratedAbsolutePoseEstimator's data on every GPU of the communicator. Collective: every process of
 * the communicator calls it; points2D / points3D are read on the process that owns rank `root` only (one
 * PCIe upload, ncclBroadcast over NVLink). root = RLB_COMM_ROOT_ALL: EVERY process holds the arrays (always
 * true for rlb_comm_create_local); each GPU then uploads 1/G of them over its own PCIe link and the slices
 * are exchanged over NVLink - the upload time divides by the number of GPUs. */
#define RLB_COMM_ROOT_ALL (-1)
int rlb_comm_scorer_create_pose(rlb_comm* comm, int n, const double* points2D, const double* points3D,
                                double fx, double fy, int root, rlb_comm_scorer** out);
int rlb_comm_scorer_destroy(rlb_comm_scorer* cs);
rlb_scorer* rlb_comm_scorer_local(rlb_comm_scorer* cs, int local_index);

/* One round over k_total hypotheses (collective). samples: ALL k_total x sample_size indices, the same
 * array on every process; rank r uploads and scores ids [r*K/G .. (r+1)*K/G). best_id is the GLOBAL
 * hypothesis id (-1: no model anywhere), identical on every rank. scores / counts (optional, k_total
 * entries): the slices of this process's GPUs are filled, the rest is left untouched. */
int rlb_comm_hypothesize_and_score(rlb_comm_scorer* cs, int precision, int solver, int k_total,
                                   const int32_t* samples, const double* sq_thr, const double* weights,
                                   double* scores, int32_t* counts, int64_t* best_id, double* best_score,
                                   double* best_model);
/* The same round with the hypotheses themselves returned (models: k_total x model_size, valid: k_total; the
 * slices of this process's GPUs are filled): what the drop-in LocallyOptimizedMSAC needs to replay
 * ransac.h:181-240 over a sharded batch (gpu::DeviceScorer::InitPoseMulti). */
int rlb_comm_hypothesize_and_score_models(rlb_comm_scorer* cs, int precision, int solver, int k_total,
                                          const int32_t* samples, const double* sq_thr, const double* weights,
                                          double* models, uint8_t* valid, double* scores, int32_t* counts,
                                          int64_t* best_id, double* best_score, double* best_model);
/* Device-pointer form of the same round for the one-process-per-GPU communicator: this rank's shard
 * (k_local hypotheses, global ids id_offset ...) is already in HBM; asynchronous on the context stream.
 * d_winner: {score, global id as f64, model[model_size]} on every rank. */
int rlb_comm_round_dev(rlb_comm_scorer* cs, int precision, int solver, int k_local, const int32_t* d_samples,
                       const double* sq_thr, const double* weights, int64_t id_offset, double* d_models,
                       uint8_t* d_valid, double* d_scores, int32_t* d_counts, double* d_winner);
/* GetInliers of the winning model: answered by the first local GPU (K = 1 passes do not shard). */
int rlb_comm_get_inliers(rlb_comm_scorer* cs, const double* model, const double* sq_thr_inl,
                         const double* sq_thr_score, const double* weights, int32_t* idx, int32_t* counts,
                         double* score);

/* ---- device-resident batched localization ------------------------------------
 * Q independent P3P LO-MSAC runs (pyransaclib.ransaclib_localization,
 * examples/ransaclib_localization.cc:58-165, one call per query image) in ONE launch: each
 * CTA of a persistent grid runs the whole of LocallyOptimizedMSAC::EstimateModel
 * (ransac.h:128-275) for one query at a time — seeded std::mt19937 sampling and shuffles,
 * P3P + 4th-point test, MSAC scoring, local optimisation with the Levenberg-Marquardt pose
 * refinement (calibrated_absolute_pose_estimator.cc:170-212) — without host round trips.
 * Queries are concatenated: query q owns matches [offsets[q], offsets[q+1]). */
typedef struct rlb_localize_options {
  uint32_t min_iters, max_iters;   /* RansacOptions (ransac.h:47-60) */
  double success_prob;
  double sq_thr;                   /* squared inlier threshold, also the P3P 4th-point test */
  int32_t lo_steps;                /* LORansacOptions (ransac.h:64-92) */
  double thr_mult;
  int32_t lsq_iters;
  int32_t min_sample_mult;         /* x4 must be <= 32 */
  int32_t non_min_sample_mult;     /* x4 must be <= 32 */
  uint32_t lo_start;
  int32_t final_lsq;
  int32_t debug_flags;             /* bit 0: force the sequential path of every shuffle (tests);
                                    * bit 1: per-phase cycle counters into the tail of the inlier region (tools) */
  /* Line LO-MSAC only (one CTA = one SM runs a problem): a run that after `handback_after` iterations
   * still needs more than `handback_min_remaining` further hypotheses is HANDED BACK to the caller:
   * result.success = 2, result.num_iterations = iterations done, nothing else written; the caller then
   * runs it where sweeps use the whole GPU (the drop-in LocallyOptimizedMSAC does, and counts it in
   * GpuRunCounters::handbacks). handback_after = 0 (the default): never hand back. */
  uint32_t handback_after;
  uint32_t handback_min_remaining;
} rlb_localize_options;

/* A handed-back run carries its state: when an inlier buffer was passed and the problem has at least
 * RLB_LOMSAC_RESUME_WORDS points, result.num_inliers = RLB_LOMSAC_RESUME_WORDS and the problem's slice of the
 * inlier buffer holds (int32 words)
 *   [0] RLB_LOMSAC_RESUME_MAGIC  [1] iterations done  [2] number_lo_iterations  [3] position of the local
 *   optimisation's std::mt19937 inside its block (0..624)  [4..11] best_model_score, *best_model (a, b, c)
 *   [12..19] score and coefficients of the best minimal model (doubles as lo, hi words)  [20..643] the 624 words
 *   of that engine's current block
 * so that the caller resumes at that iteration (the sampler is a function of seed and iteration: replay it)
 * instead of starting over. Otherwise result.num_inliers = 0: start over (same run either way). */
#define RLB_LOMSAC_RESUME_WORDS 644
#define RLB_LOMSAC_RESUME_MAGIC 0x524c4231

typedef struct rlb_localize_result {
  int32_t success;        /* 0 for queries with <= 3 matches (ransaclib_localization.cc:86-92) */
  int32_t num_inliers;    /* RansacStatistics::best_num_inliers */
  uint32_t num_iterations;
  int32_t num_lo;         /* number_lo_iterations */
  double score;           /* best_model_score */
  double inlier_ratio;
  int32_t passes;         /* driver counters: f64 passes over the matches, LM refinements */
  int32_t lsq_calls;
} rlb_localize_result;

/* ransaclib_localization's fixed options (ransaclib_localization.cc:100-107). */
void rlb_localize_default_options(rlb_localize_options* opt, double inlier_threshold_px,
                                  int number_lo_steps, uint32_t min_num_iterations,
                                  uint32_t max_num_iterations);

/* HOST pointers. offsets: num_queries + 1 entries; points2D: total x 2; points3D: total x 3;
 * fx, fy, seeds: one per query (the reference seeds from std::random_device; here the seed is an
 * argument so runs are reproducible). Outputs: poses = num_queries x 12 (column-major [R | c]),
 * results = num_queries records, inliers = `total` int32 (ascending inlier indices of query q
 * start at offsets[q]; entries past results[q].num_inliers are unspecified) or NULL. The matches go up in a
 * few chunks of queries; each chunk's grid is launched on its own stream behind that chunk's copies
 * (cudaStreamWaitEvent), so upload and compute overlap and no kernel ever waits on the host. */
int rlb_localize_batch(rlb_ctx* ctx, int num_queries, const int64_t* offsets,
                       const double* points2D, const double* points3D, const double* fx,
                       const double* fy, const uint32_t* seeds, const rlb_localize_options* opt,
                       double* poses, rlb_localize_result* results, int32_t* inliers);
/* DEVICE pointers (everything already in HBM, outputs stay in HBM); asynchronous on the
 * context stream. max_matches = largest offsets[q+1]-offsets[q] (sizes the scratch). */
int rlb_localize_batch_dev(rlb_ctx* ctx, int num_queries, int max_matches, const int64_t* d_offsets,
                           const double* d_points2D, const double* d_points3D, const double* d_fx,
                           const double* d_fy, const uint32_t* d_seeds,
                           const rlb_localize_options* opt, double* d_poses,
                           rlb_localize_result* d_results, int32_t* d_inliers);

/* ---- device-resident line LO-MSAC ---------------------------------------------
 * P independent runs of LocallyOptimizedMSAC<Vector3d, ., LineEstimator>::EstimateModel
 * (ransac.h:128-275 over examples/line_estimator.{h,cc}; BASELINE configs[0] is one such run) in
 * ONE launch: each CTA runs the whole control flow for one problem at a time - seeded
 * std::mt19937 sampling and Fisher-Yates shuffles, the 2-point solver, MSAC scoring, local
 * optimisation with the covariance line fit (line_estimator.cc:74-107) - without host round
 * trips. Options and results are the structs of the localization entry points (the fields are
 * those of LORansacOptions / RansacStatistics); min_sample_mult * 2 and non_min_sample_mult * 2
 * must be <= 32. Problems are concatenated: problem p owns points [offsets[p], offsets[p+1]). */
typedef rlb_localize_options rlb_lomsac_options;
typedef rlb_localize_result rlb_lomsac_result;

/* RansacOptions / LORansacOptions defaults (ransac.h:47-92). */
void rlb_lomsac_default_options(rlb_lomsac_options* opt);

/* HOST pointers. xy: total x 2 (x0,y0,x1,y1,... = column-major Matrix2Xd); seeds: one per problem
 * (LORansacOptions::random_seed_). lines: num_problems x 3, IN/OUT: on entry the caller's models
 * (EstimateModel leaves *best_model untouched when nothing is found, and its final least-squares
 * step reads it, ransac.h:257-259), on return the best models. results[p].success = 0 for
 * problems with fewer than 2 points. inliers: `total` int32 (ascending inlier indices of problem
 * p start at offsets[p]) or NULL. */
int rlb_line_lomsac_batch(rlb_ctx* ctx, int num_problems, const int64_t* offsets, const double* xy,
                          const uint32_t* seeds, const rlb_lomsac_options* opt, double* lines,
                          rlb_lomsac_result* results, int32_t* inliers);
/* DEVICE pointers (outputs stay in HBM); asynchronous on the context stream. max_points = largest
 * offsets[p+1]-offsets[p] (sizes the scratch). */
int rlb_line_lomsac_batch_dev(rlb_ctx* ctx, int num_problems, int max_points, const int64_t* d_offsets,
                              const double* d_xy, const uint32_t* d_seeds, const rlb_lomsac_options* opt,
                              double* d_lines, rlb_lomsac_result* d_results, int32_t* d_inliers);
/* One run over the data set of a LINE scorer (already resident): what the drop-in
 * LocallyOptimizedMSAC calls for LineEstimator. line: 3 doubles, in/out; inliers: n int32 or NULL. */
int rlb_scorer_line_lomsac(rlb_scorer* s, const rlb_lomsac_options* opt, uint32_t seed, double* line,
                           rlb_lomsac_result* result, int32_t* inliers);

/* ---- measurement helpers --------------------------------------------------- */
/* When enabled, the context brackets every sweep kernel and every inlier kernel with
 * CUDA events on its stream; rlb_ctx_kernel_time_ms waits for the LAST such kernel
 * and returns its device time (which: 0 = K x N sweep, 1 = inlier pass). */
int rlb_ctx_set_timing(rlb_ctx* ctx, int enabled);
int rlb_ctx_kernel_time_ms(rlb_ctx* ctx, int which, double* ms);
/* Plan of the LAST K x N sweep of this context: kernel variant (9 = tensor-core sweep sweep_pose_tc_kernel,
 * 0..3 = FFMA2 sweeps with 4/1/2/8 hypotheses per thread, 7/8 = f64 sweeps, -1 = generic) and the number of
 * point-axis splits. Reporting only; the planner (capi.cu, score_dev) picks both per call. */
int rlb_ctx_last_sweep_plan(rlb_ctx* ctx, int* variant, int* splits);
/* Register-resident FMA microbenchmark: returns achieved TFLOP/s of the FP32
 * (dtype 1: FFMA, 2: packed FFMA2) or FP64 (dtype 0: DFMA) pipe on this device.
 * Used by bench.py for the FMA-pipe roofline denominator ("of measured"). */
int rlb_measure_fma_peak(rlb_ctx* ctx, int dtype, double* tflops);

#ifdef __cplusplus
}
#endif

#endif /* RANSACLIB_B200_H_ */
