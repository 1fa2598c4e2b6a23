// A HybridSolver (README.md:164-208 of the reference) whose EvaluateModelOnPoint reaches the GPU as SOURCE:
// the data and the solvers of HybridLineEstimator, but scored through rlb_scorer_create_custom_hybrid - the path
// any user-defined two-data-type solver takes. A run must equal the run of the built-in HybridLineEstimator
// (tests/test_custom_functor.py), which is what makes this the parity check of that path.
#ifndef RANSACLIB_B200_EXAMPLES_HYBRID_CUSTOM_LINE_ESTIMATOR_H_
#define RANSACLIB_B200_EXAMPLES_HYBRID_CUSTOM_LINE_ESTIMATOR_H_

#include "hybrid_line_estimator.h"

namespace ransac_lib {

class HybridCustomLineEstimator : public HybridLineEstimator {
 public:
  HybridCustomLineEstimator(int n0, const double* points, int n1, const double* points_with_normals,
                            const std::vector<double>& prior_probabilities, gpu::Context* ctx = nullptr)
      : HybridLineEstimator(gpu::BorrowData(), n0, points, n1, points_with_normals, prior_probabilities, ctx) {
    // hybrid_line_estimator.cc:100-109: both data types measure the distance of (x, y) to the line; a type-1
    // row also carries its normal (unused by the residual, used by its minimal solver on the host).
    static const char* kSource =
        "__device__ double rlb_eval_t(const double* m, int type, const double* p) {\n"
        "  const double r = m[0] * p[0] + m[1] * p[1] + m[2] * 1.0;\n"
        "  return r * r;\n"
        "}\n";
    scorer_.InitCustomHybrid(ctx ? *ctx : gpu::Context::ThreadDefault(), n0, 2, points, n1, 4, points_with_normals,
                             3, kSource);
  }
};

}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_HYBRID_CUSTOM_LINE_ESTIMATOR_H_
