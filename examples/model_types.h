// Eigen-free model types of the example solvers. Eigen is not available in this image;
// the memory layouts are Eigen's, so a build that has Eigen can pass Eigen::Vector3d /
// Eigen::Matrix<double,3,4> through the same Pack/Unpack functions (their .data() is the
// same 3 / 12 contiguous doubles).
#ifndef RANSACLIB_B200_EXAMPLES_MODEL_TYPES_H_
#define RANSACLIB_B200_EXAMPLES_MODEL_TYPES_H_

#include <vector>

namespace ransac_lib {

// A 2D line a*x + b*y + c = 0 (reference: Eigen::Vector3d, line_estimator.h:57-58).
struct Line2D {
  double coeff[3];
  double& operator[](int i) { return coeff[i]; }
  const double& operator[](int i) const { return coeff[i]; }
};
typedef std::vector<Line2D> Lines2D;

namespace calibrated_absolute_pose {

// [R | c], column-major 3x4 like Eigen::Matrix<double,3,4>
// (reference calibrated_absolute_pose_estimator.h:55,68-70): R world->camera, c centre.
struct CameraPose {
  double m[12];
  double& operator()(int r, int c) { return m[r + 3 * c]; }
  const double& operator()(int r, int c) const { return m[r + 3 * c]; }
};
typedef std::vector<CameraPose> CameraPoses;

}  // namespace calibrated_absolute_pose
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_MODEL_TYPES_H_
