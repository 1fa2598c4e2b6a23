// File formats of the reference's localization tools (examples/localization.cc,
// localization_with_gt.cc, localization_gc.cc) and the pose output they write:
//   list file : one query per line   name MODEL width height <intrinsics...> [qw qx qy qz cx cy cz]
//   match file: one 2D-3D match per line   u v X Y Z      (file name = query name + postfix)
//   poses file: one line per localized query   name qw qx qy qz tx ty tz   (t = -R c)
// Parsing only; nothing here touches the device.
#ifndef RANSACLIB_B200_EXAMPLES_LOCALIZATION_IO_H_
#define RANSACLIB_B200_EXAMPLES_LOCALIZATION_IO_H_

#include <cmath>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace ransac_lib {
namespace localization_io {

struct Query {
  std::string name;
  std::string camera_model;
  int width = 0, height = 0;
  double focal_x = 0.0, focal_y = 0.0, c_x = 0.0, c_y = 0.0;
  std::vector<double> distortion;  // parsed, never applied (the reference tools do not undistort either)
  bool has_gt = false;
  double gt_q[4] = {1.0, 0.0, 0.0, 0.0};  // w x y z, world -> camera
  double gt_c[3] = {0.0, 0.0, 0.0};       // camera centre
};

// Which of the reference's three list readers to mimic.
enum ListDialect {
  kListPlain,  // localization.cc:66-115      (PINHOLE: fy is read and then overwritten with fx, :91-92)
  kListGt,     // localization_with_gt.cc:89-136 (PINHOLE keeps fy; 7 ground-truth values follow)
  kListGc      // localization_gc.cc:64-96    (SIMPLE_RADIAL and PINHOLE only)
};

// Intrinsics layout per camera model: {number of focal lengths, number of distortion parameters}.
inline bool ModelLayout(const std::string& model, ListDialect dialect, int* num_focals, int* num_dist) {
  if (model == "SIMPLE_RADIAL") {
    *num_focals = 1;
    *num_dist = 1;
    return true;
  }
  if (model == "PINHOLE") {
    *num_focals = 2;
    *num_dist = 0;
    return true;
  }
  if (dialect == kListGc) return false;
  if (model == "VSFM") {
    *num_focals = 1;
    *num_dist = 1;
    return true;
  }
  if (model == "OPENCV") {
    *num_focals = 2;
    *num_dist = 4;
    return true;
  }
  if (model == "BROWN_3_PARAMS") {
    *num_focals = 2;
    *num_dist = 3;
    return true;
  }
  return false;
}

// Returns false (with *error set) on an unreadable file, an unknown camera model or a short line: the
// reference would carry on with uninitialised intrinsics there.
inline bool LoadQueryList(const std::string& filename, ListDialect dialect, std::vector<Query>* queries,
                          std::string* error) {
  std::ifstream in(filename.c_str());
  if (!in.is_open()) {
    *error = "cannot read the image list from " + filename;
    return false;
  }
  queries->clear();
  std::string line;
  int line_no = 0;
  while (std::getline(in, line)) {
    ++line_no;
    if (line.find_first_not_of(" \t\r") == std::string::npos) continue;
    std::istringstream ss(line);
    Query q;
    ss >> q.name >> q.camera_model >> q.width >> q.height;
    int nf = 0, nd = 0;
    if (!ss || !ModelLayout(q.camera_model, dialect, &nf, &nd)) {
      *error = filename + ":" + std::to_string(line_no) + ": unknown camera model or malformed line";
      return false;
    }
    ss >> q.focal_x;
    if (nf == 2) {
      ss >> q.focal_y;
    } else {
      q.focal_y = q.focal_x;
    }
    ss >> q.c_x >> q.c_y;
    q.distortion.resize(nd);
    for (int i = 0; i < nd; ++i) ss >> q.distortion[i];
    if (dialect == kListPlain && q.camera_model == "PINHOLE") q.focal_y = q.focal_x;
    if (dialect == kListGt) {
      ss >> q.gt_q[0] >> q.gt_q[1] >> q.gt_q[2] >> q.gt_q[3] >> q.gt_c[0] >> q.gt_c[1] >> q.gt_c[2];
      q.has_gt = true;
    }
    if (!ss) {
      *error = filename + ":" + std::to_string(line_no) + ": too few values for " + q.camera_model;
      return false;
    }
    queries->push_back(q);
  }
  return true;
}

// Appends the matches of one file to p2d (u v) / p3d (X Y Z); returns their number or -1 if the file
// cannot be read. Lines without five numbers are skipped (the reference pushes whatever the failed
// extraction left behind). invert_y_z: localization.cc:142-146; the principal point is subtracted
// unless the pixels are already centred (localization.cc:231-236).
inline int LoadMatches(const std::string& filename, bool invert_y_z, bool points_centered, double c_x,
                       double c_y, std::vector<double>* p2d, std::vector<double>* p3d) {
  std::ifstream in(filename.c_str());
  if (!in.is_open()) return -1;
  std::string line;
  int count = 0;
  while (std::getline(in, line)) {
    std::istringstream ss(line);
    double u, v, X, Y, Z;
    if (!(ss >> u >> v >> X >> Y >> Z)) continue;
    if (invert_y_z) {
      Y *= -1.0;
      Z *= -1.0;
    }
    if (!points_centered) {
      u -= c_x;
      v -= c_y;
    }
    p2d->push_back(u);
    p2d->push_back(v);
    p3d->push_back(X);
    p3d->push_back(Y);
    p3d->push_back(Z);
    ++count;
  }
  return count;
}

// [R|c] (column-major 3x4) -> (qvec w,x,y,z ; tvec = -R c): what Eigen::Quaterniond(R) followed by
// normalize() yields (localization.cc:291-294, ransaclib_localization.cc:147-150).
inline void PoseToQvecTvec(const double* m, double* qvec, double* tvec) {
  double R[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) R[3 * r + c] = m[r + 3 * c];
  for (int r = 0; r < 3; ++r) tvec[r] = -(R[3 * r] * m[9] + R[3 * r + 1] * m[10] + R[3 * r + 2] * m[11]);
  double w, x, y, z;
  const double trace = R[0] + R[4] + R[8];
  if (trace > 0.0) {
    double t = std::sqrt(trace + 1.0);
    w = 0.5 * t;
    t = 0.5 / t;
    x = (R[7] - R[5]) * t;
    y = (R[2] - R[6]) * t;
    z = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    double q[3];
    q[i] = 0.5 * t;
    t = 0.5 / t;
    w = (R[3 * k + j] - R[3 * j + k]) * t;
    q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
    x = q[0];
    y = q[1];
    z = q[2];
  }
  const double nq = std::sqrt(w * w + x * x + y * y + z * z);
  if (nq > 0.0) {
    w /= nq;
    x /= nq;
    y /= nq;
    z /= nq;
  }
  qvec[0] = w;
  qvec[1] = x;
  qvec[2] = y;
  qvec[3] = z;
}

// Pose error against ground truth as localization_with_gt.cc:327-334: |c - c_gt| and the angle (degrees)
// of R^T R_gt.
inline void PoseError(const double* pose /* [R|c] column-major */, const double* gt_q, const double* gt_c,
                      double* position_error, double* orientation_error_deg) {
  const double dx = pose[9] - gt_c[0], dy = pose[10] - gt_c[1], dz = pose[11] - gt_c[2];
  *position_error = std::sqrt(dx * dx + dy * dy + dz * dz);
  double n = std::sqrt(gt_q[0] * gt_q[0] + gt_q[1] * gt_q[1] + gt_q[2] * gt_q[2] + gt_q[3] * gt_q[3]);
  if (!(n > 0.0)) n = 1.0;
  const double w = gt_q[0] / n, x = gt_q[1] / n, y = gt_q[2] / n, z = gt_q[3] / n;
  const double G[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z),     2 * (x * z + w * y),
                       2 * (x * y + w * z),     1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                       2 * (x * z - w * y),     2 * (y * z + w * x),     1 - 2 * (x * x + y * y)};
  // trace(R^T G) = sum_ij R_ij G_ij ; R_ij = pose[i + 3 j]
  double tr = 0.0;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) tr += pose[i + 3 * j] * G[3 * i + j];
  double c = 0.5 * (tr - 1.0);
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  *orientation_error_deg = std::acos(c) * 180.0 / M_PI;
}

}  // namespace localization_io
}  // namespace ransac_lib

#endif  // RANSACLIB_B200_EXAMPLES_LOCALIZATION_IO_H_
