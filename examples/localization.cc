// Visual localization from 2D-3D match files: the reference's three command-line tools
// (examples/localization.cc, localization_with_gt.cc, localization_gc.cc) as ONE program over the
// batched device-resident LO-MSAC (rlb_localize_batch): all queries of the list are parsed first,
// uploaded once, and localized by a single kernel launch instead of one EstimateModel call per image.
//
//   localization [--gt | --gc] [--dry-run] [--seed S] [--device D] <positional arguments>
//
//   default : images_with_intrinsics outfile inlier_threshold num_lo_steps invert_Y_Z points_centered
//             [match-file postfix]                                   (localization.cc:166-171)
//   --gt    : same arguments; the list carries ground-truth poses, <outfile>.stats.txt and
//             <outfile>.times.txt are written too                    (localization_with_gt.cc:187-222,355)
//   --gc    : images_with_intrinsics outfile [match-file postfix]   (localization_gc.cc:141-164)
//
// The reference seeds every query from std::random_device; so does this tool unless --seed is given
// (query i then uses S + i, which makes a run reproducible and comparable with a CPU run).
// --dry-run parses everything and prints one summary line per query without touching a GPU.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <random>
#include <string>
#include <vector>

#include <ransaclib_b200.h>

#include "localization_io.h"

namespace io = ransac_lib::localization_io;

namespace {

// What differs between the three reference tools.
struct Tool {
  io::ListDialect dialect;
  uint32_t min_iters, max_iters, lo_start;
  int final_lsq;
  double thr_mult;          // threshold_multiplier_ (default sqrt(2), ransac.h:77)
  int skip_if_matches_le;   // "Found only k matches -> skipping image"
  int skip_if_inliers_lt;   // no pose line below this many inliers
};

const Tool kPlain = {io::kListPlain, 200u, 100000u, 60u, 1, std::sqrt(2.0), 3, 0};  // localization.cc:219-252
const Tool kGt = {io::kListGt, 100u, 10000u, 60u, 1, 16.0, 4, 4};                   // localization_with_gt.cc:247-284,320
const Tool kGc = {io::kListGc, 20u, 10000u, 20u, 0, std::sqrt(2.0), 11, 12};        // localization_gc.cc:177-197,232

double Median(std::vector<double> v) {  // localization_with_gt.cc:58-68
  if (v.empty()) return 0.0;
  std::sort(v.begin(), v.end());
  const size_t n = v.size();
  return n % 2u == 1u ? v[n / 2] : 0.5 * (v[n / 2 - 1] + v[n / 2]);
}

}  // namespace

int main(int argc, char** argv) {
  const Tool* tool = &kPlain;
  bool dry_run = false, have_seed = false;
  uint32_t seed = 0;
  int device = 0;
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    const std::string a(argv[i]);
    if (a == "--gt") {
      tool = &kGt;
    } else if (a == "--gc") {
      tool = &kGc;
    } else if (a == "--dry-run") {
      dry_run = true;
    } else if (a == "--seed" && i + 1 < argc) {
      seed = static_cast<uint32_t>(std::strtoul(argv[++i], nullptr, 10));
      have_seed = true;
    } else if (a == "--device" && i + 1 < argc) {
      device = std::atoi(argv[++i]);
    } else {
      pos.push_back(a);
    }
  }
  const bool gc = tool == &kGc;
  std::cout << " usage: " << argv[0] << " [--gt|--gc] [--dry-run] [--seed S] [--device D] images_with_intrinsics outfile "
            << (gc ? "" : "inlier_threshold num_lo_steps invert_Y_Z points_centered ") << "[match-file postfix]"
            << std::endl;
  if (pos.size() < (gc ? 2u : 6u)) return -1;

  const std::string list = pos[0], outfile = pos[1];
  const double thr_px = gc ? 5.0 : std::atof(pos[2].c_str());
  const int lo_steps = gc ? 10 : std::atoi(pos[3].c_str());
  const bool invert_y_z = gc ? true : std::atoi(pos[4].c_str()) != 0;
  const bool centered = gc ? true : std::atoi(pos[5].c_str()) != 0;
  std::string postfix = ".individual_datasets.matches.txt";
  if (pos.size() > (gc ? 2u : 6u)) postfix = pos[gc ? 2 : 6];

  std::vector<io::Query> queries;
  std::string error;
  if (!io::LoadQueryList(list, tool->dialect, &queries, &error)) {
    std::cerr << " ERROR: " << error << std::endl;
    return -1;
  }
  const int num_listed = static_cast<int>(queries.size());
  std::cout << " Found " << num_listed << " query images " << std::endl;

  // Gather every usable query into one concatenated batch.
  std::vector<int> listed;  // batch index -> list index
  std::vector<int64_t> offsets(1, 0);
  std::vector<double> p2d, p3d, fx, fy;
  for (int i = 0; i < num_listed; ++i) {
    const io::Query& q = queries[i];
    const size_t keep2 = p2d.size(), keep3 = p3d.size();
    const int n = io::LoadMatches(q.name + postfix, invert_y_z, centered, q.c_x, q.c_y, &p2d, &p3d);
    if (n < 0) {
      std::cerr << "  ERROR: Could not load matches from " << q.name + postfix << std::endl;
      continue;
    }
    std::cout << " image " << q.name << " has # " << n << " matches as input to RANSAC" << std::endl;
    if (n <= tool->skip_if_matches_le) {
      std::cout << " Found only " << n << " matches for query image " << q.name << " -> skipping image" << std::endl;
      p2d.resize(keep2);
      p3d.resize(keep3);
      continue;
    }
    listed.push_back(i);
    offsets.push_back(offsets.back() + n);
    fx.push_back(q.focal_x);
    fy.push_back(q.focal_y);
  }
  const int Q = static_cast<int>(listed.size());

  if (dry_run) {
    for (int b = 0; b < Q; ++b) {
      const io::Query& q = queries[listed[b]];
      double s2 = 0.0, s3 = 0.0;
      for (int64_t j = offsets[b]; j < offsets[b + 1]; ++j) {
        s2 += p2d[2 * j] + p2d[2 * j + 1];
        s3 += p3d[3 * j] + p3d[3 * j + 1] + p3d[3 * j + 2];
      }
      std::printf("DRY %s %s %d %d %.17g %.17g %lld %.17g %.17g\n", q.name.c_str(), q.camera_model.c_str(),
                  q.width, q.height, q.focal_x, q.focal_y, static_cast<long long>(offsets[b + 1] - offsets[b]), s2, s3);
    }
    return 0;
  }

  std::ofstream ofs(outfile.c_str());
  if (!ofs.is_open()) {
    std::cerr << " ERROR: Cannot write to " << outfile << std::endl;
    return -1;
  }

  rlb_localize_options opt;
  rlb_localize_default_options(&opt, thr_px, lo_steps, tool->min_iters, tool->max_iters);
  opt.min_sample_mult = 7;
  opt.lsq_iters = 4;
  opt.lo_start = tool->lo_start;
  opt.final_lsq = tool->final_lsq;
  opt.thr_mult = tool->thr_mult;

  std::vector<uint32_t> seeds(Q);
  std::random_device rd;
  for (int b = 0; b < Q; ++b) seeds[b] = have_seed ? seed + static_cast<uint32_t>(listed[b]) : rd();

  std::vector<double> poses(static_cast<size_t>(Q) * 12);
  std::vector<rlb_localize_result> res(Q);
  double seconds = 0.0;
  if (Q > 0) {
    rlb_ctx* ctx = nullptr;
    if (rlb_ctx_create(device, &ctx) != RLB_OK) {
      std::cerr << " ERROR: " << rlb_last_error() << std::endl;  // no GPU, no CPU fallback
      return -2;
    }
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = rlb_localize_batch(ctx, Q, offsets.data(), p2d.data(), p3d.data(), fx.data(), fy.data(),
                                      seeds.data(), &opt, poses.data(), res.data(), nullptr);
    seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rc != RLB_OK) {
      std::cerr << " ERROR: " << rlb_last_error() << std::endl;
      return -2;
    }
    rlb_ctx_destroy(ctx);
  }
  std::cout << " LO-MSAC on " << Q << " images took " << seconds << " s (one batched launch)" << std::endl;

  const bool gt = tool == &kGt;
  std::ofstream ofs_times;
  if (gt) ofs_times.open((outfile + ".times.txt").c_str());
  const double kMax = std::numeric_limits<double>::max();
  std::vector<double> pos_err(num_listed, kMax), rot_err(num_listed, kMax);  // localization_with_gt.cc:224-227
  const double pos_thr[4] = {0.05, 0.03, 0.02, 0.01}, rot_thr[4] = {5.0, 3.0, 2.0, 1.0};
  int within[4] = {0, 0, 0, 0};

  for (int b = 0; b < Q; ++b) {
    const int i = listed[b];
    const io::Query& q = queries[i];
    const rlb_localize_result& r = res[b];
    std::cout << "  " << i << " " << q.name << " " << q.focal_x << " " << q.focal_y << std::endl;
    std::cout << "   ... LOMSAC found " << r.num_inliers << " inliers in " << r.num_iterations
              << " iterations with an inlier ratio of " << r.inlier_ratio << std::endl;
    std::cout << "   ... LOMSAC executed " << r.num_lo << " local optimization stages" << std::endl;
    if (gt) ofs_times << seconds / static_cast<double>(Q) << std::endl;  // amortised: the batch is one launch
    if (r.num_inliers < tool->skip_if_inliers_lt) continue;
    const double* pose = &poses[static_cast<size_t>(b) * 12];
    double qv[4], tv[3];
    io::PoseToQvecTvec(pose, qv, tv);
    if (gt) {
      io::PoseError(pose, q.gt_q, q.gt_c, &pos_err[i], &rot_err[i]);
      std::cout << "   pose error: " << pos_err[i] << ", " << rot_err[i] << std::endl;
      for (int k = 0; k < 4; ++k)
        if (pos_err[i] <= pos_thr[k] && rot_err[i] <= rot_thr[k]) within[k] += 1;
    }
    ofs << q.name << " " << qv[0] << " " << qv[1] << " " << qv[2] << " " << qv[3] << " " << tv[0] << " " << tv[1]
        << " " << tv[2] << std::endl;
  }

  if (gt) {  // localization_with_gt.cc:352-394
    std::ofstream ofs_stats((outfile + ".stats.txt").c_str());
    if (!ofs_stats.is_open()) {
      std::cerr << " ERROR: Cannot write to " << outfile << ".stats.txt" << std::endl;
      return -1;
    }
    const double mean_time = num_listed > 0 ? seconds / static_cast<double>(num_listed) : 0.0;
    const double med_pos = Median(pos_err), med_rot = Median(rot_err);
    for (std::ostream* o : {static_cast<std::ostream*>(&std::cout), static_cast<std::ostream*>(&ofs_stats)}) {
      *o << std::endl << " Mean RANSAC time: " << mean_time << " s " << std::endl;
      *o << " Median position error: " << med_pos << "m " << " Median orientation error: " << med_rot << " deg"
         << std::endl;
      for (int k = 0; k < 4; ++k)
        *o << " % images within " << pos_thr[k] * 100.0 << "cm and " << rot_thr[k] << "deg: "
           << static_cast<double>(within[k]) / static_cast<double>(num_listed) * 100.0 << std::endl;
    }
  }
  return 0;
}
