// tcs_b200/cpp/segmentation_evaluation.cpp - headless equivalent of the reference's segmentation_evaluation tool
// (src/segmentation_evaluation.cpp:17-275): loads a segmentation and a ground-truth cloud (PointXYZL), assigns
// ground-truth labels to segmentation labels (maximum-weight assignment) and counts erroneous points.  CPU only.
//   tcs_segmentation_evaluation segmentation.pcd groundtruth.pcd [--errors errors.pcd]
#include <cstdio>
#include <cstring>
#include <string>

#include "tcs/evaluation.h"
#include "tcs/pcd_io.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "Usage: %s segmentation groundtruth [--errors <pcd-file>]\n", argv[0]);
    return 1;
  }
  pcl::PointCloud<pcl::PointXYZL> seg, gt;
  try {
    if (!tcs::pcd::load_labeled(argv[1], seg) || !tcs::pcd::load_labeled(argv[2], gt)) {
      std::fprintf(stderr, "clouds need x/y/z fields\n");
      return 1;
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 1;
  }
  if (seg.size() != gt.size()) {
    std::fprintf(stderr, "Segmentation and groundtruth clouds have different number of points (%zu vs %zu)\n", seg.size(), gt.size());
    return 2;
  }
  const tcs::evaluation::Score s = tcs::evaluation::score(seg, gt);
  std::printf("Points: %zu\nErroneous points: %zu\nError: %.4f %%\n", s.finite_points, s.erroneous_points, 100.0 * s.error_rate());
  for (const auto& kv : s.assignment) std::printf("groundtruth label %u -> segmentation label %u\n", kv.first, kv.second);
  for (int i = 3; i + 1 < argc; ++i)
    if (!std::strcmp(argv[i], "--errors")) {
      pcl::PointCloud<pcl::PointXYZL> err = gt;
      for (auto& p : err.points) p.label = 0;
      for (size_t k : s.error_indices) err.points[k].label = 1;
      tcs::pcd::save_labeled(argv[i + 1], err, "binary");
    }
  return 0;
}
