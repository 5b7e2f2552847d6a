// include/tcs/evaluation.h - scoring a segmentation against ground truth, and the labelled-cloud helper of the app.
//
// Restates src/segmentation_evaluation.cpp:139-165 (overlap table between ground-truth labels and segmentation
// labels over the finite ground-truth points; a maximum-weight one-to-one assignment of ground-truth labels to
// segmentation labels, src/hungarian/hungarian.cpp:8-54; every point whose segmentation label is not the one assigned
// to its ground-truth label is erroneous; an unassigned ground-truth label maps to label 0, as the reference's
// std::map::operator[] does) and include/label_utils.h:13-30 (createLabeledCloudFromColorMap).  The reference links
// the third-party hungarian-v2.0 sources; here the assignment is the O(n^3) shortest-augmenting-path form of the
// Hungarian method with potentials, written for this file.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <map>
#include <vector>

#include "pcl_compat.h"

namespace tcs {
namespace evaluation {

// Maximum-weight assignment on an R x C non-negative weight matrix (row-major).  Returns for every row the column it
// is matched to, or -1 (min(R, C) rows get a column).
inline std::vector<int> max_weight_assignment(const std::vector<long long>& w, int R, int C) {
  std::vector<int> result(R, -1);
  if (R == 0 || C == 0) return result;
  // minimise cost = big - w on an n x m problem with n <= m (transpose when there are more rows than columns)
  const bool transpose = R > C;
  const int n = transpose ? C : R, m = transpose ? R : C;
  long long big = 0;
  for (long long v : w) big = std::max(big, v);
  auto cost = [&](int i, int j) -> long long { return big - (transpose ? w[(size_t)j * C + i] : w[(size_t)i * C + j]); };
  const long long INF = std::numeric_limits<long long>::max() / 4;
  std::vector<long long> u(n + 1, 0), v(m + 1, 0);
  std::vector<int> p(m + 1, 0), way(m + 1, 0);  // p[j] = row matched to column j (1-based, 0 = free)
  for (int i = 1; i <= n; ++i) {
    p[0] = i;
    int j0 = 0;
    std::vector<long long> minv(m + 1, INF);
    std::vector<char> used(m + 1, 0);
    do {
      used[j0] = 1;
      const int i0 = p[j0];
      long long delta = INF;
      int j1 = 0;
      for (int j = 1; j <= m; ++j) {
        if (used[j]) continue;
        const long long cur = cost(i0 - 1, j - 1) - u[i0] - v[j];
        if (cur < minv[j]) { minv[j] = cur; way[j] = j0; }
        if (minv[j] < delta) { delta = minv[j]; j1 = j; }
      }
      for (int j = 0; j <= m; ++j) {
        if (used[j]) { u[p[j]] += delta; v[j] -= delta; }
        else minv[j] -= delta;
      }
      j0 = j1;
    } while (p[j0] != 0);
    do {
      const int j1 = way[j0];
      p[j0] = p[j1];
      j0 = j1;
    } while (j0);
  }
  for (int j = 1; j <= m; ++j) {
    if (!p[j]) continue;
    if (transpose) result[j - 1] = p[j] - 1;
    else result[p[j] - 1] = j - 1;
  }
  return result;
}

struct Score {
  size_t finite_points = 0;     // finite points of the ground-truth cloud
  size_t erroneous_points = 0;  // points whose segmentation label is not the one assigned to their ground-truth label
  std::map<uint32_t, uint32_t> assignment;  // ground-truth label -> segmentation label
  std::vector<size_t> error_indices;
  double error_rate() const { return finite_points ? (double)erroneous_points / (double)finite_points : 0.0; }
};

// labels per point; `finite[i]` = the ground-truth point i has finite coordinates (others are skipped, :143)
inline Score score(const std::vector<uint32_t>& segmentation, const std::vector<uint32_t>& groundtruth,
                   const std::vector<char>& finite) {
  Score s;
  const size_t n = std::min(segmentation.size(), groundtruth.size());
  std::map<uint32_t, int> gt_id, seg_id;
  std::vector<uint32_t> gt_labels, seg_labels;
  for (size_t i = 0; i < n; ++i) {
    if (!finite.empty() && !finite[i]) continue;
    ++s.finite_points;
    if (!gt_id.count(groundtruth[i])) { gt_id[groundtruth[i]] = (int)gt_labels.size(); gt_labels.push_back(groundtruth[i]); }
    if (!seg_id.count(segmentation[i])) { seg_id[segmentation[i]] = (int)seg_labels.size(); seg_labels.push_back(segmentation[i]); }
  }
  const int R = (int)gt_labels.size(), C = (int)seg_labels.size();
  std::vector<long long> overlap((size_t)R * C, 0);
  for (size_t i = 0; i < n; ++i) {
    if (!finite.empty() && !finite[i]) continue;
    ++overlap[(size_t)gt_id[groundtruth[i]] * C + seg_id[segmentation[i]]];
  }
  const std::vector<int> match = max_weight_assignment(overlap, R, C);
  std::vector<uint32_t> assigned(R, 0u);  // unassigned ground-truth labels map to 0
  for (int r = 0; r < R; ++r) {
    if (match[r] >= 0) assigned[r] = seg_labels[match[r]];
    s.assignment[gt_labels[r]] = assigned[r];
  }
  for (size_t i = 0; i < n; ++i) {
    if (!finite.empty() && !finite[i]) continue;
    if (segmentation[i] != assigned[gt_id[groundtruth[i]]]) {
      ++s.erroneous_points;
      s.error_indices.push_back(i);
    }
  }
  return s;
}

inline Score score(const pcl::PointCloud<pcl::PointXYZL>& segmentation, const pcl::PointCloud<pcl::PointXYZL>& groundtruth) {
  const size_t n = std::min(segmentation.size(), groundtruth.size());
  std::vector<uint32_t> a(n), b(n);
  std::vector<char> fin(n);
  for (size_t i = 0; i < n; ++i) {
    a[i] = segmentation[i].label;
    b[i] = groundtruth[i].label;
    fin[i] = std::isfinite(groundtruth[i].x) && std::isfinite(groundtruth[i].y) && std::isfinite(groundtruth[i].z);
  }
  return score(a, b, fin);
}

}  // namespace evaluation

namespace labels {
// include/label_utils.h:13-30: copy the cloud, label = colour of the point's vertex, 0 for points without a vertex
template <typename PointT, typename VertexId>
typename pcl::PointCloud<pcl::PointXYZL>::Ptr createLabeledCloudFromColorMap(const pcl::PointCloud<PointT>& cloud,
                                                                             const std::vector<VertexId>& point_to_vertex_map,
                                                                             const std::vector<uint32_t>& color_map) {
  const VertexId NIL = std::numeric_limits<VertexId>::max();
  typename pcl::PointCloud<pcl::PointXYZL>::Ptr labeled(new pcl::PointCloud<pcl::PointXYZL>);
  labeled->points.resize(cloud.size());
  labeled->width = cloud.width;
  labeled->height = cloud.height;
  for (size_t i = 0; i < cloud.size(); ++i) {
    pcl::PointXYZL& q = labeled->points[i];
    q.x = cloud[i].x; q.y = cloud[i].y; q.z = cloud[i].z;
    const VertexId v = i < point_to_vertex_map.size() ? point_to_vertex_map[i] : NIL;
    q.label = (v == NIL || (size_t)v >= color_map.size()) ? 0u : color_map[v];
  }
  return labeled;
}
}  // namespace labels
}  // namespace tcs
