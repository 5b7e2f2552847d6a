// include/tcs/random_walker.h - pcl::segmentation::randomWalker(graph, ...) on the GPU.
// Same entry points as the reference's free functions (include/random_walker.h:95-137): the graph's edge weights
// and vertex colours (0 = unlabelled, >0 = seed colour) go in, the colours come back overwritten with the labels;
// the 5-argument form also returns the dense potentials and the colour -> column map.  All connected components are
// solved at once by the multi-right-hand-side Jacobi-PCG of libtcs_b200 (tcs_random_walker); components without a
// seed keep colour 0 like in the reference (impl/random_walker.hpp:191-192).
#pragma once

#include <cstdio>
#include <map>

#include "graph/point_cloud_graph.h"

namespace pcl {
namespace segmentation {

namespace detail {
template <typename GraphT>
bool randomWalkerImpl(GraphT& graph, Eigen::MatrixXf* potentials, std::map<uint32_t, size_t>* colors_to_columns_map,
                      const tcs_solver_opts* opts, tcs_solve_info* info_out, uint32_t min_colours = 0) {
  tcs_graph* h = graph.device();
  std::vector<uint32_t>& colours = graph.vertex_colors();
  uint32_t K = min_colours;
  for (size_t v = 0; v < colours.size(); ++v) K = std::max(K, colours[v]);
  tcs_solve_info info;
  std::vector<float> pot;
  if (potentials) pot.resize(colours.size() * (size_t)K);
  const int st = tcs_random_walker(h, colours.data(), (int)K, opts, potentials ? pot.data() : nullptr, &info);
  if (info_out) *info_out = info;
  if (st != TCS_OK && st != TCS_ERR_NOT_CONVERGED) graph.ctx().check(st);
  graph.colours_synced();
  if (potentials) {
    potentials->resize((long)colours.size(), (long)K);
    if (!pot.empty()) std::memcpy(potentials->data(), pot.data(), pot.size() * sizeof(float));  // both column-major
  }
  if (colors_to_columns_map) {
    colors_to_columns_map->clear();
    for (uint32_t c = 1; c <= K; ++c) (*colors_to_columns_map)[c] = c - 1;
  }
  return st == TCS_OK;  // false = "Random walker segmentation failed" (solver did not converge)
}
}  // namespace detail

template <typename GraphT>
bool randomWalker(GraphT& graph) {
  return detail::randomWalkerImpl(graph, nullptr, nullptr, nullptr, nullptr);
}

template <typename GraphT>
bool randomWalker(GraphT& graph, Eigen::MatrixXf& potentials, std::map<uint32_t, size_t>& colors_to_columns_map) {
  return detail::randomWalkerImpl(graph, &potentials, &colors_to_columns_map, nullptr, nullptr);
}

}  // namespace segmentation
}  // namespace pcl
