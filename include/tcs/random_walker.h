// include/tcs/random_walker.h - pcl::segmentation::randomWalker(graph, ...) on the GPU.
// Same entry points as the reference's free functions (include/random_walker.h:95-137, impl/random_walker.hpp:283-335):
//   randomWalker (graph)                                              - the graph's own weight / colour maps
//   randomWalker (graph, weights, colors)                             - external property maps
//   randomWalker (graph, weights, colors, potentials, colors_to_columns_map)
// Edge weights and vertex colours (0 = unlabelled, anything else = a seed of that colour) go in, the colours come back
// overwritten with the labels; the 5-argument form also returns the dense potentials (rows = vertices, one column per
// colour) and the colour -> column map.  All connected components are solved at once by the multi-right-hand-side PCG
// of libtcs_b200 (tcs_random_walker); components without a seed keep colour 0 like in the reference
// (impl/random_walker.hpp:191-192).
//
// Colours are arbitrary uint32 ids (the reference maps them through a bimap): the distinct non-zero ids are compacted
// to 1..K before the solve and mapped back afterwards.  Two documented deviations: columns are ordered by ascending
// colour id (the reference: order of first encounter while scanning the rows, impl/random_walker.hpp:153-157) - this
// only decides which of two EXACTLY equal potentials wins, and colors_to_columns_map tells the caller the order; and
// at most TCS_MAX_LABELS (64) distinct colours are supported (the reference has no limit): more makes the call fail
// (message on stderr, returns false) instead of solving.
#pragma once

#include <algorithm>
#include <cstdio>
#include <map>

#include "graph/point_cloud_graph.h"

namespace pcl {
namespace segmentation {

namespace detail {

// min_colours > 0: colours are already 1..min_colours (the class path: one column per seed label, used or not)
template <typename GraphT>
bool randomWalkerImpl(GraphT& graph, Eigen::MatrixXf* potentials, std::map<uint32_t, size_t>* colors_to_columns_map,
                      const tcs_solver_opts* opts, tcs_solve_info* info_out, uint32_t min_colours = 0) {
  tcs_graph* h = graph.device();
  std::vector<uint32_t>& colours = graph.vertex_colors();
  // distinct non-zero colour ids, ascending
  std::vector<uint32_t> ids;
  for (uint32_t c = 1; c <= min_colours; ++c) ids.push_back(c);
  {
    std::vector<uint32_t> seen;
    for (size_t v = 0; v < colours.size(); ++v)
      if (colours[v] > min_colours) seen.push_back(colours[v]);
    std::sort(seen.begin(), seen.end());
    seen.erase(std::unique(seen.begin(), seen.end()), seen.end());
    ids.insert(ids.end(), seen.begin(), seen.end());
  }
  const size_t K = ids.size();
  if (K > (size_t)TCS_MAX_LABELS) {
    std::fprintf(stderr, "[pcl::segmentation::randomWalker] %zu distinct seed colours: this implementation solves at most %d\n", K,
                 TCS_MAX_LABELS);
    return false;
  }
  bool identity = true;
  for (size_t i = 0; i < K; ++i) identity = identity && ids[i] == (uint32_t)(i + 1);
  if (!identity)
    for (size_t v = 0; v < colours.size(); ++v)
      if (colours[v]) colours[v] = (uint32_t)(std::lower_bound(ids.begin(), ids.end(), colours[v]) - ids.begin()) + 1u;
  tcs_solve_info info;
  std::vector<float> pot;
  if (potentials) pot.resize(colours.size() * K);
  const int st = tcs_random_walker(h, colours.data(), (int)K, opts, potentials ? pot.data() : nullptr, &info);
  if (info_out) *info_out = info;
  if (!identity)
    for (size_t v = 0; v < colours.size(); ++v)
      if (colours[v]) colours[v] = ids[colours[v] - 1];
  if (st != TCS_OK && st != TCS_ERR_NOT_CONVERGED) graph.ctx().check(st);
  graph.colours_synced();
  if (potentials) {
    potentials->resize((long)colours.size(), (long)K);
    if (!pot.empty()) std::memcpy(potentials->data(), pot.data(), pot.size() * sizeof(float));  // both column-major
  }
  if (colors_to_columns_map) {
    colors_to_columns_map->clear();
    for (size_t i = 0; i < K; ++i) (*colors_to_columns_map)[ids[i]] = i;
  }
  return st == TCS_OK;  // false = "Random walker segmentation failed" (solver did not converge)
}

// is `map` the graph's own storage (then nothing has to be copied)?
template <typename GraphT, typename Map> inline bool own_map(const GraphT&, const Map&) { return false; }
template <typename GraphT> inline bool own_map(const GraphT& g, const tcs::detail::edge_weight_map<GraphT>& m) { return m.graph() == &g; }
template <typename GraphT> inline bool own_map(const GraphT& g, const tcs::detail::vertex_color_map<GraphT>& m) { return m.graph() == &g; }

template <class Graph, class EdgeWeightMap, class VertexColorMap>
bool randomWalkerMaps(Graph& graph, EdgeWeightMap weights, VertexColorMap colors, Eigen::MatrixXf* potentials,
                      std::map<uint32_t, size_t>* colors_to_columns_map) {
  typedef typename Graph::edge_iterator EdgeIterator;
  if (!own_map(graph, weights)) {
    std::vector<float>& w = graph.edge_weights();
    EdgeIterator ei, ee;
    for (boost::tie(ei, ee) = edges(graph); ei != ee; ++ei) w[(*ei).idx] = (float)weights[*ei];
  }
  const size_t V = graph.num_vertices();
  const bool own_colours = own_map(graph, colors);
  if (!own_colours) {
    std::vector<uint32_t>& c = graph.vertex_colors();
    for (size_t v = 0; v < V; ++v) c[v] = (uint32_t)colors[(typename Graph::vertex_descriptor)v];
  }
  const bool ok = randomWalkerImpl(graph, potentials, colors_to_columns_map, nullptr, nullptr);
  if (!own_colours) {
    const std::vector<uint32_t>& c = graph.vertex_colors();
    for (size_t v = 0; v < V; ++v) colors[(typename Graph::vertex_descriptor)v] = c[v];
  }
  return ok;
}

}  // namespace detail

template <class Graph>
bool randomWalker(Graph& graph) {
  return detail::randomWalkerImpl(graph, nullptr, nullptr, nullptr, nullptr);
}

template <class Graph, class EdgeWeightMap, class VertexColorMap>
bool randomWalker(Graph& graph, EdgeWeightMap weights, VertexColorMap colors) {
  return detail::randomWalkerMaps(graph, weights, colors, nullptr, nullptr);
}

template <class Graph, class EdgeWeightMap, class VertexColorMap, class Color>
bool randomWalker(Graph& graph, EdgeWeightMap weights, VertexColorMap colors, Eigen::MatrixXf& potentials,
                  std::map<Color, size_t>& colors_to_columns_map) {
  std::map<uint32_t, size_t> m;
  const bool ok = detail::randomWalkerMaps(graph, weights, colors, &potentials, &m);
  colors_to_columns_map.clear();
  for (std::map<uint32_t, size_t>::const_iterator it = m.begin(); it != m.end(); ++it) colors_to_columns_map[(Color)it->first] = it->second;
  return ok;
}

// (not in the reference; kept from round 1) the graph's own maps + potentials
template <class Graph>
bool randomWalker(Graph& graph, Eigen::MatrixXf& potentials, std::map<uint32_t, size_t>& colors_to_columns_map) {
  return detail::randomWalkerImpl(graph, &potentials, &colors_to_columns_map, nullptr, nullptr);
}

}  // namespace segmentation
}  // namespace pcl
