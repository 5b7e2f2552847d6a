// include/tcs/graph/common.h - free functions of the reference's include/graph/common.h:65-204 on the GPU:
// computeNormalsAndCurvatures, computeSignedCurvatures, smoothen, createSubgraphsFromConnectedComponents.
#pragma once

#include "point_cloud_graph.h"

namespace pcl {
namespace graph {

// extension: arithmetic mode used by the stages below (TCS_MODE_PARITY reproduces the reference's float32 order)
inline int& arithmeticMode() { static thread_local int mode = TCS_MODE_PARITY; return mode; }

template <typename GraphT>
void computeNormalsAndCurvatures(GraphT& graph, bool neighborhood_1ring = false) {
  tcs_graph* h = graph.device();
  graph.ctx().check(tcs_normals(h, neighborhood_1ring ? 1 : 0, arithmeticMode()));
  graph.invalidate_vertices();
}

template <typename GraphT>
void computeSignedCurvatures(GraphT& graph) {
  tcs_graph* h = graph.device();
  graph.ctx().check(tcs_signed_curvatures(h));
  graph.invalidate_vertices();
}

template <typename GraphT>
void smoothen(GraphT& graph, float spatial_sigma, float influence_sigma) {
  tcs_graph* h = graph.device();
  graph.ctx().check(tcs_smoothen(h, spatial_sigma, influence_sigma));
  graph.invalidate_vertices();
}

// A connected component as a light view: the reference returns boost::subgraph children whose local vertex ids ascend
// with the global ids (graph/impl/common.hpp:134-149); local_to_global(i) = vertices[i].
struct ComponentView {
  std::vector<uint32_t> vertices;  // global vertex ids, ascending
  uint32_t local_to_global(size_t i) const { return vertices[i]; }
  size_t num_vertices() const { return vertices.size(); }
};

// returns the number of components; component ids are ordered by their smallest vertex id
template <typename GraphT>
size_t createSubgraphsFromConnectedComponents(GraphT& graph, std::vector<ComponentView>& subgraphs) {
  tcs_graph* h = graph.device();
  int nc = 0;
  std::vector<int32_t> comp(graph.num_vertices());
  graph.ctx().check(tcs_components(h, &nc, comp.data()));
  subgraphs.assign((size_t)nc, ComponentView());
  for (size_t v = 0; v < comp.size(); ++v) subgraphs[(size_t)comp[v]].vertices.push_back((uint32_t)v);
  return (size_t)nc;
}

}  // namespace graph
}  // namespace pcl
