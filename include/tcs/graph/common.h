// include/tcs/graph/common.h - free functions of the reference's include/graph/common.h:65-204 on the GPU:
// computeNormalsAndCurvatures, computeSignedCurvatures, smoothen, createSubgraphsFromConnectedComponents,
// createSubgraphsFromColorMap, createSubgraphsFromIndices.
#pragma once

#include <map>
#include <set>

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

// One subgraph per distinct colour, ordered by ascending colour; vertices keep ascending global order
// (graph/impl/common.hpp:151-185: a std::map from colour to subgraph, filled in vertex order).
template <typename GraphT, typename ColorMap>
size_t createSubgraphsFromColorMap(GraphT& graph, const ColorMap& color_map, std::vector<ComponentView>& subgraphs) {
  typedef typename std::decay<decltype(color_map[0])>::type Color;
  std::map<Color, ComponentView> by_colour;
  const size_t V = graph.num_vertices();
  for (size_t v = 0; v < V; ++v) by_colour[color_map[v]].vertices.push_back((uint32_t)v);
  subgraphs.clear();
  for (typename std::map<Color, ComponentView>::iterator it = by_colour.begin(); it != by_colour.end(); ++it)
    subgraphs.push_back(it->second);
  return subgraphs.size();
}

// Two subgraphs: the listed vertices (in list order), and all the others in ascending order
// (graph/impl/common.hpp:187-211).
template <typename GraphT>
void createSubgraphsFromIndices(GraphT& graph, const pcl::PointIndices& indices, std::vector<ComponentView>& subgraphs) {
  std::set<int> index_set;
  subgraphs.assign(2, ComponentView());
  for (size_t i = 0; i < indices.indices.size(); ++i) {
    index_set.insert(indices.indices[i]);
    subgraphs[0].vertices.push_back((uint32_t)indices.indices[i]);
  }
  const size_t V = graph.num_vertices();
  for (size_t v = 0; v < V; ++v)
    if (!index_set.count((int)v)) subgraphs[1].vertices.push_back((uint32_t)v);
}

// One subgraph per index list plus a last one with the vertices in none of them (graph/impl/common.hpp:213-239).
template <typename GraphT>
void createSubgraphsFromIndices(GraphT& graph, const std::vector<pcl::PointIndices>& indices, std::vector<ComponentView>& subgraphs) {
  std::set<int> index_set;
  subgraphs.assign(indices.size() + 1, ComponentView());
  for (size_t i = 0; i < indices.size(); ++i)
    for (size_t j = 0; j < indices[i].indices.size(); ++j) {
      index_set.insert(indices[i].indices[j]);
      subgraphs[i].vertices.push_back((uint32_t)indices[i].indices[j]);
    }
  const size_t V = graph.num_vertices();
  for (size_t v = 0; v < V; ++v)
    if (!index_set.count((int)v)) subgraphs.back().vertices.push_back((uint32_t)v);
}

// The edges a boost::subgraph child would hold: edge ids (ascending) of the parent whose two endpoints are both in
// the view (boost::add_vertex(global, subgraph) pulls in every parent edge between the child's vertices).
template <typename GraphT>
std::vector<uint32_t> inducedEdges(const GraphT& graph, const ComponentView& view) {
  std::vector<char> inside(graph.num_vertices(), 0);
  for (size_t i = 0; i < view.vertices.size(); ++i) inside[view.vertices[i]] = 1;
  const std::vector<uint32_t>& s = graph.edge_sources();
  const std::vector<uint32_t>& t = graph.edge_targets();
  std::vector<uint32_t> out;
  for (size_t e = 0; e < s.size(); ++e)
    if (inside[s[e]] && inside[t[e]]) out.push_back((uint32_t)e);
  return out;
}

}  // namespace graph
}  // namespace pcl
