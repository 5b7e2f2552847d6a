// include/tcs/graph/nearest_neighbors_graph_builder.h - pcl::graph::NearestNeighborsGraphBuilder on the GPU.
// Same interface as the reference (include/graph/nearest_neighbors_graph_builder.h:60-178): default 10 neighbours
// and k-nearest search, useNearestKSearch / useRadiusSearch, setNumberOfNeighbors, setRadius, compute(graph),
// getPointToVertexMap().  compute() runs tcs_build_nn_graph (grid-hash exact search, tcs_b200/csrc/knn.cu).
// setSearchMethod / getSearchMethod are not mirrored: there is no pcl::search here, the search structure is the
// builder's own.
#pragma once

#include "voxel_grid_graph_builder.h"

namespace pcl {
namespace graph {

template <typename PointT, typename GraphT>
class NearestNeighborsGraphBuilder : public GraphBuilder<PointT, GraphT> {
  using pcl::PCLBase<PointT>::input_;
  using GraphBuilder<PointT, GraphT>::point_to_vertex_map_;
  enum SearchType { KNN, RADIUS };

 public:
  NearestNeighborsGraphBuilder() : num_neighbors_(10), radius_(0.0), search_type_(KNN) {}
  void useNearestKSearch() { search_type_ = KNN; }
  void useRadiusSearch() { search_type_ = RADIUS; }
  void setNumberOfNeighbors(size_t num_neighbors) { num_neighbors_ = num_neighbors; }
  void setRadius(double radius) { radius_ = radius; }
  void setDevice(int device) { device_ = device; }

  virtual void compute(GraphT& graph) {
    std::shared_ptr<tcs::detail::Context> ctx = tcs::detail::Context::instance(device_);
    const size_t n = input_ ? input_->size() : 0;
    const unsigned char* base = n ? reinterpret_cast<const unsigned char*>(&input_->points[0]) : nullptr;
    const uint32_t* rgba = nullptr;
    if (n && tcs::has_color<PointT>::value) rgba = color_ptr(&input_->points[0], std::integral_constant<bool, tcs::has_color<PointT>::value>());
    ctx->check(tcs_cloud_upload(ctx->get(), reinterpret_cast<const float*>(base), sizeof(PointT), rgba, sizeof(PointT), n));
    tcs_graph* h = nullptr;
    ctx->check(tcs_build_nn_graph(ctx->get(), (int)num_neighbors_, radius_, search_type_ == RADIUS ? 1 : 0, &h));
    graph.adopt(ctx, h);
    std::vector<uint32_t> map(n);
    if (n) ctx->check(tcs_graph_download_point_to_vertex(h, map.data()));
    point_to_vertex_map_.resize(n);
    for (size_t i = 0; i < n; ++i)
      point_to_vertex_map_[i] = map[i] == 0xFFFFFFFFu ? std::numeric_limits<size_t>::max() : (size_t)map[i];
  }

 private:
  template <typename Q> static const uint32_t* color_ptr(const Q* p, std::true_type) { return &p->rgba; }
  template <typename Q> static const uint32_t* color_ptr(const Q*, std::false_type) { return nullptr; }
  size_t num_neighbors_;
  double radius_;
  SearchType search_type_;
  int device_ = 0;
};

}  // namespace graph
}  // namespace pcl
