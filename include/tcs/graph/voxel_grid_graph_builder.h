// include/tcs/graph/voxel_grid_graph_builder.h - pcl::graph::VoxelGridGraphBuilder on the GPU.
// Same interface as the reference (include/graph/voxel_grid_graph_builder.h:75-93, graph_builder.h:87-102):
// VoxelGridGraphBuilder(float resolution), setInputCloud, compute(graph), set/getVoxelResolution,
// getPointToVertexMap().  compute() runs tcs_build_voxel_graph (Morton sort + segmented reduce + hashed neighbours).
#pragma once

#include "point_cloud_graph.h"

namespace pcl {
namespace graph {

template <typename PointT, typename GraphT>
class GraphBuilder : public pcl::PCLBase<PointT> {
 public:
  typedef PointT PointInT;
  typedef typename GraphT::point_type PointOutT;
  virtual ~GraphBuilder() {}
  virtual void compute(GraphT& graph) = 0;
  // point index -> vertex id; points that are not in the graph (non-finite) map to the maximum value
  // (graph_builder.h:96-102)
  const std::vector<size_t>& getPointToVertexMap() const { return point_to_vertex_map_; }

 protected:
  std::vector<size_t> point_to_vertex_map_;
};

template <typename PointT, typename GraphT>
class VoxelGridGraphBuilder : public GraphBuilder<PointT, GraphT> {
  using pcl::PCLBase<PointT>::input_;
  using GraphBuilder<PointT, GraphT>::point_to_vertex_map_;

 public:
  explicit VoxelGridGraphBuilder(float voxel_resolution = 0.006f) : voxel_resolution_(voxel_resolution) {}
  void setVoxelResolution(float resolution) { voxel_resolution_ = resolution; }
  float getVoxelResolution() const { return voxel_resolution_; }
  // device / leaf order are extensions (defaults reproduce the reference)
  void setDevice(int device) { device_ = device; }
  void setLeafOrderDescending(bool d) { leaf_order_desc_ = d; }

  virtual void compute(GraphT& graph) {
    std::shared_ptr<tcs::detail::Context> ctx = tcs::detail::Context::instance(device_);
    const size_t n = input_ ? input_->size() : 0;
    // invalid input -> empty graph, no throw (graph/impl/voxel_grid_graph_builder.hpp:72-77)
    const unsigned char* base = n ? reinterpret_cast<const unsigned char*>(&input_->points[0]) : nullptr;
    const uint32_t* rgba = nullptr;
    if (n && tcs::has_color<PointT>::value) rgba = color_ptr(&input_->points[0]);
    ctx->check(tcs_cloud_upload(ctx->get(), reinterpret_cast<const float*>(base), sizeof(PointT), rgba, sizeof(PointT), n));
    tcs_graph* h = nullptr;
    ctx->check(tcs_build_voxel_graph(ctx->get(), voxel_resolution_, leaf_order_desc_ ? 1 : 0, &h));
    graph.adopt(ctx, h);
    std::vector<uint32_t> map(n);
    if (n) ctx->check(tcs_graph_download_point_to_vertex(h, map.data()));
    point_to_vertex_map_.resize(n);
    for (size_t i = 0; i < n; ++i)
      point_to_vertex_map_[i] = map[i] == 0xFFFFFFFFu ? std::numeric_limits<size_t>::max() : (size_t)map[i];
  }

 private:
  template <typename Q> static const uint32_t* color_ptr(const Q* p) { return color_ptr_impl(p, std::integral_constant<bool, tcs::has_color<Q>::value>()); }
  template <typename Q> static const uint32_t* color_ptr_impl(const Q* p, std::true_type) { return &p->rgba; }
  template <typename Q> static const uint32_t* color_ptr_impl(const Q*, std::false_type) { return nullptr; }
  float voxel_resolution_;
  int device_ = 0;
  bool leaf_order_desc_ = true;
};

}  // namespace graph
}  // namespace pcl
