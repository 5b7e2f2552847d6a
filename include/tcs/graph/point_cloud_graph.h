// include/tcs/graph/point_cloud_graph.h - pcl::graph::point_cloud_graph on top of the device graph of libtcs_b200.
//
// Mirrors the reference's BGL-compatible graph type (include/graph/point_cloud_graph.h:335-588, free functions
// :590-835) as far as the random-walker path uses it: vertices are the points of a pcl::PointCloud, edges carry a
// float weight, vertices a uint32 colour; edge ids are insertion order; out-edge lists are in insertion order.
// Storage differs by design: the graph lives on the GPU (CSR + SoA, tcs_graph) and a host mirror is materialised
// lazily when the host reads it; host edits (add_edge, g[v], colour map writes) are uploaded before the next device
// stage.  One graph belongs to one host thread (like the reference's, which is not thread-safe either).
#pragma once

#include <algorithm>
#include <utility>

#include "../../tcs_b200.h"
#include "../pcl_compat.h"

namespace boost {
#ifndef BOOST_GRAPH_PROPERTIES_HPP
// property tags the reference's call sites name (boost/graph/properties.hpp); defined here only without Boost.Graph
enum vertex_color_t { vertex_color };
enum edge_weight_t { edge_weight };
enum vertex_index_t { vertex_index };
enum edge_index_t { edge_index };
#endif
}  // namespace boost

namespace tcs {
namespace detail {

// one tcs_ctx per host thread and device
class Context {
 public:
  explicit Context(int device) {
    if (tcs_ctx_create(device, &h_) != TCS_OK)
      throw pcl::PCLException(std::string("tcs_b200: cannot create a CUDA context (no CPU fallback): ") + tcs_last_error(nullptr));
  }
  ~Context() { tcs_ctx_destroy(h_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  tcs_ctx* get() const { return h_; }
  // error convention of the reference: invalid input -> ComputeFailedException (impl/random_walker_segmentation.hpp:111)
  void check(int status) const {
    if (status == TCS_OK) return;
    std::string msg = std::string("tcs_b200: ") + tcs_last_error(h_);
    if (status == TCS_ERR_INPUT || status == TCS_ERR_STATE) throw pcl::ComputeFailedException(msg);
    throw pcl::PCLException(msg);
  }
  static tcs::shared_ptr<Context> instance(int device = 0) {
    static thread_local std::vector<std::weak_ptr<Context>> cache;
    if ((int)cache.size() <= device) cache.resize(device + 1);
    std::shared_ptr<Context> c = cache[device].lock();
    if (!c) { c = std::make_shared<Context>(device); cache[device] = c; }
    return tcs::shared_ptr<Context>(c);
  }

 private:
  tcs_ctx* h_ = nullptr;
};

struct DeviceGraph {
  std::shared_ptr<Context> ctx;
  tcs_graph* h = nullptr;
  ~DeviceGraph() { if (h) tcs_graph_destroy(h); }
};

template <typename PointT> inline void set_rgba(PointT&, uint32_t, std::false_type) {}
template <typename PointT> inline void set_rgba(PointT& p, uint32_t c, std::true_type) { p.rgba = c; }
template <typename PointT> inline uint32_t get_rgba(const PointT&, std::false_type) { return 0; }
template <typename PointT> inline uint32_t get_rgba(const PointT& p, std::true_type) { return p.rgba; }

}  // namespace detail
}  // namespace tcs

namespace pcl {
namespace graph {

template <typename PointT>
class point_cloud_graph {
 public:
  typedef PointT vertex_bundled;
  typedef PointT point_type;
  typedef pcl::PointCloud<PointT> point_cloud_type;
  typedef typename point_cloud_type::Ptr point_cloud_ptr;
  typedef typename point_cloud_type::ConstPtr point_cloud_const_ptr;
  typedef uint32_t vertex_descriptor;
  struct edge_descriptor {
    uint32_t src, tgt, idx;
    bool operator==(const edge_descriptor& o) const { return idx == o.idx; }
  };
  typedef size_t vertices_size_type;
  typedef size_t edges_size_type;
  typedef std::integral_constant<bool, tcs::has_color<PointT>::value> colored;

  point_cloud_graph() : cloud_(new point_cloud_type) {}
  // vertices = the points of `cloud`, no edges (reference ctor, point_cloud_graph.h:444)
  explicit point_cloud_graph(const point_cloud_ptr& cloud) : cloud_(cloud) { colour_.assign(cloud_->size(), 0u); }
  explicit point_cloud_graph(vertices_size_type n) : cloud_(new point_cloud_type) {
    cloud_->resize(n);
    colour_.assign(n, 0u);
  }

  // ---- sizes (valid in either state)
  vertices_size_type num_vertices() const { return dev_ ? dev_sizes().first : cloud_->size(); }
  edges_size_type num_edges() const { return dev_ && !topo_host_dirty_ ? dev_sizes().second : esrc_.size(); }

  // ---- host view -------------------------------------------------------------------------------------------
  const PointT& operator[](vertex_descriptor v) const { pull_vertices(); return cloud_->points[v]; }
  PointT& operator[](vertex_descriptor v) { pull_vertices(); bundle_host_dirty_ = true; return cloud_->points[v]; }
  point_cloud_const_ptr point_cloud() const { pull_vertices(); return cloud_; }
  const std::vector<uint32_t>& edge_sources() const { pull_topology(); return esrc_; }
  const std::vector<uint32_t>& edge_targets() const { pull_topology(); return etgt_; }
  const std::vector<float>& edge_weights() const { pull_weights(); return weight_; }
  std::vector<float>& edge_weights() { pull_weights(); weight_host_dirty_ = true; return weight_; }
  const std::vector<uint32_t>& vertex_colors() const { pull_colours(); return colour_; }
  std::vector<uint32_t>& vertex_colors() { pull_colours(); colour_host_dirty_ = true; return colour_; }
  // out-edge lists (CSR: rowptr, adjacent vertex, edge id), insertion order
  const std::vector<uint32_t>& out_rowptr() const { pull_topology(); return rowptr_; }
  const std::vector<uint32_t>& out_targets() const { pull_topology(); return col_; }
  const std::vector<uint32_t>& out_edge_ids() const { pull_topology(); return eid_; }

  // boost::add_edge(u, v, g): appended to the global edge list and to both out-edge lists
  std::pair<edge_descriptor, bool> add_edge(vertex_descriptor u, vertex_descriptor v, float w = 0.0f) {
    pull_topology(); pull_weights();
    edge_descriptor e = {u, v, (uint32_t)esrc_.size()};
    esrc_.push_back(u); etgt_.push_back(v); weight_.push_back(w);
    topo_host_dirty_ = true;
    csr_valid_ = false;
    return std::make_pair(e, true);
  }

  // ---- device side -------------------------------------------------------------------------------------------
  // adopt a graph built on the device (VoxelGridGraphBuilder)
  void adopt(const std::shared_ptr<tcs::detail::Context>& ctx, tcs_graph* h) {
    dev_ = std::make_shared<tcs::detail::DeviceGraph>();
    dev_->ctx = ctx;
    dev_->h = h;
    have_vertices_ = have_topo_ = have_weights_ = have_colours_ = false;
    bundle_host_dirty_ = topo_host_dirty_ = weight_host_dirty_ = colour_host_dirty_ = false;
  }
  // device handle with every pending host edit uploaded
  tcs_graph* device() {
    if (!dev_ || topo_host_dirty_) rebuild_device();
    if (bundle_host_dirty_) push_vertices();
    if (weight_host_dirty_) { ctx().check(tcs_graph_upload_weights(dev_->h, weight_.data())); weight_host_dirty_ = false; }
    return dev_->h;
  }
  const tcs::detail::Context& ctx() const { return *dev_->ctx; }
  std::shared_ptr<tcs::detail::Context> context() {
    if (!dev_) rebuild_device();
    return dev_->ctx;
  }
  // a device stage changed these: the host mirror is stale
  void invalidate_vertices() { have_vertices_ = false; }
  void invalidate_weights() { have_weights_ = false; }
  void invalidate_colours() { have_colours_ = false; }
  void invalidate_topology() { have_topo_ = false; have_weights_ = false; csr_valid_ = false; }
  bool colours_edited_on_host() const { return colour_host_dirty_; }
  void colours_synced() { colour_host_dirty_ = false; have_colours_ = true; }
  std::vector<uint32_t>& raw_colours() { return colour_; }

 private:
  std::pair<size_t, size_t> dev_sizes() const {
    size_t V = 0, E = 0;
    tcs_graph_sizes(dev_->h, &V, &E, nullptr, nullptr);
    return std::make_pair(V, E);
  }
  void rebuild_device() {
    std::shared_ptr<tcs::detail::Context> c = dev_ ? dev_->ctx : tcs::detail::Context::instance(0);
    if (dev_) { pull_vertices(); pull_weights(); pull_colours(); }
    const size_t V = cloud_->size();
    std::vector<float> xyz(3 * V), n4(4 * V), cv(V);
    std::vector<uint32_t> rgba(V);
    for (size_t v = 0; v < V; ++v) {
      const PointT& p = cloud_->points[v];
      xyz[3 * v] = p.x; xyz[3 * v + 1] = p.y; xyz[3 * v + 2] = p.z;
      for (int k = 0; k < 4; ++k) n4[4 * v + k] = p.data_n[k];
      cv[v] = p.curvature;
      rgba[v] = tcs::detail::get_rgba(p, colored());
    }
    tcs_graph* h = nullptr;
    c->check(tcs_graph_create(c->get(), V, xyz.data(), colored::value ? rgba.data() : nullptr, n4.data(), cv.data(),
                              esrc_.size(), esrc_.data(), etgt_.data(), weight_.data(), &h));
    dev_ = std::make_shared<tcs::detail::DeviceGraph>();
    dev_->ctx = c;
    dev_->h = h;
    have_vertices_ = have_topo_ = have_weights_ = have_colours_ = true;
    bundle_host_dirty_ = topo_host_dirty_ = weight_host_dirty_ = false;
    if (colour_.size() != V) colour_.assign(V, 0u);
  }
  void push_vertices() {
    const size_t V = cloud_->size();
    std::vector<float> xyz(3 * V), n4(4 * V), cv(V);
    for (size_t v = 0; v < V; ++v) {
      const PointT& p = cloud_->points[v];
      xyz[3 * v] = p.x; xyz[3 * v + 1] = p.y; xyz[3 * v + 2] = p.z;
      for (int k = 0; k < 4; ++k) n4[4 * v + k] = p.data_n[k];
      cv[v] = p.curvature;
    }
    ctx().check(tcs_graph_upload_xyz(dev_->h, xyz.data()));
    ctx().check(tcs_graph_upload_normals(dev_->h, n4.data(), cv.data()));
    bundle_host_dirty_ = false;
  }
  void pull_vertices() const {
    if (!dev_ || have_vertices_ || bundle_host_dirty_) return;
    const size_t V = dev_sizes().first;
    std::vector<float> xyz(3 * V), n4(4 * V), cv(V);
    std::vector<uint32_t> rgba(V);
    ctx().check(tcs_graph_download_vertices(dev_->h, xyz.data(), rgba.data(), n4.data(), cv.data()));
    cloud_->points.resize(V);
    cloud_->width = (uint32_t)V; cloud_->height = 1; cloud_->is_dense = true;
    for (size_t v = 0; v < V; ++v) {
      PointT& p = cloud_->points[v];
      p.x = xyz[3 * v]; p.y = xyz[3 * v + 1]; p.z = xyz[3 * v + 2];
      for (int k = 0; k < 4; ++k) p.data_n[k] = n4[4 * v + k];
      p.curvature = cv[v];
      tcs::detail::set_rgba(p, rgba[v], colored());
    }
    have_vertices_ = true;
  }
  void pull_topology() const {
    if (dev_ && !have_topo_ && !topo_host_dirty_) {
      const std::pair<size_t, size_t> s = dev_sizes();
      esrc_.resize(s.second); etgt_.resize(s.second);
      rowptr_.resize(s.first + 1); col_.resize(2 * s.second); eid_.resize(2 * s.second);
      ctx().check(tcs_graph_download_edges(dev_->h, esrc_.data(), etgt_.data(), nullptr));
      ctx().check(tcs_graph_download_csr(dev_->h, rowptr_.data(), col_.data(), eid_.data()));
      have_topo_ = true;
      csr_valid_ = true;
    }
    if (!csr_valid_) {  // host-built edge list: out-edge lists in insertion order
      const size_t V = cloud_->size(), E = esrc_.size();
      rowptr_.assign(V + 1, 0);
      for (size_t e = 0; e < E; ++e) { rowptr_[esrc_[e] + 1]++; rowptr_[etgt_[e] + 1]++; }
      for (size_t v = 0; v < V; ++v) rowptr_[v + 1] += rowptr_[v];
      col_.assign(2 * E, 0); eid_.assign(2 * E, 0);
      std::vector<uint32_t> fill(rowptr_.begin(), rowptr_.end() - 1);
      for (size_t e = 0; e < E; ++e) {
        col_[fill[esrc_[e]]] = etgt_[e]; eid_[fill[esrc_[e]]++] = (uint32_t)e;
        col_[fill[etgt_[e]]] = esrc_[e]; eid_[fill[etgt_[e]]++] = (uint32_t)e;
      }
      csr_valid_ = true;
    }
  }
  void pull_weights() const {
    if (!dev_ || have_weights_ || weight_host_dirty_ || topo_host_dirty_) return;
    weight_.resize(dev_sizes().second);
    ctx().check(tcs_graph_download_edges(dev_->h, nullptr, nullptr, weight_.data()));
    have_weights_ = true;
  }
  void pull_colours() const {
    if (colour_.size() != num_vertices()) colour_.resize(num_vertices(), 0u);
    if (!dev_ || have_colours_ || colour_host_dirty_) return;
    if (tcs_graph_d_colours(dev_->h)) ctx().check(tcs_graph_download_colours(dev_->h, colour_.data()));
    have_colours_ = true;
  }

  mutable point_cloud_ptr cloud_;
  mutable std::vector<uint32_t> esrc_, etgt_, colour_, rowptr_, col_, eid_;
  mutable std::vector<float> weight_;
  std::shared_ptr<tcs::detail::DeviceGraph> dev_;
  mutable bool have_vertices_ = true, have_topo_ = true, have_weights_ = true, have_colours_ = true, csr_valid_ = false;
  bool bundle_host_dirty_ = false, topo_host_dirty_ = false, weight_host_dirty_ = false, colour_host_dirty_ = false;
};

// ---- BGL-style free functions (the subset the reference's call sites use) --------------------------------------
template <typename P> inline size_t num_vertices(const point_cloud_graph<P>& g) { return g.num_vertices(); }
template <typename P> inline size_t num_edges(const point_cloud_graph<P>& g) { return g.num_edges(); }
template <typename P>
inline std::pair<typename point_cloud_graph<P>::edge_descriptor, bool> add_edge(uint32_t u, uint32_t v, point_cloud_graph<P>& g) {
  return g.add_edge(u, v);
}
template <typename P> inline uint32_t source(const typename point_cloud_graph<P>::edge_descriptor& e, const point_cloud_graph<P>&) { return e.src; }
template <typename P> inline uint32_t target(const typename point_cloud_graph<P>::edge_descriptor& e, const point_cloud_graph<P>&) { return e.tgt; }
// edge i of the global edge list (insertion order = edge id)
template <typename P> inline typename point_cloud_graph<P>::edge_descriptor edge_at(size_t i, const point_cloud_graph<P>& g) {
  typename point_cloud_graph<P>::edge_descriptor e = {g.edge_sources()[i], g.edge_targets()[i], (uint32_t)i};
  return e;
}
// adjacent vertices of v in out-edge (insertion) order
template <typename P> inline std::vector<uint32_t> adjacent_vertices(uint32_t v, const point_cloud_graph<P>& g) {
  const std::vector<uint32_t>& rp = g.out_rowptr();
  return std::vector<uint32_t>(g.out_targets().begin() + rp[v], g.out_targets().begin() + rp[v + 1]);
}
template <typename P> inline size_t out_degree(uint32_t v, const point_cloud_graph<P>& g) { return g.out_rowptr()[v + 1] - g.out_rowptr()[v]; }
template <typename P> inline typename pcl::PointCloud<P>::ConstPtr point_cloud(const point_cloud_graph<P>& g) { return g.point_cloud(); }
// property maps: get(boost::vertex_color, g)[v], get(boost::edge_weight, g)[e.idx]
template <typename P> inline std::vector<uint32_t>& get(boost::vertex_color_t, point_cloud_graph<P>& g) { return g.vertex_colors(); }
template <typename P> inline const std::vector<uint32_t>& get(boost::vertex_color_t, const point_cloud_graph<P>& g) { return g.vertex_colors(); }
template <typename P> inline std::vector<float>& get(boost::edge_weight_t, point_cloud_graph<P>& g) { return g.edge_weights(); }
template <typename P> inline const std::vector<float>& get(boost::edge_weight_t, const point_cloud_graph<P>& g) { return g.edge_weights(); }

}  // namespace graph
}  // namespace pcl
