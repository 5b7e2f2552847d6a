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
#include "graph_iterators.h"

namespace boost {
#ifndef BOOST_GRAPH_PROPERTIES_HPP
// The Boost.Graph names the reference's call sites spell out (boost/graph/properties.hpp, graph_traits.hpp,
// property_map.hpp, tuple.hpp); defined here only when Boost.Graph is not installed.
#define TCS_BOOST_COMPAT 1
enum vertex_color_t { vertex_color };
enum edge_weight_t { edge_weight };
enum vertex_index_t { vertex_index };
enum edge_index_t { edge_index };
struct undirected_tag {};
struct allow_parallel_edge_tag {};
struct vertex_list_graph_tag {};
// boost::graph_traits<G>: reads the graph's own typedefs
template <typename G>
struct graph_traits {
  typedef typename G::vertex_descriptor vertex_descriptor;
  typedef typename G::edge_descriptor edge_descriptor;
  typedef typename G::vertex_iterator vertex_iterator;
  typedef typename G::edge_iterator edge_iterator;
  typedef typename G::out_edge_iterator out_edge_iterator;
  typedef typename G::in_edge_iterator in_edge_iterator;
  typedef typename G::adjacency_iterator adjacency_iterator;
  typedef typename G::vertices_size_type vertices_size_type;
  typedef typename G::edges_size_type edges_size_type;
  typedef typename G::degree_size_type degree_size_type;
  typedef typename G::directed_category directed_category;
  typedef typename G::edge_parallel_category edge_parallel_category;
  typedef typename G::traversal_category traversal_category;
  static vertex_descriptor null_vertex() { return (vertex_descriptor)-1; }
};
// boost::property_traits<Map>
template <typename Map>
struct property_traits {
  typedef typename Map::value_type value_type;
  typedef typename Map::reference reference;
  typedef typename Map::key_type key_type;
};
// boost::tie (a, b) = pair
template <typename A, typename B>
struct tcs_tied {
  A& a; B& b;
  tcs_tied(A& a_, B& b_) : a(a_), b(b_) {}
  template <typename U, typename V> tcs_tied& operator=(const std::pair<U, V>& p) { a = p.first; b = p.second; return *this; }
};
template <typename A, typename B> inline tcs_tied<A, B> tie(A& a, B& b) { return tcs_tied<A, B>(a, b); }
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
  typedef tcs::detail::edge_descriptor edge_descriptor;
  typedef size_t vertices_size_type;
  typedef size_t edges_size_type;
  typedef size_t degree_size_type;
  // what boost::graph_traits reads (VertexListGraph, EdgeListGraph, IncidenceGraph, AdjacencyGraph; undirected, vecS)
  typedef tcs::detail::vertex_iterator vertex_iterator;
  typedef tcs::detail::edge_iterator<point_cloud_graph> edge_iterator;
  typedef tcs::detail::out_edge_iterator<point_cloud_graph> out_edge_iterator;
  typedef out_edge_iterator in_edge_iterator;
  typedef tcs::detail::adjacency_iterator<point_cloud_graph> adjacency_iterator;
  typedef boost::undirected_tag directed_category;
  typedef boost::allow_parallel_edge_tag edge_parallel_category;
  typedef boost::vertex_list_graph_tag traversal_category;
  typedef tcs::detail::edge_weight_map<point_cloud_graph> edge_weight_map;
  typedef tcs::detail::const_edge_weight_map<point_cloud_graph> const_edge_weight_map;
  typedef tcs::detail::vertex_color_map<point_cloud_graph> vertex_color_map;
  typedef tcs::detail::const_vertex_color_map<point_cloud_graph> const_vertex_color_map;
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

// ---- BGL free functions (VertexListGraph / EdgeListGraph / IncidenceGraph / AdjacencyGraph / PropertyGraph) --------
// Same names and meaning as for the reference's type (include/graph/point_cloud_graph.h:590-835); iterator pairs work
// with boost::tie and the reference's as_range (include/as_range.h).
template <typename P> inline size_t num_vertices(const point_cloud_graph<P>& g) { return g.num_vertices(); }
template <typename P> inline size_t num_edges(const point_cloud_graph<P>& g) { return g.num_edges(); }
template <typename P>
inline std::pair<typename point_cloud_graph<P>::edge_descriptor, bool> add_edge(uint32_t u, uint32_t v, point_cloud_graph<P>& g) {
  return g.add_edge(u, v);
}
template <typename P> inline uint32_t source(const tcs::detail::edge_descriptor& e, const point_cloud_graph<P>&) { return e.src; }
template <typename P> inline uint32_t target(const tcs::detail::edge_descriptor& e, const point_cloud_graph<P>&) { return e.tgt; }
template <typename P>
inline std::pair<tcs::detail::vertex_iterator, tcs::detail::vertex_iterator> vertices(const point_cloud_graph<P>& g) {
  return std::make_pair(tcs::detail::vertex_iterator(0), tcs::detail::vertex_iterator((uint32_t)g.num_vertices()));
}
template <typename P>
inline std::pair<typename point_cloud_graph<P>::edge_iterator, typename point_cloud_graph<P>::edge_iterator> edges(
    const point_cloud_graph<P>& g) {
  typedef typename point_cloud_graph<P>::edge_iterator It;
  return std::make_pair(It(&g, 0), It(&g, g.edge_sources().size()));
}
template <typename P>
inline std::pair<typename point_cloud_graph<P>::out_edge_iterator, typename point_cloud_graph<P>::out_edge_iterator> out_edges(
    uint32_t v, const point_cloud_graph<P>& g) {
  typedef typename point_cloud_graph<P>::out_edge_iterator It;
  const std::vector<uint32_t>& rp = g.out_rowptr();
  return std::make_pair(It(&g, v, rp[v]), It(&g, v, rp[v + 1]));
}
// undirected: the in-edges of a vertex are its out-edges
template <typename P>
inline std::pair<typename point_cloud_graph<P>::out_edge_iterator, typename point_cloud_graph<P>::out_edge_iterator> in_edges(
    uint32_t v, const point_cloud_graph<P>& g) {
  return out_edges(v, g);
}
template <typename P>
inline std::pair<typename point_cloud_graph<P>::adjacency_iterator, typename point_cloud_graph<P>::adjacency_iterator>
adjacent_vertices(uint32_t v, const point_cloud_graph<P>& g) {
  typedef typename point_cloud_graph<P>::adjacency_iterator It;
  const std::vector<uint32_t>& rp = g.out_rowptr();
  return std::make_pair(It(&g, rp[v]), It(&g, rp[v + 1]));
}
template <typename P> inline size_t out_degree(uint32_t v, const point_cloud_graph<P>& g) { return g.out_rowptr()[v + 1] - g.out_rowptr()[v]; }
template <typename P> inline size_t in_degree(uint32_t v, const point_cloud_graph<P>& g) { return out_degree(v, g); }
template <typename P> inline size_t degree(uint32_t v, const point_cloud_graph<P>& g) { return out_degree(v, g); }
// boost::edge (u, v, g): the first edge between u and v in u's out-edge order
template <typename P>
inline std::pair<typename point_cloud_graph<P>::edge_descriptor, bool> edge(uint32_t u, uint32_t v, const point_cloud_graph<P>& g) {
  const std::vector<uint32_t>& rp = g.out_rowptr();
  for (uint32_t a = rp[u]; a < rp[u + 1]; ++a)
    if (g.out_targets()[a] == v) {
      const uint32_t id = g.out_edge_ids()[a];
      typename point_cloud_graph<P>::edge_descriptor e = {g.edge_sources()[id], g.edge_targets()[id], id};
      return std::make_pair(e, true);
    }
  typename point_cloud_graph<P>::edge_descriptor none = {0u, 0u, 0xFFFFFFFFu};
  return std::make_pair(none, false);
}
// edge i of the global edge list (insertion order = edge id)
template <typename P> inline typename point_cloud_graph<P>::edge_descriptor edge_at(size_t i, const point_cloud_graph<P>& g) {
  typename point_cloud_graph<P>::edge_descriptor e = {g.edge_sources()[i], g.edge_targets()[i], (uint32_t)i};
  return e;
}
template <typename P> inline typename pcl::PointCloud<P>::ConstPtr point_cloud(const point_cloud_graph<P>& g) { return g.point_cloud(); }
// property maps: get (boost::vertex_color, g)[v], get (boost::edge_weight, g)[e]  (also [edge id])
template <typename P> inline typename point_cloud_graph<P>::vertex_color_map get(boost::vertex_color_t, point_cloud_graph<P>& g) {
  return typename point_cloud_graph<P>::vertex_color_map(&g);
}
template <typename P> inline typename point_cloud_graph<P>::const_vertex_color_map get(boost::vertex_color_t, const point_cloud_graph<P>& g) {
  return typename point_cloud_graph<P>::const_vertex_color_map(&g);
}
template <typename P> inline typename point_cloud_graph<P>::edge_weight_map get(boost::edge_weight_t, point_cloud_graph<P>& g) {
  return typename point_cloud_graph<P>::edge_weight_map(&g);
}
template <typename P> inline typename point_cloud_graph<P>::const_edge_weight_map get(boost::edge_weight_t, const point_cloud_graph<P>& g) {
  return typename point_cloud_graph<P>::const_edge_weight_map(&g);
}
template <typename P> inline tcs::detail::edge_index_map<point_cloud_graph<P>> get(boost::edge_index_t, const point_cloud_graph<P>&) {
  return tcs::detail::edge_index_map<point_cloud_graph<P>>();
}
template <typename P> inline tcs::detail::vertex_index_map get(boost::vertex_index_t, const point_cloud_graph<P>&) {
  return tcs::detail::vertex_index_map();
}
// get (tag, g, key) / put (tag, g, key, value)
template <typename Tag, typename P, typename Key>
inline auto get(Tag tag, const point_cloud_graph<P>& g, const Key& k) -> typename std::decay<decltype(get(tag, g)[k])>::type {
  return get(tag, g)[k];
}
template <typename Tag, typename P, typename Key, typename Value>
inline void put(Tag tag, point_cloud_graph<P>& g, const Key& k, const Value& v) { get(tag, g)[k] = v; }

}  // namespace graph
}  // namespace pcl

// get (map, key) / put (map, key, value) on the maps above (boost/property_map/property_map.hpp spelling)
namespace tcs {
namespace detail {
template <typename G> inline float& get(const edge_weight_map<G>& m, const edge_descriptor& e) { return m[e]; }
template <typename G> inline const float& get(const const_edge_weight_map<G>& m, const edge_descriptor& e) { return m[e]; }
template <typename G> inline uint32_t& get(const vertex_color_map<G>& m, uint32_t v) { return m[v]; }
template <typename G> inline const uint32_t& get(const const_vertex_color_map<G>& m, uint32_t v) { return m[v]; }
template <typename G> inline void put(const edge_weight_map<G>& m, const edge_descriptor& e, float w) { m[e] = w; }
template <typename G> inline void put(const vertex_color_map<G>& m, uint32_t v, uint32_t c) { m[v] = c; }
}  // namespace detail
}  // namespace tcs

// The reference's call sites qualify these with boost:: (boost::edges (graph), boost::source (edge, graph), ...).
namespace boost {
using pcl::graph::num_vertices;
using pcl::graph::num_edges;
using pcl::graph::add_edge;
using pcl::graph::source;
using pcl::graph::target;
using pcl::graph::vertices;
using pcl::graph::edges;
using pcl::graph::out_edges;
using pcl::graph::in_edges;
using pcl::graph::adjacent_vertices;
using pcl::graph::out_degree;
using pcl::graph::in_degree;
using pcl::graph::degree;
using pcl::graph::edge;
using pcl::graph::get;
using pcl::graph::put;
using tcs::detail::get;
using tcs::detail::put;
#ifdef TCS_BOOST_COMPAT
// boost::property_map<Graph, Tag>::type
template <typename G, typename Tag> struct property_map;
template <typename G> struct property_map<G, edge_weight_t> { typedef typename G::edge_weight_map type; typedef typename G::const_edge_weight_map const_type; };
template <typename G> struct property_map<G, vertex_color_t> { typedef typename G::vertex_color_map type; typedef typename G::const_vertex_color_map const_type; };
template <typename G> struct property_map<G, edge_index_t> { typedef tcs::detail::edge_index_map<G> type; typedef type const_type; };
template <typename G> struct property_map<G, vertex_index_t> { typedef tcs::detail::vertex_index_map type; typedef type const_type; };
#endif
}  // namespace boost
