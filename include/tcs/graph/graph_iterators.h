// include/tcs/graph/graph_iterators.h - the BGL iteration surface of pcl::graph::point_cloud_graph.
//
// The reference's graph type models VertexListGraph / EdgeListGraph / IncidenceGraph / AdjacencyGraph
// (include/graph/point_cloud_graph.h:335-588); its callers iterate `boost::vertices (g)`, `boost::edges (g)`,
// `boost::out_edges (v, g)` through iterator pairs (`boost::tie`, `as_range`: include/impl/io.hpp:33,
// include/conversions.h:140-159) and read / write `boost::get (boost::edge_weight, g)[edge]`,
// `boost::get (boost::vertex_color, g)[vertex]`.  Here the graph lives on the GPU (CSR + SoA) with a lazily
// materialised host mirror; these iterators walk that mirror in the reference's orders: vertices ascending, edges in
// insertion order (= edge id), out-edges of a vertex in insertion order.
//
// The classes are templates on the graph type so that point_cloud_graph can name them in its own typedefs (what
// boost::graph_traits reads when Boost.Graph is installed).
#pragma once

#include <cstddef>
#include <cstdint>
#include <iterator>
#include <utility>

namespace tcs {
namespace detail {

// {source, target, edge id}: what boost::source / boost::target / the edge_index map read
struct edge_descriptor {
  uint32_t src, tgt, idx;
  bool operator==(const edge_descriptor& o) const { return idx == o.idx; }
  bool operator!=(const edge_descriptor& o) const { return idx != o.idx; }
  bool operator<(const edge_descriptor& o) const { return idx < o.idx; }
};

// 0, 1, 2, ... (vertex descriptors are indices, vecS)
class vertex_iterator {
 public:
  typedef std::random_access_iterator_tag iterator_category;
  typedef uint32_t value_type;
  typedef std::ptrdiff_t difference_type;
  typedef const uint32_t* pointer;
  typedef uint32_t reference;
  vertex_iterator() : v_(0) {}
  explicit vertex_iterator(uint32_t v) : v_(v) {}
  uint32_t operator*() const { return v_; }
  uint32_t operator[](difference_type d) const { return (uint32_t)(v_ + d); }
  vertex_iterator& operator++() { ++v_; return *this; }
  vertex_iterator operator++(int) { vertex_iterator t(*this); ++v_; return t; }
  vertex_iterator& operator--() { --v_; return *this; }
  vertex_iterator& operator+=(difference_type d) { v_ = (uint32_t)(v_ + d); return *this; }
  vertex_iterator operator+(difference_type d) const { return vertex_iterator((uint32_t)(v_ + d)); }
  difference_type operator-(const vertex_iterator& o) const { return (difference_type)v_ - (difference_type)o.v_; }
  bool operator==(const vertex_iterator& o) const { return v_ == o.v_; }
  bool operator!=(const vertex_iterator& o) const { return v_ != o.v_; }
  bool operator<(const vertex_iterator& o) const { return v_ < o.v_; }

 private:
  uint32_t v_;
};

// the global edge list in insertion order
template <typename G>
class edge_iterator {
 public:
  typedef std::forward_iterator_tag iterator_category;
  typedef edge_descriptor value_type;
  typedef std::ptrdiff_t difference_type;
  typedef const edge_descriptor* pointer;
  typedef edge_descriptor reference;
  edge_iterator() : g_(nullptr), i_(0) {}
  edge_iterator(const G* g, size_t i) : g_(g), i_(i) {}
  edge_descriptor operator*() const {
    edge_descriptor e = {g_->edge_sources()[i_], g_->edge_targets()[i_], (uint32_t)i_};
    return e;
  }
  edge_iterator& operator++() { ++i_; return *this; }
  edge_iterator operator++(int) { edge_iterator t(*this); ++i_; return t; }
  bool operator==(const edge_iterator& o) const { return i_ == o.i_; }
  bool operator!=(const edge_iterator& o) const { return i_ != o.i_; }

 private:
  const G* g_;
  size_t i_;
};

// the out-edge list of one vertex in insertion order; source (e) is that vertex (BGL's out_edges contract for
// undirected graphs), whichever end it was when the edge was added
template <typename G>
class out_edge_iterator {
 public:
  typedef std::forward_iterator_tag iterator_category;
  typedef edge_descriptor value_type;
  typedef std::ptrdiff_t difference_type;
  typedef const edge_descriptor* pointer;
  typedef edge_descriptor reference;
  out_edge_iterator() : g_(nullptr), v_(0), pos_(0) {}
  out_edge_iterator(const G* g, uint32_t v, size_t pos) : g_(g), v_(v), pos_(pos) {}
  edge_descriptor operator*() const {
    edge_descriptor e = {v_, g_->out_targets()[pos_], g_->out_edge_ids()[pos_]};
    return e;
  }
  out_edge_iterator& operator++() { ++pos_; return *this; }
  out_edge_iterator operator++(int) { out_edge_iterator t(*this); ++pos_; return t; }
  bool operator==(const out_edge_iterator& o) const { return pos_ == o.pos_; }
  bool operator!=(const out_edge_iterator& o) const { return pos_ != o.pos_; }

 private:
  const G* g_;
  uint32_t v_;
  size_t pos_;
};

// the neighbours of one vertex in out-edge order
template <typename G>
class adjacency_iterator {
 public:
  typedef std::forward_iterator_tag iterator_category;
  typedef uint32_t value_type;
  typedef std::ptrdiff_t difference_type;
  typedef const uint32_t* pointer;
  typedef uint32_t reference;
  adjacency_iterator() : g_(nullptr), pos_(0) {}
  adjacency_iterator(const G* g, size_t pos) : g_(g), pos_(pos) {}
  uint32_t operator*() const { return g_->out_targets()[pos_]; }
  adjacency_iterator& operator++() { ++pos_; return *this; }
  adjacency_iterator operator++(int) { adjacency_iterator t(*this); ++pos_; return t; }
  bool operator==(const adjacency_iterator& o) const { return pos_ == o.pos_; }
  bool operator!=(const adjacency_iterator& o) const { return pos_ != o.pos_; }

 private:
  const G* g_;
  size_t pos_;
};

// ---- property maps: boost::get (boost::edge_weight, g) / boost::get (boost::vertex_color, g) ---------------------
// Lvalue property maps over the graph's own storage (writes mark the host mirror dirty: the next device stage uploads
// it).  Indexable by descriptor (the BGL way) and, for edges, by edge id.
template <typename G>
class edge_weight_map {
 public:
  typedef float value_type;
  typedef float& reference;
  typedef edge_descriptor key_type;
  explicit edge_weight_map(G* g) : g_(g) {}
  float& operator[](const edge_descriptor& e) const { return g_->edge_weights()[e.idx]; }
  float& operator[](size_t edge_id) const { return g_->edge_weights()[edge_id]; }
  size_t size() const { return g_->num_edges(); }
  G* graph() const { return g_; }

 private:
  G* g_;
};
template <typename G>
class const_edge_weight_map {
 public:
  typedef float value_type;
  typedef const float& reference;
  typedef edge_descriptor key_type;
  explicit const_edge_weight_map(const G* g) : g_(g) {}
  const float& operator[](const edge_descriptor& e) const { return g_->edge_weights()[e.idx]; }
  const float& operator[](size_t edge_id) const { return g_->edge_weights()[edge_id]; }
  size_t size() const { return g_->num_edges(); }
  const G* graph() const { return g_; }

 private:
  const G* g_;
};
template <typename G>
class vertex_color_map {
 public:
  typedef uint32_t value_type;
  typedef uint32_t& reference;
  typedef uint32_t key_type;
  explicit vertex_color_map(G* g) : g_(g) {}
  uint32_t& operator[](uint32_t v) const { return g_->vertex_colors()[v]; }
  size_t size() const { return g_->num_vertices(); }
  G* graph() const { return g_; }

 private:
  G* g_;
};
template <typename G>
class const_vertex_color_map {
 public:
  typedef uint32_t value_type;
  typedef const uint32_t& reference;
  typedef uint32_t key_type;
  explicit const_vertex_color_map(const G* g) : g_(g) {}
  const uint32_t& operator[](uint32_t v) const { return g_->vertex_colors()[v]; }
  size_t size() const { return g_->num_vertices(); }
  const G* graph() const { return g_; }

 private:
  const G* g_;
};
// edge_index: the edge id (the reference's boost::subgraph keeps such a map, random_walker_segmentation.h:128-135)
template <typename G>
class edge_index_map {
 public:
  typedef int value_type;
  typedef int reference;
  typedef edge_descriptor key_type;
  int operator[](const edge_descriptor& e) const { return (int)e.idx; }
};
// vertex_index: identity
class vertex_index_map {
 public:
  typedef uint32_t value_type;
  typedef uint32_t reference;
  typedef uint32_t key_type;
  uint32_t operator[](uint32_t v) const { return v; }
};

}  // namespace detail
}  // namespace tcs
