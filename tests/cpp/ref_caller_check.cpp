// tests/cpp/ref_caller_check.cpp - code written the way callers of the REFERENCE write it, compiled against include/tcs:
//   * the edge loop of saveGraph (include/impl/io.hpp:31-39): as_range (boost::edges (graph)), boost::source / target,
//     weights[edge] with weights = boost::get (boost::edge_weight, graph);
//   * the loops of include/conversions.h:140-159: as_range (boost::vertices (graph)), boost::tie (ei, ee) = boost::edges,
//     boost::out_degree (vertex, graph), boost::edge (u, v, graph).second, boost::add_edge;
//   * randomWalker (graph, weights, colors) and the 5-argument form (include/random_walker.h:110-137) with external maps
//     and sparse colour ids; more than 64 colours fails cleanly;
//   * RandomWalkerSegmentation<PointXYZ | PointXYZRGB | PointXYZRGBA> (src/random_walker_segmentation.cpp:44), a NaN seed;
//   * EdgeWeightComputer::setTermBalancingFunction (graph/edge_weight_computer.h:325).
// Prints "ref_caller_check ok" and returns 0 when every REQUIRE holds.  Needs a GPU (the solves run in libtcs_b200).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <map>
#include <sstream>
#include <string>

#include "tcs/random_walker_segmentation.h"

#define REQUIRE(c) do { if (!(c)) { std::fprintf(stderr, "ref_caller_check: %s:%d: REQUIRE(%s) failed\n", __FILE__, __LINE__, #c); return 1; } } while (0)

// a range over an iterator pair, as callers of the reference use with range-for (their include/as_range.h)
template <class It>
struct pair_range {
  std::pair<It, It> p;
  It begin() const { return p.first; }
  It end() const { return p.second; }
};
template <class It> pair_range<It> as_range(const std::pair<It, It>& p) { pair_range<It> r = {p}; return r; }

typedef pcl::PointXYZRGBNormal VertexPoint;
typedef pcl::graph::point_cloud_graph<VertexPoint> Graph;
typedef boost::graph_traits<Graph>::vertex_descriptor VertexId;
typedef boost::graph_traits<Graph>::edge_descriptor EdgeId;
typedef boost::graph_traits<Graph>::edge_iterator EdgeIterator;
typedef boost::graph_traits<Graph>::out_edge_iterator OutEdgeIterator;

// an external edge-weight property map (ReadablePropertyMap over edge descriptors)
struct ExternalWeights {
  typedef float value_type;
  typedef float reference;
  typedef EdgeId key_type;
  std::map<EdgeId, float> w;
  float operator[](const EdgeId& e) const { return w.find(e)->second; }
};

// ladder graph: two rails of n vertices (0..n-1 and n..2n-1) with strong rail edges and weak rungs
static void make_ladder(Graph& graph, int n, float rail, float rung) {
  for (int i = 0; i + 1 < n; ++i) {
    boost::add_edge(i, i + 1, graph);
    boost::add_edge(n + i, n + i + 1, graph);
  }
  for (int i = 0; i < n; ++i)
    if (!boost::edge(i, n + i, graph).second) boost::add_edge(i, n + i, graph);
  auto weights = boost::get(boost::edge_weight, graph);
  for (const auto& edge : as_range(boost::edges(graph))) {
    const bool is_rung = boost::target(edge, graph) == boost::source(edge, graph) + (VertexId)n;
    weights[edge] = is_rung ? rung : rail;
  }
}

template <typename PointT>
static int run_class(const char* name, bool nan_seed, std::vector<int>& cluster_of) {
  typedef pcl::segmentation::RandomWalkerSegmentation<PointT> RWS;
  typedef std::integral_constant<bool, tcs::has_color<PointT>::value> colored;
  // two coplanar patches side by side with a 2.4 cm gap in x (two components at 6 mm voxels), distinct colours
  typename pcl::PointCloud<PointT>::Ptr cloud(new pcl::PointCloud<PointT>);
  for (int iy = 0; iy < 40; ++iy)
    for (int ix = 0; ix < 80; ++ix) {
      PointT p;
      p.x = 0.003f * (ix - 40) + (ix >= 40 ? 0.024f : 0.0f);
      p.y = 0.003f * iy;
      p.z = 1.0f + 0.0005f * ((ix * 7 + iy * 3) % 5);
      // (textured: with one flat colour per patch the globally normalised RGB term would be 0 / 0, in the reference too)
      tcs::detail::set_rgba(p, (ix < 40 ? 0xFFC82828u : 0xFF2828C8u) + (uint32_t)((ix * 5 + iy * 11) % 17) * 0x000100u, colored());
      cloud->push_back(p);
    }
  pcl::PointCloud<pcl::PointXYZL>::Ptr seeds(new pcl::PointCloud<pcl::PointXYZL>);
  pcl::PointXYZL s;
  s.x = -0.06f; s.y = 0.06f; s.z = 1.0f; s.label = 7; seeds->push_back(s);
  s.x = 0.084f; s.y = 0.06f; s.z = 1.0f; s.label = 3; seeds->push_back(s);
  if (nan_seed) { s.x = std::numeric_limits<float>::quiet_NaN(); s.label = 3; seeds->push_back(s); }
  RWS rws;
  rws.setInputCloud(cloud);
  rws.setSeeds(seeds);
  std::vector<pcl::PointIndices> clusters;
  rws.segment(clusters);
  REQUIRE(clusters.size() == 3);   // labels 3, 7 (ascending) + the unlabelled rest
  cluster_of.assign(cloud->size(), -1);
  for (size_t c = 0; c < clusters.size(); ++c)
    for (size_t k = 0; k < clusters[c].indices.size(); ++k) cluster_of[clusters[c].indices[k]] = (int)c;
  // the left patch carries label 7 (cluster 1), the right one label 3 (cluster 0), nothing is left over
  size_t wrong = 0;
  for (int iy = 0; iy < 40; ++iy)
    for (int ix = 0; ix < 80; ++ix) wrong += cluster_of[iy * 80 + ix] != (ix < 40 ? 1 : 0);
  std::printf("  RandomWalkerSegmentation<%s>%s: %zu vertices, %zu clusters, %zu points off\n", name, nan_seed ? " + NaN seed" : "",
              rws.getGraph()->num_vertices(), clusters.size(), wrong);
  REQUIRE(wrong == 0);
  return 0;
}

int main() {
  try {
    // ---------------------------------------------------------------- graph surface on a host-built graph
    const int n = 6;
    Graph graph(2 * n);
    for (const auto& vertex : as_range(boost::vertices(graph))) {
      graph[vertex].x = 0.01f * (float)(vertex % n);
      graph[vertex].y = vertex < (VertexId)n ? 0.0f : 0.01f;
      graph[vertex].z = 1.0f;
    }
    make_ladder(graph, n, 1.0f, 0.01f);
    REQUIRE(boost::num_vertices(graph) == (size_t)(2 * n) && boost::num_edges(graph) == (size_t)(3 * n - 2));
    // the saveGraph loop (include/impl/io.hpp:31-39)
    std::ostringstream file;
    file << "# Edges\n";
    auto weights = boost::get(boost::edge_weight, graph);
    for (const auto& edge : as_range(boost::edges(graph))) {
      auto src = boost::source(edge, graph);
      auto tgt = boost::target(edge, graph);
      auto wgh = weights[edge];
      file << src << " " << tgt << " " << wgh << "\n";
    }
    const std::string text = file.str();
    REQUIRE(text.rfind("# Edges\n0 1 1\n6 7 1\n1 2 1\n7 8 1\n", 0) == 0);
    REQUIRE(text.find("\n0 6 0.01\n") != std::string::npos);
    // boost::tie + explicit iterators (include/conversions.h:144-157), degrees, out-edges, adjacency
    EdgeIterator ei, ee;
    size_t count = 0;
    float sum = 0.f;
    for (boost::tie(ei, ee) = boost::edges(graph); ei != ee; ++ei) {
      const VertexId& i1 = boost::source(*ei, graph);
      const VertexId& i2 = boost::target(*ei, graph);
      REQUIRE(i1 != i2);
      sum += boost::get(boost::edge_weight, graph, *ei);
      ++count;
    }
    REQUIRE(count == boost::num_edges(graph) && std::fabs(sum - (2.0f * (n - 1) + 0.01f * n)) < 1e-5f);
    REQUIRE(boost::out_degree(0, graph) == 2 && boost::out_degree(1, graph) == 3 && boost::degree(n, graph) == 2);
    OutEdgeIterator oi, oe;
    size_t seen = 0;
    for (boost::tie(oi, oe) = boost::out_edges(1, graph); oi != oe; ++oi) {
      REQUIRE(boost::source(*oi, graph) == 1u);            // out-edges are oriented away from the vertex
      const VertexId t = boost::target(*oi, graph);
      REQUIRE(t == 0u || t == 2u || t == (VertexId)(n + 1));
      ++seen;
    }
    REQUIRE(seen == 3);
    seen = 0;
    for (const auto& nb : as_range(boost::adjacent_vertices(n, graph))) { REQUIRE(nb == (VertexId)(n + 1) || nb == 0u); ++seen; }
    REQUIRE(seen == 2);
    REQUIRE(boost::edge(2, n + 2, graph).second && boost::edge(n + 2, 2, graph).second && !boost::edge(0, n + 1, graph).second);
    REQUIRE(boost::get(boost::edge_index, graph)[boost::edge(0, 1, graph).first] == 0);

    // ---------------------------------------------------------------- randomWalker, 3- and 5-argument forms
    // external maps; sparse colour ids: seed 1000 on the lower rail's left end, seed 5 on the upper rail's right end
    ExternalWeights ext;
    for (const auto& edge : as_range(boost::edges(graph))) ext.w[edge] = weights[edge];
    std::vector<uint32_t> colors(2 * n, 0u);
    colors[0] = 1000u;
    colors[2 * n - 1] = 5u;
    REQUIRE(pcl::segmentation::randomWalker(graph, ext, &colors[0]));
    for (int i = 0; i < n; ++i) REQUIRE(colors[i] == 1000u && colors[n + i] == 5u);   // weak rungs: each rail follows its seed
    // the graph's own maps give the same answer
    auto own_colors = boost::get(boost::vertex_color, graph);
    for (const auto& vertex : as_range(boost::vertices(graph))) own_colors[vertex] = 0u;
    own_colors[0] = 1000u;
    own_colors[2 * n - 1] = 5u;
    REQUIRE(pcl::segmentation::randomWalker(graph, boost::get(boost::edge_weight, graph), own_colors));
    for (int i = 0; i < 2 * n; ++i) REQUIRE(boost::get(boost::vertex_color, graph)[i] == colors[i]);
    // 5 arguments: potentials (rows = vertices, one column per colour) and the colour -> column map
    std::vector<uint32_t> colors2(2 * n, 0u);
    colors2[0] = 1000u;
    colors2[2 * n - 1] = 5u;
    Eigen::MatrixXf potentials;
    std::map<uint32_t, size_t> colors_to_columns_map;
    REQUIRE(pcl::segmentation::randomWalker(graph, ext, &colors2[0], potentials, colors_to_columns_map));
    REQUIRE(colors_to_columns_map.size() == 2 && colors_to_columns_map[5u] == 0 && colors_to_columns_map[1000u] == 1);
    REQUIRE(potentials.rows() == 2 * n && potentials.cols() == 2);
    REQUIRE(potentials(0, 1) == 1.0f && potentials(0, 0) == 0.0f && potentials(2 * n - 1, 0) == 1.0f);   // seeds are one-hot
    for (int v = 0; v < 2 * n; ++v) {
      REQUIRE(std::fabs(potentials(v, 0) + potentials(v, 1) - 1.0f) < 1e-4f);
      REQUIRE((potentials(v, 1) > potentials(v, 0)) == (colors2[v] == 1000u));
    }
    // more distinct colours than the implementation solves: fails cleanly (false, message on stderr), no exception
    {
      Graph big(200);
      for (int i = 0; i + 1 < 200; ++i) boost::add_edge(i, i + 1, big);
      auto w = boost::get(boost::edge_weight, big);
      for (const auto& edge : as_range(boost::edges(big))) w[edge] = 1.0f;
      auto c = boost::get(boost::vertex_color, big);
      for (int i = 0; i < 200; i += 2) c[i] = 10u + (uint32_t)i;   // 100 colours
      REQUIRE(!pcl::segmentation::randomWalker(big));
    }

    // ---------------------------------------------------------------- the class for every instantiated point type
    std::vector<int> c_rgb, c_rgba, c_xyz, c_nan;
    if (run_class<pcl::PointXYZRGB>("PointXYZRGB", false, c_rgb)) return 1;
    if (run_class<pcl::PointXYZRGBA>("PointXYZRGBA", false, c_rgba)) return 1;
    if (run_class<pcl::PointXYZ>("PointXYZ", false, c_xyz)) return 1;
    if (run_class<pcl::PointXYZRGB>("PointXYZRGB", true, c_nan)) return 1;
    REQUIRE(c_rgb == c_rgba && c_rgb == c_nan);

    // ---------------------------------------------------------------- setTermBalancingFunction
    {
      typedef pcl::graph::EdgeWeightComputer<Graph> EWC;
      float ref_w = 0.f;
      for (int variant = 0; variant < 3; ++variant) {
        Graph g2(2 * n);
        for (const auto& vertex : as_range(boost::vertices(g2))) {
          g2[vertex].x = 0.01f * (float)(vertex % n);
          g2[vertex].y = vertex < (VertexId)n ? 0.0f : 0.01f;
          g2[vertex].z = 1.0f;
        }
        make_ladder(g2, n, 1.0f, 1.0f);
        EWC wc;
        wc.addTerm<pcl::graph::terms::XYZ>(1e-4f);           // d = |dp|^2 = 1e-4 on every edge -> value / influence = 1
        if (variant == 1) wc.setTermBalancingFunction(EWC::BALANCING_INVERSE);
        if (variant == 2) wc.setTermBalancingFunction(pcl::graph::balancing::Linear());
        wc.compute(g2);
        const float w0 = boost::get(boost::edge_weight, g2)[0];
        if (variant == 0) { ref_w = w0; REQUIRE(std::fabs(w0 - std::exp(-1.0f)) < 1e-5f); }
        if (variant == 1) REQUIRE(std::fabs(w0 - 0.5f) < 1e-5f && w0 != ref_w);
        if (variant == 2) REQUIRE(std::fabs(w0) < 1e-5f);
      }
      EWC wc;
      bool threw = false;
      try { wc.setTermBalancingFunction([](float v, float s) { return v * s; }); } catch (const pcl::PCLException&) { threw = true; }
      REQUIRE(threw);   // an arbitrary host callable cannot run on the device
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "ref_caller_check: exception: %s\n", e.what());
    return 3;
  }
  std::printf("ref_caller_check ok\n");
  return 0;
}
