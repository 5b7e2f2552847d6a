// tests/cpp/graph_io_check.cpp - saveGraph / loadGraph (include/tcs/io.h) and the createSubgraphsFrom* helpers
// (include/tcs/graph/common.h) on a host-built graph; with the argument "device" also on a graph that lives on the GPU
// (voxel builder + edge weights, saved, loaded back, segmented from the same seeds: same colours).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "tcs/graph/common.h"
#include "tcs/graph/edge_weight_computer.h"
#include "tcs/graph/voxel_grid_graph_builder.h"
#include "tcs/io.h"
#include "tcs/random_walker.h"

typedef pcl::PointXYZRGBNormal PointT;
typedef pcl::graph::point_cloud_graph<PointT> Graph;

#define REQUIRE(c) do { if (!(c)) { std::fprintf(stderr, "graph_io_check: %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static bool close_rel(float a, float b, float rel) { return std::fabs(a - b) <= rel * std::fmax(std::fabs(a), std::fabs(b)) + 1e-30f; }

int main(int argc, char** argv) {
  const std::string dir = argc > 1 ? argv[1] : ".";
  const bool device = argc > 2 && std::string(argv[2]) == "device";
  // ---- host graph: 7 vertices, 8 edges, awkward values
  pcl::PointCloud<PointT>::Ptr cloud(new pcl::PointCloud<PointT>);
  cloud->resize(7);
  for (int v = 0; v < 7; ++v) {
    PointT& p = cloud->points[v];
    p.x = 0.1f * v + 1e-7f; p.y = -3.25f + v * v; p.z = 1.0f / (v + 3);
    p.data_n[0] = 0.6f; p.data_n[1] = -0.8f * (v & 1); p.data_n[2] = 1e-9f * v; p.data_n[3] = 0.f;
    p.curvature = v == 3 ? -0.125f : 0.01f * v;
    p.rgba = 0xFF000000u | (uint32_t)(v * 0x102030);
  }
  Graph g(cloud);
  const int es[8] = {0, 1, 2, 3, 4, 5, 6, 2}, et[8] = {1, 2, 3, 4, 5, 6, 0, 5};
  const float ew[8] = {1.f, 0.5f, 1e-5f, 0.333333343f, 123456.789f, 7.25e-12f, 0.f, 0.999999f};
  for (int e = 0; e < 8; ++e) g.add_edge(es[e], et[e], ew[e]);
  const std::string path = dir + "/graph.pcd";
  REQUIRE(saveGraph(path, g));
  Graph h;
  REQUIRE(loadGraph(path, h));
  REQUIRE(h.num_vertices() == 7 && h.num_edges() == 8);
  for (int v = 0; v < 7; ++v) {
    const PointT &a = g[v], &b = h[v];
    REQUIRE(a.x == b.x && a.y == b.y && a.z == b.z && a.curvature == b.curvature && a.rgba == b.rgba);
    for (int k = 0; k < 3; ++k) REQUIRE(a.data_n[k] == b.data_n[k]);
  }
  for (int e = 0; e < 8; ++e) {
    REQUIRE(h.edge_sources()[e] == (uint32_t)es[e] && h.edge_targets()[e] == (uint32_t)et[e]);
    REQUIRE(close_rel(h.edge_weights()[e], ew[e], 1e-5f));  // operator<< prints 6 significant digits
  }
  REQUIRE(h.out_rowptr() == g.out_rowptr() && h.out_targets() == g.out_targets() && h.out_edge_ids() == g.out_edge_ids());
  REQUIRE(!loadGraph(dir + "/does_not_exist.pcd", h));
  // a hand-written tail: comments and blank lines are skipped (impl/io.hpp:84-85)
  {
    FILE* f = std::fopen((dir + "/tail.pcd").c_str(), "w");
    std::fputs("VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 3\nHEIGHT 1\nPOINTS 3\nDATA ascii\n"
               "0 0 1\n1 0 1\n0 1 1\n# Edges\n0 1 0.5\n\n# a comment\n1 2 0.25\n", f);
    std::fclose(f);
    Graph t;
    REQUIRE(loadGraph(dir + "/tail.pcd", t));
    REQUIRE(t.num_vertices() == 3 && t.num_edges() == 2 && t.edge_weights()[1] == 0.25f && t.edge_targets()[1] == 2u);
  }
  // ---- subgraph helpers
  using namespace pcl::graph;
  std::vector<ComponentView> sub;
  const std::vector<uint32_t> colours = {5, 2, 5, 9, 2, 2, 5};
  REQUIRE(createSubgraphsFromColorMap(g, colours, sub) == 3);
  REQUIRE((sub[0].vertices == std::vector<uint32_t>{1, 4, 5}) && (sub[1].vertices == std::vector<uint32_t>{0, 2, 6}) &&
          (sub[2].vertices == std::vector<uint32_t>{3}));
  REQUIRE((inducedEdges(g, sub[0]) == std::vector<uint32_t>{4}));       // edge 4 = (4, 5)
  REQUIRE((inducedEdges(g, sub[1]) == std::vector<uint32_t>{6}));       // edge 6 = (6, 0)
  REQUIRE(inducedEdges(g, sub[2]).empty());
  pcl::PointIndices pi;
  pi.indices = {6, 1, 3};
  createSubgraphsFromIndices(g, pi, sub);
  REQUIRE(sub.size() == 2 && (sub[0].vertices == std::vector<uint32_t>{6, 1, 3}) && (sub[1].vertices == std::vector<uint32_t>{0, 2, 4, 5}));
  std::vector<pcl::PointIndices> many(2);
  many[0].indices = {2, 0};
  many[1].indices = {5};
  createSubgraphsFromIndices(g, many, sub);
  REQUIRE(sub.size() == 3 && (sub[0].vertices == std::vector<uint32_t>{2, 0}) && (sub[1].vertices == std::vector<uint32_t>{5}) &&
          (sub[2].vertices == std::vector<uint32_t>{1, 3, 4, 6}));
  many.clear();
  createSubgraphsFromIndices(g, many, sub);
  REQUIRE(sub.size() == 1 && sub[0].vertices.size() == 7);

  if (device) {
    // a bumpy sheet, voxelised and weighted on the GPU; saved, loaded, and both graphs segmented from the same seeds
    pcl::PointCloud<pcl::PointXYZRGB>::Ptr in(new pcl::PointCloud<pcl::PointXYZRGB>);
    for (int y = 0; y < 120; ++y)
      for (int x = 0; x < 160; ++x) {
        pcl::PointXYZRGB p;
        p.z = 1.0f + 0.05f * std::sin(0.07f * x) * std::cos(0.05f * y) + (x > 80 ? 0.15f : 0.f);
        p.x = (x - 80) * 0.002f * p.z; p.y = (y - 60) * 0.002f * p.z;
        p.rgba = x > 80 ? 0xFFC02020u : 0xFF2020C0u;
        in->push_back(p);
      }
    Graph dg;
    VoxelGridGraphBuilder<pcl::PointXYZRGB, Graph> gb(0.006f);
    gb.setInputCloud(in);
    gb.compute(dg);
    computeNormalsAndCurvatures(dg);
    computeSignedCurvatures(dg);
    typedef EdgeWeightComputer<Graph> EWC;
    EWC wc;
    wc.addTerm<terms::XYZ>(3.0f, EWC::NORMALIZATION_GLOBAL);
    wc.addTerm<terms::Normal>(0.01f, 0.0f);
    wc.addTerm<terms::RGB>(3.0f, EWC::NORMALIZATION_GLOBAL);
    wc.setSmallWeightThreshold(1e-5f);
    wc.setSmallWeightPolicy(EWC::SMALL_WEIGHT_COERCE_TO_THRESHOLD);
    wc.compute(dg);
    REQUIRE(saveGraph(dir + "/device_graph.pcd", dg));
    Graph lg;
    REQUIRE(loadGraph(dir + "/device_graph.pcd", lg));
    REQUIRE(lg.num_vertices() == dg.num_vertices() && lg.num_edges() == dg.num_edges() && lg.num_edges() > 1000);
    REQUIRE(lg.edge_sources() == dg.edge_sources() && lg.edge_targets() == dg.edge_targets());
    for (size_t v = 0; v < dg.num_vertices(); ++v) REQUIRE(dg[v].x == lg[v].x && dg[v].normal_z == lg[v].normal_z && dg[v].rgba == lg[v].rgba);
    const size_t V = dg.num_vertices();
    std::vector<uint32_t>& c1 = dg.vertex_colors();
    std::vector<uint32_t>& c2 = lg.vertex_colors();
    for (size_t v = 0; v < V; ++v) c1[v] = c2[v] = 0;
    size_t left = 0, right = 0;
    for (size_t v = 0; v < V; ++v) {
      if (dg[v].rgba == 0xFF2020C0u && left < 3 && v % 50 == 7) { c1[v] = c2[v] = 1; ++left; }
      if (dg[v].rgba == 0xFFC02020u && right < 3 && v % 50 == 11) { c1[v] = c2[v] = 2; ++right; }
    }
    REQUIRE(left && right);
    REQUIRE(pcl::segmentation::randomWalker(dg));
    REQUIRE(pcl::segmentation::randomWalker(lg));
    size_t diff = 0;
    for (size_t v = 0; v < V; ++v) diff += dg.vertex_colors()[v] != lg.vertex_colors()[v];
    std::printf("graph_io_check device: V=%zu E=%zu differing labels after the 6-digit weight round trip: %zu\n", V, dg.num_edges(), diff);
    REQUIRE(diff <= V / 200);
  }
  std::printf("graph_io_check ok\n");
  return 0;
}
