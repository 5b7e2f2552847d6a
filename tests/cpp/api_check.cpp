// tests/cpp/api_check.cpp - exercises the C++ drop-in API (include/tcs/*.h) the way a caller of the reference would:
//   mode "class": RandomWalkerSegmentation<PointXYZRGB>::setInputCloud / setSeeds / segment (+ getPotentials)
//   mode "cli"  : the app pipeline of src/random_walker_segmentation_app.cpp:122-134 + :216-222 written against the
//                 same names (builder, normals, smoothen, normals, signs, EdgeWeightComputer, setInputGraph, segment)
// in.bin : u64 n, f32 xyz[3n], u32 rgba[n], u64 ns, f32 seed_xyz[3ns], u32 seed_label[ns]
// out.bin: i32 cluster_of_point_or_vertex[m], u64 V, u32 vertex_colour[V], u64 K, f32 potentials[V*K] (col-major)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "tcs/random_walker_segmentation.h"

typedef pcl::PointXYZRGB PointT;
typedef pcl::segmentation::RandomWalkerSegmentation<PointT> RWS;

static void die(const char* m) { std::fprintf(stderr, "api_check: %s\n", m); std::exit(2); }

int main(int argc, char** argv) {
  if (argc < 4) die("usage: api_check class|cli in.bin out.bin [resolution]");
  const std::string mode = argv[1];
  const float resolution = argc > 4 ? (float)std::atof(argv[4]) : 0.006f;
  FILE* f = std::fopen(argv[2], "rb");
  if (!f) die("cannot open input");
  uint64_t n = 0, ns = 0;
  if (std::fread(&n, 8, 1, f) != 1) die("short input");
  std::vector<float> xyz(3 * n);
  std::vector<uint32_t> rgba(n);
  if (n && (std::fread(xyz.data(), 12, n, f) != n || std::fread(rgba.data(), 4, n, f) != n)) die("short input");
  if (std::fread(&ns, 8, 1, f) != 1) die("short input");
  std::vector<float> sxyz(3 * ns);
  std::vector<uint32_t> slabel(ns);
  if (ns && (std::fread(sxyz.data(), 12, ns, f) != ns || std::fread(slabel.data(), 4, ns, f) != ns)) die("short input");
  std::fclose(f);

  pcl::PointCloud<PointT>::Ptr cloud(new pcl::PointCloud<PointT>);
  cloud->resize(n);
  for (size_t i = 0; i < n; ++i) {
    PointT& p = cloud->points[i];
    p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2];
    p.rgba = rgba[i];
  }
  pcl::PointCloud<pcl::PointXYZL>::Ptr seeds(new pcl::PointCloud<pcl::PointXYZL>);
  seeds->resize(ns);
  for (size_t i = 0; i < ns; ++i) {
    pcl::PointXYZL& s = seeds->points[i];
    s.x = sxyz[3 * i]; s.y = sxyz[3 * i + 1]; s.z = sxyz[3 * i + 2];
    s.label = slabel[i];
  }

  std::vector<pcl::PointIndices> clusters;
  // TCS_API_POTENTIALS=0: do not keep the potentials (the reference's default); TCS_API_REPEAT=n: time n further
  // segment() calls of the class path on fresh objects (profiles/: e2e through the C++ class)
  const bool keep_potentials = !std::getenv("TCS_API_POTENTIALS") || std::atoi(std::getenv("TCS_API_POTENTIALS")) != 0;
  const int repeat = std::getenv("TCS_API_REPEAT") ? std::atoi(std::getenv("TCS_API_REPEAT")) : 0;
  RWS rws(keep_potentials);
  size_t m = n;
  try {
    if (mode == "class") {
      rws.setVoxelResolution(resolution);
      rws.setInputCloud(cloud);
      rws.setSeeds(seeds);
      rws.segment(clusters);
      double best = 1e300, sum = 0.0;
      for (int r = 0; r < repeat; ++r) {
        std::vector<pcl::PointIndices> again;
        const std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
        RWS timed(keep_potentials);
        timed.setVoxelResolution(resolution);
        timed.setInputCloud(cloud);
        timed.setSeeds(seeds);
        timed.segment(again);
        const double ms = 1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        best = ms < best ? ms : best;
        sum += ms;
      }
      if (repeat > 0)
        std::printf("api_check class timing: %d x (construct + setInputCloud + setSeeds + segment(clusters)), host cloud in, clusters on the "
                    "host out: best %.2f ms, mean %.2f ms, potentials %s\n", repeat, best, sum / repeat, keep_potentials ? "kept" : "not kept");
    } else {
      using namespace pcl::graph;
      typedef RWS::Graph Graph;
      RWS::GraphPtr g(new Graph);
      VoxelGridGraphBuilder<PointT, Graph> gb(resolution);
      gb.setInputCloud(cloud);
      gb.compute(*g);
      computeNormalsAndCurvatures(*g);
      smoothen(*g, 0.012f, 0.0012f);
      computeNormalsAndCurvatures(*g);
      computeSignedCurvatures(*g);
      typedef EdgeWeightComputer<Graph> EWC;
      EWC wc;  // factory defaults, include/factory/edge_weight_computer_factory.h:41-46
      wc.addTerm<terms::XYZ>(3.0f, EWC::NORMALIZATION_GLOBAL);
      wc.addTerm<terms::Normal>(0.01f, 0.01f);
      wc.addTerm<terms::Curvature>(0.0001f, 0.01f);
      wc.addTerm<terms::RGB>(3.0f, EWC::NORMALIZATION_LOCAL);
      wc.setSmallWeightThreshold(1e-5f);
      wc.setSmallWeightPolicy(EWC::SMALL_WEIGHT_COERCE_TO_THRESHOLD);
      wc.compute(*g);
      rws.setInputGraph(g);
      rws.setSeeds(seeds);
      rws.segment(clusters);
      m = g->num_vertices();
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "api_check: exception: %s\n", e.what());
    return 3;
  }
  std::vector<int32_t> cluster_of(m, -1);
  for (size_t c = 0; c < clusters.size(); ++c)
    for (size_t k = 0; k < clusters[c].indices.size(); ++k) cluster_of[clusters[c].indices[k]] = (int32_t)c;
  const RWS::Graph& g = *rws.getGraph();
  const std::vector<uint32_t>& colours = g.vertex_colors();
  const Eigen::MatrixXf& pot = rws.getPotentials();
  const uint64_t V = colours.size(), K = (uint64_t)pot.cols();
  FILE* o = std::fopen(argv[3], "wb");
  if (!o) die("cannot open output");
  std::fwrite(cluster_of.data(), 4, m, o);
  std::fwrite(&V, 8, 1, o);
  std::fwrite(colours.data(), 4, V, o);
  std::fwrite(&K, 8, 1, o);
  std::fwrite(pot.data(), 4, V * K, o);
  std::fclose(o);
  const tcs_solve_info& info = rws.getSolveInfo();
  std::printf("api_check %s: n=%zu V=%zu E=%zu clusters=%zu iterations=%d converged=%d\n", mode.c_str(), (size_t)n,
              (size_t)V, g.num_edges(), clusters.size(), info.iterations, info.converged);
  return info.converged ? 0 : 4;
}
