// tcs_b200/cpp/random_walker_segmentation_app.cpp - headless equivalent of the reference's
// app_random_walker_segmentation (src/random_walker_segmentation_app.cpp:37-296) on the GPU path.
//
//   tcs_random_walker_segmentation <pcd-file> --load-seeds <pcd-file> [--save] [--save-clusters] [--potential]
//       [--builder vg|nnk|nnr] [-v res] [--nn k] [--radius r] [-z xyz] [-n normal] [-b curvature] [-c colour] [--weight-threshold t] [--convex-multiplier m]
//       [--out segmentation.pcd] [--format ascii|binary|binary_compressed]
//
// Same pipeline and option names / defaults as the reference (factory/edge_weight_computer_factory.h:20-46,
// factory/graph_builder_factory.h:22-45): voxel-grid graph, normals, smoothen(0.012, 0.0012), normals, curvature
// signs, edge weights {XYZ global, Normal, Curvature, RGB local; coerce small weights}, random walker from the loaded
// seeds; --save writes segmentation.pcd (labels = vertex colours, 0 for points without a vertex), --save-clusters
// writes cluster%i.pcd.  There is no viewer on this box: interactive seed selection (and therefore --save-seeds) is
// unavailable, the tool always behaves like --tv-no-gui, which the reference only allows together with --load-seeds
// (:86-90).  --builder selects the voxel grid (default), the k-nearest (--nn, default 14) or the radius (--radius,
// default 0.006) graph builder, as factory/graph_builder_factory.h:22-60 does.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <memory>
#include <string>

#include "tcs/evaluation.h"
#include "tcs/graph/nearest_neighbors_graph_builder.h"
#include "tcs/pcd_io.h"
#include "tcs/random_walker_segmentation.h"

typedef pcl::PointXYZRGB PointT;
typedef pcl::segmentation::RandomWalkerSegmentation<PointT> RWS;

static bool has_switch(int argc, char** argv, const char* name) {
  for (int i = 1; i < argc; ++i)
    if (!std::strcmp(argv[i], name)) return true;
  return false;
}
static const char* value_of(int argc, char** argv, const char* name) {
  for (int i = 1; i + 1 < argc; ++i)
    if (!std::strcmp(argv[i], name)) return argv[i + 1];
  return nullptr;
}
static float float_opt(int argc, char** argv, const char* name, float dflt) {
  const char* v = value_of(argc, argv, name);
  return v ? (float)std::atof(v) : dflt;
}

struct Stopwatch {  // the reference's MEASURE_RUNTIME (include/measure_runtime.h:7-14)
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  void lap(const char* what) {
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%s[%.0f ms]\n", what, ms);
    t0 = std::chrono::steady_clock::now();
  }
};

int main(int argc, char** argv) {
  if (argc < 2 || has_switch(argc, argv, "--help")) {
    std::fprintf(stderr,
                 "Usage: %s <pcd-file>\n--load-seeds <pcd-file>\n--save\n--save-clusters\n--potential\n"
                 "--out <pcd-file> (default segmentation.pcd)\n--format ascii|binary|binary_compressed\n"
                 "Edge Weight Computer: -z <xyz influence 3.0> -n <normal influence 0.01> -b <curvature influence 0.0001>\n"
                 "  -c <color influence 3.0> --weight-threshold <1e-5> --convex-multiplier <0.01>\n"
                 "Graph Builder: --builder vg|nnk|nnr -v <voxel resolution 0.006> --nn <number of neighbors 14>\n"
                 "  --radius <sphere radius 0.006>\n",
                 argv[0]);
    return 1;
  }
  const char* seeds_file = value_of(argc, argv, "--load-seeds");
  if (!seeds_file) {
    std::fprintf(stderr, "No GUI mode can only be used with --load-seeds option.\n");
    return 2;
  }
  const std::string builder = value_of(argc, argv, "--builder") ? value_of(argc, argv, "--builder") : "vg";
  if (builder != "vg" && builder != "nnk" && builder != "nnr") {
    std::fprintf(stderr, "unknown builder type '%s' (vg, nnk, nnr)\n", builder.c_str());
    return 2;
  }
  const int number_of_neighbors = (int)float_opt(argc, argv, "--nn", 14.0f);
  const float radius = float_opt(argc, argv, "--radius", 0.006f);
  const bool option_save = has_switch(argc, argv, "--save"), option_save_clusters = has_switch(argc, argv, "--save-clusters");
  const bool mode_potential = has_switch(argc, argv, "--potential");
  const std::string out_name = value_of(argc, argv, "--out") ? value_of(argc, argv, "--out") : "segmentation.pcd";
  const std::string format = value_of(argc, argv, "--format") ? value_of(argc, argv, "--format") : "ascii";
  const float resolution = float_opt(argc, argv, "-v", 0.006f);
  const float xyz = float_opt(argc, argv, "-z", 3.0f), normal = float_opt(argc, argv, "-n", 0.01f);
  const float curvature = float_opt(argc, argv, "-b", 0.0001f), color = float_opt(argc, argv, "-c", 3.0f);
  const float threshold = float_opt(argc, argv, "--weight-threshold", 0.00001f);
  const float convex = float_opt(argc, argv, "--convex-multiplier", 0.01f);

  pcl::PointCloud<PointT>::Ptr cloud(new pcl::PointCloud<PointT>);
  pcl::PointCloud<pcl::PointXYZL>::Ptr seeds(new pcl::PointCloud<pcl::PointXYZL>);
  bool input_has_color = false;
  try {
    if (!tcs::pcd::load_cloud(argv[1], *cloud, &input_has_color)) {
      std::fprintf(stderr, "%s has no x/y/z fields\n", argv[1]);
      return 1;
    }
    if (!tcs::pcd::load_labeled(seeds_file, *seeds)) {
      std::fprintf(stderr, "%s has no x/y/z fields\n", seeds_file);
      return 1;
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 1;
  }
  std::printf("Edge Weight Computer: xyz influence %g, normal influence %g, curvature influence %g, color influence %g, "
              "weight threshold %g, convex multiplier %g\n", xyz, normal, curvature, color, threshold, convex);
  if (builder == "vg") std::printf("Graph Builder: VOXEL GRID, voxel resolution %g\n", resolution);
  else if (builder == "nnk") std::printf("Graph Builder: NEAREST NEIGHBORS KNN, number of neighbors %d\n", number_of_neighbors);
  else std::printf("Graph Builder: NEAREST NEIGHBORS RADIUS, sphere radius %g\n", radius);

  try {
    using namespace pcl::graph;
    typedef RWS::Graph Graph;
    RWS::GraphPtr g(new Graph);
    std::unique_ptr<GraphBuilder<PointT, Graph>> gbp;
    if (builder == "vg") {
      gbp.reset(new VoxelGridGraphBuilder<PointT, Graph>(resolution));
    } else {
      NearestNeighborsGraphBuilder<PointT, Graph>* nn = new NearestNeighborsGraphBuilder<PointT, Graph>;
      if (builder == "nnk") { nn->setNumberOfNeighbors(number_of_neighbors); nn->useNearestKSearch(); }
      else { nn->setRadius(radius); nn->useRadiusSearch(); }
      gbp.reset(nn);
    }
    GraphBuilder<PointT, Graph>& gb = *gbp;
    gb.setInputCloud(cloud);
    Stopwatch sw;
    gb.compute(*g);                         sw.lap("Building graph... ");
    computeNormalsAndCurvatures(*g);        sw.lap("Computing normals... ");
    smoothen(*g, 0.012f, 0.0012f);          sw.lap("Smoothening graph... ");
    computeNormalsAndCurvatures(*g);        sw.lap("Computing normals... ");
    computeSignedCurvatures(*g);            sw.lap("Computing curvature signs... ");
    typedef EdgeWeightComputer<Graph> EWC;
    EWC wc;
    wc.addTerm<terms::XYZ>(xyz, EWC::NORMALIZATION_GLOBAL);
    wc.addTerm<terms::Normal>(normal, convex);
    wc.addTerm<terms::Curvature>(curvature, convex);
    if (input_has_color) wc.addTerm<terms::RGB>(color, EWC::NORMALIZATION_LOCAL);
    wc.setSmallWeightThreshold(threshold);
    wc.setSmallWeightPolicy(EWC::SMALL_WEIGHT_COERCE_TO_THRESHOLD);
    wc.compute(*g);                         sw.lap("Computing edge weights... ");
    std::printf("Built a graph with %zu vertices and %zu edges\n", g->num_vertices(), g->num_edges());

    RWS rws(mode_potential);
    rws.setInputGraph(g);
    rws.setSeeds(seeds);
    std::vector<pcl::PointIndices> clusters;
    rws.segment(clusters);                  sw.lap("Random walker segmentation... ");
    const tcs_solve_info& info = rws.getSolveInfo();
    std::printf("Segmented into %zu clusters (%d labels, %d PCG iterations, converged %d)\n", clusters.size(), info.ncolours,
                info.iterations, info.converged);

    // with setInputGraph the clusters index graph vertices; map them back to points for the files
    const std::vector<size_t>& p2v = gb.getPointToVertexMap();
    const std::vector<uint32_t>& colours = g->vertex_colors();
    if (option_save_clusters) {
      const size_t K = clusters.empty() ? 0 : clusters.size() - 1;
      std::vector<pcl::PointCloud<PointT>> parts(K);
      for (size_t i = 0; i < cloud->size(); ++i) {
        const size_t v = p2v[i];
        if (v == std::numeric_limits<size_t>::max() || v >= colours.size() || colours[v] == 0 || colours[v] > K) continue;
        parts[colours[v] - 1].push_back((*cloud)[i]);
      }
      for (size_t c = 0; c < K; ++c)
        if (parts[c].size()) tcs::pcd::save_colored("cluster" + std::to_string(c) + ".pcd", parts[c], format);
    }
    if (option_save) {
      pcl::PointCloud<pcl::PointXYZL>::Ptr labeled = tcs::labels::createLabeledCloudFromColorMap(*cloud, p2v, colours);
      tcs::pcd::save_labeled(out_name, *labeled, format);
      std::printf("Saved %s\n", out_name.c_str());
    }
    return info.converged ? 0 : 4;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
}
