// include/tcs/random_walker_segmentation.h - pcl::segmentation::RandomWalkerSegmentation<PointT> on the GPU.
// Drop-in for the reference class (include/random_walker_segmentation.h:87-291, impl/...hpp:53-241): same public
// methods, same defaults (voxel resolution 0.006, weight terms of preComputeGraph :120-131), same cluster layout
// (one cluster per seed label in ascending label order, plus a last cluster of unlabelled points).
// Everything below the signatures runs in libtcs_b200.so; without a CUDA device construction throws.
#pragma once

#include <set>

#include "graph/common.h"
#include "graph/edge_weight_computer.h"
#include "graph/voxel_grid_graph_builder.h"
#include "random_walker.h"

namespace pcl {
namespace segmentation {

template <typename PointT>
class RandomWalkerSegmentation : public pcl::PCLBase<PointT> {
  using pcl::PCLBase<PointT>::input_;
  using pcl::PCLBase<PointT>::initCompute;

 public:
  typedef PointT Point;
  typedef pcl::PointCloud<Point> PointCloud;
  typedef typename PointCloud::Ptr PointCloudPtr;
  typedef typename PointCloud::ConstPtr PointCloudConstPtr;
  typedef typename tcs::with_normal<PointT>::type PointWithNormal;
  typedef pcl::PointCloud<PointWithNormal> PointWithNormalCloud;
  typedef pcl::graph::point_cloud_graph<PointWithNormal> Graph;
  typedef tcs::shared_ptr<Graph> GraphPtr;
  typedef tcs::shared_ptr<const Graph> GraphConstPtr;
  typedef pcl::graph::ComponentView GraphRef;

  explicit RandomWalkerSegmentation(bool store_potentials = false)
      : input_as_cloud_(true), graph_builder_(0.006f), store_potentials_(store_potentials), mode_(TCS_MODE_PARITY) {
    tcs_solver_opts_default(&solver_opts_);
    std::memset(&info_, 0, sizeof info_);
  }
  virtual ~RandomWalkerSegmentation() {}

  virtual void setInputCloud(const PointCloudConstPtr& cloud) {
    pcl::PCLBase<PointT>::setInputCloud(cloud);
    input_as_cloud_ = true;
    graph_components_.clear();
  }
  void setInputGraph(const GraphPtr& graph) {
    graph_ = graph;
    input_as_cloud_ = false;
    graph_components_.clear();
  }
  void setSeeds(const typename pcl::PointCloud<PointXYZL>::ConstPtr& seeds) {
    seeds_ = seeds;
    std::set<uint32_t> labels;
    for (size_t i = 0; i < seeds_->size(); ++i) labels.insert(seeds_->at(i).label);
    label_color_.clear();
    for (std::set<uint32_t>::iterator it = labels.begin(); it != labels.end(); ++it) {
      const uint32_t c = (uint32_t)label_color_.size() + 1;
      label_color_[*it] = c;
    }
  }

  void preComputeGraph() {
    if (graph_components_.size()) return;
    if (input_as_cloud_) {
      if (!initCompute()) throw pcl::ComputeFailedException("[pcl::segmentation::RandomWalkerSegmentation::preComputeGraph] init failed");
      graph_.reset(new Graph);
      graph_builder_.setInputCloud(input_);
      graph_builder_.compute(*graph_);
      const int saved = pcl::graph::arithmeticMode();
      pcl::graph::arithmeticMode() = mode_;
      pcl::graph::computeNormalsAndCurvatures(*graph_);
      pcl::graph::computeSignedCurvatures(*graph_);
      pcl::graph::arithmeticMode() = saved;
      typedef pcl::graph::EdgeWeightComputer<Graph> EWC;
      EWC computer;
      computer.setArithmeticMode(mode_);
      computer.template addTerm<pcl::graph::terms::XYZ>(3.0f, EWC::NORMALIZATION_GLOBAL);
      computer.template addTerm<pcl::graph::terms::Normal>(0.01f, 0.0f);
      computer.template addTerm<pcl::graph::terms::Curvature>(0.0001f, 0.0f);
      computer.template addTerm<pcl::graph::terms::RGB>(3.0f, EWC::NORMALIZATION_GLOBAL);
      computer.setSmallWeightThreshold(1e-5f);
      computer.setSmallWeightPolicy(EWC::SMALL_WEIGHT_COERCE_TO_THRESHOLD);
      computer.compute(*graph_);
    }
    pcl::graph::createSubgraphsFromConnectedComponents(*graph_, graph_components_);
  }

  void segment(std::vector<pcl::PointIndices>& clusters) {
    if (!seeds_) throw pcl::ComputeFailedException("[pcl::segmentation::RandomWalkerSegmentation::segment] no seeds");
    preComputeGraph();
    const size_t V = graph_->num_vertices();
    const size_t K = label_color_.size();
    std::vector<uint32_t>& colors = graph_->vertex_colors();
    std::fill(colors.begin(), colors.end(), 0u);
    // seeds paint their nearest vertex, later seeds overwrite (impl:149-152)
    const size_t ns = seeds_->size();
    if (ns && V) {
      std::vector<float> sxyz(3 * ns);
      for (size_t i = 0; i < ns; ++i) { sxyz[3 * i] = seeds_->at(i).x; sxyz[3 * i + 1] = seeds_->at(i).y; sxyz[3 * i + 2] = seeds_->at(i).z; }
      std::vector<int32_t> nearest(ns);
      graph_->ctx().check(tcs_nearest_vertices(graph_->device(), sxyz.data(), ns, nearest.data()));
      // a seed with NaN / Inf coordinates has no nearest vertex (-1): it paints nothing, like k_paint_seeds in the C ABI
      for (size_t i = 0; i < ns; ++i)
        if (nearest[i] >= 0 && (size_t)nearest[i] < V) colors[nearest[i]] = label_color_[seeds_->at(i).label];
    }
    tcs_solver_opts opts = solver_opts_;
    bool ok = true;
    if (V && K) {
      if (store_potentials_) {
        std::map<uint32_t, size_t> col;
        ok = detail::randomWalkerImpl(*graph_, &potentials_, &col, &opts, &info_, (uint32_t)K);
      } else {
        ok = detail::randomWalkerImpl(*graph_, (Eigen::MatrixXf*)nullptr, (std::map<uint32_t, size_t>*)nullptr, &opts, &info_, (uint32_t)K);
      }
    }
    if (!ok) std::fprintf(stderr, "[pcl::segmentation::RandomWalkerSegmentation::segment] Random walker segmentation failed to converge\n");
    // cluster assembly (impl:198-223)
    clusters.clear();
    clusters.resize(K + 1);
    const std::vector<uint32_t>& out = graph_->vertex_colors();
    if (input_as_cloud_) {
      const std::vector<size_t>& p2v = graph_builder_.getPointToVertexMap();
      for (size_t i = 0; i < input_->size(); ++i) {
        const size_t v = p2v[i];
        if (v >= V || out[v] == 0) clusters.back().indices.push_back((int)i);
        else clusters[out[v] - 1].indices.push_back((int)i);
      }
    } else {
      for (size_t v = 0; v < V; ++v) {
        if (out[v] == 0) clusters.back().indices.push_back((int)v);
        else clusters[out[v] - 1].indices.push_back((int)v);
      }
    }
  }

  GraphConstPtr getGraph() const { return graph_; }
  const Eigen::MatrixXf& getPotentials() const {
    if (!store_potentials_)
      std::fprintf(stderr, "[pcl::segmentation::RandomWalkerSegmentation::getPotentials] potentials were disabled at construction time\n");
    return potentials_;
  }

  // ---- extensions (the reference fixes these at 0.006 / its own arithmetic) ----------------------------------
  void setVoxelResolution(float r) { graph_builder_.setVoxelResolution(r); graph_components_.clear(); }
  float getVoxelResolution() const { return graph_builder_.getVoxelResolution(); }
  void setArithmeticMode(int mode) { mode_ = mode; }
  void setSolverOptions(const tcs_solver_opts& o) { solver_opts_ = o; }
  const tcs_solve_info& getSolveInfo() const { return info_; }

 private:
  bool input_as_cloud_;
  GraphPtr graph_;
  std::vector<GraphRef> graph_components_;
  typename pcl::PointCloud<PointXYZL>::ConstPtr seeds_;
  pcl::graph::VoxelGridGraphBuilder<PointT, Graph> graph_builder_;
  std::map<uint32_t, uint32_t> label_color_;
  bool store_potentials_;
  Eigen::MatrixXf potentials_;
  int mode_;
  tcs_solver_opts solver_opts_;
  tcs_solve_info info_;
};

}  // namespace segmentation
}  // namespace pcl
