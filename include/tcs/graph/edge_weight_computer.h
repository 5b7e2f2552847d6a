// include/tcs/graph/edge_weight_computer.h - pcl::graph::EdgeWeightComputer on the GPU.
// Same interface as the reference (include/graph/edge_weight_computer.h:259-328): addTerm<terms::X>(influence
// [, convex_influence_multiplier][, normalization]), setSmallWeightPolicy, setSmallWeightThreshold, compute(graph).
// The reference's user-supplied balancing std::function is not supported on the device: the Gaussian
// exp(-d / influence) of edge_weight_computer.h:333-337 is what compute() evaluates (setTermBalancingFunction throws).
#pragma once

#include "point_cloud_graph.h"

namespace pcl {
namespace graph {

namespace terms {
// tag types of include/graph/edge_weight_computer_terms.h:55-140
struct XYZ { static const int kind = TCS_TERM_XYZ; };
struct Normal { static const int kind = TCS_TERM_NORMAL; };
struct Curvature { static const int kind = TCS_TERM_CURVATURE; };
struct RGB { static const int kind = TCS_TERM_RGB; };
}  // namespace terms

template <typename GraphT>
class EdgeWeightComputer {
 public:
  typedef tcs::shared_ptr<EdgeWeightComputer> Ptr;
  enum NormalizationType { NORMALIZATION_NONE = TCS_NORM_NONE, NORMALIZATION_GLOBAL = TCS_NORM_GLOBAL, NORMALIZATION_LOCAL = TCS_NORM_LOCAL };
  enum SmallWeightPolicy { SMALL_WEIGHT_IGNORE = TCS_SMALL_IGNORE, SMALL_WEIGHT_COERCE_TO_THRESHOLD = TCS_SMALL_COERCE, SMALL_WEIGHT_REMOVE_EDGE = TCS_SMALL_REMOVE };

  EdgeWeightComputer() : policy_(SMALL_WEIGHT_IGNORE), threshold_(0.0f), mode_(TCS_MODE_PARITY) {}

  template <typename TermT>
  void addTerm(float influence, float convex_influence_multiplier, NormalizationType normalization = NORMALIZATION_NONE) {
    tcs_term t = {TermT::kind, influence, convex_influence_multiplier, (int)normalization};
    terms_.push_back(t);
  }
  template <typename TermT>
  void addTerm(float influence, NormalizationType normalization = NORMALIZATION_NONE) {
    addTerm<TermT>(influence, 1.0f, normalization);
  }
  void setSmallWeightPolicy(SmallWeightPolicy policy) { policy_ = policy; }
  void setSmallWeightThreshold(float threshold) { threshold_ = threshold; }
  template <typename F> void setTermBalancingFunction(F) {
    throw pcl::PCLException("tcs_b200: custom term balancing functions cannot run on the device");
  }
  // extension: TCS_MODE_PARITY (reference summation order, default) or TCS_MODE_FAST
  void setArithmeticMode(int mode) { mode_ = mode; }

  void compute(GraphT& graph) {
    tcs_graph* h = graph.device();
    graph.ctx().check(tcs_edge_weights(h, terms_.data(), (int)terms_.size(), threshold_, (int)policy_, mode_));
    if (policy_ == SMALL_WEIGHT_REMOVE_EDGE) graph.invalidate_topology();
    graph.invalidate_weights();
  }

 private:
  std::vector<tcs_term> terms_;
  SmallWeightPolicy policy_;
  float threshold_;
  int mode_;
};

}  // namespace graph
}  // namespace pcl
