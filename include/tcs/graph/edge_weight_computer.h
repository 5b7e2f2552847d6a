// include/tcs/graph/edge_weight_computer.h - pcl::graph::EdgeWeightComputer on the GPU.
// Same interface as the reference (include/graph/edge_weight_computer.h:259-328): addTerm<terms::X>(influence
// [, convex_influence_multiplier][, normalization]), setSmallWeightPolicy, setSmallWeightThreshold, compute(graph).
// setTermBalancingFunction selects one of a closed set of device functions (the reference's Gaussian
// exp(-d / influence), edge_weight_computer.h:333-337, is the default); an arbitrary host callable throws.
#pragma once

#include <cmath>

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

namespace balancing {
// host callables with the reference's TermBalancingFunction signature that the device also implements
struct Gaussian { static const int kind = TCS_BALANCE_GAUSSIAN; float operator()(float v, float s) const { return s > 0.f ? std::exp(-v / s) : 1.f; } };
struct Inverse { static const int kind = TCS_BALANCE_INVERSE; float operator()(float v, float s) const { return s > 0.f ? 1.f / (1.f + v / s) : 1.f; } };
struct GaussianSquared { static const int kind = TCS_BALANCE_GAUSSIAN_SQUARED; float operator()(float v, float s) const { return s > 0.f ? std::exp(-(v / s) * (v / s)) : 1.f; } };
struct Linear { static const int kind = TCS_BALANCE_LINEAR; float operator()(float v, float s) const { return s > 0.f ? std::max(0.f, 1.f - v / s) : 1.f; } };
}  // namespace balancing

template <typename GraphT>
class EdgeWeightComputer {
 public:
  typedef tcs::shared_ptr<EdgeWeightComputer> Ptr;
  enum NormalizationType { NORMALIZATION_NONE = TCS_NORM_NONE, NORMALIZATION_GLOBAL = TCS_NORM_GLOBAL, NORMALIZATION_LOCAL = TCS_NORM_LOCAL };
  enum SmallWeightPolicy { SMALL_WEIGHT_IGNORE = TCS_SMALL_IGNORE, SMALL_WEIGHT_COERCE_TO_THRESHOLD = TCS_SMALL_COERCE, SMALL_WEIGHT_REMOVE_EDGE = TCS_SMALL_REMOVE };

  EdgeWeightComputer() : policy_(SMALL_WEIGHT_IGNORE), threshold_(0.0f), mode_(TCS_MODE_PARITY), balancing_(TCS_BALANCE_GAUSSIAN) {}

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
  // setTermBalancingFunction (edge_weight_computer.h:207,325).  The reference takes any
  // boost::function<float (float value, float influence)>; a host callable cannot run on the device, so the choice
  // is a closed set: by enum, or by one of the functor types of pcl::graph::balancing (which are also plain host
  // callables with the reference's signature).  Any other callable throws.
  enum TermBalancing {
    BALANCING_GAUSSIAN = TCS_BALANCE_GAUSSIAN,                  // exp (-value / influence): the reference's default
    BALANCING_INVERSE = TCS_BALANCE_INVERSE,                    // 1 / (1 + value / influence)
    BALANCING_GAUSSIAN_SQUARED = TCS_BALANCE_GAUSSIAN_SQUARED,  // exp (-(value / influence)^2)
    BALANCING_LINEAR = TCS_BALANCE_LINEAR                       // max (0, 1 - value / influence)
  };
  void setTermBalancingFunction(TermBalancing f) { balancing_ = (int)f; }
  template <typename F> void setTermBalancingFunction(F f) { setBalancingImpl(f, 0); }
  // extension: TCS_MODE_PARITY (reference summation order, default) or TCS_MODE_FAST
  void setArithmeticMode(int mode) { mode_ = mode; }

  void compute(GraphT& graph) {
    tcs_graph* h = graph.device();
    graph.ctx().check(tcs_edge_weights_ex(h, terms_.data(), (int)terms_.size(), threshold_, (int)policy_, mode_, balancing_));
    if (policy_ == SMALL_WEIGHT_REMOVE_EDGE) graph.invalidate_topology();
    graph.invalidate_weights();
  }

 private:
  template <typename F> auto setBalancingImpl(F, int) -> decltype((void)F::kind, void()) { balancing_ = F::kind; }
  template <typename F> void setBalancingImpl(F, long) {
    throw pcl::PCLException("tcs_b200: this term balancing function cannot run on the device; use one of pcl::graph::balancing::*");
  }
  std::vector<tcs_term> terms_;
  SmallWeightPolicy policy_;
  float threshold_;
  int mode_;
  int balancing_;
};

}  // namespace graph
}  // namespace pcl
