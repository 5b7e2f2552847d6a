// include/tcs/pcl_compat.h - the few PCL / Eigen names the reference's public API is written in.
//
// The reference's boundary is a C++ template API over PCL point clouds (include/random_walker_segmentation.h:87-291).
// When PCL is installed its own headers are used; in images without PCL (this one) the minimal layout-compatible
// definitions below stand in, so code written against the reference's signatures compiles unchanged:
//   pcl::PointXYZ / PointXYZRGB / PointXYZRGBA / PointXYZL / PointNormal / PointXYZRGBNormal (PCL 1.7 memory layout),
//   pcl::PointCloud<T> (+ ::Ptr / ::ConstPtr), pcl::PointIndices, pcl::PCLBase<T>, pcl::ComputeFailedException,
//   Eigen::MatrixXf (column-major float matrix) for getPotentials().
#pragma once

#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#if defined(__has_include)
#if __has_include(<pcl/point_types.h>) && __has_include(<pcl/point_cloud.h>)
#define TCS_HAVE_PCL 1
#endif
#endif

#ifdef TCS_HAVE_PCL
#include <pcl/PointIndices.h>
#include <pcl/exceptions.h>
#include <pcl/pcl_base.h>
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include <Eigen/Core>
namespace tcs {
template <typename T> using shared_ptr = boost::shared_ptr<T>;
}
#else

namespace tcs {
template <typename T> using shared_ptr = std::shared_ptr<T>;
}

namespace pcl {

// 16-byte aligned records exactly as PCL 1.7 lays them out (32 B, 32 B, 48 B)
struct alignas(16) PointXYZ {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
};
struct alignas(16) PointXYZRGBA {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
  union { uint32_t rgba; struct { uint8_t b, g, r, a; }; };
  float _pad2[3] = {0, 0, 0};
  PointXYZRGBA() : rgba(0xFF000000u) {}
};
struct alignas(16) PointXYZRGB {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
  union { uint32_t rgba; float rgb; struct { uint8_t b, g, r, a; }; };
  float _pad2[3] = {0, 0, 0};
  PointXYZRGB() : rgba(0xFF000000u) {}
};
struct alignas(16) PointXYZL {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
  uint32_t label = 0;
  float _pad2[3] = {0, 0, 0};
};
struct alignas(16) PointNormal {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
  union { float data_n[4]; float normal[3]; struct { float normal_x, normal_y, normal_z; }; };
  float curvature = 0;
  float _pad2[3] = {0, 0, 0};
  PointNormal() { data_n[0] = data_n[1] = data_n[2] = data_n[3] = 0; }
};
struct alignas(16) PointXYZRGBNormal {
  float x = 0, y = 0, z = 0, _pad = 1.0f;
  union { float data_n[4]; float normal[3]; struct { float normal_x, normal_y, normal_z; }; };
  union { uint32_t rgba; float rgb; struct { uint8_t b, g, r, a; }; };
  float curvature = 0;
  float _pad2[2] = {0, 0};
  PointXYZRGBNormal() : rgba(0xFF000000u) { data_n[0] = data_n[1] = data_n[2] = data_n[3] = 0; }
};
static_assert(sizeof(PointXYZ) == 16 && sizeof(PointXYZRGB) == 32 && sizeof(PointXYZRGBA) == 32 &&
                  sizeof(PointXYZL) == 32 && sizeof(PointNormal) == 48 && sizeof(PointXYZRGBNormal) == 48,
              "PCL point layouts");

template <typename PointT>
class PointCloud {
 public:
  typedef PointT PointType;
  typedef tcs::shared_ptr<PointCloud<PointT>> Ptr;
  typedef tcs::shared_ptr<const PointCloud<PointT>> ConstPtr;
  std::vector<PointT> points;
  uint32_t width = 0, height = 0;
  bool is_dense = true;
  size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void resize(size_t n) { points.resize(n); width = (uint32_t)n; height = 1; }
  void push_back(const PointT& p) { points.push_back(p); width = (uint32_t)points.size(); height = 1; }
  void clear() { points.clear(); width = height = 0; }
  PointT& operator[](size_t i) { return points[i]; }
  const PointT& operator[](size_t i) const { return points[i]; }
  PointT& at(size_t i) { return points.at(i); }
  const PointT& at(size_t i) const { return points.at(i); }
  typename std::vector<PointT>::iterator begin() { return points.begin(); }
  typename std::vector<PointT>::iterator end() { return points.end(); }
  typename std::vector<PointT>::const_iterator begin() const { return points.begin(); }
  typename std::vector<PointT>::const_iterator end() const { return points.end(); }
};

struct PointIndices {
  std::vector<int> indices;
  typedef tcs::shared_ptr<PointIndices> Ptr;
  typedef tcs::shared_ptr<const PointIndices> ConstPtr;
};

class PCLException : public std::runtime_error {
 public:
  explicit PCLException(const std::string& what) : std::runtime_error(what) {}
};
class ComputeFailedException : public PCLException {
 public:
  explicit ComputeFailedException(const std::string& what) : PCLException(what) {}
};

// pcl::PCLBase: input cloud + optional index subset (all points when none is set)
template <typename PointT>
class PCLBase {
 public:
  typedef typename PointCloud<PointT>::ConstPtr PointCloudConstPtr;
  virtual ~PCLBase() {}
  virtual void setInputCloud(const PointCloudConstPtr& cloud) { input_ = cloud; }
  PointCloudConstPtr getInputCloud() const { return input_; }

 protected:
  bool initCompute() { return input_ && !input_->empty(); }
  bool deinitCompute() { return true; }
  PointCloudConstPtr input_;
};

}  // namespace pcl

namespace Eigen {
// column-major float matrix, the subset of Eigen::MatrixXf that getPotentials() callers use
class MatrixXf {
 public:
  MatrixXf() {}
  MatrixXf(long r, long c) { resize(r, c); }
  void resize(long r, long c) { rows_ = r; cols_ = c; data_.assign((size_t)r * (size_t)c, 0.f); }
  long rows() const { return rows_; }
  long cols() const { return cols_; }
  float& operator()(long r, long c) { return data_[(size_t)c * rows_ + r]; }
  float operator()(long r, long c) const { return data_[(size_t)c * rows_ + r]; }
  float* data() { return data_.data(); }
  const float* data() const { return data_.data(); }

 private:
  long rows_ = 0, cols_ = 0;
  std::vector<float> data_;
};
}  // namespace Eigen

#endif  // TCS_HAVE_PCL

namespace tcs {
// trait: does the point type carry colour (EdgeWeightComputer::addTerm drops RGB terms otherwise)
template <typename PointT> struct has_color { static const bool value = false; };
template <> struct has_color<pcl::PointXYZRGB> { static const bool value = true; };
template <> struct has_color<pcl::PointXYZRGBA> { static const bool value = true; };
template <> struct has_color<pcl::PointXYZRGBNormal> { static const bool value = true; };
// point type augmented with normal + curvature (the reference's PointTypeAssociations map,
// include/random_walker_segmentation.h:101-111)
template <typename PointT> struct with_normal;
template <> struct with_normal<pcl::PointXYZ> { typedef pcl::PointNormal type; };
template <> struct with_normal<pcl::PointNormal> { typedef pcl::PointNormal type; };
template <> struct with_normal<pcl::PointXYZRGB> { typedef pcl::PointXYZRGBNormal type; };
template <> struct with_normal<pcl::PointXYZRGBA> { typedef pcl::PointXYZRGBNormal type; };
template <> struct with_normal<pcl::PointXYZRGBNormal> { typedef pcl::PointXYZRGBNormal type; };
}  // namespace tcs
