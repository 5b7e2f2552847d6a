// include/tcs/io.h - saveGraph / loadGraph: the reference's graph interchange format (include/impl/io.hpp:13-101).
//
// A graph file is the PCD file of the graph's point cloud (ASCII, what pcl::io::savePCDFile writes by default) with a
// text tail appended: the line "# Edges" followed by one "source target weight" line per edge, in edge order
// (:30-40; weights go through operator<< of a std::ofstream, i.e. 6 significant digits).  loadGraph reads the cloud,
// makes one vertex per point, then adds the edges of the tail in file order, skipping empty lines and lines that
// start with '#' (:72-91).  The PCD part is written with float fields at 9 significant digits (exact round trip) and
// the colour as an unsigned `rgba` field, which PCL's reader matches to the rgb of PointXYZRGBNormal.
#pragma once

#include <fstream>
#include <sstream>
#include <string>

#include "graph/point_cloud_graph.h"
#include "pcd_io.h"

template <typename PointT>
bool saveGraph(const std::string& filename, const pcl::graph::point_cloud_graph<PointT>& graph) {
  typedef pcl::graph::point_cloud_graph<PointT> Graph;
  typename Graph::point_cloud_const_ptr cloud = graph.point_cloud();
  const size_t V = cloud->size();
  tcs::pcd::Table t;
  t.n = V;
  t.width = (uint32_t)V;
  t.height = 1;
  const PointT* p0 = V ? &cloud->points[0] : nullptr;
  std::vector<uint32_t> rgba(V);
  for (size_t v = 0; v < V; ++v) rgba[v] = tcs::detail::get_rgba(cloud->points[v], typename Graph::colored());
  try {
    tcs::pcd::detail::add_field(t, "x", 'F', p0 ? &p0->x : nullptr, sizeof(PointT), V);
    tcs::pcd::detail::add_field(t, "y", 'F', p0 ? &p0->y : nullptr, sizeof(PointT), V);
    tcs::pcd::detail::add_field(t, "z", 'F', p0 ? &p0->z : nullptr, sizeof(PointT), V);
    if (Graph::colored::value) tcs::pcd::detail::add_field(t, "rgba", 'U', rgba.data(), 4, V);
    tcs::pcd::detail::add_field(t, "normal_x", 'F', p0 ? &p0->data_n[0] : nullptr, sizeof(PointT), V);
    tcs::pcd::detail::add_field(t, "normal_y", 'F', p0 ? &p0->data_n[1] : nullptr, sizeof(PointT), V);
    tcs::pcd::detail::add_field(t, "normal_z", 'F', p0 ? &p0->data_n[2] : nullptr, sizeof(PointT), V);
    tcs::pcd::detail::add_field(t, "curvature", 'F', p0 ? &p0->curvature : nullptr, sizeof(PointT), V);
    tcs::pcd::write(filename, t, "ascii");
  } catch (const std::exception&) {
    return false;
  }
  std::ofstream file(filename, std::fstream::app);
  if (!file.is_open()) return false;
  file << "# Edges\n";
  const std::vector<uint32_t>& src = graph.edge_sources();
  const std::vector<uint32_t>& tgt = graph.edge_targets();
  const std::vector<float>& w = graph.edge_weights();
  for (size_t e = 0; e < src.size(); ++e) file << src[e] << " " << tgt[e] << " " << w[e] << "\n";
  file.close();
  return true;
}

template <typename PointT>
bool loadGraph(const std::string& filename, pcl::graph::point_cloud_graph<PointT>& graph) {
  typedef pcl::graph::point_cloud_graph<PointT> Graph;
  typename Graph::point_cloud_ptr cloud(new typename Graph::point_cloud_type);
  try {
    const tcs::pcd::Table t = tcs::pcd::read(filename);
    const int fx = t.find("x"), fy = t.find("y"), fz = t.find("z");
    if (fx < 0 || fy < 0 || fz < 0) return false;
    int fc = t.find("rgba");
    if (fc < 0) fc = t.find("rgb");
    const int fn[3] = {t.find("normal_x"), t.find("normal_y"), t.find("normal_z")};
    const int fk = t.find("curvature");
    cloud->points.resize(t.n);
    cloud->width = t.width ? t.width : (uint32_t)t.n;
    cloud->height = t.width ? t.height : 1;
    for (size_t i = 0; i < t.n; ++i) {
      PointT& p = cloud->points[i];
      p.x = tcs::pcd::detail::get_f32(t, fx, i);
      p.y = tcs::pcd::detail::get_f32(t, fy, i);
      p.z = tcs::pcd::detail::get_f32(t, fz, i);
      for (int k = 0; k < 3; ++k)
        if (fn[k] >= 0) p.data_n[k] = tcs::pcd::detail::get_f32(t, fn[k], i);
      if (fk >= 0) p.curvature = tcs::pcd::detail::get_f32(t, fk, i);
      if (fc >= 0) tcs::detail::set_rgba(p, tcs::pcd::detail::get_bits32(t, fc, i), typename Graph::colored());
    }
  } catch (const std::exception&) {
    return false;
  }
  std::ifstream stream(filename);
  if (!stream) return false;
  graph = Graph(cloud);
  std::string line;
  bool edges_section = false;
  while (std::getline(stream, line)) {
    if (!edges_section) {
      if (line == "# Edges") edges_section = true;
    } else {
      if (!line.size() || line[0] == '#') continue;
      int src = 0, tgt = 0;
      float wgh = 0.f;
      std::stringstream(line) >> src >> tgt >> wgh;
      graph.add_edge((uint32_t)src, (uint32_t)tgt, wgh);
    }
  }
  return true;
}
