// oracle/orc_build.cpp - CPU ORACLE (test infrastructure, see orc.h): voxel-grid graph builder.
// Restates pcl::graph::VoxelGridGraphBuilder::compute (graph/impl/voxel_grid_graph_builder.hpp:69-143) and
// the PCL 1.7.2 octree / getMinMax3D / CentroidPoint semantics it relies on.
#include <algorithm>
#include <cfloat>
#include <unordered_map>

#include "orc_internal.hpp"

namespace {

// pcl::octree::OctreePointCloud bounding-box state (appendix A.2)
struct OctreeBox {
  double res;
  double mn[3], mx[3];
  unsigned depth = 0;
  size_t leaf_count = 0;

  // OctreePointCloud::getKeyBitSize()
  void get_key_bit_size() {
    const float min_value = std::numeric_limits<float>::epsilon();
    unsigned mk[3];
    for (int a = 0; a < 3; ++a) mk[a] = static_cast<unsigned>((mx[a] - mn[a]) / res);
    unsigned max_voxels = std::max(std::max(std::max(mk[0], mk[1]), mk[2]), 2u);
    // OctreeBase::Log2(n) = log(n) / log(2.0)
    double l2 = std::log((double)max_voxels) / std::log(2.0);
    depth = std::max(std::min(64u, static_cast<unsigned>(std::ceil(l2 - min_value))), 0u);
    double side = static_cast<double>(1ull << depth) * res - min_value;
    if (leaf_count == 0) {
      for (int a = 0; a < 3; ++a) {
        double over = (side - (mx[a] - mn[a])) / 2.0;
        mn[a] -= over;
        mx[a] += over;
      }
    } else {
      for (int a = 0; a < 3; ++a) mx[a] = mn[a] + side;
    }
  }
};

inline uint64_t morton3(uint32_t kx, uint32_t ky, uint32_t kz, int depth) {
  // child index at a level = x<<2 | y<<1 | z, most significant level first
  uint64_t m = 0;
  for (int b = depth - 1; b >= 0; --b)
    m = (m << 3) | (uint64_t)(((kx >> b) & 1u) << 2 | ((ky >> b) & 1u) << 1 | ((kz >> b) & 1u));
  return m;
}

}  // namespace

extern "C" orc_graph* orc_build_voxel_graph(const float* xyz, const uint32_t* rgba, size_t n, float resolution,
                                            int leaf_order_desc, int* status) {
  if (status) *status = ORC_OK;
  orc_graph* g = new orc_graph;
  g->N = n;
  g->has_color = rgba != nullptr;
  g->p2v.assign(n, 0xFFFFFFFFu);
  // initCompute() fails on an empty cloud -> empty graph (voxel_grid_graph_builder.hpp:72-77)
  if (n == 0 || !(resolution > 0.0f)) { if (status && n) *status = ORC_ERR_INPUT; return g; }

  // :79-87  x/=z, y/=z, z=log z on a copy
  std::vector<float> tx(n), ty(n), tz(n);
  for (size_t i = 0; i < n; ++i) {
    float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
    tx[i] = px / pz;
    ty[i] = py / pz;
    tz[i] = orc_logf(pz);
  }
  // :89-90 pcl::getMinMax3D over finite points (A.1)
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  size_t nfinite = 0;
  for (size_t i = 0; i < n; ++i) {
    if (!std::isfinite(tx[i]) || !std::isfinite(ty[i]) || !std::isfinite(tz[i])) continue;
    ++nfinite;
    mn[0] = std::min(mn[0], tx[i]); mn[1] = std::min(mn[1], ty[i]); mn[2] = std::min(mn[2], tz[i]);
    mx[0] = std::max(mx[0], tx[i]); mx[1] = std::max(mx[1], ty[i]); mx[2] = std::max(mx[2], tz[i]);
  }
  if (nfinite == 0) return g;

  // :93-97 octree(resolution), defineBoundingBox, addPointsFromInputCloud (A.2)
  OctreeBox box;
  box.res = (double)resolution;
  for (int a = 0; a < 3; ++a) {
    box.mn[a] = std::min((double)mn[a], (double)mx[a]);
    box.mx[a] = std::max((double)mn[a], (double)mx[a]);
  }
  box.get_key_bit_size();

  std::vector<uint32_t> kx(n), ky(n), kz(n);
  std::vector<size_t> inserted;  // finite point indices in insertion order
  inserted.reserve(nfinite);
  for (size_t i = 0; i < n; ++i) {
    if (!std::isfinite(tx[i]) || !std::isfinite(ty[i]) || !std::isfinite(tz[i])) continue;
    const double p[3] = {(double)tx[i], (double)ty[i], (double)tz[i]};
    // OctreePointCloud::adoptBoundingBoxToPoint: grow the tree while the point is outside [min, max)
    for (;;) {
      bool low[3], up[3], any = false;
      for (int a = 0; a < 3; ++a) {
        low[a] = p[a] < box.mn[a];
        up[a] = p[a] >= box.mx[a];
        any = any || low[a] || up[a];
      }
      if (!any) break;
      if (box.depth >= 31) { if (status) *status = ORC_ERR_DEPTH; return g; }
      // the old root becomes child ((!upX)<<2 | (!upY)<<1 | !upZ) of a new root: existing keys gain a top bit
      const uint32_t bit = 1u << box.depth;
      for (size_t j : inserted) {
        if (!up[0]) kx[j] |= bit;
        if (!up[1]) ky[j] |= bit;
        if (!up[2]) kz[j] |= bit;
      }
      double side = static_cast<double>(1ull << box.depth) * box.res;
      for (int a = 0; a < 3; ++a) if (!up[a]) box.mn[a] -= side;
      box.depth++;
      side = static_cast<double>(1ull << box.depth) * box.res - std::numeric_limits<float>::epsilon();
      for (int a = 0; a < 3; ++a) box.mx[a] = box.mn[a] + side;
      box.leaf_count = inserted.size();
      box.get_key_bit_size();
    }
    // OctreePointCloud::genOctreeKeyforPoint
    kx[i] = static_cast<unsigned>((p[0] - box.mn[0]) / box.res);
    ky[i] = static_cast<unsigned>((p[1] - box.mn[1]) / box.res);
    kz[i] = static_cast<unsigned>((p[2] - box.mn[2]) / box.res);
    inserted.push_back(i);
    box.leaf_count = inserted.size();
  }
  g->depth = (int)box.depth;
  for (int a = 0; a < 3; ++a) g->bbmin[a] = box.mn[a];
  if (box.depth > 21) { if (status) *status = ORC_ERR_DEPTH; return g; }

  // leaves in LeafNodeIterator order (A.2): depth first, children visited 7..0 (PCL <= 1.8) => descending
  // Morton code; point indices inside a leaf ascend (push_back in insertion order)
  std::vector<std::pair<uint64_t, uint32_t>> order(inserted.size());
  for (size_t j = 0; j < inserted.size(); ++j) {
    size_t i = inserted[j];
    order[j] = std::make_pair(morton3(kx[i], ky[i], kz[i], (int)box.depth), (uint32_t)i);
  }
  if (leaf_order_desc)
    std::sort(order.begin(), order.end(), [](const std::pair<uint64_t, uint32_t>& a, const std::pair<uint64_t, uint32_t>& b) {
      return a.first != b.first ? a.first > b.first : a.second < b.second; });
  else
    std::sort(order.begin(), order.end());

  size_t V = 0;
  for (size_t j = 0; j < order.size(); ++j) if (j == 0 || order[j].first != order[j - 1].first) ++V;
  g->alloc_vertices(V);  // :99  graph = GraphT(leafCount); PointXYZRGBNormal() = zeros
  g->keys.assign(3 * V, 0);

  struct KeyHash { size_t operator()(uint64_t k) const { return std::hash<uint64_t>()(k); } };
  std::unordered_map<uint64_t, uint32_t, KeyHash> key_to_vertex;
  key_to_vertex.reserve(V * 2);
  auto pack = [](uint32_t a, uint32_t b, uint32_t c) { return ((uint64_t)a << 42) | ((uint64_t)b << 21) | (uint64_t)c; };

  size_t j = 0;
  for (uint32_t v = 0; j < order.size(); ++v) {
    // Step 1 (:110-119): pcl::CentroidPoint over the ORIGINAL points of the leaf (A.3)
    size_t j0 = j;
    float sx = 0.f, sy = 0.f, sz = 0.f, sr = 0.f, sg = 0.f, sb = 0.f, sa = 0.f;
    while (j < order.size() && order[j].first == order[j0].first) {
      uint32_t i = order[j].second;
      sx += xyz[3 * i]; sy += xyz[3 * i + 1]; sz += xyz[3 * i + 2];
      if (rgba) {
        uint32_t c = rgba[i];
        sr += static_cast<float>((c >> 16) & 0xFF);
        sg += static_cast<float>((c >> 8) & 0xFF);
        sb += static_cast<float>(c & 0xFF);
        sa += static_cast<float>((c >> 24) & 0xFF);
      }
      g->p2v[i] = v;
      ++j;
    }
    float cnt = static_cast<float>(j - j0);  // Eigen: vector / size_t -> divide by (float)n
    g->x[v] = sx / cnt; g->y[v] = sy / cnt; g->z[v] = sz / cnt;
    if (rgba)
      g->rgba[v] = static_cast<uint32_t>(sa / cnt) << 24 | static_cast<uint32_t>(sr / cnt) << 16 |
                   static_cast<uint32_t>(sg / cnt) << 8 | static_cast<uint32_t>(sb / cnt);
    // Step 2 (:121-123): key -> vertex
    uint32_t i0 = order[j0].second;
    uint32_t key[3] = {kx[i0], ky[i0], kz[i0]};
    g->keys[3 * v] = key[0]; g->keys[3 * v + 1] = key[1]; g->keys[3 * v + 2] = key[2];
    key_to_vertex[pack(key[0], key[1], key[2])] = v;
    // Step 3 (:125-141): edges to already visited neighbours, (dx,dy,dz) lexicographic
    for (int dx = (key[0] > 0) ? -1 : 0; dx <= 1; ++dx)
      for (int dy = (key[1] > 0) ? -1 : 0; dy <= 1; ++dy)
        for (int dz = (key[2] > 0) ? -1 : 0; dz <= 1; ++dz) {
          auto f = key_to_vertex.find(pack(key[0] + dx, key[1] + dy, key[2] + dz));
          if (f != key_to_vertex.end() && f->second != v) {
            g->esrc.push_back(v);
            g->etgt.push_back(f->second);
          }
        }
  }
  g->ew.assign(g->esrc.size(), 0.f);
  g->build_out_edges();
  return g;
}

extern "C" orc_graph* orc_graph_from_edges(size_t V, const float* xyz, const uint32_t* rgba, const float* normals4,
                                           const float* curvature, size_t E, const uint32_t* src,
                                           const uint32_t* tgt, const float* weight) {
  orc_graph* g = new orc_graph;
  g->alloc_vertices(V);
  g->has_color = rgba != nullptr;
  for (size_t v = 0; v < V; ++v) {
    if (xyz) { g->x[v] = xyz[3 * v]; g->y[v] = xyz[3 * v + 1]; g->z[v] = xyz[3 * v + 2]; }
    if (rgba) g->rgba[v] = rgba[v];
    if (normals4) { g->nx[v] = normals4[4 * v]; g->ny[v] = normals4[4 * v + 1]; g->nz[v] = normals4[4 * v + 2]; g->nw[v] = normals4[4 * v + 3]; }
    if (curvature) g->curv[v] = curvature[v];
  }
  g->esrc.assign(src, src + E);
  g->etgt.assign(tgt, tgt + E);
  if (weight) g->ew.assign(weight, weight + E); else g->ew.assign(E, 0.f);
  g->build_out_edges();
  return g;
}

// ---- NearestNeighborsGraphBuilder::compute, graph/impl/nearest_neighbors_graph_builder.hpp:49-119 ------------------
// Vertices are the finite points in index order (:58-72; copyPointCloud of the finite indices).  The copied cloud has
// height 1, so the default search is always pcl::search::KdTree (:75-83), an exact FLANN search: per vertex the
// num_neighbors + 1 nearest points (the first being the query itself, :95-97), or for RADIUS the nearest
// num_neighbors + 1 among the points with d^2 < (float)(radius * radius) (:100-103; FLANN's RadiusResultSet keeps
// dist < radius, and max_nn bounds the sorted result).  Distances are FLANN's L2_Simple<float>: ((dx*dx) + dy*dy) + dz*dz
// in float32.  Restated here as a brute-force search; neighbours are ordered by (d^2, index) and the query vertex itself
// is skipped (FLANN's order among equal distances is unspecified - with duplicated points the reference may return a
// duplicate instead of the query as "first neighbour"; this restatement fixes the deterministic choice).  Edges are
// added in the reference's loop order with its `if (!boost::edge(i, nb).second)` test (:105-107).
extern "C" orc_graph* orc_build_nn_graph(const float* xyz, const uint32_t* rgba, size_t n, int num_neighbors,
                                         double radius, int use_radius, int* status) {
  if (status) *status = ORC_OK;
  orc_graph* g = new orc_graph;
  g->N = n;
  g->has_color = rgba != nullptr;
  g->p2v.assign(n, 0xFFFFFFFFu);
  std::vector<uint32_t> finite;
  for (size_t i = 0; i < n; ++i)
    if (std::isfinite(xyz[3 * i]) && std::isfinite(xyz[3 * i + 1]) && std::isfinite(xyz[3 * i + 2])) finite.push_back((uint32_t)i);
  const size_t V = finite.size();
  g->alloc_vertices(V);
  for (size_t v = 0; v < V; ++v) {
    const size_t i = finite[v];
    g->x[v] = xyz[3 * i]; g->y[v] = xyz[3 * i + 1]; g->z[v] = xyz[3 * i + 2];
    if (rgba) g->rgba[v] = rgba[i];
    g->p2v[i] = (uint32_t)v;   // :112-116
  }
  const size_t k = num_neighbors < 0 ? 0 : (size_t)num_neighbors;
  const float r2 = (float)(radius * radius);
  std::vector<std::vector<uint32_t>> nbr(V);
  std::vector<std::pair<float, uint32_t>> cand;
  for (size_t i = 0; i < V; ++i) {
    cand.clear();
    for (size_t j = 0; j < V; ++j) {
      if (j == i) continue;
      const float dx = g->x[i] - g->x[j], dy = g->y[i] - g->y[j], dz = g->z[i] - g->z[j];
      const float d = dx * dx + dy * dy + dz * dz;
      if (use_radius && !(d < r2)) continue;
      cand.emplace_back(d, (uint32_t)j);
    }
    const size_t m = std::min(k, cand.size());
    std::partial_sort(cand.begin(), cand.begin() + m, cand.end());
    for (size_t j = 0; j < m; ++j) nbr[i].push_back(cand[j].second);
  }
  // boost::edge(i, nb, graph).second on an undirected graph: the edge exists if either endpoint added it before
  std::vector<std::vector<uint32_t>> adj(V);
  for (size_t i = 0; i < V; ++i)
    for (uint32_t nb : nbr[i]) {
      if (std::find(adj[i].begin(), adj[i].end(), nb) != adj[i].end()) continue;
      adj[i].push_back(nb);
      adj[nb].push_back((uint32_t)i);
      g->esrc.push_back((uint32_t)i);
      g->etgt.push_back(nb);
    }
  g->ew.assign(g->esrc.size(), 0.f);
  g->build_out_edges();
  return g;
}
