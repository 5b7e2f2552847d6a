// oracle/orc_internal.hpp - internal data structures of the CPU oracle (TEST INFRASTRUCTURE, see orc.h).
#ifndef TCS_ORACLE_INTERNAL_HPP
#define TCS_ORACLE_INTERNAL_HPP

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "orc.h"

// Transcendentals: the reference calls the float overloads of libm (std::log / std::exp / std::atan2 /
// std::cos / std::sin on float arguments).  Their last bit depends on the libm version; oracle and GPU
// both DEFINE them as the float64 function rounded once to float32, which is
// the correctly rounded float function in all but ~2^-29 of the cases.
static inline float orc_logf(float v) { return (float)std::log((double)v); }
static inline float orc_expf(float v) { return (float)std::exp((double)v); }
static inline float orc_atan2f(float a, float b) { return (float)std::atan2((double)a, (double)b); }
static inline float orc_cosf(float v) { return (float)std::cos((double)v); }
static inline float orc_sinf(float v) { return (float)std::sin((double)v); }

// Eigen 3.2 reduction orders (appendix A.8): a non-vectorised 3-vector reduces as e0 + (e1 + e2);
// a Vector4f (SSE3 haddps) reduces as (e0 + e1) + (e2 + e3).
static inline float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}
static inline float sqn3(float a0, float a1, float a2) { return a0 * a0 + (a1 * a1 + a2 * a2); }

struct orc_graph {
  size_t V = 0, E = 0, N = 0;
  bool has_color = false;
  // vertex bundle (PointXYZRGBNormal / PointNormal fields)
  std::vector<float> x, y, z, nx, ny, nz, nw, curv;
  std::vector<uint32_t> rgba;
  // boost::adjacency_list<vecS, vecS, undirectedS, ..., listS>: global edge list in insertion order
  std::vector<uint32_t> esrc, etgt;
  std::vector<float> ew;
  // out-edge vectors (insertion order) flattened: rowptr[V+1], col = target, eid = edge id
  std::vector<uint32_t> rowptr, col, eid;
  // voxel builder outputs
  std::vector<uint32_t> keys;  // 3 per vertex
  std::vector<uint32_t> p2v;   // per input point, nil = 0xFFFFFFFF
  double bbmin[3] = {0, 0, 0};
  int depth = 0;

  void alloc_vertices(size_t n) {
    V = n;
    x.assign(n, 0.f); y.assign(n, 0.f); z.assign(n, 0.f);
    nx.assign(n, 0.f); ny.assign(n, 0.f); nz.assign(n, 0.f); nw.assign(n, 0.f);
    curv.assign(n, 0.f);
    rgba.assign(n, 0u);
  }
  // rebuild rowptr/col/eid from the edge list: add_edge(u, v) appends to out-edges of u and of v
  void build_out_edges() {
    E = esrc.size();
    rowptr.assign(V + 1, 0);
    for (size_t e = 0; e < E; ++e) { rowptr[esrc[e] + 1]++; rowptr[etgt[e] + 1]++; }
    for (size_t v = 0; v < V; ++v) rowptr[v + 1] += rowptr[v];
    col.assign(2 * E, 0); eid.assign(2 * E, 0);
    std::vector<uint32_t> fill(rowptr.begin(), rowptr.end() - 1);
    for (size_t e = 0; e < E; ++e) {
      uint32_t s = esrc[e], t = etgt[e];
      col[fill[s]] = t; eid[fill[s]++] = (uint32_t)e;
      col[fill[t]] = s; eid[fill[t]++] = (uint32_t)e;
    }
  }
};

// orc_solve.cpp
int orc_components_impl(const orc_graph* g, std::vector<int32_t>& comp);

#endif
