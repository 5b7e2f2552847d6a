// oracle/orc_api.cpp - CPU ORACLE (test infrastructure, see orc.h): accessors, the class-level pipeline
// (impl/random_walker_segmentation.hpp:82-224, src/random_walker_segmentation_app.cpp:122-134) and the
// evaluation score (src/segmentation_evaluation.cpp:139-165).
#include <algorithm>
#include <chrono>
#include <climits>
#include <map>
#include <set>

#include "orc_internal.hpp"

extern "C" {

void orc_graph_free(orc_graph* g) { delete g; }
size_t orc_num_vertices(const orc_graph* g) { return g->V; }
size_t orc_num_edges(const orc_graph* g) { return g->E; }
size_t orc_num_points(const orc_graph* g) { return g->N; }
int orc_depth(const orc_graph* g) { return g->depth; }
void orc_get_bbox_min(const orc_graph* g, double* m) { for (int a = 0; a < 3; ++a) m[a] = g->bbmin[a]; }

void orc_get_vertices(const orc_graph* g, float* xyz, uint32_t* rgba, float* normals4, float* curvature) {
  for (size_t v = 0; v < g->V; ++v) {
    if (xyz) { xyz[3 * v] = g->x[v]; xyz[3 * v + 1] = g->y[v]; xyz[3 * v + 2] = g->z[v]; }
    if (rgba) rgba[v] = g->rgba[v];
    if (normals4) { normals4[4 * v] = g->nx[v]; normals4[4 * v + 1] = g->ny[v]; normals4[4 * v + 2] = g->nz[v]; normals4[4 * v + 3] = g->nw[v]; }
    if (curvature) curvature[v] = g->curv[v];
  }
}
void orc_get_keys(const orc_graph* g, uint32_t* k) { if (!g->keys.empty()) std::memcpy(k, g->keys.data(), g->keys.size() * 4); }
void orc_get_edges(const orc_graph* g, uint32_t* s, uint32_t* t, float* w) {
  if (s) std::copy(g->esrc.begin(), g->esrc.end(), s);
  if (t) std::copy(g->etgt.begin(), g->etgt.end(), t);
  if (w) std::copy(g->ew.begin(), g->ew.end(), w);
}
void orc_get_csr(const orc_graph* g, uint32_t* rowptr, uint32_t* col, uint32_t* eid) {
  if (rowptr) std::copy(g->rowptr.begin(), g->rowptr.end(), rowptr);
  if (col) std::copy(g->col.begin(), g->col.end(), col);
  if (eid) std::copy(g->eid.begin(), g->eid.end(), eid);
}
void orc_get_point_to_vertex(const orc_graph* g, uint32_t* map) { std::copy(g->p2v.begin(), g->p2v.end(), map); }
void orc_set_xyz(orc_graph* g, const float* xyz) {
  for (size_t v = 0; v < g->V; ++v) { g->x[v] = xyz[3 * v]; g->y[v] = xyz[3 * v + 1]; g->z[v] = xyz[3 * v + 2]; }
}
void orc_set_normals(orc_graph* g, const float* n4, const float* curvature) {
  for (size_t v = 0; v < g->V; ++v) {
    if (n4) { g->nx[v] = n4[4 * v]; g->ny[v] = n4[4 * v + 1]; g->nz[v] = n4[4 * v + 2]; g->nw[v] = n4[4 * v + 3]; }
    if (curvature) g->curv[v] = curvature[v];
  }
}
void orc_set_weights(orc_graph* g, const float* w) { std::copy(w, w + g->E, g->ew.begin()); }

orc_graph* orc_segment_cloud(const float* xyz, const uint32_t* rgba, size_t n, float resolution, int preset,
                             int mode, const float* seed_xyz, const uint32_t* seed_label, size_t ns, int solver,
                             int32_t* point_cluster, uint32_t* vertex_colour_out, int* ncolours_out,
                             double* stage_ms, int* status) {
  typedef std::chrono::steady_clock clk;
  auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
  double t[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  int st = ORC_OK;
  // preComputeGraph (:103-136) / app pipeline (app.cpp:129-134); stage split = MEASURE_RUNTIME sites
  auto t0 = clk::now();
  orc_graph* g = orc_build_voxel_graph(xyz, rgba, n, resolution, 1, &st);
  auto t1 = clk::now(); t[0] = ms(t0, t1);
  if (st == ORC_OK && g->V > 0) {
    orc_normals(g, 0, mode);
    auto t2 = clk::now(); t[1] = ms(t1, t2);
    if (preset == 1) {
      orc_smoothen(g, 0.012f, 0.0012f);
      auto t3 = clk::now(); t[2] = ms(t2, t3);
      orc_normals(g, 0, mode);
      t2 = clk::now(); t[3] = ms(t3, t2);
    }
    orc_signed_curvatures(g);
    auto t4 = clk::now(); t[4] = ms(t2, t4);
    orc_term terms[4];
    if (preset == 0) {  // impl/random_walker_segmentation.hpp:124-129
      terms[0] = {ORC_TERM_XYZ, 3.0f, 1.0f, ORC_NORM_GLOBAL};
      terms[1] = {ORC_TERM_NORMAL, 0.01f, 0.0f, ORC_NORM_NONE};
      terms[2] = {ORC_TERM_CURVATURE, 0.0001f, 0.0f, ORC_NORM_NONE};
      terms[3] = {ORC_TERM_RGB, 3.0f, 1.0f, ORC_NORM_GLOBAL};
    } else {            // factory/edge_weight_computer_factory.h:20-25,41-46
      terms[0] = {ORC_TERM_XYZ, 3.0f, 1.0f, ORC_NORM_GLOBAL};
      terms[1] = {ORC_TERM_NORMAL, 0.01f, 0.01f, ORC_NORM_NONE};
      terms[2] = {ORC_TERM_CURVATURE, 0.0001f, 0.01f, ORC_NORM_NONE};
      terms[3] = {ORC_TERM_RGB, 3.0f, 1.0f, ORC_NORM_LOCAL};
    }
    orc_edge_weights(g, terms, 4, 1e-5f, ORC_SMALL_COERCE, mode);
    auto t5 = clk::now(); t[5] = ms(t4, t5);
    std::vector<int32_t> comp;
    orc_components_impl(g, comp);
    auto t6 = clk::now(); t[6] = ms(t5, t6);

    // setSeeds (:82-101): unique labels ascending -> colours 1..K ; segment (:144-152): paint nearest vertices
    std::set<uint32_t> labels(seed_label, seed_label + ns);
    std::map<uint32_t, uint32_t> label_colour;
    for (uint32_t l : labels) { uint32_t c = (uint32_t)label_colour.size() + 1; label_colour[l] = c; }
    const int K = (int)labels.size();
    if (ncolours_out) *ncolours_out = K;
    std::vector<uint32_t> colours(g->V, 0u);
    std::vector<int32_t> nearest(ns);
    orc_nearest_vertices(g, seed_xyz, ns, nearest.data());
    for (size_t i = 0; i < ns; ++i) colours[nearest[i]] = label_colour[seed_label[i]];
    auto t7 = clk::now(); t[7] = ms(t6, t7);
    int rs = orc_random_walker(g, colours.data(), K, solver, nullptr, nullptr);
    if (rs != ORC_OK) st = rs;
    auto t8 = clk::now(); t[8] = ms(t7, t8);
    // cluster assembly (:198-223): clusters[colour-1], nil / unlabelled -> last cluster (index K)
    if (point_cluster)
      for (size_t i = 0; i < n; ++i) {
        uint32_t v = g->p2v[i];
        point_cluster[i] = (v >= g->V || colours[v] == 0) ? K : (int32_t)colours[v] - 1;
      }
    if (vertex_colour_out) std::copy(colours.begin(), colours.end(), vertex_colour_out);
    t[9] = ms(t8, clk::now());
  } else if (point_cluster) {
    std::set<uint32_t> labels(seed_label, seed_label + ns);
    if (ncolours_out) *ncolours_out = (int)labels.size();
    for (size_t i = 0; i < n; ++i) point_cluster[i] = (int32_t)labels.size();
  }
  if (stage_ms) std::copy(t, t + 10, stage_ms);
  if (status) *status = st;
  return g;
}

// #erroneous points: overlap table -> maximum-weight assignment -> points outside assigned pairs.
// The reference delegates the assignment to the vendored Kuhn-Munkres solver (src/hungarian/hungarian.cpp:8-54);
// any optimal assignment gives the same matched weight, so a small exact O(n^3) solver is used here.
long orc_segmentation_error(const uint32_t* gt, const uint32_t* seg, size_t n) {
  std::map<uint32_t, int> gi, si;
  for (size_t i = 0; i < n; ++i) {
    if (!gi.count(gt[i])) { int k = (int)gi.size(); gi[gt[i]] = k; }
    if (!si.count(seg[i])) { int k = (int)si.size(); si[seg[i]] = k; }
  }
  int R = (int)gi.size(), C = (int)si.size(), m = std::max(R, C);
  std::vector<std::vector<long>> w(m, std::vector<long>(m, 0));
  for (size_t i = 0; i < n; ++i) w[gi[gt[i]]][si[seg[i]]]++;
  long big = 0;
  for (auto& r : w) for (long x : r) big = std::max(big, x);
  // Hungarian (potentials form) on cost = big - w, square m x m
  std::vector<long> u(m + 1, 0), v(m + 1, 0);
  std::vector<int> p(m + 1, 0), way(m + 1, 0);
  for (int i = 1; i <= m; ++i) {
    p[0] = i;
    int j0 = 0;
    std::vector<long> minv(m + 1, LONG_MAX);
    std::vector<char> used(m + 1, 0);
    do {
      used[j0] = 1;
      int i0 = p[j0], j1 = 0;
      long delta = LONG_MAX;
      for (int j = 1; j <= m; ++j)
        if (!used[j]) {
          long cur = (big - w[i0 - 1][j - 1]) - u[i0] - v[j];
          if (cur < minv[j]) { minv[j] = cur; way[j] = j0; }
          if (minv[j] < delta) { delta = minv[j]; j1 = j; }
        }
      for (int j = 0; j <= m; ++j)
        if (used[j]) { u[p[j]] += delta; v[j] -= delta; } else minv[j] -= delta;
      j0 = j1;
    } while (p[j0] != 0);
    do { int j1 = way[j0]; p[j0] = p[j1]; j0 = j1; } while (j0);
  }
  long matched = 0;
  for (int j = 1; j <= m; ++j) matched += w[p[j] - 1][j - 1];
  return (long)n - matched;
}

}  // extern "C"
