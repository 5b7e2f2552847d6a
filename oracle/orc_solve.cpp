// oracle/orc_solve.cpp - CPU ORACLE (test infrastructure, see orc.h): connected components, seed lookup,
// Grady random walker per component.  Restates graph/impl/common.hpp:134-149, impl/random_walker.hpp:95-243
// and impl/random_walker_segmentation.hpp:144-196.
#include <cfloat>
#include <map>

#include "orc_internal.hpp"
#include "orc_ldlt.hpp"

// boost::connected_components: component ids in order of the smallest member vertex (appendix A.7)
int orc_components_impl(const orc_graph* g, std::vector<int32_t>& comp) {
  const size_t V = g->V;
  comp.assign(V, -1);
  int nc = 0;
  std::vector<uint32_t> stack;
  for (size_t s = 0; s < V; ++s) {
    if (comp[s] >= 0) continue;
    comp[s] = nc;
    stack.assign(1, (uint32_t)s);
    while (!stack.empty()) {
      uint32_t v = stack.back();
      stack.pop_back();
      for (uint32_t a = g->rowptr[v]; a < g->rowptr[v + 1]; ++a) {
        uint32_t u = g->col[a];
        if (comp[u] < 0) { comp[u] = nc; stack.push_back(u); }
      }
    }
    ++nc;
  }
  return nc;
}

extern "C" int orc_components(const orc_graph* g, int32_t* component) {
  std::vector<int32_t> comp;
  int nc = orc_components_impl(g, comp);
  if (component) std::memcpy(component, comp.data(), comp.size() * sizeof(int32_t));
  return nc;
}

// pcl::getApproximateIndices -> KdTreeFLANN exact 1-NN, squared L2 on xyz accumulated x, y, z (A.6);
// ties resolved to the lowest vertex id (FLANN leaves it unspecified)
extern "C" void orc_nearest_vertices(const orc_graph* g, const float* seed_xyz, size_t ns, int32_t* out) {
  for (size_t s = 0; s < ns; ++s) {
    float sx = seed_xyz[3 * s], sy = seed_xyz[3 * s + 1], sz = seed_xyz[3 * s + 2];
    float best = std::numeric_limits<float>::infinity();
    int32_t bi = -1;
    for (size_t v = 0; v < g->V; ++v) {
      float dx = sx - g->x[v], dy = sy - g->y[v], dz = sz - g->z[v];
      float d = 0.f;
      d += dx * dx; d += dy * dy; d += dz * dz;
      if (d < best) { best = d; bi = (int32_t)v; }
    }
    out[s] = bi;
  }
}

namespace {

// Jacobi-preconditioned CG in float64 on the reduced system of one component (ground truth for large cases)
bool pcg_f64(int n, const std::vector<double>& diag, const std::vector<int>& ti, const std::vector<int>& tj,
             const std::vector<double>& tv, const double* b, double* x) {
  std::vector<double> r(b, b + n), z(n), p(n), ap(n);
  std::fill(x, x + n, 0.0);
  double bnorm = 0;
  for (int i = 0; i < n; ++i) bnorm += b[i] * b[i];
  if (bnorm == 0) return true;
  double rz = 0;
  for (int i = 0; i < n; ++i) { z[i] = r[i] / diag[i]; p[i] = z[i]; rz += r[i] * z[i]; }
  for (int it = 0; it < 20 * n + 100; ++it) {
    for (int i = 0; i < n; ++i) ap[i] = diag[i] * p[i];
    for (size_t k = 0; k < ti.size(); ++k) { ap[ti[k]] += tv[k] * p[tj[k]]; ap[tj[k]] += tv[k] * p[ti[k]]; }
    double pap = 0;
    for (int i = 0; i < n; ++i) pap += p[i] * ap[i];
    double alpha = rz / pap, rr = 0, rz2 = 0;
    for (int i = 0; i < n; ++i) {
      x[i] += alpha * p[i];
      r[i] -= alpha * ap[i];
      z[i] = r[i] / diag[i];
      rr += r[i] * r[i];
      rz2 += r[i] * z[i];
    }
    if (rr <= 1e-24 * bnorm) return true;
    double beta = rz2 / rz;
    rz = rz2;
    for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
  }
  return false;
}

}  // namespace

// detail::RandomWalker::segment() per connected component
extern "C" int orc_random_walker(orc_graph* g, uint32_t* colours, int ncolours, int solver, float* potentials,
                                 double* residuals) {
  const size_t V = g->V;
  int status = ORC_OK;
  if (potentials) std::fill(potentials, potentials + V * (size_t)ncolours, 0.0f);
  if (residuals) std::fill(residuals, residuals + ncolours, 0.0);
  std::vector<int32_t> comp;
  int nc = orc_components_impl(g, comp);
  std::vector<std::vector<uint32_t>> members(nc);
  for (size_t v = 0; v < V; ++v) members[comp[v]].push_back((uint32_t)v);  // local ids ascend with global ids

  // computeVertexDegrees (:103-114): edge loop scatter-add == per-vertex sum in out-edge order
  std::vector<float> degree(V, 0.f);
  for (size_t e = 0; e < g->E; ++e) {
    float w = g->ew[e];
    degree[g->esrc[e]] += w;
    degree[g->etgt[e]] += w;
  }

  std::vector<int> row_of(V, -1);
  for (int c = 0; c < nc; ++c) {
    const std::vector<uint32_t>& mem = members[c];
    // buildLinearSystem (:116-184)
    std::vector<uint32_t> rows;              // L_vertex_bimap: row -> vertex
    std::vector<uint32_t> col_colour;        // B_color_bimap: column -> colour (first-encounter order)
    std::map<uint32_t, int> colour_col;
    std::vector<uint32_t> seeds;
    std::vector<float> diag;
    std::vector<int> ti, tj; std::vector<float> tv;          // strictly lower L entries
    std::vector<int> bi, bj; std::vector<float> bv;          // B triplets
    for (uint32_t v : mem) {
      if (colours[v]) seeds.push_back(v);
      if (colours[v] || std::fabs(degree[v]) < std::numeric_limits<float>::epsilon()) continue;
      int r = (int)rows.size();
      rows.push_back(v);
      row_of[v] = r;
      diag.push_back(degree[v]);
      for (uint32_t a = g->rowptr[v]; a < g->rowptr[v + 1]; ++a) {
        float w = g->ew[g->eid[a]];
        uint32_t t = g->col[a];
        uint32_t colour = colours[t];
        if (colour) {
          auto f = colour_col.find(colour);
          int column;
          if (f == colour_col.end()) {
            column = (int)col_colour.size();
            colour_col[colour] = column;
            col_colour.push_back(colour);
          } else {
            column = f->second;
          }
          bi.push_back(r); bj.push_back(column); bv.push_back(w);
        } else if (row_of[t] >= 0 && row_of[t] != r) {
          // "only if a row for this vertex already exists" -> lower triangle only (:165-171)
          ti.push_back(r); tj.push_back(row_of[t]); tv.push_back(-w);
        }
      }
    }
    const int n = (int)rows.size(), k = (int)col_colour.size();
    std::vector<float> X((size_t)n * k, 0.f);
    // solveLinearSystem (:186-207)
    if (n > 0 && k > 0) {
      std::vector<double> Bd((size_t)n * k, 0.0);
      std::vector<float> Bf((size_t)n * k, 0.f);
      for (size_t q = 0; q < bi.size(); ++q) Bf[(size_t)bj[q] * n + bi[q]] += bv[q];
      for (size_t q = 0; q < Bf.size(); ++q) Bd[q] = (double)Bf[q];
      std::vector<double> dd(diag.begin(), diag.end()), tvd(tv.begin(), tv.end());
      bool ok = true;
      if (solver == ORC_SOLVE_LDLT_F32) {
        orc::SimplicialLDLT<float> ldlt;
        ok = ldlt.compute(n, diag, ti, tj, tv);
        if (ok) for (int j = 0; j < k; ++j) ldlt.solve(&Bf[(size_t)j * n], &X[(size_t)j * n]);
      } else if (solver == ORC_SOLVE_LDLT_F64) {
        orc::SimplicialLDLT<double> ldlt;
        ok = ldlt.compute(n, dd, ti, tj, tvd);
        std::vector<double> xd(n);
        if (ok) for (int j = 0; j < k; ++j) {
          ldlt.solve(&Bd[(size_t)j * n], xd.data());
          for (int i = 0; i < n; ++i) X[(size_t)j * n + i] = (float)xd[i];
        }
      } else {
        std::vector<double> xd(n);
        for (int j = 0; j < k; ++j) {
          ok = pcg_f64(n, dd, ti, tj, tvd, &Bd[(size_t)j * n], xd.data()) && ok;
          for (int i = 0; i < n; ++i) X[(size_t)j * n + i] = (float)xd[i];
        }
      }
      if (!ok) status = ORC_ERR_SOLVE;
      if (residuals) {
        std::vector<double> ax(n);
        for (int j = 0; j < k; ++j) {
          const float* xj = &X[(size_t)j * n];
          for (int i = 0; i < n; ++i) ax[i] = dd[i] * (double)xj[i];
          for (size_t q = 0; q < ti.size(); ++q) {
            ax[ti[q]] += tvd[q] * (double)xj[tj[q]];
            ax[tj[q]] += tvd[q] * (double)xj[ti[q]];
          }
          double rn = 0, bn = 0;
          for (int i = 0; i < n; ++i) {
            double b = Bd[(size_t)j * n + i];
            rn += (ax[i] - b) * (ax[i] - b);
            bn += b * b;
          }
          double rel = bn > 0 ? std::sqrt(rn / bn) : 0.0;
          int cc = (int)col_colour[j] - 1;
          if (cc >= 0 && cc < ncolours) residuals[cc] = std::max(residuals[cc], rel);
        }
      }
      // assignColors (:209-222): first maximal coefficient of the row
      for (int i = 0; i < n; ++i) {
        int best = 0;
        float bvx = X[i];
        for (int j = 1; j < k; ++j)
          if (X[(size_t)j * n + i] > bvx) { bvx = X[(size_t)j * n + i]; best = j; }
        colours[rows[i]] = col_colour[best];
      }
    }
    // getPotentials (:224-243) -> potentials_(global vertex, colour - 1) (random_walker_segmentation.hpp:178-186)
    if (potentials) {
      for (int j = 0; j < k; ++j) {
        int cc = (int)col_colour[j] - 1;
        if (cc < 0 || cc >= ncolours) continue;
        for (int i = 0; i < n; ++i) potentials[(size_t)cc * V + rows[i]] = X[(size_t)j * n + i];
      }
      for (uint32_t s : seeds) {
        int cc = (int)colours[s] - 1;
        if (cc >= 0 && cc < ncolours) potentials[(size_t)cc * V + s] = 1.0f;
      }
    }
    for (uint32_t v : rows) row_of[v] = -1;
  }
  return status;
}
