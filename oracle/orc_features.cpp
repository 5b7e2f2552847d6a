// oracle/orc_features.cpp - CPU ORACLE (test infrastructure, see orc.h): normals/curvature, smoothing,
// curvature signs, edge weights.  Restates graph/impl/common.hpp:53-132,241-279 and
// graph/impl/edge_weight_computer.hpp:110-188 (+ edge_weight_computer.h:333-469, edge_weight_computer_terms.h).
#include <cfloat>

#include "orc_internal.hpp"

namespace {

// pcl::computeRoots (A.4): smallest root of the characteristic polynomial of a scaled symmetric 3x3 matrix
float smallest_root(const float m[3][3]) {
  float c0 = m[0][0] * m[1][1] * m[2][2] + 2.0f * m[0][1] * m[0][2] * m[1][2] - m[0][0] * m[1][2] * m[1][2] -
             m[1][1] * m[0][2] * m[0][2] - m[2][2] * m[0][1] * m[0][1];
  float c1 = m[0][0] * m[1][1] - m[0][1] * m[0][1] + m[0][0] * m[2][2] - m[0][2] * m[0][2] + m[1][1] * m[2][2] -
             m[1][2] * m[1][2];
  float c2 = m[0][0] + m[1][1] + m[2][2];
  if (std::fabs(c0) < FLT_EPSILON) return 0.0f;  // computeRoots2: roots(0) = 0
  const float s_inv3 = (float)(1.0 / 3.0);
  const float s_sqrt3 = std::sqrt(3.0f);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > 0.0f) a_over_3 = 0.0f;
  float half_b = 0.5f * (c0 + c2_over_3 * (2.0f * c2_over_3 * c2_over_3 - c1));
  float q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > 0.0f) q = 0.0f;
  float rho = std::sqrt(-a_over_3);
  float theta = orc_atan2f(std::sqrt(-q), half_b) * s_inv3;
  float cos_theta = orc_cosf(theta);
  float sin_theta = orc_sinf(theta);
  float r0 = c2_over_3 + 2.0f * rho * cos_theta;
  float r1 = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  float r2 = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  if (r0 >= r1) std::swap(r0, r1);
  if (r1 >= r2) {
    std::swap(r1, r2);
    if (r0 >= r1) std::swap(r0, r1);
  }
  if (r0 <= 0.0f) return 0.0f;  // "eigenval ... can not be negative! Set it to 0" -> computeRoots2
  return r0;
}

// pcl::eigen33(mat, eigenvalue, eigenvector) (A.4)
void eigen33(const float c[3][3], float& eigenvalue, float ev[3]) {
  float scale = 0.0f;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) scale = std::max(scale, std::fabs(c[i][j]));
  if (scale <= FLT_MIN) scale = 1.0f;
  float s[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) s[i][j] = c[i][j] / scale;
  float root = smallest_root(s);
  eigenvalue = root * scale;
  s[0][0] -= root; s[1][1] -= root; s[2][2] -= root;
  auto cross = [](const float* a, const float* b, float* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
  };
  float v1[3], v2[3], v3[3];
  cross(s[0], s[1], v1); cross(s[0], s[2], v2); cross(s[1], s[2], v3);
  float l1 = sqn3(v1[0], v1[1], v1[2]), l2 = sqn3(v2[0], v2[1], v2[2]), l3 = sqn3(v3[0], v3[1], v3[2]);
  const float* best; float len;
  if (l1 >= l2 && l1 >= l3) { best = v1; len = l1; }
  else if (l2 >= l1 && l2 >= l3) { best = v2; len = l2; }
  else { best = v3; len = l3; }
  float sl = std::sqrt(len);
  for (int i = 0; i < 3; ++i) ev[i] = best[i] / sl;
}

}  // namespace

// graph/impl/common.hpp:53-103
extern "C" void orc_normals(orc_graph* g, int one_ring, int mode) {
  const size_t V = g->V;
  std::vector<float> onx(V), ony(V), onz(V), onw(V), ocv(V);
  std::vector<uint32_t> nb;
  for (size_t v = 0; v < V; ++v) {
    // :64-76 neighbourhood multiset [v] + sum_u ([u] + adj(u)), duplicates kept
    nb.assign(1, (uint32_t)v);
    for (uint32_t a = g->rowptr[v]; a < g->rowptr[v + 1]; ++a) {
      uint32_t u = g->col[a];
      nb.push_back(u);
      if (!one_ring)
        for (uint32_t b = g->rowptr[u]; b < g->rowptr[u + 1]; ++b) nb.push_back(g->col[b]);
    }
    float n[4] = {0, 0, 0, 0};
    float curvature;
    if (nb.size() < 3) {
      // pcl::computePointNormal: fewer than 3 indices -> NaN plane, NaN curvature
      n[0] = n[1] = n[2] = n[3] = std::numeric_limits<float>::quiet_NaN();
      curvature = std::numeric_limits<float>::quiet_NaN();
    } else {
      // pcl::computeMeanAndCovarianceMatrix: single pass, nine float accumulators, no centring
      float accu[9];
      if (mode == ORC_MODE_PARITY) {
        for (int k = 0; k < 9; ++k) accu[k] = 0.f;
        for (uint32_t i : nb) {
          float px = g->x[i], py = g->y[i], pz = g->z[i];
          accu[0] += px * px; accu[1] += px * py; accu[2] += px * pz;
          accu[3] += py * py; accu[4] += py * pz; accu[5] += pz * pz;
          accu[6] += px; accu[7] += py; accu[8] += pz;
        }
      } else {
        // fast-mode definition: float products, order-free float64 sums, rounded once to float
        double d[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (uint32_t i : nb) {
          float px = g->x[i], py = g->y[i], pz = g->z[i];
          d[0] += (double)(px * px); d[1] += (double)(px * py); d[2] += (double)(px * pz);
          d[3] += (double)(py * py); d[4] += (double)(py * pz); d[5] += (double)(pz * pz);
          d[6] += (double)px; d[7] += (double)py; d[8] += (double)pz;
        }
        for (int k = 0; k < 9; ++k) accu[k] = (float)d[k];
      }
      // accu /= (float)count : Eigen 3.2 DenseBase::operator/= multiplies by the reciprocal (A.8)
      float inv = 1.0f / static_cast<float>(nb.size());
      for (int k = 0; k < 9; ++k) accu[k] *= inv;
      float c[3][3];
      c[0][0] = accu[0] - accu[6] * accu[6];
      c[0][1] = accu[1] - accu[6] * accu[7];
      c[0][2] = accu[2] - accu[6] * accu[8];
      c[1][1] = accu[3] - accu[7] * accu[7];
      c[1][2] = accu[4] - accu[7] * accu[8];
      c[2][2] = accu[5] - accu[8] * accu[8];
      c[1][0] = c[0][1]; c[2][0] = c[0][2]; c[2][1] = c[1][2];
      // pcl::solvePlaneParameters
      float lambda, ev[3];
      eigen33(c, lambda, ev);
      n[0] = ev[0]; n[1] = ev[1]; n[2] = ev[2];
      float eig_sum = c[0][0] + c[1][1] + c[2][2];
      curvature = (eig_sum != 0) ? std::fabs(lambda / eig_sum) : 0.0f;
    }
    // :81-82 normal[3] = 0; normal.normalize()  (Vector4f: (x^2+y^2)+(z^2+0); v *= 1/norm)
    n[3] = 0.0f;
    {
      float nrm = std::sqrt((n[0] * n[0] + n[1] * n[1]) + (n[2] * n[2] + n[3] * n[3]));
      float inv = 1.0f / nrm;
      for (int k = 0; k < 4; ++k) n[k] *= inv;
    }
    // :87-99 orientation
    float gx = g->nx[v], gy = g->ny[v], gz = g->nz[v], gw = g->nw[v];
    if (std::isfinite(gx) && std::isfinite(gy) && std::isfinite(gz) && std::isfinite(gw) &&
        (gx != 0.f || gy != 0.f || gz != 0.f || gw != 0.f)) {
      float d = (gx * n[0] + gy * n[1]) + (gz * n[2] + gw * n[3]);
      if (d < 0) for (int k = 0; k < 4; ++k) n[k] *= -1.0f;
    } else {
      // pcl::flipNormalTowardsViewpoint(point, 0, 0, 0, normal) (A.5)
      float vx = 0.0f - g->x[v], vy = 0.0f - g->y[v], vz = 0.0f - g->z[v];
      float cos_theta = (vx * n[0] + vy * n[1]) + (vz * n[2] + 0.0f * n[3]);
      if (cos_theta < 0) {
        for (int k = 0; k < 4; ++k) n[k] *= -1.0f;
        n[3] = 0.0f;
        n[3] = -1.0f * ((n[0] * g->x[v] + n[1] * g->y[v]) + (n[2] * g->z[v] + n[3] * 1.0f));
      }
    }
    onx[v] = n[0]; ony[v] = n[1]; onz[v] = n[2]; onw[v] = n[3];
    ocv[v] = std::isnan(curvature) ? 0.0f : curvature;  // :101
  }
  // the reference writes vertex v before visiting v+1, but only xyz of other vertices is read -> same result
  g->nx.swap(onx); g->ny.swap(ony); g->nz.swap(onz); g->nw.swap(onw); g->curv.swap(ocv);
}

// graph/impl/common.hpp:241-279
extern "C" void orc_smoothen(orc_graph* g, float spatial_sigma, float influence_sigma) {
  const size_t V = g->V;
  std::vector<float> K(V, 0.f), P0(V, 0.f), P1(V, 0.f), P2(V, 0.f);
  for (size_t e = 0; e < g->E; ++e) {
    uint32_t s = g->esrc[e], t = g->etgt[e];
    float xx = g->x[s] - g->x[t], xy = g->y[s] - g->y[t], xz = g->z[s] - g->z[t];  // x = p - q
    float d1 = dot3(g->nx[t], g->ny[t], g->nz[t], xx, xy, xz);                     // nq . x
    float d2 = dot3(g->nx[s], g->ny[s], g->nz[s], -xx, -xy, -xz);                  // np . (-x)
    // std::pow(float, int) and the literal 2 promote to double; std::exp(double); result stored in float
    double nrm = (double)std::sqrt(sqn3(xx, xy, xz));
    double ss2 = 2 * ((double)spatial_sigma * (double)spatial_sigma);
    double si2 = 2 * ((double)influence_sigma * (double)influence_sigma);
    float ws = (float)std::exp(-(nrm * nrm) / ss2);
    float wi1 = (float)std::exp(-((double)d1 * (double)d1) / si2);
    float wi2 = (float)std::exp(-((double)d2 * (double)d2) / si2);
    float w1 = ws * wi1, w2 = ws * wi2;
    K[s] += w1;
    K[t] += w2;
    P0[s] += g->nx[t] * d1 * w1; P1[s] += g->ny[t] * d1 * w1; P2[s] += g->nz[t] * d1 * w1;
    P0[t] += g->nx[s] * d2 * w2; P1[t] += g->ny[s] * d2 * w2; P2[t] += g->nz[s] * d2 * w2;
  }
  for (size_t v = 0; v < V; ++v)
    if (K[v] > 0.01) {
      g->x[v] -= P0[v] / K[v];
      g->y[v] -= P1[v] / K[v];
      g->z[v] -= P2[v] / K[v];
    }
}

// graph/impl/common.hpp:105-132
extern "C" void orc_signed_curvatures(orc_graph* g) {
  std::vector<float> conv(g->V, 0.0f);
  for (size_t e = 0; e < g->E; ++e) {
    uint32_t v1 = g->esrc[e], v2 = g->etgt[e];
    float dx = g->x[v2] - g->x[v1], dy = g->y[v2] - g->y[v1], dz = g->z[v2] - g->z[v1];
    float n1x = g->nx[v1], n1y = g->ny[v1], n1z = g->nz[v1];
    float n2x = g->nx[v2], n2y = g->ny[v2], n2z = g->nz[v2];
    float dn = dot3(dx, dy, dz, n1x, n1y, n1z);
    float rx = dx - dn * n1x, ry = dy - dn * n1y, rz = dz - dn * n1z;
    float sgn = dot3(rx, ry, rz, n2x, n2y, n2z) > 0 ? 1.0f : -1.0f;
    float c = sgn * sqn3(n1x - n2x, n1y - n2y, n1z - n2z);
    conv[v1] += c;
    conv[v2] += c;
  }
  for (size_t v = 0; v < g->V; ++v) g->curv[v] = std::copysign(g->curv[v], conv[v]);
}

namespace {

// graph/edge_weight_computer_terms.h:55-140
float term_distance(const orc_graph* g, int kind, uint32_t a, uint32_t b) {
  switch (kind) {
    case ORC_TERM_XYZ:
      return sqn3(g->x[b] - g->x[a], g->y[b] - g->y[a], g->z[b] - g->z[a]);
    case ORC_TERM_NORMAL:
      return 0.5f * sqn3(g->nx[a] - g->nx[b], g->ny[a] - g->ny[b], g->nz[a] - g->nz[b]);
    case ORC_TERM_CURVATURE:
      return std::fabs(g->curv[a]) * std::fabs(g->curv[b]);
    case ORC_TERM_RGB: {
      // getBGRVector3cMap().cast<float>(): (b, g, r)
      uint32_t ca = g->rgba[a], cb = g->rgba[b];
      float d0 = (float)(ca & 0xFF) - (float)(cb & 0xFF);
      float d1 = (float)((ca >> 8) & 0xFF) - (float)((cb >> 8) & 0xFF);
      float d2 = (float)((ca >> 16) & 0xFF) - (float)((cb >> 16) & 0xFF);
      return std::sqrt(sqn3(d0, d1, d2)) / 255.0f;
    }
  }
  return 0.f;
}

// EdgeWeightComputer::gaussian (edge_weight_computer.h:333-337)
inline float gaussian(float val, float influence) { return influence > 0.0f ? orc_expf(-val / influence) : 1.0f; }
// setTermBalancingFunction (edge_weight_computer.h:207,325): the closed set the device implements (ORC_BALANCE_*);
// 0 = the reference's default Gaussian
inline float balance(int kind, float val, float influence) {
  if (kind == 0) return gaussian(val, influence);
  if (!(influence > 0.0f)) return 1.0f;
  const float q = val / influence;
  if (kind == 1) return 1.0f / (1.0f + q);
  if (kind == 2) return orc_expf(-(q * q));
  const float l = 1.0f - q;
  return l > 0.0f ? l : 0.0f;
}

}  // namespace

// graph/impl/edge_weight_computer.hpp:110-188
extern "C" int orc_edge_weights(orc_graph* g, const orc_term* terms, int nterms, float threshold, int policy,
                                int mode) {
  return orc_edge_weights_ex(g, terms, nterms, threshold, policy, mode, 0);
}

extern "C" int orc_edge_weights_ex(orc_graph* g, const orc_term* terms, int nterms, float threshold, int policy,
                                   int mode, int balancing) {
  const size_t E = g->E, V = g->V;
  // terms are stored per normalisation class and evaluated plain -> global -> local (:148-154)
  std::vector<int> plain, glob, loc;
  for (int i = 0; i < nterms; ++i) {
    if ((terms[i].kind == ORC_TERM_RGB) && !g->has_color) continue;  // addTermImpl(false): incompatible point type
    if (terms[i].normalization == ORC_NORM_GLOBAL) glob.push_back(i);
    else if (terms[i].normalization == ORC_NORM_LOCAL) loc.push_back(i);
    else plain.push_back(i);
  }
  std::vector<std::vector<float>> gd(glob.size()), ld(loc.size()), lsum(loc.size());
  std::vector<float> gavg(glob.size(), 0.f);
  std::vector<std::vector<uint32_t>> ldeg(loc.size());
  if (!glob.empty() || !loc.empty()) {
    std::vector<float> gtot(glob.size(), 0.f);
    std::vector<double> gtot64(glob.size(), 0.0);
    for (size_t i = 0; i < glob.size(); ++i) gd[i].assign(E, 0.f);
    for (size_t i = 0; i < loc.size(); ++i) { ld[i].assign(E, 0.f); lsum[i].assign(V, 0.f); ldeg[i].assign(V, 0); }
    for (size_t e = 0; e < E; ++e) {
      uint32_t v1 = g->esrc[e], v2 = g->etgt[e];
      for (size_t i = 0; i < glob.size(); ++i) {  // GloballyNormalizedTerm::round1
        float w = term_distance(g, terms[glob[i]].kind, v1, v2);
        gd[i][e] = w;
        gtot[i] += w;
        gtot64[i] += (double)w;
      }
      for (size_t i = 0; i < loc.size(); ++i) {  // LocallyNormalizedTerm::round1
        float w = term_distance(g, terms[loc[i]].kind, v1, v2);
        ld[i][e] = w;
        lsum[i][v1] += w;  // per-vertex sums stay sequential float in both modes (cheap to reproduce)
        lsum[i][v2] += w;
        ++ldeg[i][v1];
        ++ldeg[i][v2];
      }
    }
    for (size_t i = 0; i < glob.size(); ++i) {  // extract(): average = total / size()
      float tot = mode == ORC_MODE_PARITY ? gtot[i] : (float)gtot64[i];
      gavg[i] = tot / static_cast<float>(E);
    }
    for (size_t i = 0; i < loc.size(); ++i)
      for (size_t v = 0; v < V; ++v)
        if (ldeg[i][v]) lsum[i][v] /= static_cast<float>(ldeg[i][v]);
  }
  for (size_t e = 0; e < E; ++e) {
    uint32_t v1 = g->esrc[e], v2 = g->etgt[e];
    bool convex = g->curv[v1] > 0.0f && g->curv[v2] > 0.0f;  // detail::IsConvex (:95-102)
    float a = 1.0f;
    auto influence = [&](int t, bool cvx) { return terms[t].influence * (cvx ? terms[t].convex_mult : 1.0f); };
    for (int t : plain) a *= balance(balancing, term_distance(g, terms[t].kind, v1, v2), influence(t, convex));
    for (size_t i = 0; i < glob.size(); ++i) a *= balance(balancing, gd[i][e] / gavg[i], influence(glob[i], convex));
    for (size_t i = 0; i < loc.size(); ++i) {
      // LocallyNormalizedTerm::round2 (edge_weight_computer.h:451-460)
      float n = (lsum[i][v1] + lsum[i][v2]) / 2.0f;
      float w = (n > 0.0f && influence(loc[i], false) > 0.0f) ? ld[i][e] / n : 0.0f;
      a *= balance(balancing, w, influence(loc[i], convex));
    }
    g->ew[e] = a;
  }
  if (policy == ORC_SMALL_COERCE) {
    for (size_t e = 0; e < E; ++e) if (g->ew[e] < threshold) g->ew[e] = threshold;
  } else if (policy == ORC_SMALL_REMOVE) {
    // remove_edge_if: surviving edges keep their relative order in the edge list and in every out-edge list
    std::vector<uint32_t> s2, t2; std::vector<float> w2;
    for (size_t e = 0; e < E; ++e)
      if (!(g->ew[e] < threshold)) { s2.push_back(g->esrc[e]); t2.push_back(g->etgt[e]); w2.push_back(g->ew[e]); }
    g->esrc.swap(s2); g->etgt.swap(t2); g->ew.swap(w2);
    g->build_out_edges();
  }
  return ORC_OK;
}
