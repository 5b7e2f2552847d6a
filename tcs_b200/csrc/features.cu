// tcs_b200/csrc/features.cu - per-vertex normals/curvature, bilateral smoothing, curvature signs.
// Replaces pcl::graph::computeNormalsAndCurvatures / smoothen / computeSignedCurvatures
// (graph/impl/common.hpp:53-132, 241-279) and the PCL routines they call.
// Compiled with -fmad=false: float arithmetic is the reference's operation sequence, no contraction.
#include <cfloat>

#include "common.cuh"

namespace {

// Eigen 3.2 reduction orders (appendix A.8): 3-vector e0 + (e1 + e2); Vector4f (e0 + e1) + (e2 + e3)
__device__ __forceinline__ float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}
__device__ __forceinline__ float sqn3(float a0, float a1, float a2) { return a0 * a0 + (a1 * a1 + a2 * a2); }




// pcl::computeRoots (appendix A.4): smallest root of the characteristic polynomial
__device__ float smallest_root(const float m[3][3]) {
  float c0 = m[0][0] * m[1][1] * m[2][2] + 2.0f * m[0][1] * m[0][2] * m[1][2] - m[0][0] * m[1][2] * m[1][2] -
             m[1][1] * m[0][2] * m[0][2] - m[2][2] * m[0][1] * m[0][1];
  float c1 = m[0][0] * m[1][1] - m[0][1] * m[0][1] + m[0][0] * m[2][2] - m[0][2] * m[0][2] + m[1][1] * m[2][2] -
             m[1][2] * m[1][2];
  float c2 = m[0][0] + m[1][1] + m[2][2];
  if (fabsf(c0) < FLT_EPSILON) return 0.0f;
  const float s_inv3 = (float)(1.0 / 3.0);
  const float s_sqrt3 = sqrtf(3.0f);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > 0.0f) a_over_3 = 0.0f;
  float half_b = 0.5f * (c0 + c2_over_3 * (2.0f * c2_over_3 * c2_over_3 - c1));
  float q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > 0.0f) q = 0.0f;
  float rho = sqrtf(-a_over_3);
  float theta = (float)atan2((double)sqrtf(-q), (double)half_b) * s_inv3;
  float cos_theta = (float)cos((double)theta);
  float sin_theta = (float)sin((double)theta);
  float r0 = c2_over_3 + 2.0f * rho * cos_theta;
  float r1 = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  float r2 = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  float t;
  if (r0 >= r1) { t = r0; r0 = r1; r1 = t; }
  if (r1 >= r2) {
    t = r1; r1 = r2; r2 = t;
    if (r0 >= r1) { t = r0; r0 = r1; r1 = t; }
  }
  if (r0 <= 0.0f) return 0.0f;
  return r0;
}

// pcl::eigen33(mat, eigenvalue, eigenvector): smallest eigenpair
__device__ void eigen33(const float c[3][3], float& eigenvalue, float ev[3]) {
  float scale = 0.0f;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) scale = fmaxf(scale, fabsf(c[i][j]));
  if (scale <= FLT_MIN) scale = 1.0f;
  float s[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) s[i][j] = c[i][j] / scale;
  float root = smallest_root(s);
  eigenvalue = root * scale;
  s[0][0] -= root; s[1][1] -= root; s[2][2] -= root;
  float v1[3], v2[3], v3[3];
#define CROSS(a, b, o)                          \
  o[0] = a[1] * b[2] - a[2] * b[1];             \
  o[1] = a[2] * b[0] - a[0] * b[2];             \
  o[2] = a[0] * b[1] - a[1] * b[0];
  CROSS(s[0], s[1], v1) CROSS(s[0], s[2], v2) CROSS(s[1], s[2], v3)
#undef CROSS
  float l1 = sqn3(v1[0], v1[1], v1[2]), l2 = sqn3(v2[0], v2[1], v2[2]), l3 = sqn3(v3[0], v3[1], v3[2]);
  const float* best;
  float len;
  if (l1 >= l2 && l1 >= l3) { best = v1; len = l1; }
  else if (l2 >= l1 && l2 >= l3) { best = v2; len = l2; }
  else { best = v3; len = l3; }
  float sl = sqrtf(len);
#pragma unroll
  for (int i = 0; i < 3; ++i) ev[i] = best[i] / sl;
}

// everything after the nine moment sums (common.hpp:80-101 + pcl::computePointNormal tail)
__device__ void finish_normal(float accu[9], uint32_t count, float px, float py, float pz, float gx, float gy,
                              float gz, float gw, float out_n[4], float& out_curv) {
  float n[4] = {0.f, 0.f, 0.f, 0.f};
  float curvature;
  if (count < 3) {
    n[0] = n[1] = n[2] = n[3] = __int_as_float(0x7FC00000);
    curvature = __int_as_float(0x7FC00000);
  } else {
    float inv = 1.0f / (float)count;  // Eigen 3.2 operator/= multiplies by the reciprocal
#pragma unroll
    for (int k = 0; k < 9; ++k) accu[k] *= inv;
    float c[3][3];
    c[0][0] = accu[0] - accu[6] * accu[6];
    c[0][1] = accu[1] - accu[6] * accu[7];
    c[0][2] = accu[2] - accu[6] * accu[8];
    c[1][1] = accu[3] - accu[7] * accu[7];
    c[1][2] = accu[4] - accu[7] * accu[8];
    c[2][2] = accu[5] - accu[8] * accu[8];
    c[1][0] = c[0][1]; c[2][0] = c[0][2]; c[2][1] = c[1][2];
    float lambda, ev[3];
    eigen33(c, lambda, ev);
    n[0] = ev[0]; n[1] = ev[1]; n[2] = ev[2];
    float eig_sum = c[0][0] + c[1][1] + c[2][2];
    curvature = (eig_sum != 0) ? fabsf(lambda / eig_sum) : 0.0f;
  }
  n[3] = 0.0f;
  {
    float nrm = sqrtf((n[0] * n[0] + n[1] * n[1]) + (n[2] * n[2] + n[3] * n[3]));
    float inv = 1.0f / nrm;
#pragma unroll
    for (int k = 0; k < 4; ++k) n[k] *= inv;
  }
  if (isfinite(gx) && isfinite(gy) && isfinite(gz) && isfinite(gw) &&
      (gx != 0.f || gy != 0.f || gz != 0.f || gw != 0.f)) {
    float d = (gx * n[0] + gy * n[1]) + (gz * n[2] + gw * n[3]);
    if (d < 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) n[k] *= -1.0f;
    }
  } else {
    // pcl::flipNormalTowardsViewpoint(point, 0, 0, 0, normal) (appendix A.5)
    float vx = 0.0f - px, vy = 0.0f - py, vz = 0.0f - pz;
    float cos_theta = (vx * n[0] + vy * n[1]) + (vz * n[2] + 0.0f * n[3]);
    if (cos_theta < 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) n[k] *= -1.0f;
      n[3] = 0.0f;
      n[3] = -1.0f * ((n[0] * px + n[1] * py) + (n[2] * pz + n[3] * 1.0f));
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) out_n[k] = n[k];
  out_curv = isnan(curvature) ? 0.0f : curvature;
}

#define ACCUM(px, py, pz)                                              \
  accu[0] += px * px; accu[1] += px * py; accu[2] += px * pz;          \
  accu[3] += py * py; accu[4] += py * pz; accu[5] += pz * pz;          \
  accu[6] += px; accu[7] += py; accu[8] += pz;

// PARITY mode: one thread per vertex walks the multiset [v] + sum_u([u] + adj(u)) in the reference's order with
// nine sequential float accumulators (pcl::computeMeanAndCovarianceMatrix).
__global__ void k_normals_parity(size_t V, int one_ring, const uint32_t* __restrict__ rowptr,
                                 const uint32_t* __restrict__ col, const float* __restrict__ x,
                                 const float* __restrict__ y, const float* __restrict__ z,
                                 const float* __restrict__ gnx, const float* __restrict__ gny,
                                 const float* __restrict__ gnz, const float* __restrict__ gnw,
                                 float* __restrict__ onx, float* __restrict__ ony, float* __restrict__ onz,
                                 float* __restrict__ onw, float* __restrict__ ocv) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  float accu[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  uint32_t count = 1;
  const float vx = x[v], vy = y[v], vz = z[v];
  ACCUM(vx, vy, vz)
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    float px = x[u], py = y[u], pz = z[u];
    ACCUM(px, py, pz)
    ++count;
    if (!one_ring) {
      uint32_t b1 = rowptr[u + 1];
      for (uint32_t b = rowptr[u]; b < b1; ++b) {
        uint32_t t = col[b];
        float qx = x[t], qy = y[t], qz = z[t];
        ACCUM(qx, qy, qz)
        ++count;
      }
    }
  }
  float n[4], cv;
  finish_normal(accu, count, vx, vy, vz, gnx[v], gny[v], gnz[v], gnw[v], n, cv);
  onx[v] = n[0]; ony[v] = n[1]; onz[v] = n[2]; onw[v] = n[3]; ocv[v] = cv;
}
#undef ACCUM

// FAST mode: the 2-ring multiset factorises, sum_{M(v)} m = m(v) + sum_{u in adj(v)} S1(u) with
// S1(u) = m(u) + sum_{t in adj(u)} m(t), so two 1-ring gathers (~12 neighbours each) replace ~165 point gathers.
// Moments are float32 products accumulated in float64 (effectively exact => order-free) and rounded once.
// S1 rows: 10 doubles {xx, xy, xz, yy, yz, zz, x, y, z, count}.
__global__ void k_moments_ring1(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                                const float* __restrict__ x, const float* __restrict__ y,
                                const float* __restrict__ z, double* __restrict__ S1) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  double d[9];
  {
    float px = x[v], py = y[v], pz = z[v];
    d[0] = (double)(px * px); d[1] = (double)(px * py); d[2] = (double)(px * pz);
    d[3] = (double)(py * py); d[4] = (double)(py * pz); d[5] = (double)(pz * pz);
    d[6] = (double)px; d[7] = (double)py; d[8] = (double)pz;
  }
  uint32_t a0 = rowptr[v], a1 = rowptr[v + 1];
  for (uint32_t a = a0; a < a1; ++a) {
    uint32_t u = col[a];
    float px = x[u], py = y[u], pz = z[u];
    d[0] += (double)(px * px); d[1] += (double)(px * py); d[2] += (double)(px * pz);
    d[3] += (double)(py * py); d[4] += (double)(py * pz); d[5] += (double)(pz * pz);
    d[6] += (double)px; d[7] += (double)py; d[8] += (double)pz;
  }
  double2* o = (double2*)(S1 + 10 * v);
  o[0] = make_double2(d[0], d[1]); o[1] = make_double2(d[2], d[3]); o[2] = make_double2(d[4], d[5]);
  o[3] = make_double2(d[6], d[7]); o[4] = make_double2(d[8], (double)(1 + a1 - a0));
}

__global__ void k_normals_fast(size_t V, int one_ring, const uint32_t* __restrict__ rowptr,
                               const uint32_t* __restrict__ col, const float* __restrict__ x,
                               const float* __restrict__ y, const float* __restrict__ z,
                               const double* __restrict__ S1, const float* __restrict__ gnx,
                               const float* __restrict__ gny, const float* __restrict__ gnz,
                               const float* __restrict__ gnw, float* __restrict__ onx, float* __restrict__ ony,
                               float* __restrict__ onz, float* __restrict__ onw, float* __restrict__ ocv) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  const float vx = x[v], vy = y[v], vz = z[v];
  double d[10];
  if (one_ring) {
    const double2* s = (const double2*)(S1 + 10 * v);
#pragma unroll
    for (int k = 0; k < 5; ++k) { double2 t = s[k]; d[2 * k] = t.x; d[2 * k + 1] = t.y; }
  } else {
    d[0] = (double)(vx * vx); d[1] = (double)(vx * vy); d[2] = (double)(vx * vz);
    d[3] = (double)(vy * vy); d[4] = (double)(vy * vz); d[5] = (double)(vz * vz);
    d[6] = (double)vx; d[7] = (double)vy; d[8] = (double)vz; d[9] = 1.0;
    for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
      const double2* s = (const double2*)(S1 + 10 * (size_t)col[a]);
#pragma unroll
      for (int k = 0; k < 5; ++k) { double2 t = s[k]; d[2 * k] += t.x; d[2 * k + 1] += t.y; }
    }
  }
  float accu[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) accu[k] = (float)d[k];
  float n[4], cv;
  finish_normal(accu, (uint32_t)d[9], vx, vy, vz, gnx[v], gny[v], gnz[v], gnw[v], n, cv);
  onx[v] = n[0]; ony[v] = n[1]; onz[v] = n[2]; onw[v] = n[3]; ocv[v] = cv;
}

}  // namespace

extern "C" int tcs_normals(tcs_graph* g, int one_ring, int mode) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V;
  if (V == 0) return TCS_OK;
  CK(cudaSetDevice(ctx->device));
  DBuf<float> onx, ony, onz, onw, ocv;
  TRY(onx.alloc(ctx, V)); TRY(ony.alloc(ctx, V)); TRY(onz.alloc(ctx, V)); TRY(onw.alloc(ctx, V)); TRY(ocv.alloc(ctx, V));
  if (mode == TCS_MODE_PARITY) {
    LAUNCH(ctx, k_normals_parity, div_up(V, 128), 128, 0, V, one_ring, g->rowptr.p, g->col.p, g->x.p, g->y.p, g->z.p,
           g->nx.p, g->ny.p, g->nz.p, g->nw.p, onx.p, ony.p, onz.p, onw.p, ocv.p);
  } else {
    DBuf<double> S1;
    TRY(S1.alloc(ctx, 10 * V));
    LAUNCH(ctx, k_moments_ring1, div_up(V, 128), 128, 0, V, g->rowptr.p, g->col.p, g->x.p, g->y.p, g->z.p, S1.p);
    LAUNCH(ctx, k_normals_fast, div_up(V, 128), 128, 0, V, one_ring, g->rowptr.p, g->col.p, g->x.p, g->y.p, g->z.p,
           S1.p, g->nx.p, g->ny.p, g->nz.p, g->nw.p, onx.p, ony.p, onz.p, onw.p, ocv.p);
  }
  CK(cudaGetLastError());
  g->nx.swap(onx); g->ny.swap(ony); g->nz.swap(onz); g->nw.swap(onw); g->curv.swap(ocv);
  return TCS_OK;
}

// ---- smoothen (graph/impl/common.hpp:241-279) ---------------------------------------------------------------
// The reference scatters per edge into K[v], P[v]; every vertex receives its contributions in ascending edge id,
// which is exactly its out-edge order, so a per-vertex gather over the CSR row reproduces the float sums.
namespace {

__global__ void k_smoothen(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                           const uint32_t* __restrict__ eid, const uint32_t* __restrict__ esrc,
                           const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z,
                           const float* __restrict__ nx, const float* __restrict__ ny, const float* __restrict__ nz,
                           float spatial_sigma, float influence_sigma, float* __restrict__ ox,
                           float* __restrict__ oy, float* __restrict__ oz) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  const double ss2 = 2 * ((double)spatial_sigma * (double)spatial_sigma);
  const double si2 = 2 * ((double)influence_sigma * (double)influence_sigma);
  const float pvx = x[v], pvy = y[v], pvz = z[v];
  const float nvx = nx[v], nvy = ny[v], nvz = nz[v];
  float K = 0.f, P0 = 0.f, P1 = 0.f, P2 = 0.f;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    bool v_is_source = esrc[eid[a]] == (uint32_t)v;
    float pux = x[u], puy = y[u], puz = z[u];
    float nux = nx[u], nuy = ny[u], nuz = nz[u];
    // x = p(source) - q(target)
    float xx, xy, xz;
    if (v_is_source) { xx = pvx - pux; xy = pvy - puy; xz = pvz - puz; }
    else { xx = pux - pvx; xy = puy - pvy; xz = puz - pvz; }
    double nrm = (double)sqrtf(sqn3(xx, xy, xz));
    float ws = (float)exp(-(nrm * nrm) / ss2);
    float d;
    if (v_is_source) d = dot3(nux, nuy, nuz, xx, xy, xz);      // d1 = nq . x      -> contributes to the source
    else d = dot3(nux, nuy, nuz, -xx, -xy, -xz);               // d2 = np . (-x)   -> contributes to the target
    float wi = (float)exp(-((double)d * (double)d) / si2);
    float w = ws * wi;
    K += w;
    P0 += nux * d * w; P1 += nuy * d * w; P2 += nuz * d * w;
  }
  float rx = pvx, ry = pvy, rz = pvz;
  if (K > 0.01) { rx -= P0 / K; ry -= P1 / K; rz -= P2 / K; }
  ox[v] = rx; oy[v] = ry; oz[v] = rz;
  (void)nvx; (void)nvy; (void)nvz;
}

// computeSignedCurvatures (graph/impl/common.hpp:105-132)
__global__ void k_signed_curvature(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                                   const uint32_t* __restrict__ eid, const uint32_t* __restrict__ esrc,
                                   const float* __restrict__ x, const float* __restrict__ y,
                                   const float* __restrict__ z, const float* __restrict__ nx,
                                   const float* __restrict__ ny, const float* __restrict__ nz,
                                   const float* __restrict__ curv, float* __restrict__ ocurv) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  const float pvx = x[v], pvy = y[v], pvz = z[v];
  const float nvx = nx[v], nvy = ny[v], nvz = nz[v];
  float conv = 0.0f;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    bool v_is_source = esrc[eid[a]] == (uint32_t)v;
    float pux = x[u], puy = y[u], puz = z[u];
    float nux = nx[u], nuy = ny[u], nuz = nz[u];
    // edge (v1 = source, v2 = target): d = p2 - p1
    float dx, dy, dz, n1x, n1y, n1z, n2x, n2y, n2z;
    if (v_is_source) {
      dx = pux - pvx; dy = puy - pvy; dz = puz - pvz;
      n1x = nvx; n1y = nvy; n1z = nvz; n2x = nux; n2y = nuy; n2z = nuz;
    } else {
      dx = pvx - pux; dy = pvy - puy; dz = pvz - puz;
      n1x = nux; n1y = nuy; n1z = nuz; n2x = nvx; n2y = nvy; n2z = nvz;
    }
    float dn = dot3(dx, dy, dz, n1x, n1y, n1z);
    float rx = dx - dn * n1x, ry = dy - dn * n1y, rz = dz - dn * n1z;
    float sgn = dot3(rx, ry, rz, n2x, n2y, n2z) > 0 ? 1.0f : -1.0f;
    conv += sgn * sqn3(n1x - n2x, n1y - n2y, n1z - n2z);
  }
  ocurv[v] = copysignf(curv[v], conv);
}

}  // namespace

extern "C" int tcs_smoothen(tcs_graph* g, float spatial_sigma, float influence_sigma) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V;
  if (V == 0) return TCS_OK;
  CK(cudaSetDevice(ctx->device));
  DBuf<float> ox, oy, oz;
  TRY(ox.alloc(ctx, V)); TRY(oy.alloc(ctx, V)); TRY(oz.alloc(ctx, V));
  LAUNCH(ctx, k_smoothen, div_up(V, 128), 128, 0, V, g->rowptr.p, g->col.p, g->eid.p, g->esrc.p, g->x.p, g->y.p,
         g->z.p, g->nx.p, g->ny.p, g->nz.p, spatial_sigma, influence_sigma, ox.p, oy.p, oz.p);
  CK(cudaGetLastError());
  g->x.swap(ox); g->y.swap(oy); g->z.swap(oz);
  return TCS_OK;
}

extern "C" int tcs_signed_curvatures(tcs_graph* g) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V;
  if (V == 0) return TCS_OK;
  CK(cudaSetDevice(ctx->device));
  DBuf<float> oc;
  TRY(oc.alloc(ctx, V));
  LAUNCH(ctx, k_signed_curvature, div_up(V, 128), 128, 0, V, g->rowptr.p, g->col.p, g->eid.p, g->esrc.p, g->x.p,
         g->y.p, g->z.p, g->nx.p, g->ny.p, g->nz.p, g->curv.p, oc.p);
  CK(cudaGetLastError());
  g->curv.swap(oc);
  return TCS_OK;
}
