// tcs_b200/csrc/weights.cu - edge weights on the device.
// Replaces pcl::graph::EdgeWeightComputer::compute (graph/impl/edge_weight_computer.hpp:110-188) with the terms
// of graph/edge_weight_computer_terms.h:55-140 and the normalisation classes of
// graph/edge_weight_computer.h:333-469.  Compiled with -fmad=false.
#include <cfloat>

#include "common.cuh"

namespace {

#define MAX_TERMS 8

struct TermSet {
  int n;                   // terms in evaluation order: plain..., global..., local... (impl:148-154)
  int kind[MAX_TERMS];
  float influence[MAX_TERMS];
  float cmult[MAX_TERMS];
  int norm[MAX_TERMS];
  int slot[MAX_TERMS];     // index of the raw-distance array for normalised terms
  int nslots;
  int balancing;           // TCS_BALANCE_*: how a (normalised) distance becomes a factor of the weight
};

struct VertexView {
  const float *x, *y, *z, *nx, *ny, *nz, *curv;
  const uint32_t* rgba;
};

__device__ __forceinline__ float sqn3(float a0, float a1, float a2) { return a0 * a0 + (a1 * a1 + a2 * a2); }

// graph/edge_weight_computer_terms.h:55-140; a = source, b = target
__device__ float term_distance(const VertexView& g, int kind, uint32_t a, uint32_t b) {
  switch (kind) {
    case TCS_TERM_XYZ:
      return sqn3(g.x[b] - g.x[a], g.y[b] - g.y[a], g.z[b] - g.z[a]);
    case TCS_TERM_NORMAL:
      return 0.5f * sqn3(g.nx[a] - g.nx[b], g.ny[a] - g.ny[b], g.nz[a] - g.nz[b]);
    case TCS_TERM_CURVATURE:
      return fabsf(g.curv[a]) * fabsf(g.curv[b]);
    case TCS_TERM_RGB: {
      uint32_t ca = g.rgba[a], cb = g.rgba[b];
      float d0 = (float)(ca & 0xFF) - (float)(cb & 0xFF);
      float d1 = (float)((ca >> 8) & 0xFF) - (float)((cb >> 8) & 0xFF);
      float d2 = (float)((ca >> 16) & 0xFF) - (float)((cb >> 16) & 0xFF);
      return sqrtf(sqn3(d0, d1, d2)) / 255.0f;
    }
  }
  return 0.f;
}

// EdgeWeightComputer::gaussian (edge_weight_computer.h:333-337); std::exp(float) := float64 exp rounded once
__device__ __forceinline__ float gaussian(float val, float influence) {
  return influence > 0.0f ? (float)exp((double)(-val / influence)) : 1.0f;
}
// EdgeWeightComputer::setTermBalancingFunction (edge_weight_computer.h:325): the reference takes any
// boost::function<float (float, float)>; a host callable cannot run here, so the device offers a closed set
// (TCS_BALANCE_*), the reference's default Gaussian first.  Every function is 1 for influence <= 0 like the default.
__device__ __forceinline__ float balance(int kind, float val, float influence) {
  if (kind == TCS_BALANCE_GAUSSIAN) return gaussian(val, influence);
  if (!(influence > 0.0f)) return 1.0f;
  const float q = val / influence;
  if (kind == TCS_BALANCE_INVERSE) return 1.0f / (1.0f + q);
  if (kind == TCS_BALANCE_GAUSSIAN_SQUARED) return (float)exp((double)(-(q * q)));
  const float l = 1.0f - q;   // TCS_BALANCE_LINEAR
  return l > 0.0f ? l : 0.0f;
}

// round 1: raw distances of the normalised terms, raw[slot * E + e]
__global__ void k_edge_raw(size_t E, const uint32_t* __restrict__ esrc, const uint32_t* __restrict__ etgt,
                           VertexView g, TermSet ts, float* __restrict__ raw) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= E) return;
  uint32_t v1 = esrc[e], v2 = etgt[e];
  for (int t = 0; t < ts.n; ++t)
    if (ts.norm[t] != TCS_NORM_NONE) raw[(size_t)ts.slot[t] * E + e] = term_distance(g, ts.kind[t], v1, v2);
}

// GloballyNormalizedTerm: total_weight_ is one float accumulator over the edges in order (PARITY) - a dependent chain
// of float additions, whose floor is the FADD latency (~4 cycles per edge).  The chain must not wait for memory: all 256
// threads stream the values into shared memory one chunk ahead (coalesced, double-buffered) and thread 0 adds the
// current chunk from shared memory with 16-byte loads that do not depend on the sum.  (One warp loading 32 values per
// step and adding them through shuffles spent 15 ns per edge waiting for its own global load.)  What is left is the chain itself:
// 289 k edges take 0.88 ms whether the adds are FADD or FFMA(v, 1, total) - four cycles each at the ~1.3 GHz a
// single-CTA kernel gets right after an idle period.
constexpr int kSumChunk = 4096;
__global__ void __launch_bounds__(256) k_sum_sequential(const float* __restrict__ d, size_t E, float* __restrict__ out) {
  __shared__ __align__(16) float buf[2][kSumChunk];
  constexpr int kPer = kSumChunk / 256;
  float total = 0.0f;
  for (int j = 0; j < kPer; ++j) {
    const size_t e = (size_t)j * 256 + threadIdx.x;
    buf[0][j * 256 + threadIdx.x] = e < E ? d[e] : 0.0f;
  }
  __syncthreads();
  const size_t chunks = (E + kSumChunk - 1) / kSumChunk;
  for (size_t c = 0; c < chunks; ++c) {
    float r[kPer];
    const size_t next = (c + 1) * kSumChunk;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {   // next chunk -> registers (in flight while thread 0 adds)
      const size_t e = next + (size_t)j * 256 + threadIdx.x;
      r[j] = e < E ? d[e] : 0.0f;
    }
    if (threadIdx.x == 0) {
      const float* b = buf[c & 1];
      const int cnt = (int)min((size_t)kSumChunk, E - c * kSumChunk);
      int k = 0;
      if (cnt >= 16) {   // sixteen values ahead in registers: the shared-memory latency stays off the chain
        const float4* b4 = (const float4*)b;
        float4 a0 = b4[0], a1 = b4[1], a2 = b4[2], a3 = b4[3];
        for (; k + 32 <= cnt; k += 16) {
          const float4 n0 = b4[k / 4 + 4], n1 = b4[k / 4 + 5], n2 = b4[k / 4 + 6], n3 = b4[k / 4 + 7];
          total += a0.x; total += a0.y; total += a0.z; total += a0.w;
          total += a1.x; total += a1.y; total += a1.z; total += a1.w;
          total += a2.x; total += a2.y; total += a2.z; total += a2.w;
          total += a3.x; total += a3.y; total += a3.z; total += a3.w;
          a0 = n0; a1 = n1; a2 = n2; a3 = n3;
        }
        total += a0.x; total += a0.y; total += a0.z; total += a0.w;
        total += a1.x; total += a1.y; total += a1.z; total += a1.w;
        total += a2.x; total += a2.y; total += a2.z; total += a2.w;
        total += a3.x; total += a3.y; total += a3.z; total += a3.w;
        k += 16;
      }
      for (; k < cnt; ++k) total += b[k];
    }
#pragma unroll
    for (int j = 0; j < kPer; ++j) buf[(c + 1) & 1][j * 256 + threadIdx.x] = r[j];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = total;
}

// FAST: float64 accumulation in parallel, rounded once
__global__ void k_sum_f64(const float* __restrict__ d, size_t E, double* __restrict__ out) {
  double s = 0.0;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < E; e += (size_t)gridDim.x * blockDim.x)
    s += (double)d[e];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
  if ((threadIdx.x & 31) == 0 && s != 0.0) atomicAdd(out, s);
}

__global__ void k_finish_avg(const float* sum32, const double* sum64, int mode, size_t E, float* avg) {
  float tot = mode == TCS_MODE_PARITY ? *sum32 : (float)*sum64;
  *avg = tot / (float)E;
}

// LocallyNormalizedTerm::extract: per-vertex mean of the raw distances of its incident edges (edge order)
__global__ void k_local_mean(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ eid,
                             const float* __restrict__ d, float* __restrict__ mean) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t a0 = rowptr[v], a1 = rowptr[v + 1];
  float s = 0.0f;
  for (uint32_t a = a0; a < a1; ++a) s += d[eid[a]];
  if (a1 > a0) s /= (float)(a1 - a0);
  mean[v] = s;
}

// round 2 + small-weight coercion
__global__ void k_edge_weight(size_t E, const uint32_t* __restrict__ esrc, const uint32_t* __restrict__ etgt,
                              VertexView g, TermSet ts, const float* __restrict__ raw,
                              const float* __restrict__ gavg, const float* __restrict__ lmean, size_t V,
                              float threshold, int coerce, float* __restrict__ ew) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= E) return;
  uint32_t v1 = esrc[e], v2 = etgt[e];
  bool convex = g.curv[v1] > 0.0f && g.curv[v2] > 0.0f;  // detail::IsConvex (impl:95-102)
  float a = 1.0f;
  for (int t = 0; t < ts.n; ++t) {
    float infl = ts.influence[t] * (convex ? ts.cmult[t] : 1.0f);
    if (ts.norm[t] == TCS_NORM_NONE) {
      a *= balance(ts.balancing, term_distance(g, ts.kind[t], v1, v2), infl);
    } else if (ts.norm[t] == TCS_NORM_GLOBAL) {
      a *= balance(ts.balancing, raw[(size_t)ts.slot[t] * E + e] / gavg[ts.slot[t]], infl);
    } else {
      const float* m = lmean + (size_t)ts.slot[t] * V;
      float n = (m[v1] + m[v2]) / 2.0f;
      float w = (n > 0.0f && ts.influence[t] > 0.0f) ? raw[(size_t)ts.slot[t] * E + e] / n : 0.0f;
      a *= balance(ts.balancing, w, infl);
    }
  }
  if (coerce && a < threshold) a = threshold;
  ew[e] = a;
}

// SMALL_WEIGHT_REMOVE_EDGE: keep[e] = !(w < threshold)
__global__ void k_keep_flags(const float* __restrict__ ew, size_t E, float threshold, uint32_t* __restrict__ keep) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e < E) keep[e] = !(ew[e] < threshold) ? 1u : 0u;
}
__global__ void k_compact_edges(size_t E, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos,
                                const uint32_t* __restrict__ s, const uint32_t* __restrict__ t,
                                const float* __restrict__ w, uint32_t* __restrict__ s2, uint32_t* __restrict__ t2,
                                float* __restrict__ w2) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= E || !keep[e]) return;
  uint32_t p = pos[e];
  s2[p] = s[e]; t2[p] = t[e]; w2[p] = w[e];
}

}  // namespace

extern "C" int tcs_edge_weights(tcs_graph* g, const tcs_term* terms, int nterms, float threshold, int policy,
                                int mode) {
  return tcs_edge_weights_ex(g, terms, nterms, threshold, policy, mode, TCS_BALANCE_GAUSSIAN);
}

extern "C" int tcs_edge_weights_ex(tcs_graph* g, const tcs_term* terms, int nterms, float threshold, int policy,
                                   int mode, int balancing) {
  if (!g || (nterms && !terms) || nterms < 0) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (balancing < TCS_BALANCE_GAUSSIAN || balancing > TCS_BALANCE_LINEAR)
    return tcs_fail(ctx, TCS_ERR_INPUT, "unknown term balancing function %d", balancing);
  if (nterms > MAX_TERMS) return tcs_fail(ctx, TCS_ERR_INPUT, "at most %d edge-weight terms", MAX_TERMS);
  CK(cudaSetDevice(ctx->device));
  const size_t E = g->E, V = g->V;
  if (E == 0) return TCS_OK;
  TermSet ts;
  memset(&ts, 0, sizeof ts);
  ts.balancing = balancing;
  for (int pass = 0; pass < 3; ++pass)
    for (int i = 0; i < nterms; ++i) {
      if (terms[i].kind < 0 || terms[i].kind > 3 || terms[i].normalization < 0 || terms[i].normalization > 2)
        return tcs_fail(ctx, TCS_ERR_INPUT, "bad term %d", i);
      if (terms[i].kind == TCS_TERM_RGB && !g->has_color) continue;  // addTermImpl(false): point type has no colour
      if (terms[i].normalization != pass) continue;
      int t = ts.n++;
      ts.kind[t] = terms[i].kind;
      ts.influence[t] = terms[i].influence;
      ts.cmult[t] = terms[i].convex_mult;
      ts.norm[t] = pass;
      ts.slot[t] = pass == TCS_NORM_NONE ? -1 : ts.nslots++;
    }
  VertexView vv = {g->x.p, g->y.p, g->z.p, g->nx.p, g->ny.p, g->nz.p, g->curv.p, g->rgba.p};
  DBuf<float> raw, gavg, lmean, s32;
  DBuf<double> s64;
  if (ts.nslots) {
    TRY(raw.alloc(ctx, (size_t)ts.nslots * E));
    TRY(gavg.alloc(ctx, MAX_TERMS)); TRY(s32.alloc(ctx, 1)); TRY(s64.alloc(ctx, 1));
    TRY(lmean.alloc(ctx, (size_t)ts.nslots * V));
    LAUNCH(ctx, k_edge_raw, div_up(E, 256), 256, 0, E, g->esrc.p, g->etgt.p, vv, ts, raw.p);
    for (int t = 0; t < ts.n; ++t) {
      const float* d = raw.p + (size_t)ts.slot[t] * E;
      if (ts.norm[t] == TCS_NORM_GLOBAL) {
        if (mode == TCS_MODE_PARITY) {
          LAUNCH(ctx, k_sum_sequential, 1, 256, 0, d, E, s32.p);
        } else {
          CK(cudaMemsetAsync(s64.p, 0, 8, ctx->stream));
          LAUNCH(ctx, k_sum_f64, std::min<unsigned>(div_up(E, 256), ctx->sm_count * 4), 256, 0, d, E, s64.p);
        }
        LAUNCH(ctx, k_finish_avg, 1, 1, 0, s32.p, s64.p, mode, E, gavg.p + ts.slot[t]);
      } else if (ts.norm[t] == TCS_NORM_LOCAL) {
        LAUNCH(ctx, k_local_mean, div_up(V, 128), 128, 0, V, g->rowptr.p, g->eid.p, d, lmean.p + (size_t)ts.slot[t] * V);
      }
    }
  }
  LAUNCH(ctx, k_edge_weight, div_up(E, 256), 256, 0, E, g->esrc.p, g->etgt.p, vv, ts, raw.p, gavg.p, lmean.p, V,
         threshold, policy == TCS_SMALL_COERCE ? 1 : 0, g->ew.p);
  CK(cudaGetLastError());
  if (policy == TCS_SMALL_REMOVE) {
    // remove_edge_if: surviving edges keep their relative order in the edge list and in every out-edge list
    DBuf<uint32_t> keep, pos;
    TRY(keep.alloc(ctx, E + 1)); TRY(pos.alloc(ctx, E + 1));
    LAUNCH(ctx, k_keep_flags, div_up(E, 256), 256, 0, g->ew.p, E, threshold, keep.p);
    TRY(tcs_impl_scan_u32(ctx, keep.p, pos.p, E));
    uint32_t E2 = 0;
    TRY(tcs_impl_read_u32(ctx, pos.p + E, &E2));
    DBuf<uint32_t> s2, t2;
    DBuf<float> w2;
    TRY(s2.alloc(ctx, E2)); TRY(t2.alloc(ctx, E2)); TRY(w2.alloc(ctx, E2));
    LAUNCH(ctx, k_compact_edges, div_up(E, 256), 256, 0, E, keep.p, pos.p, g->esrc.p, g->etgt.p, g->ew.p, s2.p, t2.p,
           w2.p);
    g->esrc.swap(s2); g->etgt.swap(t2); g->ew.swap(w2);
    g->E = E2;
    TRY(tcs_impl_build_out_edges(g));
    g->ncomp = -1;
  }
  return TCS_OK;
}
