// tcs_b200/csrc/amg.cuh - multilevel preconditioner of the random-walker PCG (included by pcg.cu).
//
// Jacobi-PCG needs thousands of iterations on these systems (edge weights span five orders of magnitude: a scene is a
// set of almost disconnected membranes).  This file builds an aggregation hierarchy of the graph Laplacian ON THE
// DEVICE and applies it as   z = D^-1 r + Z0 V(Z0^T r):   Jacobi on the fine level plus one V(1,1) cycle on the coarse
// levels (additive on top, so the fine level keeps its single SpMM per iteration; multiplicative below, where the
// operators are small).  Symmetric positive definite, so plain PCG stays valid; what is accepted as converged is
// still decided by the float64 true residual of the fine system - the hierarchy only changes how fast we get there
// (640x480 scene: 2200 -> ~35 iterations; C4: 8050 Jacobi / 1850 two-level -> see profiles/r02_*).
//
//   coarsening   : pairwise heavy-edge matching by handshaking (every unmatched row picks its strongest unmatched
//                  neighbour, s_ij = w_ij^2 / (d_i d_j); mutual picks pair up; kMatchRounds rounds), then rows left
//                  over join the aggregate of their strongest matched neighbour.  Two such passes per level
//                  (aggregates of ~4-5 rows, operator complexity ~1.2).  Aggregates are numbered by their smallest
//                  row, so coarse rows keep the Morton locality of the fine ones.
//   Galerkin     : A_c = Z^T A Z for piecewise-constant Z: cross-aggregate entries get the key (agg_i, agg_j), a
//                  stable radix sort groups them, one thread sums a run in float64.  The diagonal is NOT differenced:
//                  every level carries g = diagonal excess (the coupling to the seeds) and d = sum |offdiag| + g, so
//                  all levels are weakly diagonally dominant M-matrices by construction (no cancellation).
//   cycle        : pre-smooth x = w D^-1 r, residual, restrict (children gather: deterministic), recurse,
//                  x += over * Z e (over-correction 1.5: piecewise-constant interpolation underestimates the
//                  correction), post-smooth.  Rows without off-diagonal entries (collapsed components) use w = 1:
//                  they are solved exactly.
//   bottom       : at most kBottomMax rows: dense float64 inverse by Gauss-Jordan in the shared memory of one CTA.
#pragma once

#include <cstdlib>
#include <vector>

#include "common.cuh"

namespace amg {

constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr int kBottomMax = 160;   // 160 x 160 float64 = 200 KB of shared memory
constexpr int kMaxLevels = 24;

static inline double env_double(const char* name, double dflt) { const char* v = getenv(name); return v ? atof(v) : dflt; }
static inline int env_i(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }

// CSR view of one operator: off-diagonal entries (negative values) + diagonal
struct Csr {
  int n = 0;
  long long nnz = 0;
  const uint32_t* rowptr = nullptr;
  const int* col = nullptr;
  const float* val = nullptr;
  const float* diag = nullptr;
};

struct Level {
  Csr A;                         // level 0: borrowed from the assembled system
  DBuf<uint32_t> rowptr;         // owned storage of levels >= 1
  DBuf<int> col;
  DBuf<float> val, diag;
  DBuf<float> winvd;             // smoother weight / diagonal (levels >= 1)
  DBuf<float> gdiag;             // diagonal excess in float32: the cycle's SpMM runs in difference form (levels >= 1)
  DBuf<double> g;                // diagonal excess (set-up only)
  // transfer to the next level
  int nc = 0;
  DBuf<uint32_t> agg;            // n: row of the next level
  DBuf<uint32_t> cptr, cidx;     // children of every row of the next level (nc + 1, n)
  // cycle vectors at the full working width (levels >= 1): right-hand side, iterate, scratch / result
  DBuf<float> r, x, t;
};

struct Hierarchy {
  bool on = false;
  int nlev = 0;                  // levels incl. the fine one; level nlev - 1 is the bottom
  Level L[kMaxLevels];
  bool dense = false;            // bottom solved by the dense inverse (else: by its smoother alone)
  int m = 0;                     // bottom rows
  DBuf<double> Einv;             // m x m float64 (the bottom operator is as ill-conditioned as the whole system)
  float over = 1.5f;
  float setup_ms = 0.f;
  long long launches_per_apply = 0;
};

// ---------------------------------------------------------------------------------------------------- set-up kernels
__global__ void k_excess(int n, int keep_sign, const uint32_t* __restrict__ rowptr, const float* __restrict__ val,
                         const float* __restrict__ diag, double* __restrict__ g) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = (double)diag[i];
  for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) s += (double)val[a];
  g[i] = (s > 0.0 || keep_sign) ? s : 0.0;
}

__global__ void k_fill_u32(size_t n, uint32_t v, uint32_t* __restrict__ p) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// strongest unmatched neighbour of every unmatched row: s_ij = w_ij^2 / (d_i d_j), d_i common to the row
__global__ void k_match_pick(int n, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                             const float* __restrict__ val, const float* __restrict__ diag,
                             const uint32_t* __restrict__ matched, uint32_t* __restrict__ best) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t bj = kNone;
  if (matched[i] == kNone) {
    float bs = 0.f;
    for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) {
      const uint32_t j = (uint32_t)col[a];
      if (matched[j] != kNone) continue;
      const float w = val[a], dj = diag[j];
      if (!(dj > 0.f)) continue;
      const float s = w * w / dj;
      if (s > bs || (s == bs && s > 0.f && j < bj)) { bs = s; bj = j; }
    }
  }
  best[i] = bj;
}

__global__ void k_match_commit(int n, const uint32_t* __restrict__ best, uint32_t* __restrict__ matched) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t b = best[i];
  if (b != kNone && best[b] == (uint32_t)i) matched[i] = b;
}

// rep[i] = smallest row of i's aggregate: a pair's smaller row, or - for a row left unmatched - the pair of its
// strongest matched neighbour; head[i] = 1 for the aggregate's representative
__global__ void k_aggregate(int n, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                            const float* __restrict__ val, const float* __restrict__ diag,
                            const uint32_t* __restrict__ matched, uint32_t* __restrict__ rep,
                            uint32_t* __restrict__ head) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t r = (uint32_t)i;
  const uint32_t mi = matched[i];
  if (mi != kNone) r = min(r, mi);
  else {
    float bs = 0.f;
    uint32_t bj = kNone;
    for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) {
      const uint32_t j = (uint32_t)col[a];
      if (matched[j] == kNone) continue;
      const float w = val[a], dj = diag[j];
      if (!(dj > 0.f)) continue;
      const float s = w * w / dj;
      if (s > bs || (s == bs && s > 0.f && j < bj)) { bs = s; bj = j; }
    }
    if (bj != kNone) r = min(bj, matched[bj]);
  }
  rep[i] = r;
  head[i] = r == (uint32_t)i ? 1u : 0u;
}

__global__ void k_agg_ids(int n, const uint32_t* __restrict__ rep, const uint32_t* __restrict__ id,
                          uint32_t* __restrict__ agg) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) agg[i] = id[rep[i]];
}

__global__ void k_cross_count(int n, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                              const uint32_t* __restrict__ agg, uint32_t* __restrict__ cnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ai = agg[i];
  uint32_t c = 0;
  for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) c += agg[col[a]] != ai ? 1u : 0u;
  cnt[i] = c;
}

__global__ void k_cross_emit(int n, int shift, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                             const float* __restrict__ val, const uint32_t* __restrict__ agg,
                             const uint32_t* __restrict__ pos, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ai = agg[i];
  uint32_t p = pos[i];
  for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) {
    const uint32_t aj = agg[col[a]];
    if (aj != ai) {
      keys[p] = ((uint64_t)ai << shift) | (uint64_t)aj;
      vals[p] = __float_as_uint(val[a]);
      ++p;
    }
  }
}

__global__ void k_run_heads(size_t total, const uint64_t* __restrict__ keys, uint32_t* __restrict__ head) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e < total) head[e] = (e == 0 || keys[e] != keys[e - 1]) ? 1u : 0u;
}

// one thread per run of equal keys: float64 sum in sorted (= stable, CSR) order
__global__ void k_run_reduce(size_t total, int shift, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                             const uint32_t* __restrict__ head, const uint32_t* __restrict__ uid, int* __restrict__ ccol,
                             float* __restrict__ cval, uint32_t* __restrict__ crow) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= total || !head[e]) return;
  double s = 0.0;
  size_t f = e;
  do { s += (double)__uint_as_float(vals[f]); ++f; } while (f < total && !head[f]);
  const uint32_t u = uid[e];
  const uint64_t k = keys[e];
  ccol[u] = (int)(k & ((1ull << shift) - 1ull));
  cval[u] = (float)s;
  crow[u] = (uint32_t)(k >> shift);
}

// ptr[r] = first u with row[u] >= r, for r = 0..m (row[] ascending, cnt entries)
__global__ void k_ptr_from_sorted_rows(size_t cnt, int m, const uint32_t* __restrict__ row, uint32_t* __restrict__ ptr) {
  size_t u = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (u > cnt) return;
  const int lo = u == 0 ? 0 : (int)row[u - 1] + 1;
  const int hi = u == cnt ? m : (int)row[u];
  for (int r = lo; r <= hi; ++r) ptr[r] = (uint32_t)u;
}

__global__ void k_sum_excess(int n, const uint32_t* __restrict__ agg, const double* __restrict__ g, double* __restrict__ gc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicAdd(&gc[agg[i]], g[i]);
}

// d = sum |offdiag| + g ; smoother weight w / d (1 / d for rows without off-diagonal entries, 0 for empty rows)
__global__ void k_level_diag(int m, float w, const uint32_t* __restrict__ rowptr, const float* __restrict__ val,
                             const double* __restrict__ g, float* __restrict__ diag, float* __restrict__ winvd,
                             float* __restrict__ gdiag) {
  int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  double s = g[a];
  gdiag[a] = (float)s;
  const uint32_t e0 = rowptr[a], e1 = rowptr[a + 1];
  for (uint32_t e = e0; e < e1; ++e) s -= (double)val[e];
  diag[a] = (float)s;
  winvd[a] = s > 1e-30 ? (float)((e1 > e0 ? (double)w : 1.0) / s) : 0.f;
}

__global__ void k_iota(int n, uint32_t* __restrict__ p) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (uint32_t)i;
}

__global__ void k_compose(int n, const uint32_t* __restrict__ a1, const uint32_t* __restrict__ a2, uint32_t* __restrict__ out,
                          uint32_t* __restrict__ idx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { out[i] = a2[a1[i]]; idx[i] = (uint32_t)i; }
}

// ---------------------------------------------------------------------------------------------------- bottom level
__global__ void k_dense_build(int m, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                              const float* __restrict__ val, const double* __restrict__ g, double* __restrict__ E) {
  int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= m) return;
  for (int b = 0; b < m; ++b) E[(size_t)a * m + b] = 0.0;
  double d = g[a];   // diagonal = excess + sum |offdiag| in float64: the near-null modes survive
  for (uint32_t e = rowptr[a]; e < rowptr[a + 1]; ++e) d -= (double)val[e];
  // an empty row (seedless component collapsed to one node) gets a unit diagonal: its right-hand side is zero
  E[(size_t)a * m + a] = d > 1e-30 ? d * (1.0 + 1e-9) : 1.0;
  for (uint32_t e = rowptr[a]; e < rowptr[a + 1]; ++e) E[(size_t)a * m + col[e]] += (double)val[e];
}

// in-place Gauss-Jordan inverse of a symmetric positive definite m x m matrix in shared memory (one CTA)
__global__ void __launch_bounds__(1024) k_dense_invert(int m, const double* __restrict__ E, double* __restrict__ Einv) {
  extern __shared__ double sE[];
  __shared__ double s_piv;
  double* colk = sE + (size_t)m * m;  // copy of the pivot column
  for (int t = threadIdx.x; t < m * m; t += blockDim.x) sE[t] = E[t];
  __syncthreads();
  for (int k = 0; k < m; ++k) {
    if (threadIdx.x == 0) s_piv = 1.0 / sE[(size_t)k * m + k];
    for (int i = threadIdx.x; i < m; i += blockDim.x) colk[i] = sE[(size_t)i * m + k];
    __syncthreads();
    const double p = s_piv;
    // row k scaled by p first (each thread reads the unscaled pivot row through rowk = E[k][j])
    for (int t = threadIdx.x; t < m * m; t += blockDim.x) {
      const int i = t / m, j = t - i * m;
      if (i == k) continue;
      const double f = colk[i];
      if (j == k) sE[t] = -f * p;
      else sE[t] -= f * sE[(size_t)k * m + j] * p;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < m; j += blockDim.x) sE[(size_t)k * m + j] = j == k ? p : sE[(size_t)k * m + j] * p;
    __syncthreads();
  }
  for (int t = threadIdx.x; t < m * m; t += blockDim.x) Einv[t] = sE[t];
}

// e = E^-1 r on the bottom level (m <= kBottomMax).  One CTA per row a: thread (part, q) sums its share of the b range
// in float64, the shares are added in a fixed order (deterministic).  (One thread per (a, q) with a serial loop over b
// took 46 us for m = 126: four CTAs of dependent loads.)
__global__ void __launch_bounds__(128) k_dense_apply(int m, int Q, const double* __restrict__ Einv, const float4* __restrict__ r,
                                                     float4* __restrict__ e) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  __shared__ double sm[128 * 4];
  const int a = blockIdx.x;
  const int parts = blockDim.x / Q;
  const int part = threadIdx.x / Q, q = threadIdx.x - part * Q;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  if (part < parts)
    for (int b = part; b < m; b += parts) {
      const double f = Einv[(size_t)a * m + b];
      const float4 v = r[(size_t)b * Q + q];
      s0 += f * v.x; s1 += f * v.y; s2 += f * v.z; s3 += f * v.w;
    }
  sm[threadIdx.x * 4 + 0] = s0; sm[threadIdx.x * 4 + 1] = s1; sm[threadIdx.x * 4 + 2] = s2; sm[threadIdx.x * 4 + 3] = s3;
  __syncthreads();
  if (threadIdx.x < Q) {
    double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
    for (int p = 0; p < parts; ++p) {
      const double* v = sm + (size_t)(p * Q + q) * 4;
      t0 += v[0]; t1 += v[1]; t2 += v[2]; t3 += v[3];
    }
    e[(size_t)a * Q + q] = make_float4((float)t0, (float)t1, (float)t2, (float)t3);
  }
}

// ---------------------------------------------------------------------------------------------------- cycle kernels
// restriction (children gather, fixed order) fused with the pre-smoothing of the coarse level:
//   rc[a] = sum_{i in a} rf[i] ;  xc[a] = winvd[a] * rc[a]
__global__ void __launch_bounds__(256) k_restrict(int nc, int Q, const uint32_t* __restrict__ cptr,
                                                  const uint32_t* __restrict__ cidx, const float4* __restrict__ rf,
                                                  const float* __restrict__ winvd, float4* __restrict__ rc,
                                                  float4* __restrict__ xc) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)nc * Q) return;
  const int a = (int)(t / Q), q = (int)(t - (size_t)a * Q);
  const uint32_t c0 = cptr[a], c1 = cptr[a + 1];
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  uint32_t c = c0;
  for (; c + 4 <= c1; c += 4) {  // four children in flight
    const uint32_t i0 = cidx[c], i1 = cidx[c + 1], i2 = cidx[c + 2], i3 = cidx[c + 3];
    const float4 v0 = rf[(size_t)i0 * Q + q], v1 = rf[(size_t)i1 * Q + q], v2 = rf[(size_t)i2 * Q + q], v3 = rf[(size_t)i3 * Q + q];
    s.x += v0.x; s.y += v0.y; s.z += v0.z; s.w += v0.w;
    s.x += v1.x; s.y += v1.y; s.z += v1.z; s.w += v1.w;
    s.x += v2.x; s.y += v2.y; s.z += v2.z; s.w += v2.w;
    s.x += v3.x; s.y += v3.y; s.z += v3.z; s.w += v3.w;
  }
  for (; c < c1; ++c) {
    const float4 v = rf[(size_t)cidx[c] * Q + q];
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  rc[t] = s;
  if (xc) {
    const float wd = winvd[a];
    xc[t] = make_float4(wd * s.x, wd * s.y, wd * s.z, wd * s.w);
  }
}

// y = A x in difference form (g_i x_i + sum_j val_ij (x_j - x_i), val = -w: the corrections are dominated by the
// near-null modes - constants on weakly coupled segments, |x| ~ |r| / lambda_min - and d_i x_i - sum w_ij x_j would
// drown the residual of exactly those modes in float32 cancellation noise), then
//   SMOOTH = 0: out = rhs - y                       (residual of the pre-smoothed iterate)
//   SMOOTH = 1: out = x + winvd * (rhs - y)         (post-smoothing; DOT: dots[c] += rhs . out per column, float64)
template <int SMOOTH, int DOT>
__global__ void __launch_bounds__(256, 6) k_level_spmm(int n, int Q, const uint32_t* __restrict__ rowptr,
                                                    const int* __restrict__ col, const float* __restrict__ val,
                                                    const float* __restrict__ gdiag, const float* __restrict__ winvd,
                                                    const float4* __restrict__ X, const float4* __restrict__ RHS,
                                                    float4* __restrict__ OUT, double* __restrict__ dots) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  if (active) {
    for (int row = blockIdx.x * rows_per_block + local; row < n; row += gridDim.x * rows_per_block) {
      uint32_t a = rowptr[row];
      const uint32_t a1 = rowptr[row + 1];
      const size_t i = (size_t)row * Q + q;
      const float4 x = X[i];
      const float4 b = RHS[i];
      const float d = gdiag[row];
      const F4p xp = f4_pack(x);
      F4p yy = f4_pack(make_float4(d * x.x, d * x.y, d * x.z, d * x.w));
      for (; a + 4 <= a1; a += 4) {
        const int c0 = col[a], c1 = col[a + 1], c2 = col[a + 2], c3 = col[a + 3];
        const float w0 = val[a], w1 = val[a + 1], w2 = val[a + 2], w3 = val[a + 3];
        const float4 t0 = X[(size_t)c0 * Q + q], t1 = X[(size_t)c1 * Q + q];
        const float4 t2 = X[(size_t)c2 * Q + q], t3 = X[(size_t)c3 * Q + q];
        f4_diff_fma(yy, w0, t0, xp);
        f4_diff_fma(yy, w1, t1, xp);
        f4_diff_fma(yy, w2, t2, xp);
        f4_diff_fma(yy, w3, t3, xp);
      }
      for (; a < a1; ++a) {
        const int c = col[a];
        const float w = val[a];
        const float4 t = X[(size_t)c * Q + q];
        f4_diff_fma(yy, w, t, xp);
      }
      const float4 y = f4_unpack(yy);
      float4 o = make_float4(b.x - y.x, b.y - y.y, b.z - y.z, b.w - y.w);
      if (SMOOTH) {
        const float wd = winvd[row];
        o = make_float4(x.x + wd * o.x, x.y + wd * o.y, x.z + wd * o.z, x.w + wd * o.w);
      }
      OUT[i] = o;
      if (DOT) {
        acc[0] += (double)b.x * (double)o.x; acc[1] += (double)b.y * (double)o.y;
        acc[2] += (double)b.z * (double)o.z; acc[3] += (double)b.w * (double)o.w;
      }
    }
  }
  if (DOT) reduce_columns(acc, q, Q, active, sm, dots, true);
}

// x[i] += over * e[agg[i]]
__global__ void __launch_bounds__(256) k_prolong(int n, int Q, float over, const uint32_t* __restrict__ agg,
                                                 const float4* __restrict__ e, float4* __restrict__ x) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)n * Q) return;
  const int i = (int)(t / Q), q = (int)(t - (size_t)i * Q);
  const float4 c = e[(size_t)agg[i] * Q + q];
  float4 v = x[t];
  v.x += over * c.x; v.y += over * c.y; v.z += over * c.z; v.w += over * c.w;
  x[t] = v;
}

// dots[c] += a . b per column (float64): the coarse part of r.z when level 1 is the dense bottom
__global__ void __launch_bounds__(256) k_dot(int n, int Q, const float4* __restrict__ a, const float4* __restrict__ b,
                                             double* __restrict__ dots) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  if (active)
    for (int row = blockIdx.x * rows_per_block + local; row < n; row += gridDim.x * rows_per_block) {
      const float4 u = a[(size_t)row * Q + q], v = b[(size_t)row * Q + q];
      acc[0] += (double)u.x * v.x; acc[1] += (double)u.y * v.y; acc[2] += (double)u.z * v.z; acc[3] += (double)u.w * v.w;
    }
  reduce_columns(acc, q, Q, active, sm, dots, true);
}

// ---------------------------------------------------------------------------------------------------- set-up (host)
// One pairwise pass: aggregates of A (agg, count m) and the Galerkin operator on them (storage in C, view in *out).
// g / gc: diagonal excess of the rows of A / of the aggregates.
struct PassOut {
  int m = 0;
  DBuf<uint32_t> agg;
  DBuf<uint32_t> rowptr;
  DBuf<int> col;
  DBuf<float> val, diag, winvd, gdiag;
  DBuf<double> g;
  long long nnz = 0;
  Csr view() const {
    Csr c;
    c.n = m; c.nnz = nnz; c.rowptr = rowptr.p; c.col = col.p; c.val = val.p; c.diag = diag.p;
    return c;
  }
};

inline int coarsen_pass(tcs_ctx* ctx, const Csr& A, const double* g, float w, int rounds, PassOut& out) {
  const int n = A.n;
  DBuf<uint32_t> matched, best, rep, head, id;
  TRY(matched.alloc(ctx, n)); TRY(best.alloc(ctx, n)); TRY(rep.alloc(ctx, n)); TRY(head.alloc(ctx, (size_t)n + 1));
  TRY(id.alloc(ctx, (size_t)n + 1)); TRY(out.agg.alloc(ctx, n));
  const unsigned gb = div_up(n, 256);
  LAUNCH(ctx, k_fill_u32, gb, 256, 0, (size_t)n, kNone, matched.p);
  for (int r = 0; r < rounds; ++r) {
    LAUNCH(ctx, k_match_pick, gb, 256, 0, n, A.rowptr, A.col, A.val, A.diag, matched.p, best.p);
    LAUNCH(ctx, k_match_commit, gb, 256, 0, n, best.p, matched.p);
  }
  LAUNCH(ctx, k_aggregate, gb, 256, 0, n, A.rowptr, A.col, A.val, A.diag, matched.p, rep.p, head.p);
  TRY(tcs_impl_scan_u32(ctx, head.p, id.p, n));
  LAUNCH(ctx, k_agg_ids, gb, 256, 0, n, rep.p, id.p, out.agg.p);
  // cross-aggregate entries -> (agg_i, agg_j) keys
  uint32_t* cnt = best.p;   // reuse
  uint32_t* pos = head.p;   // reuse (n + 1)
  LAUNCH(ctx, k_cross_count, gb, 256, 0, n, A.rowptr, A.col, out.agg.p, cnt);
  TRY(tcs_impl_scan_u32(ctx, cnt, pos, n));
  // one synchronising read for both counts
  CK(cudaMemcpyAsync((uint32_t*)ctx->h_pin, id.p + n, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync((uint32_t*)ctx->h_pin + 1, pos + n, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  const uint32_t m = ((uint32_t*)ctx->h_pin)[0], total = ((uint32_t*)ctx->h_pin)[1];
  out.m = (int)m;
  int shift = 1;
  while ((1u << shift) < m && shift < 31) ++shift;
  TRY(out.rowptr.alloc(ctx, (size_t)m + 1 + 8));
  TRY(out.diag.alloc(ctx, (size_t)m + 8)); TRY(out.winvd.alloc(ctx, (size_t)m + 8)); TRY(out.gdiag.alloc(ctx, (size_t)m + 8));
  TRY(out.g.alloc(ctx, (size_t)m + 1));
  CK(cudaMemsetAsync(out.g.p, 0, ((size_t)m + 1) * 8, ctx->stream));
  LAUNCH(ctx, k_sum_excess, gb, 256, 0, n, out.agg.p, g, out.g.p);
  uint32_t nnzc = 0;
  if (total) {
    DBuf<uint64_t> k0, k1;
    DBuf<uint32_t> v0, v1, rh, uid, crow;
    TRY(k0.alloc(ctx, total)); TRY(k1.alloc(ctx, total)); TRY(v0.alloc(ctx, total)); TRY(v1.alloc(ctx, total));
    TRY(rh.alloc(ctx, (size_t)total + 1)); TRY(uid.alloc(ctx, (size_t)total + 1));
    LAUNCH(ctx, k_cross_emit, gb, 256, 0, n, shift, A.rowptr, A.col, A.val, out.agg.p, pos, k0.p, v0.p);
    TRY(tcs_impl_sort_pairs_u64(ctx, k0.p, k1.p, v0.p, v1.p, total, 2 * shift));
    LAUNCH(ctx, k_run_heads, div_up(total, 256), 256, 0, (size_t)total, k1.p, rh.p);
    TRY(tcs_impl_scan_u32(ctx, rh.p, uid.p, total));
    TRY(tcs_impl_read_u32(ctx, uid.p + total, &nnzc));
    TRY(out.col.alloc(ctx, (size_t)nnzc + 8)); TRY(out.val.alloc(ctx, (size_t)nnzc + 16)); TRY(crow.alloc(ctx, (size_t)nnzc + 1));
    LAUNCH(ctx, k_run_reduce, div_up(total, 256), 256, 0, (size_t)total, shift, k1.p, v1.p, rh.p, uid.p, out.col.p, out.val.p, crow.p);
    LAUNCH(ctx, k_ptr_from_sorted_rows, div_up((size_t)nnzc + 1, 256), 256, 0, (size_t)nnzc, (int)m, crow.p, out.rowptr.p);
  } else {
    TRY(out.col.alloc(ctx, 8)); TRY(out.val.alloc(ctx, 16));
    CK(cudaMemsetAsync(out.rowptr.p, 0, ((size_t)m + 1) * 4, ctx->stream));
  }
  out.nnz = nnzc;
  LAUNCH(ctx, k_level_diag, div_up(m, 256), 256, 0, (int)m, w, out.rowptr.p, out.val.p, out.g.p, out.diag.p, out.winvd.p, out.gdiag.p);
  return TCS_OK;
}

// Builds the hierarchy of the assembled system S.  Leaves H.on = false (plain Jacobi-PCG) when the system is too small
// to be worth it or does not coarsen.
inline int setup(tcs_ctx* ctx, tcs_system& S, int Kp0, Hierarchy& H) {
  H.on = false;
  H.nlev = 0;
  const int n = S.n;
  if (env_i("TCS_AMG", 1) == 0 || n < env_i("TCS_AMG_MIN_ROWS", 1024) || S.nnz_off == 0) return TCS_OK;
  const bool dbg = getenv("TCS_DEBUG") != nullptr;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaEventRecord(e0, ctx->stream));
  const float w = (float)env_double("TCS_AMG_OMEGA", 0.9);
  const int rounds = env_i("TCS_AMG_ROUNDS", 2);
  const int passes = std::max(1, env_i("TCS_AMG_PASSES", 1));
  // TCS_AMG_SCHEDULE = "p0:p1:..." (or commas): pairwise passes of level 0 -> 1, 1 -> 2, ... (the last entry repeats)
  std::vector<int> schedule;
  if (const char* sc = getenv("TCS_AMG_SCHEDULE"))
    for (const char* c = sc; *c;) {
      schedule.push_back(std::max(1, atoi(c)));
      while (*c && *c != ',' && *c != ':') ++c;
      if (*c) ++c;
    }
  H.over = (float)env_double("TCS_AMG_OVER", 1.3);
  Level& L0 = H.L[0];
  L0.A.n = n; L0.A.nnz = S.nnz_off; L0.A.rowptr = S.rowptr.p; L0.A.col = S.col.p; L0.A.val = S.val.p; L0.A.diag = S.diag.p;
  TRY(L0.g.alloc(ctx, n));
  LAUNCH(ctx, k_excess, div_up(n, 256), 256, 0, n, env_i("TCS_AMG_SIGNED", 0), S.rowptr.p, S.val.p, S.diag.p, L0.g.p);
  int lev = 0;
  while (lev + 1 < kMaxLevels) {
    Level& F = H.L[lev];
    if (F.A.n <= kBottomMax || F.A.nnz == 0) break;
    PassOut P[4];
    DBuf<uint32_t> total;  // composition of the passes' maps
    Csr cur = F.A;
    const double* gcur = F.g.p;
    int np = 0;
    const int level_passes = schedule.empty() ? passes : schedule[std::min((size_t)lev, schedule.size() - 1)];
    for (; np < std::min(level_passes, 4); ++np) {
      TRY(coarsen_pass(ctx, cur, gcur, w, rounds, P[np]));
      if (np == 0) total.swap(P[0].agg);
      else {
        DBuf<uint32_t> t2, idx;
        TRY(t2.alloc(ctx, F.A.n)); TRY(idx.alloc(ctx, F.A.n));
        LAUNCH(ctx, k_compose, div_up(F.A.n, 256), 256, 0, F.A.n, total.p, P[np].agg.p, t2.p, idx.p);
        total.swap(t2);
      }
      cur = P[np].view();
      gcur = P[np].g.p;
      if (cur.n <= kBottomMax || cur.nnz == 0) { ++np; break; }
    }
    PassOut& last = P[np - 1];
    if (last.m >= F.A.n) break;  // nothing left to aggregate
    Level& C = H.L[lev + 1];
    F.nc = last.m;
    F.agg.swap(total);
    // children lists of the coarse rows: rows sorted by aggregate (stable => ascending row inside an aggregate)
    {
      DBuf<uint32_t> k0, k1, v0;
      TRY(k0.alloc(ctx, F.A.n)); TRY(k1.alloc(ctx, F.A.n)); TRY(v0.alloc(ctx, F.A.n)); TRY(F.cidx.alloc(ctx, F.A.n));
      TRY(F.cptr.alloc(ctx, (size_t)F.nc + 1));
      CK(cudaMemcpyAsync(k0.p, F.agg.p, (size_t)F.A.n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
      LAUNCH(ctx, k_iota, div_up(F.A.n, 256), 256, 0, F.A.n, v0.p);
      int bits = 1;
      while ((1u << bits) < (uint32_t)F.nc && bits < 32) ++bits;
      TRY(tcs_impl_sort_pairs_u32(ctx, k0.p, k1.p, v0.p, F.cidx.p, F.A.n, bits));
      LAUNCH(ctx, k_ptr_from_sorted_rows, div_up((size_t)F.A.n + 1, 256), 256, 0, (size_t)F.A.n, F.nc, k1.p, F.cptr.p);
    }
    C.rowptr.swap(last.rowptr); C.col.swap(last.col); C.val.swap(last.val); C.diag.swap(last.diag);
    C.winvd.swap(last.winvd); C.gdiag.swap(last.gdiag); C.g.swap(last.g);
    C.A.n = last.m; C.A.nnz = last.nnz; C.A.rowptr = C.rowptr.p; C.A.col = C.col.p; C.A.val = C.val.p; C.A.diag = C.diag.p;
    const size_t nk = (size_t)C.A.n * Kp0;
    TRY(C.r.alloc(ctx, nk)); TRY(C.x.alloc(ctx, nk)); TRY(C.t.alloc(ctx, nk));
    F.g.release();
    ++lev;
  }
  H.nlev = lev + 1;
  if (H.nlev < 2) {
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (dbg) fprintf(stderr, "[tcs] amg off: the system does not coarsen\n");
    return TCS_OK;
  }
  // bottom: dense inverse when it fits the shared memory of one CTA; otherwise its (diagonal-dominated) smoother
  Level& B = H.L[H.nlev - 1];
  H.m = B.A.n;
  H.dense = H.m <= kBottomMax && B.A.nnz > 0;
  if (H.dense) {
    DBuf<double> E;
    TRY(E.alloc(ctx, (size_t)H.m * H.m)); TRY(H.Einv.alloc(ctx, (size_t)H.m * H.m));
    LAUNCH(ctx, k_dense_build, div_up(H.m, 128), 128, 0, H.m, B.A.rowptr, B.A.col, B.A.val, B.g.p, E.p);
    const size_t smem = ((size_t)H.m * H.m + H.m) * sizeof(double);
    // function attributes are per device: set on every plan (cheap), not once per process
    CK(cudaFuncSetAttribute(k_dense_invert, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((size_t)kBottomMax * kBottomMax + kBottomMax) * sizeof(double))));
    LAUNCH(ctx, k_dense_invert, 1, 1024, smem, H.m, E.p, H.Einv.p);
    CK(cudaStreamSynchronize(ctx->stream));  // E is a local buffer
  }
  for (int l = 0; l < H.nlev; ++l) H.L[l].g.release();
  H.on = true;
  CK(cudaEventRecord(e1, ctx->stream));
  CK(cudaEventSynchronize(e1));
  CK(cudaEventElapsedTime(&H.setup_ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (dbg) {
    fprintf(stderr, "[tcs] amg: %d levels, rows", H.nlev);
    long long tot = 0;
    for (int l = 0; l < H.nlev; ++l) { fprintf(stderr, " %d", H.L[l].A.n); tot += H.L[l].A.nnz + H.L[l].A.n; }
    fprintf(stderr, "; operator complexity %.3f; bottom %s; set-up %.2f ms\n",
            (double)tot / (double)(H.L[0].A.nnz + H.L[0].A.n), H.dense ? "dense" : "smoother", H.setup_ms);
  }
  return TCS_OK;
}

// grid of the grid-stride level kernels: exactly the resident CTAs (6 per SM, see the launch bounds of k_level_spmm)
constexpr unsigned kLevelCtasPerSm = 6;
inline unsigned spmm_grid(tcs_ctx* ctx, int n, int Q) {
  const int rows_per_block = 256 / Q;
  const unsigned need = (unsigned)((n + rows_per_block - 1) / rows_per_block);
  return std::max(1u, std::min(need, (unsigned)ctx->sm_count * kLevelCtasPerSm));
}

template <int SMOOTH, int DOT>
inline void launch_level_spmm(tcs_ctx* ctx, const Level& C, int Q, const float4* X, const float4* RHS, float4* OUT,
                              double* dots) {
  const size_t dsm = (size_t)4 * Q * sizeof(double);
  LAUNCH_PDL(ctx, C.A.n, (k_level_spmm<SMOOTH, DOT>), spmm_grid(ctx, C.A.n, Q), 256, dsm, C.A.n, Q, C.A.rowptr, C.A.col, C.A.val,
           C.gdiag.p, C.winvd.p, X, RHS, OUT, dots);
}

// One V(1,1) cycle on levels >= 1 for the fine residual R (n x 4Q): afterwards result(H) holds e_1 (the coarse
// correction at level 1, to be prolonged by the caller as p += Z0 e_1) and rz_target[c] += r_1 . e_1 per column.
inline int apply(tcs_ctx* ctx, Hierarchy& H, int Q, const float4* R4, double* rz_target) {
  const int last = H.nlev - 1;
  const size_t dsm = (size_t)4 * Q * sizeof(double);
  // down
  for (int l = 1; l <= last; ++l) {
    Level& C = H.L[l];
    Level& F = H.L[l - 1];
    const float4* src = l == 1 ? R4 : (const float4*)F.t.p;
    const bool bottom = l == last;
    LAUNCH_PDL(ctx, C.A.n, k_restrict, div_up((size_t)C.A.n * Q, 256), 256, 0, C.A.n, Q, F.cptr.p, F.cidx.p, src, C.winvd.p,
           (float4*)C.r.p, bottom && H.dense ? (float4*)nullptr : (float4*)C.x.p);
    if (bottom) break;
    launch_level_spmm<0, 0>(ctx, C, Q, (const float4*)C.x.p, (const float4*)C.r.p, (float4*)C.t.p, nullptr);
  }
  // bottom
  {
    Level& B = H.L[last];
    if (H.dense) {
      LAUNCH_PDL(ctx, H.m, k_dense_apply, (unsigned)H.m, 128, 0, H.m, Q, H.Einv.p, (const float4*)B.r.p, (float4*)B.t.p);
    } else {
      // no dense inverse (many collapsed components): the smoother is exact on rows without off-diagonal entries
      launch_level_spmm<1, 0>(ctx, B, Q, (const float4*)B.x.p, (const float4*)B.r.p, (float4*)B.t.p, nullptr);
    }
    if (last == 1)
      LAUNCH_PDL(ctx, B.A.n, k_dot, std::min(spmm_grid(ctx, B.A.n, Q), 64u), 256, dsm, B.A.n, Q, (const float4*)B.r.p, (const float4*)B.t.p, rz_target);
  }
  // up
  for (int l = last - 1; l >= 1; --l) {
    Level& C = H.L[l];
    Level& N = H.L[l + 1];
    LAUNCH_PDL(ctx, C.A.n, k_prolong, div_up((size_t)C.A.n * Q, 256), 256, 0, C.A.n, Q, H.over, C.agg.p, (const float4*)N.t.p, (float4*)C.x.p);
    if (l == 1)
      launch_level_spmm<1, 1>(ctx, C, Q, (const float4*)C.x.p, (const float4*)C.r.p, (float4*)C.t.p, rz_target);
    else
      launch_level_spmm<1, 0>(ctx, C, Q, (const float4*)C.x.p, (const float4*)C.r.p, (float4*)C.t.p, nullptr);
  }
  return TCS_OK;
}

inline const uint32_t* fine_agg(const Hierarchy& H) { return H.L[0].agg.p; }
inline const float4* result(const Hierarchy& H) { return (const float4*)H.L[1].t.p; }

}  // namespace amg
