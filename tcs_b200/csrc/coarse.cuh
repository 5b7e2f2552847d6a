// tcs_b200/csrc/coarse.cuh - second level of the PCG preconditioner (large systems only; TCS_COARSE=0 turns it off).
//
// Jacobi-PCG needs thousands of iterations on these systems because the slow modes are (nearly) constant on strongly
// coupled parts of the graph: the edge weights span many orders of magnitude, so a scene falls into segments that
// are almost disconnected from each other, and the big segments are smooth membranes on top of that.  The second level
// removes exactly those modes with a piecewise-constant coarse space that FOLLOWS THE STRONG EDGES:
//   aggregates = connected components of the graph restricted to strong edges (w_ij^2 >= theta^2 d_i d_j), cut into
//                pieces of at most `blk` consecutive rows (rows are in Morton order, so a piece is spatially compact);
//   z = D^-1 r + Z (Z^T A Z)^-1 Z^T r      (additive two-level preconditioner, symmetric positive definite).
// Geometric aggregates alone (Morton blocks that straddle segments) make the iteration count worse; see DESIGN.md.
// The coarse operator E = Z^T A Z is small (m of a few thousand): it is assembled densely in float64, inverted once
// with cuSOLVER (potrf + potri, a plain library call outside the iteration loop) and kept in float32; one
// application is a restriction (shared-memory segment sums per block of rows), a dense m x m by m x K product and a
// gather inside k_pcg_dir.  r.z gains the coarse term (Z^T r).(E^-1 Z^T r), computed from coarse quantities alone.
// The solution the solver returns is checked by the float64 true residual exactly as before: the preconditioner only
// changes how fast the tolerance is reached, never what is accepted.
#pragma once
#include <cusolverDn.h>

#include <cstdlib>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace coarse {

struct Level {
  bool on = false;
  int n = 0, m = 0, blk = 0, nblk = 0;
  size_t restrict_smem = 0;
  DBuf<uint32_t> agg;    // n: aggregate of every row
  DBuf<uint32_t> first;  // nblk + 1: first aggregate id of every block of `blk` rows (ids are contiguous per block)
  DBuf<float> Einv;      // m x ld (ld = m rounded up to a multiple of 4, zero padded)
  int ld = 0;
  DBuf<float> rc, y;     // m x Kp0: Z^T r and E^-1 Z^T r
  float setup_ms = 0.f;
};

__global__ void k_cc_init(int n, uint32_t* __restrict__ lab) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lab[i] = (uint32_t)i;
}

// min-label hooking over strong edges; lab[] always points at a row with a label <= its own
__global__ void k_cc_hook(int n, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                          const float* __restrict__ val, const float* __restrict__ diag, float theta2,
                          uint32_t* lab, int* __restrict__ changed) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t l = lab[i];
  uint32_t mn = l;
  const float di = diag[i] * theta2;
  for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) {
    const int j = col[a];
    const float w = val[a];
    if (w * w >= di * diag[j]) mn = min(mn, lab[j]);
  }
  if (mn < l) {
    atomicMin(&lab[l], mn);
    atomicMin(&lab[i], mn);
    *changed = 1;
  }
}

__global__ void k_cc_jump(int n, uint32_t* lab) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t l = lab[i];
  while (true) {
    const uint32_t p = lab[l];
    if (p == l) break;
    l = p;
  }
  lab[i] = l;
}

__global__ void k_agg_keys(int n, int blk, const uint32_t* __restrict__ lab, uint64_t* __restrict__ keys,
                           uint32_t* __restrict__ idx) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (uint64_t)(i / blk) * (uint64_t)n + lab[i];
  idx[i] = (uint32_t)i;
}

__global__ void k_key_heads(int n, const uint64_t* __restrict__ keys, uint32_t* __restrict__ flag) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > n) return;
  flag[j] = (j < n && (j == 0 || keys[j] != keys[j - 1])) ? 1u : 0u;
}

__global__ void k_assign_agg(int n, int blk, const uint32_t* __restrict__ flag, const uint32_t* __restrict__ scan,
                             const uint32_t* __restrict__ idx, uint32_t* __restrict__ agg, uint32_t* __restrict__ first) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t a = scan[j] - (flag[j] ? 0u : 1u);
  const uint32_t i = idx[j];
  agg[i] = a;
  atomicMin(&first[i / blk], a);
}

// E = Z^T A Z (dense, float64) and the diagonal mass of every aggregate (for the regularisation of singular blocks)
__global__ void k_assemble_E(int n, int m, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                             const float* __restrict__ val, const float* __restrict__ diag,
                             const uint32_t* __restrict__ agg, double* __restrict__ E, double* __restrict__ dsum) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t a = agg[i];
  double self = (double)diag[i];
  for (uint32_t e = rowptr[i]; e < rowptr[i + 1]; ++e) {
    const uint32_t b = agg[col[e]];
    if (b == a) self += (double)val[e];
    else atomicAdd(&E[(size_t)a * m + b], (double)val[e]);
  }
  atomicAdd(&E[(size_t)a * m + a], self);
  atomicAdd(&dsum[a], (double)diag[i]);
}

__global__ void k_regularise(int m, double* __restrict__ E, const double* __restrict__ dsum, double eps) {
  int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < m) E[(size_t)a * m + a] += eps * dsum[a];
}

// potri leaves the inverse in one triangle (column-major lower = row-major upper): expand to a full float32 matrix
// (rows of ld = m rounded up to 4 floats, zero padded: the product kernel reads them with 16-byte loads)
__global__ void k_einv_f32(int m, int ld, const double* __restrict__ Ed, float* __restrict__ Ef) {
  size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)m * ld) return;
  const int a = (int)(t / ld), b = (int)(t - (size_t)a * ld);
  Ef[t] = b < m ? (float)Ed[(size_t)min(a, b) * m + max(a, b)] : 0.f;
}

// rc = Z^T R: one CTA per block of `blk` rows.  A thread keeps one column group and walks rows local, local + rows_per,
// ...; it sums in registers while the aggregate stays the same (a block holds one or two aggregates as a rule) and
// only flushes a finished run into the shared-memory segment sums.
__global__ void __launch_bounds__(256) k_restrict(int n, int Q, int blk, const float4* __restrict__ R,
                                                  const uint32_t* __restrict__ agg, const uint32_t* __restrict__ first,
                                                  float* __restrict__ rc, float* __restrict__ y) {
  extern __shared__ float acc[];
  const int b = blockIdx.x, W = 4 * Q;
  const uint32_t f0 = first[b];
  const int na = (int)(first[b + 1] - f0);
  for (int t = threadIdx.x; t < na * W; t += blockDim.x) acc[t] = 0.f;
  __syncthreads();
  const int rows_per = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const int row0 = b * blk, rows = min(blk, n - row0);
  if (local < rows_per) {
    int cur = -1;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r0 = local; r0 < rows; r0 += 8 * rows_per) {
      int a4[8];
      float4 v4[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {  // eight rows in flight
        const int r = r0 + u * rows_per;
        a4[u] = r < rows ? (int)(agg[row0 + r] - f0) : -1;
        v4[u] = r < rows ? R[(size_t)(row0 + r) * Q + q] : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (a4[u] < 0) continue;
        if (a4[u] != cur) {
          if (cur >= 0) {
            float* d = acc + (size_t)cur * W + 4 * q;
            atomicAdd(d + 0, s.x); atomicAdd(d + 1, s.y); atomicAdd(d + 2, s.z); atomicAdd(d + 3, s.w);
          }
          cur = a4[u];
          s = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        s.x += v4[u].x; s.y += v4[u].y; s.z += v4[u].z; s.w += v4[u].w;
      }
    }
    if (cur >= 0) {
      float* d = acc + (size_t)cur * W + 4 * q;
      atomicAdd(d + 0, s.x); atomicAdd(d + 1, s.y); atomicAdd(d + 2, s.z); atomicAdd(d + 3, s.w);
    }
  }
  __syncthreads();
  // y of the block's aggregates is cleared here: k_coarse_gemm adds its partial products into it
  for (int t = threadIdx.x; t < na * W; t += blockDim.x) { rc[(size_t)f0 * W + t] = acc[t]; y[(size_t)f0 * W + t] = 0.f; }
}

// y += E^-1[:, b-range] rc[b-range]  (dense m x m by m x 4Q product, split over the b range: grid.y parts add their
// partial sums into the zeroed y).  A CTA owns TA = 4 * rows_per aggregates; a thread owns 4 aggregates x 4 columns in
// registers.  E^-1 is symmetric, so the tile E^-1[a-range][b] is read as E^-1[b][a-range]: coalesced, and already in
// the [b][a] order the float4 shared-memory reads want.
#define TCS_COARSE_TB 32
__global__ void __launch_bounds__(256) k_coarse_gemm(int m, int ld, int Q, int rows_per, int b_per_part,
                                                     const float* __restrict__ Einv, const float* __restrict__ rc,
                                                     float* __restrict__ y) {
  extern __shared__ float4 sm4[];
  // thread = (ks, local, q): group `local` owns 4 aggregates, q a column group, and the KS = blockDim / (rows_per Q)
  // copies of that layout share the b values of a tile (narrow working sets would otherwise leave a CTA one warp)
  const int TA = 4 * rows_per, W = 4 * Q, G = rows_per * Q;
  const int KS = blockDim.x / G;
  float* Es = (float*)sm4;                        // [TB][TA]
  float* Rs = Es + TCS_COARSE_TB * TA;            // [TB][W]
  const int ks = threadIdx.x / G, g = threadIdx.x - ks * G;
  const int local = g / Q, q = g - local * Q;
  const int a0 = blockIdx.x * TA;
  const int b_begin = blockIdx.y * b_per_part, b_end = min(m, b_begin + b_per_part);
  float4 acc[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int b0 = b_begin; b0 < b_end; b0 += TCS_COARSE_TB) {
    const int nb = min(TCS_COARSE_TB, b_end - b0);
    // 16-byte loads, up to four in flight per thread (the tile loop has a run-time trip count: without the explicit
    // batches every load waits for the shared-memory store of the one before).  a0 and ld are multiples of 4.
    const int TA4 = TA / 4;
    for (int t0 = threadIdx.x; t0 < TCS_COARSE_TB * TA4; t0 += 4 * blockDim.x) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = t0 + u * blockDim.x;
        const int bb = t / TA4, a = 4 * (t - bb * TA4);
        v[u] = (t < TCS_COARSE_TB * TA4 && bb < nb && a0 + a < ld)
                   ? *(const float4*)(Einv + (size_t)(b0 + bb) * ld + a0 + a) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = t0 + u * blockDim.x;
        if (t < TCS_COARSE_TB * TA4) ((float4*)Es)[t] = v[u];
      }
    }
    for (int t = threadIdx.x; t < TCS_COARSE_TB * W; t += blockDim.x) Rs[t] = (t / W < nb) ? rc[(size_t)b0 * W + t] : 0.f;
    __syncthreads();
    if (ks < KS) {
      for (int bb = ks; bb < TCS_COARSE_TB; bb += KS) {
        const float4 e = *(const float4*)(Es + bb * TA + 4 * local);
        const float4 r = *(const float4*)(Rs + bb * W + 4 * q);
        acc[0].x += e.x * r.x; acc[0].y += e.x * r.y; acc[0].z += e.x * r.z; acc[0].w += e.x * r.w;
        acc[1].x += e.y * r.x; acc[1].y += e.y * r.y; acc[1].z += e.y * r.z; acc[1].w += e.y * r.w;
        acc[2].x += e.z * r.x; acc[2].y += e.z * r.y; acc[2].z += e.z * r.z; acc[2].w += e.z * r.w;
        acc[3].x += e.w * r.x; acc[3].y += e.w * r.y; acc[3].z += e.w * r.z; acc[3].w += e.w * r.w;
      }
    }
    __syncthreads();
  }
  if (ks < KS) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int a = a0 + 4 * local + i;
      if (a < m) {
        float* d = y + (size_t)a * W + 4 * q;
        atomicAdd(d + 0, acc[i].x); atomicAdd(d + 1, acc[i].y); atomicAdd(d + 2, acc[i].z); atomicAdd(d + 3, acc[i].w);
      }
    }
  }
}

// rz[c] += sum_a rc[a][c] y[a][c]
__global__ void __launch_bounds__(256) k_coarse_dot(int m, int Q, const float4* __restrict__ rc,
                                                    const float4* __restrict__ y, double* __restrict__ rz) {
  extern __shared__ double dots[];  // 4Q
  const int W = 4 * Q;
  for (int t = threadIdx.x; t < W; t += blockDim.x) dots[t] = 0.0;
  __syncthreads();
  const int rows_per = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  if (local < rows_per) {
    double d[4] = {0, 0, 0, 0};
    for (int a = blockIdx.x * rows_per + local; a < m; a += gridDim.x * rows_per) {
      const float4 r = rc[(size_t)a * Q + q], v = y[(size_t)a * Q + q];
      d[0] += (double)r.x * v.x; d[1] += (double)r.y * v.y; d[2] += (double)r.z * v.z; d[3] += (double)r.w * v.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(&dots[4 * q + j], d[j]);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < W; c += blockDim.x)
    if (dots[c] != 0.0) atomicAdd(&rz[c], dots[c]);
}

// p += Z y (after k_pcg_init, whose p is D^-1 b)
__global__ void k_add_coarse_p(int n, int Q, float4* __restrict__ P, const uint32_t* __restrict__ agg,
                               const float4* __restrict__ y) {
  size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= (size_t)n * Q) return;
  const int row = (int)(t / Q), q = (int)(t - (size_t)row * Q);
  const float4 c = y[(size_t)agg[row] * Q + q];
  float4 p = P[t];
  p.x += c.x; p.y += c.y; p.z += c.z; p.w += c.w;
  P[t] = p;
}

static inline double env_double(const char* name, double dflt) { const char* v = getenv(name); return v ? atof(v) : dflt; }
static inline int env_i(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }

// rc = Z^T R, y = E^-1 rc, rz += rc . y   (R at working width 4 Q)
inline int apply(tcs_ctx* ctx, Level& Lv, int Q, const float4* R4, double* rz_target) {
  const int W = 4 * Q;
  LAUNCH(ctx, k_restrict, Lv.nblk, 256, Lv.restrict_smem, Lv.n, Q, Lv.blk, R4, Lv.agg.p, Lv.first.p, Lv.rc.p, Lv.y.p);
  const int rows_per = std::min(256 / Q, 32);
  const int TA = 4 * rows_per;
  const unsigned gx = div_up(Lv.m, TA);
  unsigned parts = std::max(1u, std::min(32u, ((unsigned)ctx->sm_count * 4 + gx - 1) / gx));
  int b_per_part = (int)div_up(Lv.m, parts);
  b_per_part = (b_per_part + TCS_COARSE_TB - 1) / TCS_COARSE_TB * TCS_COARSE_TB;
  parts = div_up(Lv.m, b_per_part);
  const size_t smem = (size_t)TCS_COARSE_TB * (TA + W) * sizeof(float);
  const int KS = std::max(1, 256 / (rows_per * Q));
  LAUNCH(ctx, k_coarse_gemm, dim3(gx, parts), KS * rows_per * Q, smem, Lv.m, Lv.ld, Q, rows_per, b_per_part, Lv.Einv.p,
         Lv.rc.p, Lv.y.p);
  LAUNCH(ctx, k_coarse_dot, 48, 256, (size_t)W * sizeof(double), Lv.m, Q, (const float4*)Lv.rc.p, (const float4*)Lv.y.p,
         rz_target);
  return TCS_OK;
}

// Builds the aggregates and E^-1.  Leaves Lv.on = false (and returns TCS_OK) whenever the level is not worth having
// or cannot be built; the solver then runs plain Jacobi-PCG.
inline int setup(tcs_ctx* ctx, tcs_system& S, int Kp0, Level& Lv) {
  Lv.on = false;
  const int n = S.n;
  if (env_i("TCS_COARSE", 1) == 0 || n < env_i("TCS_COARSE_MIN_ROWS", 150000)) return TCS_OK;
  const bool dbg = getenv("TCS_DEBUG") != nullptr;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaEventRecord(e0, ctx->stream));
  const float theta = (float)env_double("TCS_COARSE_THETA", 0.02);
  const int max_m = env_i("TCS_COARSE_MAX_AGGREGATES", 12000);

  // ---- strong-edge components
  DBuf<uint32_t> lab;
  DBuf<int> changed;
  TRY(lab.alloc(ctx, n)); TRY(changed.alloc(ctx, 1));
  LAUNCH(ctx, k_cc_init, div_up(n, 256), 256, 0, n, lab.p);
  int rounds = 0;
  for (; rounds < 500; ++rounds) {
    CK(cudaMemsetAsync(changed.p, 0, 4, ctx->stream));
    LAUNCH(ctx, k_cc_hook, div_up(n, 256), 256, 0, n, S.rowptr.p, S.col.p, S.val.p, S.diag.p, theta * theta, lab.p, changed.p);
    LAUNCH(ctx, k_cc_jump, div_up(n, 256), 256, 0, n, lab.p);
    uint32_t c = 0;
    TRY(tcs_impl_read_u32(ctx, (const uint32_t*)changed.p, &c));
    if (!c) break;
  }

  // ---- aggregates = (block of rows, component), numbered block-major
  int blk = env_i("TCS_COARSE_BLOCK", 0);
  if (blk <= 0) { blk = 256; while (blk < 4096 && (long long)blk * 1500 < n) blk <<= 1; }
  DBuf<uint64_t> k0, k1;
  DBuf<uint32_t> i0, i1, flag, scan;
  TRY(k0.alloc(ctx, n)); TRY(k1.alloc(ctx, n)); TRY(i0.alloc(ctx, n)); TRY(i1.alloc(ctx, n));
  TRY(flag.alloc(ctx, (size_t)n + 1)); TRY(scan.alloc(ctx, (size_t)n + 1));
  TRY(Lv.agg.alloc(ctx, n));
  uint32_t m = 0;
  int nblk = 0;
  for (;; blk <<= 1) {
    nblk = (n + blk - 1) / blk;
    LAUNCH(ctx, k_agg_keys, div_up(n, 256), 256, 0, n, blk, lab.p, k0.p, i0.p);
    int bits = 1;
    while (bits < 64 && ((uint64_t)nblk * (uint64_t)n) >> bits) ++bits;
    TRY(tcs_impl_sort_pairs_u64(ctx, k0.p, k1.p, i0.p, i1.p, n, bits));
    LAUNCH(ctx, k_key_heads, div_up((size_t)n + 1, 256), 256, 0, n, k1.p, flag.p);
    TRY(tcs_impl_scan_u32(ctx, flag.p, scan.p, n));
    TRY(tcs_impl_read_u32(ctx, scan.p + n, &m));
    if ((int)m <= max_m || blk >= 8192) break;
  }
  if ((int)m > max_m || m < 2) {
    if (dbg) fprintf(stderr, "[tcs] coarse level off: %u aggregates (limit %d)\n", m, max_m);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return TCS_OK;
  }
  TRY(Lv.first.alloc(ctx, (size_t)nblk + 1));
  CK(cudaMemsetAsync(Lv.first.p, 0xFF, ((size_t)nblk + 1) * 4, ctx->stream));
  LAUNCH(ctx, k_assign_agg, div_up(n, 256), 256, 0, n, blk, flag.p, scan.p, i1.p, Lv.agg.p, Lv.first.p);
  CK(cudaMemcpyAsync(Lv.first.p + nblk, scan.p + n, 4, cudaMemcpyDeviceToDevice, ctx->stream));
  std::vector<uint32_t> hfirst((size_t)nblk + 1);
  CK(cudaMemcpyAsync(hfirst.data(), Lv.first.p, ((size_t)nblk + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  uint32_t max_na = 0;
  for (int b = 0; b < nblk; ++b) max_na = std::max(max_na, hfirst[b + 1] - hfirst[b]);
  Lv.restrict_smem = (size_t)max_na * Kp0 * sizeof(float);
  if (Lv.restrict_smem > 200 * 1024) {
    if (dbg) fprintf(stderr, "[tcs] coarse level off: %u aggregates in one block\n", max_na);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return TCS_OK;
  }
  if (Lv.restrict_smem > 48 * 1024)
    CK(cudaFuncSetAttribute(k_restrict, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Lv.restrict_smem));
  k0.release(); k1.release(); i0.release(); i1.release(); flag.release(); scan.release(); lab.release();

  // ---- E = Z^T A Z, inverted once (float64), kept in float32
  DBuf<double> E, dsum, work;
  DBuf<int> info;
  TRY(E.alloc(ctx, (size_t)m * m)); TRY(dsum.alloc(ctx, m)); TRY(info.alloc(ctx, 2));
  CK(cudaMemsetAsync(E.p, 0, (size_t)m * m * 8, ctx->stream));
  CK(cudaMemsetAsync(dsum.p, 0, (size_t)m * 8, ctx->stream));
  LAUNCH(ctx, k_assemble_E, div_up(n, 256), 256, 0, n, (int)m, S.rowptr.p, S.col.p, S.val.p, S.diag.p, Lv.agg.p, E.p, dsum.p);
  LAUNCH(ctx, k_regularise, div_up(m, 256), 256, 0, (int)m, E.p, dsum.p, env_double("TCS_COARSE_EPS", 1e-6));
  // one cuSOLVER handle per device for the life of the process (creating one costs tens of milliseconds and device
  // memory); the mutex covers enqueue + synchronise, the handle's workspace is not shared between streams in flight
  static std::mutex mu;
  static cusolverDnHandle_t handles[64] = {};
  int hinfo[2] = {-1, -1};
  bool ok = ctx->device >= 0 && ctx->device < 64;
  if (ok) {
    std::lock_guard<std::mutex> lock(mu);
    cusolverDnHandle_t& h = handles[ctx->device];
    if (!h && cusolverDnCreate(&h) != CUSOLVER_STATUS_SUCCESS) { h = nullptr; ok = false; }
    ok = ok && cusolverDnSetStream(h, ctx->stream) == CUSOLVER_STATUS_SUCCESS;
    int lw1 = 0, lw2 = 0;
    ok = ok && cusolverDnDpotrf_bufferSize(h, CUBLAS_FILL_MODE_LOWER, (int)m, E.p, (int)m, &lw1) == CUSOLVER_STATUS_SUCCESS;
    ok = ok && cusolverDnDpotri_bufferSize(h, CUBLAS_FILL_MODE_LOWER, (int)m, E.p, (int)m, &lw2) == CUSOLVER_STATUS_SUCCESS;
    if (ok) {
      TRY(work.alloc(ctx, (size_t)std::max(lw1, lw2) + 1));
      ok = cusolverDnDpotrf(h, CUBLAS_FILL_MODE_LOWER, (int)m, E.p, (int)m, work.p, lw1, info.p) == CUSOLVER_STATUS_SUCCESS;
      ok = ok && cusolverDnDpotri(h, CUBLAS_FILL_MODE_LOWER, (int)m, E.p, (int)m, work.p, lw2, info.p + 1) == CUSOLVER_STATUS_SUCCESS;
      CK(cudaMemcpyAsync(ctx->h_pin, info.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      memcpy(hinfo, ctx->h_pin, 8);
    }
  }
  if (!ok || hinfo[0] != 0 || hinfo[1] != 0) {
    if (dbg) fprintf(stderr, "[tcs] coarse level off: factorisation failed (potrf %d, potri %d)\n", hinfo[0], hinfo[1]);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return TCS_OK;
  }
  Lv.ld = ((int)m + 3) & ~3;
  TRY(Lv.Einv.alloc(ctx, (size_t)m * Lv.ld));
  LAUNCH(ctx, k_einv_f32, div_up((size_t)m * Lv.ld, 256), 256, 0, (int)m, Lv.ld, E.p, Lv.Einv.p);
  TRY(Lv.rc.alloc(ctx, (size_t)m * Kp0)); TRY(Lv.y.alloc(ctx, (size_t)m * Kp0));
  Lv.n = n; Lv.m = (int)m; Lv.blk = blk; Lv.nblk = nblk;
  Lv.on = true;
  CK(cudaEventRecord(e1, ctx->stream));
  CK(cudaEventSynchronize(e1));
  CK(cudaEventElapsedTime(&Lv.setup_ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (dbg)
    fprintf(stderr, "[tcs] coarse level: %d rows -> %u aggregates (blocks of %d rows, at most %u per block, theta %.3g, %d "
            "component rounds), set-up %.1f ms\n", n, m, blk, max_na, theta, rounds, Lv.setup_ms);
  return TCS_OK;
}

}  // namespace coarse
