// tcs_b200/csrc/pcg.cu - the random-walker linear solve: multi-right-hand-side Jacobi-preconditioned CG.
// Replaces RandomWalker::solveLinearSystem / assignColors / getPotentials (impl/random_walker.hpp:186-243):
// Eigen::SimplicialCholesky<SparseMatrix<float>, Lower> becomes PCG over all label columns at once.
//
// Data layout (HBM): CSR of the off-diagonal part of L (int32 row indices, fp32 values = -w), fp32 diagonal,
// dense blocks P, R, AP as n x Kp fp32 row-major (Kp = K rounded up to 4 => one row = Kp/4 float4), X as n x Kp
// fp64.  Precision recipe: fp32 matrix and direction/residual vectors, fp64 dot products, fp64
// solution, float64 true-residual replacement every `check_every` iterations.
#include <cfloat>
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <chrono>

#include "common.cuh"

namespace {

__device__ __forceinline__ bool is_pow2(int q) { return (q & (q - 1)) == 0; }

// Block-level reduction of per-thread partial sums of 4 columns (column group q of Q) into gout[Kp].
// packed = the thread layout of the vector kernels (thread = row * Q + q): the lanes of a warp that hold the same
// column group sit Q apart also when Q is not a power of two, so they are summed with strided shuffles first
// (without this a 12-column block makes every thread hit twelve shared-memory words with fp64 CAS loops: 15 us on C4).
__device__ __forceinline__ void reduce_columns(double acc[4], int q, int Q, bool active, double* sm, double* gout,
                                               bool packed = false) {
  const int Kp = 4 * Q;
  for (int i = threadIdx.x; i < Kp; i += blockDim.x) sm[i] = 0.0;
  __syncthreads();
  if (packed && !is_pow2(Q) && Q < 32) {
    const int lane = threadIdx.x & 31;
    for (int o = Q; o < 32; o <<= 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const double v = __shfl_down_sync(0xFFFFFFFFu, acc[j], o);
        if (lane + o < 32) acc[j] += v;
      }
    }
    if (lane < Q) {   // lane l < Q now holds the sum over lanes l, l + Q, l + 2 Q, ... (all of its column group)
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(&sm[4 * q + j], acc[j]);
    }
  } else if (is_pow2(Q) && Q <= 32) {
    for (int o = 16; o >= Q; o >>= 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] += __shfl_xor_sync(0xFFFFFFFFu, acc[j], o);
    }
    if ((threadIdx.x & 31) < Q) {
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(&sm[4 * q + j], acc[j]);
    }
  } else if (active) {
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(&sm[4 * q + j], acc[j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < Kp; i += blockDim.x)
    if (sm[i] != 0.0) atomicAdd(&gout[i], sm[i]);
}

}  // namespace
#include "spmm.cuh"
#include "spmm_ring.cuh"
#include "amg.cuh"
namespace {

// ---- the CG SpMM: AP = L P over all label columns, fused p.Ap partial dots -------------------------------------
// All SpMM kernels evaluate the Laplacian in DIFFERENCE FORM:  (L p)_i = g_i p_i + sum_j w_ij (p_i - p_j),  where
// g_i = d_i - sum_j w_ij is the diagonal excess (the coupling to the seeds; rounding noise elsewhere) - the same
// operator as d_i p_i - sum_j w_ij p_j, but the float32 rounding error is proportional to |p_i - p_j| instead of
// |p_i|: the directions are smooth, so the noise the recurrence residual accumulates (eps32 |A| |x| sqrt(n), the floor
// under the true residual of one refinement stage) drops by the smoothness factor.  `val` holds -w_ij.
// Q = Kp/4 consecutive threads own one row (one float4 of the row each): the gather of a neighbour's P row is one
// contiguous 16*Q-byte read per row group.  Grid-stride over row groups; fp64 dot accumulation per thread.
__global__ void __launch_bounds__(256) k_spmm(int n, int Q, const uint32_t* __restrict__ rowptr,
                                              const int* __restrict__ col, const float* __restrict__ val,
                                              const float* __restrict__ gdiag, const float4* __restrict__ P,
                                              float4* __restrict__ AP, double* __restrict__ pAp) {
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
      const float4 p = P[(size_t)row * Q + q];
      const float d = gdiag[row];
      const F4p pp = f4_pack(p);
      F4p yy = f4_pack(make_float4(d * p.x, d * p.y, d * p.z, d * p.w));
      for (; a + 4 <= a1; a += 4) {
        int c0 = col[a], c1 = col[a + 1], c2 = col[a + 2], c3 = col[a + 3];
        float w0 = val[a], w1 = val[a + 1], w2 = val[a + 2], w3 = val[a + 3];
        float4 t0 = P[(size_t)c0 * Q + q], t1 = P[(size_t)c1 * Q + q];
        float4 t2 = P[(size_t)c2 * Q + q], t3 = P[(size_t)c3 * Q + q];
        f4_diff_fma(yy, w0, t0, pp);
        f4_diff_fma(yy, w1, t1, pp);
        f4_diff_fma(yy, w2, t2, pp);
        f4_diff_fma(yy, w3, t3, pp);
      }
      for (; a < a1; ++a) {
        int c = col[a];
        float w = val[a];
        float4 t = P[(size_t)c * Q + q];
        f4_diff_fma(yy, w, t, pp);
      }
      const float4 y = f4_unpack(yy);
      AP[(size_t)row * Q + q] = y;
      acc[0] += (double)p.x * (double)y.x; acc[1] += (double)p.y * (double)y.y;
      acc[2] += (double)p.z * (double)y.z; acc[3] += (double)p.w * (double)y.w;
    }
  }
  reduce_columns(acc, q, Q, active, sm, pAp, true);
}

// g_i = d_i - sum_j w_ij, evaluated in float64 and rounded once (val = -w)
__global__ void k_diag_excess(int n, const uint32_t* __restrict__ rowptr, const float* __restrict__ val,
                              const float* __restrict__ diag, float* __restrict__ gdiag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = (double)diag[i];
  for (uint32_t a = rowptr[i]; a < rowptr[i + 1]; ++a) s += (double)val[a];
  gdiag[i] = (float)s;
}

// x = 0, r = b, p = D^-1 b, rz = sum b^2 / d, bb = sum b^2 ; also 1/d
__global__ void __launch_bounds__(256) k_pcg_init(int n, int Q, const float4* __restrict__ B,
                                                  const float* __restrict__ diag, float* __restrict__ invd,
                                                  float4* __restrict__ R, float4* __restrict__ P, double* __restrict__ X,
                                                  float4* __restrict__ Xd, double* __restrict__ rz,
                                                  double* __restrict__ bb) {
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double a_rz[4] = {0, 0, 0, 0}, a_bb[4] = {0, 0, 0, 0};
  if (active) {
    for (int row = blockIdx.x * rows_per_block + local; row < n; row += gridDim.x * rows_per_block) {
      const size_t i = (size_t)row * Q + q;
      const float4 b = B[i];
      const float inv = 1.0f / diag[row];
      if (q == 0) invd[row] = inv;
      R[i] = b;
      P[i] = make_float4(b.x * inv, b.y * inv, b.z * inv, b.w * inv);
      double2* x = (double2*)(X + 4 * i);
      x[0] = make_double2(0.0, 0.0);
      x[1] = make_double2(0.0, 0.0);
      Xd[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      const double di = (double)inv;
      a_bb[0] += (double)b.x * b.x; a_bb[1] += (double)b.y * b.y; a_bb[2] += (double)b.z * b.z; a_bb[3] += (double)b.w * b.w;
      a_rz[0] += (double)b.x * b.x * di; a_rz[1] += (double)b.y * b.y * di;
      a_rz[2] += (double)b.z * b.z * di; a_rz[3] += (double)b.w * b.w * di;
    }
  }
  reduce_columns(a_rz, q, Q, active, sm, rz, true);
  __syncthreads();
  reduce_columns(a_bb, q, Q, active, sm, bb, true);
}

// Step length of column c: alpha = r.z / p.Ap, or the exact line search p.r / p.Ap on the iteration after a residual
// replacement (pr != NULL, see k_true_residual).  A column that reached the tolerance is frozen (host flag), and so
// is one whose recurrence residual fell far below it between two checks (rz <= done_factor * rz_initial): iterating
// on it further only amplifies rounding noise, and once r underflows in float32 rz / pAp is no longer a step length.
__device__ __forceinline__ float pcg_alpha(int c, const double* __restrict__ rz, const double* __restrict__ pAp,
                                           const int* __restrict__ frozen, const double* __restrict__ rz0,
                                           double done_factor, const double* __restrict__ pr) {
  const double den = pAp[c], num = pr ? pr[c] : rz[c];
  const bool live = den > 0.0 && num > done_factor * rz0[c] && !frozen[c];
  const double ratio = live ? num / den : 0.0;
  return ratio < 1e30 ? (float)ratio : 0.0f;
}

// r -= alpha Ap ; rz_next += r . D^-1 r     (x += alpha p happens in k_pcg_dir, which reads p anyway)
__global__ void __launch_bounds__(256) k_pcg_update(int n, int Q, const float4* __restrict__ AP, float4* __restrict__ R,
                                                    const float* __restrict__ invd, const double* __restrict__ rz,
                                                    const double* __restrict__ pAp, double* __restrict__ rz_next,
                                                    const int* __restrict__ frozen, const double* __restrict__ rz0,
                                                    double done_factor, const double* __restrict__ pr,
                                                    double* __restrict__ rr_next) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  // rr_next (may be NULL): ||r||^2 of the recurrence residual per column - what the host probes between two float64
  // true-residual checks when the multilevel preconditioner is on
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double acc[4] = {0, 0, 0, 0}, acc2[4] = {0, 0, 0, 0};
  if (active) {
    float al[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) al[j] = pcg_alpha(4 * q + j, rz, pAp, frozen, rz0, done_factor, pr);
    const int stride = gridDim.x * rows_per_block;
    int row = blockIdx.x * rows_per_block + local;
    // four rows per step: eight independent 16-byte loads in flight per thread
    for (; row + 3 * stride < n; row += 4 * stride) {
      size_t i[4];
      float4 ap[4], r[4];
      float d[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) i[u] = (size_t)(row + u * stride) * Q + q;
#pragma unroll
      for (int u = 0; u < 4; ++u) { ap[u] = AP[i[u]]; r[u] = R[i[u]]; d[u] = invd[row + u * stride]; }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        // a dead column (alpha = 0) is left untouched: its Ap may hold inf or nan after float32 underflow
        if (al[0] != 0.f) r[u].x -= al[0] * ap[u].x;
        if (al[1] != 0.f) r[u].y -= al[1] * ap[u].y;
        if (al[2] != 0.f) r[u].z -= al[2] * ap[u].z;
        if (al[3] != 0.f) r[u].w -= al[3] * ap[u].w;
        R[i[u]] = r[u];
        // float64 products: r can sit far below sqrt(FLT_MIN) on quickly converging columns
        const double e = (double)d[u];
        const double sx = (double)r[u].x * (double)r[u].x, sy = (double)r[u].y * (double)r[u].y;
        const double sz = (double)r[u].z * (double)r[u].z, sw = (double)r[u].w * (double)r[u].w;
        acc[0] += sx * e; acc[1] += sy * e; acc[2] += sz * e; acc[3] += sw * e;
        acc2[0] += sx; acc2[1] += sy; acc2[2] += sz; acc2[3] += sw;
      }
    }
    for (; row < n; row += stride) {
      const size_t i = (size_t)row * Q + q;
      const float4 ap = AP[i];
      float4 r = R[i];
      const float d = invd[row];
      if (al[0] != 0.f) r.x -= al[0] * ap.x;
      if (al[1] != 0.f) r.y -= al[1] * ap.y;
      if (al[2] != 0.f) r.z -= al[2] * ap.z;
      if (al[3] != 0.f) r.w -= al[3] * ap.w;
      R[i] = r;
      const double e = (double)d;
      const double sx = (double)r.x * (double)r.x, sy = (double)r.y * (double)r.y;
      const double sz = (double)r.z * (double)r.z, sw = (double)r.w * (double)r.w;
      acc[0] += sx * e; acc[1] += sy * e; acc[2] += sz * e; acc[3] += sw * e;
      acc2[0] += sx; acc2[1] += sy; acc2[2] += sz; acc2[3] += sw;
    }
  }
  reduce_columns(acc, q, Q, active, sm, rz_next, true);
  if (rr_next) {
    __syncthreads();
    reduce_columns(acc2, q, Q, active, sm, rr_next, true);
  }
}

// out[c] = sum over rows [row0, row0 + n) of B[row][c]^2 (float64): ||b||^2 of one scene of a block-diagonal batch
__global__ void __launch_bounds__(256) k_col_sumsq(int row0, int n, int Q, const float4* __restrict__ B, double* __restrict__ out) {
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double acc[4] = {0, 0, 0, 0};
  if (active)
    for (int lrow = blockIdx.x * rows_per_block + local; lrow < n; lrow += gridDim.x * rows_per_block) {
      const float4 b = B[(size_t)(row0 + lrow) * Q + q];
      acc[0] += (double)b.x * b.x; acc[1] += (double)b.y * b.y; acc[2] += (double)b.z * b.z; acc[3] += (double)b.w * b.w;
    }
  reduce_columns(acc, q, Q, active, sm, out, true);
}

// X += xd ; xd = 0  (before every float64 true residual)
__global__ void k_flush_x(size_t count, double* __restrict__ X, float* __restrict__ Xd) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= count) return;
  X[i] += (double)Xd[i];
  Xd[i] = 0.f;
}

// column compaction: out[row, k] = in[row, src[k]] (0 where src[k] < 0); Kin / Kout are the row widths
template <typename T>
__global__ void k_repack_columns(size_t n, int Kin, int Kout, const int* __restrict__ src, const T* __restrict__ in,
                                 T* __restrict__ out) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n * Kout) return;
  const size_t row = i / Kout;
  const int k = (int)(i - row * Kout), sk = src[k];
  out[i] = sk >= 0 ? in[row * Kin + sk] : (T)0;
}
// X_full[row, dst[k]] = X_work[row, k] for the working columns with dst[k] >= 0
__global__ void k_scatter_columns(size_t n, int Kw, int Kfull, const int* __restrict__ dst, const double* __restrict__ Xw,
                                  double* __restrict__ Xfull) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n * Kw) return;
  const size_t row = i / Kw;
  const int dk = dst[(int)(i - row * Kw)];
  if (dk >= 0) Xfull[row * Kfull + dk] = Xw[i];
}

// xd += alpha p ; beta = rz_next / rz ; p = D^-1 r + beta p ; block 0 clears the accumulators of the coming iterations.
// x is carried as a float64 base X plus a float32 increment xd that k_flush_x folds into X before every float64
// true-residual evaluation (an fp32 x alone stalls the true residual at 1e-4; the increment of
// `check_every` iterations is small next to x, so its float32 rounding is far below the tolerance) - 8 instead of
// 16 bytes of x traffic per element and iteration, and the update lives here because this kernel reads p anyway.
__global__ void __launch_bounds__(256) k_pcg_dir(int n, int Q, const float4* __restrict__ R, float4* __restrict__ P,
                                                 float4* __restrict__ Xd, const float* __restrict__ invd,
                                                 const double* __restrict__ rz, const double* __restrict__ pAp,
                                                 const double* __restrict__ rz_next, double* __restrict__ clear_a,
                                                 double* __restrict__ clear_b, double* __restrict__ clear_c,
                                                 const int* __restrict__ frozen,
                                                 const double* __restrict__ rz0, double done_factor,
                                                 const double* __restrict__ pr, const uint32_t* __restrict__ agg,
                                                 const float4* __restrict__ yc, int restart, double* __restrict__ X64) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  // X64 != NULL: x is advanced in float64 directly (x += alpha p with float64 products) instead of through the float32
  // increment Xd - with tens of iterations per solve the increment IS the solution, and its float32 rounding is the
  // other source of the white noise that costs the second refinement stage its iterations
  // restart: the step after a residual replacement starts a fresh direction sequence (beta = 0): the refinement
  // stages of the multilevel solve are independent inner solves
  // agg / yc: coarse levels of the preconditioner (amg.cuh), z = D^-1 r + Z0 y; NULL = Jacobi alone
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < 4 * Q; i += blockDim.x) { clear_a[i] = 0.0; clear_b[i] = 0.0; if (clear_c) clear_c[i] = 0.0; }
  if (local >= rows_per_block) return;
  float be[4], al[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    al[j] = pcg_alpha(4 * q + j, rz, pAp, frozen, rz0, done_factor, pr);  // the step k_pcg_update just took
    const double den = rz[4 * q + j], num = rz_next[4 * q + j];
    // after a replacement whose p is not even a descent direction for the new r (p.r <= 0): restart from z
    const bool live = den > done_factor * rz0[4 * q + j] && den > 0.0 && !frozen[4 * q + j] && !(pr && (restart || pr[4 * q + j] <= 0.0));
    const double ratio = live ? num / den : 0.0;
    be[j] = ratio < 1e30 ? (float)ratio : 0.0f;
  }
  const int stride = gridDim.x * rows_per_block;
  int row = blockIdx.x * rows_per_block + local;
  for (; row + stride < n; row += 2 * stride) {
    const size_t i0 = (size_t)row * Q + q, i1 = (size_t)(row + stride) * Q + q;
    const float4 r0 = R[i0], r1 = R[i1];
    float4 p0 = P[i0], p1 = P[i1];
    float4 x0 = make_float4(0.f, 0.f, 0.f, 0.f), x1 = x0;
    if (!X64) { x0 = Xd[i0]; x1 = Xd[i1]; }
    const float d0 = invd[row], d1 = invd[row + stride];
    float4 c0 = make_float4(0.f, 0.f, 0.f, 0.f), c1 = c0;
    if (agg) { c0 = yc[(size_t)agg[row] * Q + q]; c1 = yc[(size_t)agg[row + stride] * Q + q]; }
    // alpha == 0 / beta == 0 never touch the old p (which may hold inf / nan on a dead column)
    if (X64) {
      double2* xa = (double2*)(X64 + 4 * i0);
      double2* xb = (double2*)(X64 + 4 * i1);
      double2 a0 = xa[0], a1 = xa[1], b0 = xb[0], b1 = xb[1];
      if (al[0] != 0.f) { a0.x += (double)al[0] * p0.x; b0.x += (double)al[0] * p1.x; }
      if (al[1] != 0.f) { a0.y += (double)al[1] * p0.y; b0.y += (double)al[1] * p1.y; }
      if (al[2] != 0.f) { a1.x += (double)al[2] * p0.z; b1.x += (double)al[2] * p1.z; }
      if (al[3] != 0.f) { a1.y += (double)al[3] * p0.w; b1.y += (double)al[3] * p1.w; }
      xa[0] = a0; xa[1] = a1; xb[0] = b0; xb[1] = b1;
    } else {
      if (al[0] != 0.f) { x0.x += al[0] * p0.x; x1.x += al[0] * p1.x; }
      if (al[1] != 0.f) { x0.y += al[1] * p0.y; x1.y += al[1] * p1.y; }
      if (al[2] != 0.f) { x0.z += al[2] * p0.z; x1.z += al[2] * p1.z; }
      if (al[3] != 0.f) { x0.w += al[3] * p0.w; x1.w += al[3] * p1.w; }
    }
    p0.x = r0.x * d0 + c0.x + (be[0] != 0.f ? be[0] * p0.x : 0.f); p0.y = r0.y * d0 + c0.y + (be[1] != 0.f ? be[1] * p0.y : 0.f);
    p0.z = r0.z * d0 + c0.z + (be[2] != 0.f ? be[2] * p0.z : 0.f); p0.w = r0.w * d0 + c0.w + (be[3] != 0.f ? be[3] * p0.w : 0.f);
    p1.x = r1.x * d1 + c1.x + (be[0] != 0.f ? be[0] * p1.x : 0.f); p1.y = r1.y * d1 + c1.y + (be[1] != 0.f ? be[1] * p1.y : 0.f);
    p1.z = r1.z * d1 + c1.z + (be[2] != 0.f ? be[2] * p1.z : 0.f); p1.w = r1.w * d1 + c1.w + (be[3] != 0.f ? be[3] * p1.w : 0.f);
    P[i0] = p0; P[i1] = p1;
    if (!X64) { Xd[i0] = x0; Xd[i1] = x1; }
  }
  for (; row < n; row += stride) {
    const size_t i = (size_t)row * Q + q;
    const float4 r = R[i];
    float4 p = P[i], x = make_float4(0.f, 0.f, 0.f, 0.f);
    const float d = invd[row];
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (agg) c = yc[(size_t)agg[row] * Q + q];
    if (X64) {
      double2* xa = (double2*)(X64 + 4 * i);
      double2 a0 = xa[0], a1 = xa[1];
      if (al[0] != 0.f) a0.x += (double)al[0] * p.x;
      if (al[1] != 0.f) a0.y += (double)al[1] * p.y;
      if (al[2] != 0.f) a1.x += (double)al[2] * p.z;
      if (al[3] != 0.f) a1.y += (double)al[3] * p.w;
      xa[0] = a0; xa[1] = a1;
    } else {
      x = Xd[i];
      if (al[0] != 0.f) x.x += al[0] * p.x;
      if (al[1] != 0.f) x.y += al[1] * p.y;
      if (al[2] != 0.f) x.z += al[2] * p.z;
      if (al[3] != 0.f) x.w += al[3] * p.w;
    }
    p.x = r.x * d + c.x + (be[0] != 0.f ? be[0] * p.x : 0.f); p.y = r.y * d + c.y + (be[1] != 0.f ? be[1] * p.y : 0.f);
    p.z = r.z * d + c.z + (be[2] != 0.f ? be[2] * p.z : 0.f); p.w = r.w * d + c.w + (be[3] != 0.f ? be[3] * p.w : 0.f);
    P[i] = p;
    if (!X64) Xd[i] = x;
  }
}

// float64 true residual r = b - L x (replaces the recurrence residual), rr = ||r||^2, rz = r . D^-1 r, pr = p . r.
// The replaced r is not the vector the current direction p was built for (it differs by the rounding the recurrence
// and the float32 x increments accumulated), so the step after a replacement uses the exact line search
// alpha = p.r / p.Ap instead of r.z / p.Ap: with r.z the step overshoots by (|r_true| / |r_recurrence|)^2 and a
// column that is almost converged diverges.
// Rows [row0, row0 + n) of the system (one scene of a block-diagonal batch, or all rows).
__global__ void __launch_bounds__(256) k_true_residual(int row0, int n, int Q, const uint32_t* __restrict__ rowptr,
                                                       const int* __restrict__ col, const float* __restrict__ val,
                                                       const float* __restrict__ diag, const float4* __restrict__ B,
                                                       const double* __restrict__ X, const float4* __restrict__ P,
                                                       float4* __restrict__ R, double* __restrict__ rr,
                                                       double* __restrict__ rz, double* __restrict__ pr) {
  extern __shared__ double sm[];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double a_rr[4] = {0, 0, 0, 0}, a_rz[4] = {0, 0, 0, 0}, a_pr[4] = {0, 0, 0, 0};
  if (active) {
    for (int lrow = blockIdx.x * rows_per_block + local; lrow < n; lrow += gridDim.x * rows_per_block) {
      const int row = row0 + lrow;
      const size_t i = (size_t)row * Q + q;
      const double d = (double)diag[row];
      const double2* xp = (const double2*)(X + 4 * i);
      double2 x0 = xp[0], x1 = xp[1];
      double y[4] = {d * x0.x, d * x0.y, d * x1.x, d * x1.y};
      for (uint32_t a = rowptr[row]; a < rowptr[row + 1]; ++a) {
        const double w = (double)val[a];
        const double2* t = (const double2*)(X + 4 * ((size_t)col[a] * Q + q));
        double2 t0 = t[0], t1 = t[1];
        y[0] += w * t0.x; y[1] += w * t0.y; y[2] += w * t1.x; y[3] += w * t1.y;
      }
      const float4 b = B[i];
      double r[4] = {(double)b.x - y[0], (double)b.y - y[1], (double)b.z - y[2], (double)b.w - y[3]};
#pragma unroll
      for (int j = 0; j < 4; ++j) a_rr[j] += r[j] * r[j];
      float4 rf = make_float4((float)r[0], (float)r[1], (float)r[2], (float)r[3]);
      R[i] = rf;
      const double di = 1.0 / d;
      a_rz[0] += (double)rf.x * rf.x * di; a_rz[1] += (double)rf.y * rf.y * di;
      a_rz[2] += (double)rf.z * rf.z * di; a_rz[3] += (double)rf.w * rf.w * di;
      const float4 pv = P[i];
      a_pr[0] += (double)pv.x * rf.x; a_pr[1] += (double)pv.y * rf.y;
      a_pr[2] += (double)pv.z * rf.z; a_pr[3] += (double)pv.w * rf.w;
    }
  }
  reduce_columns(a_rr, q, Q, active, sm, rr, true);
  __syncthreads();
  reduce_columns(a_rz, q, Q, active, sm, rz, true);
  __syncthreads();
  reduce_columns(a_pr, q, Q, active, sm, pr, true);
}

// ---- small systems: `chunk` whole PCG iterations in ONE persistent launch -------------------------------------------
// A 640 x 480 scene gives ~40 k rows.  With one launch per kernel an iteration costs ~20 us although it moves only a few
// megabytes: every kernel is a chain of dependent L2 round trips (rowptr -> col -> P, scalars -> rows) on a handful of
// rows per thread, plus three launches.  k_pcg_chunk keeps the system ON CHIP for a whole chunk of iterations: a thread
// owns the same (at most M) rows throughout and holds their r, x increment, p and Ap in registers; the CSR slice of
// the CTA's rows sits in shared memory; the only global traffic of an iteration is the gather of the neighbours' p
// (read through L2, ld.global.cg), the store of the thread's own new p, and the per-column dot products.  The three
// phases of an iteration (SpMM, residual update, direction update) are separated by grid-wide barriers (all CTAs are
// co-resident: cooperative launch).  Same arithmetic, the same order of operations per row and the same fp64
// accumulators as k_spmm / k_pcg_update / k_pcg_dir.
__constant__ unsigned c_coop_sleep = 0;
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    // release: this CTA's writes (ordered before by the bar.sync) become visible before the count; acquire on the poll
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
    unsigned seen;
    while (true) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
      if (seen >= target) break;
      if (c_coop_sleep) __nanosleep(c_coop_sleep);   // hundreds of CTAs poll one L2 line: leave room for the arrivals
    }
  }
  __syncthreads();
}

// Per-column dot products without atomics: every CTA reduces its threads' partial sums (fixed order: rows of a warp
// lane-strided, then an xor tree) and stores them at part[blockIdx][column]; after the grid barrier every CTA adds the
// gridDim.x partials of each column in the same fixed order.  Distinct addresses (no hot L2 line), and the sums - hence
// the whole solve - are bitwise reproducible from run to run.
__device__ __forceinline__ void chunk_block_partials(const double acc[4], int Q, double* arr, double* part) {
  const int Kp = 4 * Q, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, rows_per_block = blockDim.x / Q;
#pragma unroll
  for (int j = 0; j < 4; ++j) arr[threadIdx.x * 4 + j] = acc[j];
  __syncthreads();
  for (int c = warp; c < Kp; c += (int)(blockDim.x >> 5)) {
    const int q = c >> 2, jj = c & 3;
    double v = 0.0;
    for (int l = lane; l < rows_per_block; l += 32) v += arr[(l * Q + q) * 4 + jj];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    if (lane == 0) part[(size_t)blockIdx.x * Kp + c] = v;
  }
}
__device__ __forceinline__ double chunk_column_total(const double* part, int Kp, int c) {
  const int lane = threadIdx.x & 31;
  // four independent loads in flight per lane (an L2 round trip each); fixed order of additions
  double v0 = 0.0, v1 = 0.0, v2 = 0.0, v3 = 0.0;
  const int G = (int)gridDim.x;
  int b = lane;
  for (; b + 96 < G; b += 128) {
    const double a0 = __ldcg(&part[(size_t)b * Kp + c]), a1 = __ldcg(&part[(size_t)(b + 32) * Kp + c]);
    const double a2 = __ldcg(&part[(size_t)(b + 64) * Kp + c]), a3 = __ldcg(&part[(size_t)(b + 96) * Kp + c]);
    v0 += a0; v1 += a1; v2 += a2; v3 += a3;
  }
  for (; b < G; b += 32) v0 += __ldcg(&part[(size_t)b * Kp + c]);
  double v = (v0 + v1) + (v2 + v3);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

// CTA b owns the row tiles b, b + gridDim.x, ... (M of them, rows_per_block rows each); `cap` = entries of the largest
// tile.  Dynamic shared memory: M * cap {col, val} pairs, the block-reduction scratch (blockDim.x * 4 doubles), and
// per column: rz (two generations), p.r of the replacement, alpha, beta.
// part: 2 * gridDim.x * Kp doubles (p.Ap partials, then r.z partials).
template <int M>
__global__ void __launch_bounds__(256) k_pcg_chunk(int n, int Q, int chunk, int it0, int kstride, int cap,
                                                   const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                                                   const float* __restrict__ val, const float* __restrict__ diag,
                                                   const float* __restrict__ invd, float4* P, const float4* R,
                                                   float4* Xd, double* pAp3, double* rz3, const double* rz0,
                                                   const double* pr_first, const int* frozen, double done_factor,
                                                   unsigned* bar, double* part, long long* dbg) {
  extern __shared__ __align__(16) unsigned char chunk_smem[];
  const int Kp = 4 * Q;
  int2* ent = (int2*)chunk_smem;                                  // [M][cap] {col, val bits}
  double* arr = (double*)(chunk_smem + (size_t)M * cap * 8);      // [blockDim.x][4]
  double* s_rz = arr + blockDim.x * 4;                            // [2][Kp]
  double* s_pr = s_rz + 2 * Kp;                                   // [Kp]
  double* s_rz0 = s_pr + Kp;                                      // [Kp] done_factor * rz of the initial residual
  float* s_al = (float*)(s_rz0 + Kp);                             // [Kp]
  float* s_be = s_al + Kp;                                        // [Kp]
  int* s_frozen = (int*)(s_be + Kp);                              // [Kp]
  double* partA = part;
  double* partB = part + (size_t)gridDim.x * Kp;
  long long tk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#define TCS_TICK(slot) do { if (dbg) { const long long _t = clock64(); tk[slot] += _t - tprev; tprev = _t; } } while (0)
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const bool active = local < rows_per_block;
  int row[M], e0[M], e1[M];
  float4 r[M], xd[M], p[M], ap[M];
  float dg[M], iv[M];
  for (int c = threadIdx.x; c < Kp; c += blockDim.x) {
    s_rz[c] = __ldcg(&rz3[(it0 % 3) * kstride + c]);
    s_pr[c] = pr_first ? __ldcg(&pr_first[c]) : 0.0;
    s_rz0[c] = done_factor * __ldcg(&rz0[c]);
    s_frozen[c] = __ldcg(&frozen[c]);
  }
#pragma unroll
  for (int m = 0; m < M; ++m) {
    const int tile_row0 = (blockIdx.x + m * gridDim.x) * rows_per_block;
    row[m] = tile_row0 + local;
    const bool ok = active && row[m] < n;
    if (!ok) row[m] = -1;
    const int tile_row1 = min(n, tile_row0 + rows_per_block);
    const uint32_t base = tile_row0 < n ? rowptr[tile_row0] : 0u;
    const uint32_t end = tile_row0 < n ? rowptr[tile_row1] : 0u;
    for (uint32_t a = base + threadIdx.x; a < end; a += blockDim.x)
      ent[m * cap + (a - base)] = make_int2(col[a], __float_as_int(val[a]));
    e0[m] = e1[m] = 0;
    r[m] = xd[m] = p[m] = ap[m] = make_float4(0.f, 0.f, 0.f, 0.f);
    dg[m] = iv[m] = 0.f;
    if (ok) {
      e0[m] = m * cap + (int)(rowptr[row[m]] - base);
      e1[m] = m * cap + (int)(rowptr[row[m] + 1] - base);
      const size_t i = (size_t)row[m] * Q + q;
      r[m] = R[i]; xd[m] = Xd[i]; p[m] = __ldcg(&P[i]);
      dg[m] = diag[row[m]]; iv[m] = invd[row[m]];
    }
  }
  __syncthreads();
  unsigned target = 0;
  int cur = 0;  // generation of s_rz holding r.z of the current iteration
  for (int j = 0; j < chunk; ++j) {
    const bool after_replacement = j == 0 && pr_first != nullptr;
    // ---- AP = L P, p . Ap
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int m = 0; m < M; ++m) {
      if (row[m] < 0) continue;
      float4 y = make_float4(dg[m] * p[m].x, dg[m] * p[m].y, dg[m] * p[m].z, dg[m] * p[m].w);
      int a = e0[m];
      const int a1 = e1[m];
      // eight gathers in flight (an L2 round trip each); the tail batch is padded with zero-weight reads of the own
      // row, so the sum keeps CSR order and every load is issued before the first FMA of its batch
      const size_t own = (size_t)row[m] * Q + q;
      for (; a < a1; a += 8) {
        float4 t[8];
        float w[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const bool in = a + u < a1;
          const int2 e = in ? ent[a + u] : make_int2(0, 0);
          w[u] = __int_as_float(e.y);
          t[u] = __ldcg(&P[in ? (size_t)e.x * Q + q : own]);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (a + u < a1) { y.x += w[u] * (t[u].x - p[m].x); y.y += w[u] * (t[u].y - p[m].y); y.z += w[u] * (t[u].z - p[m].z); y.w += w[u] * (t[u].w - p[m].w); }
        }
      }
      ap[m] = y;
      acc[0] += (double)p[m].x * (double)y.x; acc[1] += (double)p[m].y * (double)y.y;
      acc[2] += (double)p[m].z * (double)y.z; acc[3] += (double)p[m].w * (double)y.w;
    }
    TCS_TICK(0);
    chunk_block_partials(acc, Q, arr, partA);
    TCS_TICK(1);
    grid_barrier(bar, target);
    TCS_TICK(2);
    // ---- alpha (pcg_alpha), r -= alpha Ap, r . D^-1 r
    for (int c = warp; c < Kp; c += nwarps) {
      const double den = chunk_column_total(partA, Kp, c);
      if (lane == 0) {
        const double num = after_replacement ? s_pr[c] : s_rz[cur * Kp + c];
        const bool live = den > 0.0 && num > s_rz0[c] && !s_frozen[c];
        const double ratio = live ? num / den : 0.0;
        s_al[c] = ratio < 1e30 ? (float)ratio : 0.0f;
      }
    }
    __syncthreads();
    float al[4] = {0.f, 0.f, 0.f, 0.f};
    acc[0] = acc[1] = acc[2] = acc[3] = 0.0;
    if (active) {
#pragma unroll
      for (int c = 0; c < 4; ++c) al[c] = s_al[4 * q + c];
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
      if (row[m] < 0) continue;
      if (al[0] != 0.f) r[m].x -= al[0] * ap[m].x;
      if (al[1] != 0.f) r[m].y -= al[1] * ap[m].y;
      if (al[2] != 0.f) r[m].z -= al[2] * ap[m].z;
      if (al[3] != 0.f) r[m].w -= al[3] * ap[m].w;
      const double e = (double)iv[m];
      acc[0] += (double)r[m].x * (double)r[m].x * e; acc[1] += (double)r[m].y * (double)r[m].y * e;
      acc[2] += (double)r[m].z * (double)r[m].z * e; acc[3] += (double)r[m].w * (double)r[m].w * e;
    }
    TCS_TICK(3);
    chunk_block_partials(acc, Q, arr, partB);
    TCS_TICK(4);
    grid_barrier(bar, target);
    TCS_TICK(5);
    // ---- beta, xd += alpha p, p = D^-1 r + beta p
    for (int c = warp; c < Kp; c += nwarps) {
      const double num = chunk_column_total(partB, Kp, c);
      if (lane == 0) {
        const double den = s_rz[cur * Kp + c];
        const bool live = den > s_rz0[c] && den > 0.0 && !s_frozen[c] && !(after_replacement && s_pr[c] <= 0.0);
        const double ratio = live ? num / den : 0.0;
        s_be[c] = ratio < 1e30 ? (float)ratio : 0.0f;
        s_rz[(cur ^ 1) * Kp + c] = num;
      }
    }
    __syncthreads();
    cur ^= 1;
    if (active) {
      float be[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) be[c] = s_be[4 * q + c];
#pragma unroll
      for (int m = 0; m < M; ++m) {
        if (row[m] < 0) continue;
        float4 x = xd[m], pv = p[m];
        const float d = iv[m];
        if (al[0] != 0.f) x.x += al[0] * pv.x;
        if (al[1] != 0.f) x.y += al[1] * pv.y;
        if (al[2] != 0.f) x.z += al[2] * pv.z;
        if (al[3] != 0.f) x.w += al[3] * pv.w;
        pv.x = r[m].x * d + (be[0] != 0.f ? be[0] * pv.x : 0.f); pv.y = r[m].y * d + (be[1] != 0.f ? be[1] * pv.y : 0.f);
        pv.z = r[m].z * d + (be[2] != 0.f ? be[2] * pv.z : 0.f); pv.w = r[m].w * d + (be[3] != 0.f ? be[3] * pv.w : 0.f);
        xd[m] = x; p[m] = pv;
        P[(size_t)row[m] * Q + q] = pv;
      }
    }
    TCS_TICK(6);
    grid_barrier(bar, target);
    TCS_TICK(7);
  }
  // r is not written back: every chunk is followed by the float64 residual replacement, which rewrites R (and r.z).
  // The global accumulators of the per-kernel path stay cleared.
#pragma unroll
  for (int m = 0; m < M; ++m)
    if (row[m] >= 0) Xd[(size_t)row[m] * Q + q] = xd[m];
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < 3 * kstride; i += blockDim.x) { pAp3[i] = 0.0; rz3[i] = 0.0; }
  if (dbg && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1))
    for (int i = 0; i < 8; ++i) dbg[(blockIdx.x ? 8 : 0) + i] = tk[i];
#undef TCS_TICK
}

// assignColors (impl/random_walker.hpp:209-222): first maximal potential of the row; rows whose potentials are all
// zero (seedless components, :191-192) stay unlabelled.  Also scatters potentials (getPotentials, :224-243).
__global__ void k_argmax(int n, int K, int Kp, int k0, int k1, const double* __restrict__ X,
                         const uint32_t* __restrict__ row_vertex, uint32_t* __restrict__ colours,
                         float* __restrict__ potentials, size_t V, unsigned long long* __restrict__ near_ties,
                         float* __restrict__ best_val) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const double* x = X + (size_t)r * Kp;
  const uint32_t v = row_vertex[r];
  if (potentials)
    for (int k = 0; k < K; ++k) potentials[(size_t)k * V + v] = (float)x[k];
  // the maximum runs over the solved columns [k0, k1) only (all of them unless this is a label-column shard)
  float best = (float)x[k0], second = -INFINITY;
  int bi = k0;
  bool any = best != 0.0f;
  for (int k = k0 + 1; k < k1; ++k) {
    float f = (float)x[k];
    any = any || f != 0.0f;
    if (f > best) { second = best; best = f; bi = k; }
    else if (f > second) second = f;
  }
  if (any) {
    colours[v] = (uint32_t)bi + 1u;
    best_val[v] = best;
    if (k1 - k0 > 1 && best - second < 1e-6f) atomicAdd(near_ties, 1ull);
  }
}
// best potential of the vertices that are not rows: seeds carry potential 1 for their colour, the rest 0
__global__ void k_best_init(size_t V, const uint32_t* __restrict__ colours, float* __restrict__ best_val) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v < V) best_val[v] = colours[v] ? 1.0f : 0.0f;
}

__global__ void k_seed_potentials(size_t V, const uint32_t* __restrict__ seed_colour, float* __restrict__ potentials) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t c = seed_colour[v];
  if (c) potentials[(size_t)(c - 1) * V + v] = 1.0f;
}

}  // namespace

extern "C" void tcs_solver_opts_default(tcs_solver_opts* o) {
  if (!o) return;
  o->tolerance = 1e-6;
  o->max_iterations = 200000;
  o->check_every = 50;
  o->store_potentials = 0;
  o->time_spmm = 0;
  o->column_begin = 0;
  o->column_count = 0;
}

namespace {

struct PcgLaunch {
  unsigned grid, block;
  size_t smem;
  // grid-stride kernels with their own register budgets: exactly the resident CTAs (a grid of 8 CTAs per SM leaves a
  // 78-register kernel with 2.67 waves: the last one runs a third empty)
  unsigned grid_update, grid_dir;
};

// resident CTAs of `fn` on the whole device, capped by the CTAs the rows need
template <typename F>
unsigned resident_grid(tcs_ctx* ctx, F fn, int block, size_t smem, unsigned need, unsigned fallback_per_sm) {
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, block, smem) != cudaSuccess || nb < 1) {
    cudaGetLastError();
    nb = (int)fallback_per_sm;
  }
  return std::max(1u, std::min(need, (unsigned)ctx->sm_count * (unsigned)nb));
}

__global__ void k_tile_max_nnz(int n, int R, const uint32_t* __restrict__ rowptr, uint32_t* __restrict__ out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  int r0 = t * R;
  if (r0 >= n) return;
  int r1 = min(n, r0 + R);
  atomicMax(out, rowptr[r1] - rowptr[r0]);
}

int spmm_far_lists(tcs_ctx* ctx, tcs_system& S, int R, int H, uint32_t* nfar_out, uint32_t* nuniq_out, uint32_t* max_far_out);
int spmm_plan_ring(tcs_ctx* ctx, tcs_system& S);

// Choose the SpMM kernel and its shape for this system: k_spmm_ring for large systems, k_spmm_tiled (two CTAs of 256
// threads per SM share the 227 KB) for mid-sized ones, k_spmm for tiny or very dense ones.  TCS_SPMM=v1|tiled forces.
int spmm_plan(tcs_ctx* ctx, tcs_system& S) {
  S.tiled = false;
  S.ring = false;
  S.plan_stale = false;
  const char* force = getenv("TCS_SPMM");
  if (force && !strcmp(force, "v1")) return TCS_OK;
  if (!(force && !strcmp(force, "tiled"))) {
    TRY(spmm_plan_ring(ctx, S));
    if (S.ring) return TCS_OK;
  }
  const int n = S.n, Q = S.Kp / 4;
  if (n < 8192 || S.nnz_off == 0) return TCS_OK;  // tiny systems are launch-bound, not bandwidth-bound
  const int rows_per_pass = 256 / Q;
  int R = getenv("TCS_TILE_R") ? atoi(getenv("TCS_TILE_R")) : 256;
  R = std::min(508, std::max(64, (R + 3) & ~3));  // multiple of 4: the diagonal slice is a 16-byte aligned bulk copy;
                                                   // <= 508: the rowptr prefetch holds two values per thread
  (void)rows_per_pass;
  uint32_t* d_max = (uint32_t*)ctx->d_scal + 16;
  CK(cudaMemsetAsync(d_max, 0, 8, ctx->stream));
  const int ntiles = (n + R - 1) / R;
  LAUNCH(ctx, k_tile_max_nnz, div_up(ntiles, 256), 256, 0, n, R, S.rowptr.p, d_max);
  uint32_t mx = 0;
  TRY(tcs_impl_read_u32(ctx, d_max, &mx));
  const int budget = (getenv("TCS_TILE_SMEM_KB") ? atoi(getenv("TCS_TILE_SMEM_KB")) : 110) * 1024;
  const int rowbytes = 16 * Q;
  SpmmTile T;
  T.R = R;
  T.cap = (int)((mx + 1u) & ~1u) + 4;
  T.farcap = getenv("TCS_TILE_FAR") ? atoi(getenv("TCS_TILE_FAR")) : std::max(64, R);
  T.off_rp = 16;
  T.off_diag = (T.off_rp + (R + 3) * 4 + 15) & ~15;
  T.off_cv = (T.off_diag + R * 4 + 16 + 15) & ~15;
  T.off_red = T.off_cv + T.cap * 8;
  T.off_p = (T.off_red + S.Kp * 8 + 127) & ~127;
  int left = budget - T.off_p - (R + T.farcap) * rowbytes;
  if (left < 0) return TCS_OK;  // very dense rows (user graphs): keep the generic kernel
  int H = left / (2 * rowbytes);
  if (getenv("TCS_TILE_H")) H = std::min(H, atoi(getenv("TCS_TILE_H")));
  T.H = std::min(H, 4096);

  uint32_t nfar = 0, nuniq = 0, max_far = 0;
  TRY(spmm_far_lists(ctx, S, T.R, T.H, &nfar, &nuniq, &max_far));
  if ((int)max_far > T.farcap) T.farcap = (int)((max_far + 15u) & ~15u);
  T.total = T.off_p + (R + 2 * T.H + T.farcap) * rowbytes;
  if (T.total > 113 * 1024 && T.total > budget) {
    if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] spmm tiled: far buffer of %u rows does not fit, using v1\n", max_far);
    return TCS_OK;
  }
  // function attributes are per device: set on every plan (cheap), not once per process
  CK(cudaFuncSetAttribute(k_spmm_tiled, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  S.tile = T;
  S.tiled = true;
  if (getenv("TCS_DEBUG"))
    fprintf(stderr, "[tcs] spmm tiled: R=%d H=%d cap=%d farcap=%d smem=%d B (max tile nnz %u, far entries %u = %.1f%%, distinct far rows %u, max per tile %u)\n",
            T.R, T.H, T.cap, T.farcap, T.total, mx, nfar, 100.0 * nfar / (double)S.nnz_off, nuniq, max_far);
  return TCS_OK;
}

// Resolve the far entries of a tiling (R rows per tile, halo H): cv = {col or -1 - far slot, weight}, per-tile
// de-duplicated far-row lists.  Shared by k_spmm_tiled and k_spmm_ring.
int spmm_far_lists(tcs_ctx* ctx, tcs_system& S, int R, int H, uint32_t* nfar_out, uint32_t* nuniq_out, uint32_t* max_far_out) {
  const int n = S.n;
  const int ntiles = (n + R - 1) / R;
  const size_t nnz = (size_t)S.nnz_off;
  uint32_t* d_max = (uint32_t*)ctx->d_scal + 16;
  CK(cudaMemsetAsync(d_max, 0, 8, ctx->stream));
  TRY(S.cv.alloc(ctx, nnz + 8));
  TRY(S.tile_far_ptr.alloc(ctx, (size_t)ntiles + 2));
  DBuf<uint32_t> flags, pos;
  TRY(flags.alloc(ctx, nnz + 1)); TRY(pos.alloc(ctx, nnz + 1));
  CK(cudaMemsetAsync(S.cv.p + nnz, 0, 8 * sizeof(int2), ctx->stream));
  LAUNCH(ctx, k_far_flags, div_up(n, 128), 128, 0, n, R, H, S.rowptr.p, S.col.p, S.val.p, S.cv.p, flags.p);
  TRY(tcs_impl_scan_u32(ctx, flags.p, pos.p, nnz));
  uint32_t nfar = 0, max_far = 0, nuniq = 0;
  TRY(tcs_impl_read_u32(ctx, pos.p + nnz, &nfar));
  if (nfar) {
    DBuf<unsigned long long> k0, k1;
    DBuf<uint32_t> e0, e1, head, uidx;
    TRY(k0.alloc(ctx, nfar)); TRY(k1.alloc(ctx, nfar)); TRY(e0.alloc(ctx, nfar)); TRY(e1.alloc(ctx, nfar));
    TRY(head.alloc(ctx, (size_t)nfar + 1)); TRY(uidx.alloc(ctx, (size_t)nfar + 1));
    LAUNCH(ctx, k_far_scatter, div_up(n, 128), 128, 0, n, R, S.rowptr.p, S.col.p, flags.p, pos.p, k0.p, e0.p);
    int tbits = 1;
    while ((1 << tbits) < ntiles + 1) ++tbits;
    TRY(tcs_impl_sort_pairs_u64(ctx, (uint64_t*)k0.p, (uint64_t*)k1.p, e0.p, e1.p, nfar, 32 + tbits));
    LAUNCH(ctx, k_far_heads, div_up(nfar, 256), 256, 0, k1.p, nfar, head.p);
    TRY(tcs_impl_scan_u32(ctx, head.p, uidx.p, nfar));
    LAUNCH(ctx, k_far_tile_ptr, div_up(ntiles + 1, 256), 256, 0, ntiles, k1.p, nfar, uidx.p, S.tile_far_ptr.p);
    TRY(tcs_impl_read_u32(ctx, uidx.p + nfar, &nuniq));
    TRY(S.far_rows.alloc(ctx, (size_t)nuniq + 4));
    LAUNCH(ctx, k_far_assign, div_up(nfar, 256), 256, 0, nfar, k1.p, e1.p, head.p, uidx.p, S.tile_far_ptr.p, S.cv.p,
           S.far_rows.p, d_max + 1);
    TRY(tcs_impl_read_u32(ctx, d_max + 1, &max_far));
    // tile_far_ptr[ntiles + 1] is read by the ring producer's prefetch
    CK(cudaMemcpyAsync(S.tile_far_ptr.p + ntiles + 1, S.tile_far_ptr.p + ntiles, 4, cudaMemcpyDeviceToDevice, ctx->stream));
  } else {
    CK(cudaMemsetAsync(S.tile_far_ptr.p, 0, ((size_t)ntiles + 2) * 4, ctx->stream));
    TRY(S.far_rows.alloc(ctx, 4));
  }
  *nfar_out = nfar; *nuniq_out = nuniq; *max_far_out = max_far;
  return TCS_OK;
}

static inline int env_int(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }

// Shape of k_spmm_ring for this system (one CTA of 17 warps per SM owns the whole 227 KB).
int spmm_plan_ring(tcs_ctx* ctx, tcs_system& S) {
  S.ring = false;
  const int n = S.n, Q = S.Kp / 4;
  if (Q < 1 || Q > 8 || S.nnz_off == 0) return TCS_OK;
  SpmmRing T;
  memset(&T, 0, sizeof T);
  const int rowbytes = 16 * Q;
  // 48-byte rows (9..12 columns) have no conflict-free placement (18 M bank conflicts on C4, 190 us).  TCS_RING_PAD=1
  // pads them to a 64-byte pitch in shared memory and runs them like 64-byte rows with one idle lane per row group;
  // a bulk copy cannot pad, and filling the ring with 16-byte cp.async instead costs more (+28 us measured on 64-byte
  // rows) than the conflicts it removes (196 us), so it stays off.
  // What works for them (TCS_RING_Q3PAIR=1, default): four lanes per row group (one idle), so a quarter warp gathers
  // exactly two rows, and a second ring copy shifted by 48 bytes: for every pair of row starts one combination of
  // copies is bank-conflict-free, and the plan picks it per pair of entries (ring_pair_choice).
  const bool pad = Q == 3 && env_int("TCS_RING_PAD", 0);
  const bool q3pair = Q == 3 && !pad && env_int("TCS_RING_Q3PAIR", 1) && env_int("TCS_RING_DUP", 1);
  T.Qp = (pad || q3pair) ? 4 : Q;
  T.pitch = pad ? 64 : rowbytes;
  T.fill_async = pad ? 1 : 0;
  T.G = std::min(16, 32 / T.Qp);  // 16-byte rows: 16 rows per warp step keep the tile (and its stage buffers) at 256 rows
  T.R = kRingWarps * T.G;
  T.S = std::min(8, std::max(2, env_int("TCS_RING_S", 3)));
  T.dup = ((T.pitch == 64 || q3pair) && env_int("TCS_RING_DUP", 1)) ? 1 : 0;
  T.dup_shift = q3pair ? 48 : 64;
  const int ntiles = (n + T.R - 1) / T.R;
  // small systems are launch-bound and stay with the other kernels (TCS_RING_MIN_TILES=1 forces the ring, for tests)
  if (ntiles < env_int("TCS_RING_MIN_TILES", 4 * ctx->sm_count)) return TCS_OK;
  T.nctas = std::min(ntiles, ctx->sm_count);
  // row order + sliced entry layout of every tile (independent of the ring size)
  DBuf<uint32_t> tile_entries, tile_start;
  const int metabytes = ring_meta_bytes(T.R);
  TRY(S.ring_meta.alloc(ctx, (size_t)ntiles * metabytes + 64));
  TRY(tile_entries.alloc(ctx, (size_t)ntiles + 1));
  TRY(tile_start.alloc(ctx, (size_t)ntiles + 1));
  LAUNCH(ctx, k_ring_perm, ntiles, T.R, (size_t)T.R * 8, n, T.R, T.G, env_int("TCS_RING_SORT", 1), S.rowptr.p, S.gdiag.p,
         S.ring_meta.p, tile_entries.p);
  TRY(tcs_impl_scan_u32(ctx, tile_entries.p, tile_start.p, ntiles));
  uint32_t total_entries = 0, mx = 0;
  TRY(tcs_impl_read_u32(ctx, tile_start.p + ntiles, &total_entries));
  {
    uint32_t* d_max = (uint32_t*)ctx->d_scal + 16;
    CK(cudaMemsetAsync(d_max, 0, 8, ctx->stream));
    LAUNCH(ctx, k_tile_max_nnz, div_up(ntiles, 256), 256, 0, ntiles, 1, tile_start.p, d_max);
    TRY(tcs_impl_read_u32(ctx, d_max, &mx));
  }
  if (mx > 60000u) return TCS_OK;  // slice starts are 16-bit
  T.cap = (int)mx;
  const int budget = 232448;  // 227 KB: the per-block opt-in maximum of sm_100
  auto layout = [&](int farcap, int ring_rows) {
    T.farcap = farcap;
    T.ring_rows = ring_rows;
    T.off_bar = 0;
    T.off_red = 256;
    T.off_ids = (T.off_red + S.Kp * 8 + 15) & ~15;
    T.ids_bytes = (farcap + 8) * 4;
    T.off_stage0 = (T.off_ids + 2 * T.ids_bytes + 127) & ~127;
    T.off_meta = 0;
    T.off_cv = (metabytes + 15) & ~15;
    T.off_far = T.off_cv + T.cap * 6;
    T.stage_bytes = (T.off_far + farcap * rowbytes + 127) & ~127;
    T.off_ring = T.off_stage0 + T.S * T.stage_bytes;
    T.ring_bytes = ring_rows * T.pitch;
    T.total = T.off_ring + T.ring_bytes * (T.dup ? 2 : 1) + (T.dup ? 128 : 0);
    return T.total;
  };
  int farcap = env_int("TCS_RING_FAR", std::max(64, T.R));
  int ring_rows = 1 << 16;
  while (ring_rows > 0 && layout(farcap, ring_rows) > budget) ring_rows >>= 1;
  if (getenv("TCS_RING_ROWS")) ring_rows = std::min(ring_rows, atoi(getenv("TCS_RING_ROWS")));
  uint32_t nfar = 0, nuniq = 0, max_far = 0;
  for (;; ring_rows >>= 1) {  // a smaller ring means a smaller halo and more far rows: shrink until everything fits
    T.H = (ring_rows - T.S * T.R) / 2;
    if (T.H < T.R / 2) return TCS_OK;  // very dense rows (user graphs): keep the generic kernels
    TRY(spmm_far_lists(ctx, S, T.R, T.H, &nfar, &nuniq, &max_far));
    if ((int)max_far > farcap) farcap = (int)((max_far + 7u) & ~7u);
    if (layout(farcap, ring_rows) <= budget) break;
    if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] spmm ring: %d ring rows + %u far rows do not fit, shrinking\n", ring_rows, max_far);
  }
  TRY(S.ring_desc.alloc(ctx, (size_t)ntiles + 4));
  LAUNCH(ctx, k_ring_desc, div_up(ntiles + 4, 256), 256, 0, ntiles, tile_start.p, S.tile_far_ptr.p, S.ring_desc.p);
  TRY(S.ring_blob.alloc(ctx, (size_t)total_entries * 6 + 64));
  CK(cudaMemsetAsync(S.ring_blob.p, 0, (size_t)total_entries * 6 + 64, ctx->stream));
  LAUNCH(ctx, k_ring_fill, div_up((size_t)ntiles * T.R, 128), 128, 0, n, Q, T, ntiles, S.rowptr.p, S.ring_meta.p,
         tile_start.p, S.cv.p, S.ring_blob.p);
  S.cv.release();  // plan-time scratch only: the ring kernel reads ring_meta + ring_blob
  S.ring_entries = total_entries;
  // function attributes are per device: set on every plan (cheap), not once per process
  CK(cudaFuncSetAttribute(k_spmm_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, budget));
  S.rg = T;
  // Contiguous tile ranges of equal work.  First guess: a tile costs its entries + far rows + a per-row prologue.
  // Then the kernel itself is the cost model: it reports every CTA's cycles, the tiles of a CTA are re-weighted by
  // measured / modelled time, and the ranges are cut again (the iteration matrix is launched thousands of times).
  std::vector<uint4> hdesc((size_t)ntiles);
  CK(cudaMemcpyAsync(hdesc.data(), S.ring_desc.p, (size_t)ntiles * sizeof(uint4), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  std::vector<double> w((size_t)ntiles);
  for (int t = 0; t < ntiles; ++t)
    w[t] = (double)hdesc[t].y + 2.0 * (hdesc[t].w - hdesc[t].z) + 2.0 * T.R;
  std::vector<int> cta_ptr((size_t)T.nctas + 1);
  auto cut = [&]() {
    std::vector<double> pre((size_t)ntiles + 1, 0.0);
    for (int t = 0; t < ntiles; ++t) pre[t + 1] = pre[t] + w[t];
    int t = 0;
    for (int b = 0; b < T.nctas; ++b) {
      cta_ptr[b] = t;
      const double goal = pre[ntiles] * (b + 1) / T.nctas;
      const int limit = ntiles - (T.nctas - 1 - b);  // leave one tile for every later CTA
      ++t;
      // take tile t while that brings the range closer to its goal
      while (t < limit && pre[t + 1] - goal <= goal - pre[t]) ++t;
      t = std::min(t, limit);
    }
    cta_ptr[T.nctas] = ntiles;
  };
  TRY(S.ring_cta_ptr.alloc(ctx, (size_t)T.nctas + 1));
  TRY(S.ring_clk.alloc(ctx, (size_t)T.nctas));
  const int rounds = (S.P.p && S.AP.p) ? env_int("TCS_RING_BALANCE", 3) : 0;
  std::vector<long long> clk((size_t)T.nctas);
  double imbalance = 0.0;
  for (int round = 0;; ++round) {
    cut();
    CK(cudaMemcpyAsync(S.ring_cta_ptr.p, cta_ptr.data(), ((size_t)T.nctas + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (round >= rounds) break;
    DBuf<double> scratch;
    TRY(scratch.alloc(ctx, S.Kp));
    for (int rep = 0; rep < 2; ++rep)  // the second launch runs with warm instruction / L2 state
      LAUNCH(ctx, k_spmm_ring, (unsigned)T.nctas, kRingThreads, (size_t)T.total, n, Q, T, S.ring_meta.p, S.ring_blob.p, S.far_rows.p,
             S.ring_desc.p, S.ring_cta_ptr.p, (const float4*)S.P.p, (float4*)S.AP.p, scratch.p, S.ring_clk.p);
    CK(cudaMemcpyAsync(clk.data(), S.ring_clk.p, (size_t)T.nctas * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    long long mx_clk = 0, sum_clk = 0;
    for (int b = 0; b < T.nctas; ++b) {
      double model = 0.0;
      for (int t = cta_ptr[b]; t < cta_ptr[b + 1]; ++t) model += w[t];
      const double f = model > 0.0 ? (double)clk[b] / model : 1.0;
      for (int t = cta_ptr[b]; t < cta_ptr[b + 1]; ++t) w[t] *= f;
      mx_clk = std::max(mx_clk, clk[b]);
      sum_clk += clk[b];
    }
    imbalance = (double)mx_clk * T.nctas / (double)std::max(1ll, sum_clk);
    if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] spmm ring balance round %d: max/mean CTA cycles = %.3f (max %lld)\n", round, imbalance, mx_clk);
  }
  S.ring = true;
  if (getenv("TCS_DEBUG"))
    fprintf(stderr, "[tcs] spmm ring: R=%d H=%d S=%d ring=%d rows dup=%d cap=%d farcap=%d stage=%d B smem=%d B ctas=%d (max tile entries %u, padding %.1f%%, far entries %u = %.1f%%, distinct far rows %u, max per tile %u)\n",
            T.R, T.H, T.S, T.ring_rows, T.dup, T.cap, T.farcap, T.stage_bytes, T.total, T.nctas, mx,
            100.0 * ((double)total_entries / (double)S.nnz_off - 1.0), nfar,
            100.0 * nfar / (double)S.nnz_off, nuniq, max_far);
  return TCS_OK;
}

PcgLaunch pcg_launch_shape(tcs_ctx* ctx, int n, int Q, bool vector_kernels = true);

// AP = L P, pAp += p . Ap per column
void launch_spmm(tcs_ctx* ctx, tcs_system& S, const float4* P4, float4* AP4, double* pAp) {
  const int n = S.n, Q = S.Kp / 4;
  if (S.ring) {
    const SpmmRing& T = S.rg;
    LAUNCH_PDL(ctx, n, k_spmm_ring, (unsigned)T.nctas, kRingThreads, (size_t)T.total, n, Q, T, S.ring_meta.p, S.ring_blob.p, S.far_rows.p,
           S.ring_desc.p, S.ring_cta_ptr.p, P4, AP4, pAp, (long long*)nullptr);
  } else if (S.tiled) {
    const SpmmTile& T = S.tile;
    const int ntiles = (n + T.R - 1) / T.R;
    const unsigned grid = (unsigned)std::min(ntiles, 2 * ctx->sm_count);
    LAUNCH(ctx, k_spmm_tiled, grid, 256, (size_t)T.total, n, Q, T, S.rowptr.p, S.cv.p, S.far_rows.p, S.tile_far_ptr.p,
           S.gdiag.p, P4, AP4, pAp);
  } else {
    const PcgLaunch L = pcg_launch_shape(ctx, n, Q, false);
    LAUNCH_PDL(ctx, n, k_spmm, L.grid, L.block, L.smem, n, Q, S.rowptr.p, S.col.p, S.val.p, S.gdiag.p, P4, AP4, pAp);
  }
}

PcgLaunch pcg_launch_shape(tcs_ctx* ctx, int n, int Q, bool vector_kernels) {
  PcgLaunch L;
  L.block = 256;
  const int rows_per_block = 256 / Q;
  unsigned need = (unsigned)((n + rows_per_block - 1) / rows_per_block);
  // persistent-style grid: a multiple of the SM count, 8 resident CTAs of 256 threads per SM
  L.grid = std::max(1u, std::min(need, (unsigned)ctx->sm_count * 8u));
  L.smem = (size_t)4 * Q * sizeof(double);
  L.grid_update = L.grid_dir = L.grid;
  if (vector_kernels) {   // two occupancy queries: once per solve, not per launch
    L.grid_update = resident_grid(ctx, k_pcg_update, 256, L.smem, need, 4);
    L.grid_dir = resident_grid(ctx, k_pcg_dir, 256, 0, need, 3);
  }
  return L;
}

}  // namespace

#include "pcg_batch.cuh"

// Multi-right-hand-side PCG on an assembled system S (S.B in, S.X out as n x S.Kp float64).  `seg` (may be NULL)
// describes a block-diagonal system of several scenes: the float64 convergence test is then applied per scene and
// column (a column is done when EVERY scene's relative residual has reached the tolerance).
int tcs_impl_pcg_solve(tcs_ctx* ctx, tcs_system& S, int K, const tcs_solver_opts& opts, const tcs_segments* seg,
                       tcs_pcg_result& res) {
  const int n = S.n, Kp = S.Kp, Q = Kp / 4;
  res = tcs_pcg_result();
  if (n <= 0 || K <= 0) return TCS_OK;
  TRY(S.gdiag.alloc(ctx, (size_t)n + 8));
  LAUNCH(ctx, k_diag_excess, div_up(n, 256), 256, 0, n, S.rowptr.p, S.val.p, S.diag.p, S.gdiag.p);
  if (seg && seg->count() > 1 && env_int("TCS_BATCH_PCG", 1) != 0) {
    // a batch of independent scenes: every (scene, column) pair runs its own recurrence (pcg_batch.cuh)
    bool ran = false;
    TRY(pcg_solve_batched(ctx, S, K, opts, *seg, res, &ran));
    if (ran) return TCS_OK;
  }
  int iterations = 0, converged = 1, kernel_at_start = 0;
  double rel[TCS_MAX_LABELS] = {0};
  float spmm_sum_ms = 0.f;
  int spmm_timed = 0, spmm_launches = 0;

  // Working set: the label columns still iterating, packed to a multiple of 4.  wmap[kw] = original column of working
  // column kw (-1 = padding).  Columns are independent systems (one right-hand side each): when enough of them
  // have reached the tolerance to drop a float4 of columns, the survivors are re-packed and the SpMM re-planned
  // for the narrower rows, so converged columns stop costing bandwidth (C4: 6 of 16 columns finish within 700 of
  // 8050 iterations).  TCS_COMPACT=0 keeps all columns (they are then frozen instead).
  const int Kp0 = Kp;
  int Kw = Kp0, Qw = Q;
  std::vector<int> wmap(Kw);
  for (int k = 0; k < Kw; ++k) wmap[k] = k < K ? k : -1;
  std::vector<char> done(K, 0);
  bool compact_enabled = env_int("TCS_COMPACT", 1) != 0 && !(seg && seg->count() > 1);
  bool compacted = false;
  bool dropped_columns = false;  // converged columns have left the working set (SpMM timing samples stop there)
  const size_t nk = (size_t)n * Kp0;
  DBuf<double> Xw;   // float64 solution of the working columns (becomes S.X when no compaction happens)
  DBuf<float> Xd;    // float32 increment since the last flush
  DBuf<float> Bw;    // right-hand sides of the working columns (S.B itself until the first compaction)
  S.X.release();
  TRY(Xw.alloc(ctx, nk)); TRY(Xd.alloc(ctx, nk));
  TRY(S.P.alloc(ctx, nk)); TRY(S.R.alloc(ctx, nk)); TRY(S.AP.alloc(ctx, nk));
  DBuf<float> invd;
  TRY(invd.alloc(ctx, n));
  // scalar accumulators, row stride Kp0: pAp[3], rz[3], rr, bb, rz of the initial residual, p.r, recurrence rr[3]
  const int kScalRows = 13;
  DBuf<double> scal;
  TRY(scal.alloc(ctx, (size_t)kScalRows * Kp0));
  CK(cudaMemsetAsync(scal.p, 0, (size_t)kScalRows * Kp0 * 8, ctx->stream));
  double* d_pAp = scal.p;
  double* d_rz = scal.p + 3 * Kp0;
  double* d_rr = scal.p + 6 * Kp0;
  double* d_bb = scal.p + 7 * Kp0;
  double* d_rz0 = scal.p + 8 * Kp0;
  double* d_pr = scal.p + 9 * Kp0;   // p . r of the last residual replacement
  double* d_rr3 = scal.p + 10 * Kp0; // ||r||^2 of the recurrence residual (three generations, like rz)
  // block-diagonal batch: per-scene ||r||^2 and ||b||^2 (scene-major, stride Kp0), agg / worst ratio per column
  const int nseg = (seg && seg->count() > 1) ? seg->count() : 0;
  DBuf<double> seg_rr;
  std::vector<double> h_seg_rr, h_seg_bb, seg_scale((size_t)Kp0, 1.0);
  if (nseg) {
    TRY(seg_rr.alloc(ctx, 2 * (size_t)nseg * Kp0));
    CK(cudaMemsetAsync(seg_rr.p, 0, 2 * (size_t)nseg * Kp0 * 8, ctx->stream));
    h_seg_rr.resize((size_t)nseg * Kp0);
    h_seg_bb.resize((size_t)nseg * Kp0);
  }
  // device-side "done": recurrence residual four decades below the tolerance (in the D^-1 norm, relative to b) -
  // far from the tolerance itself (the two norms differ), far above float32 underflow of r
  const double done_factor = 1e-8 * opts.tolerance * opts.tolerance;
  DBuf<int> frozen, d_cols;
  TRY(frozen.alloc(ctx, Kp0)); TRY(d_cols.alloc(ctx, 2 * (size_t)Kp0));
  CK(cudaMemsetAsync(frozen.p, 0, (size_t)Kp0 * 4, ctx->stream));
  int* h_frozen = (int*)((char*)ctx->h_pin + 2048);
  for (int k = 0; k < Kp0; ++k) h_frozen[k] = 0;
  PcgLaunch L = pcg_launch_shape(ctx, n, Qw);
  // small systems: whole chunks of iterations in one persistent cooperative launch (k_pcg_chunk) instead of three
  // launches per iteration; no SpMM plan is needed then (the measurement hooks plan on demand)
  DBuf<unsigned> coop_bar;
  DBuf<double> coop_part;     // per-CTA partial dot products
  DBuf<long long> coop_dbg;   // TCS_DEBUG: cycles per phase of the first and the last CTA
  TRY(coop_bar.alloc(ctx, 1));
  if (getenv("TCS_DEBUG")) TRY(coop_dbg.alloc(ctx, 16));
  unsigned coop_grid = 0;
  bool coop_disabled = false;
  int coop_M = 0, coop_cap = 0;
  size_t coop_smem = 0;
  const void* coop_fn = nullptr;
  auto plan_iteration = [&]() -> int {
    coop_grid = 0;
    const int rows_per_block = 256 / Qw;
    const int ntiles = (n + rows_per_block - 1) / rows_per_block;
    // at most 4 row tiles per CTA, all CTAs co-resident: n * Q <= ~1.2 M (a 640 x 480 scene with 16 labels is 0.2 M)
    if (!coop_disabled && env_int("TCS_COOP", 1) != 0 && !opts.time_spmm && !getenv("TCS_SPMM") && ntiles <= 4 * 8 * ctx->sm_count &&
        n <= env_int("TCS_COOP_MAX_ROWS", 100000)) {
      uint32_t* d_max = (uint32_t*)ctx->d_scal + 16;
      CK(cudaMemsetAsync(d_max, 0, 4, ctx->stream));
      LAUNCH(ctx, k_tile_max_nnz, div_up(ntiles, 256), 256, 0, n, rows_per_block, S.rowptr.p, d_max);
      uint32_t mx = 0;
      TRY(tcs_impl_read_u32(ctx, d_max, &mx));
      const int cap = (int)((mx + 1u) & ~1u);
      const void* fns[4] = {(const void*)k_pcg_chunk<1>, (const void*)k_pcg_chunk<2>, (const void*)k_pcg_chunk<3>,
                            (const void*)k_pcg_chunk<4>};
      const int max_ctas = env_int("TCS_COOP_CTAS", 8);
      for (int M = 1; M <= 4; ++M) {
        const size_t smem = (size_t)M * cap * 8 + 256 * 4 * 8 + (size_t)Kw * (4 * 8 + 3 * 4);
        if (smem > 200 * 1024) break;
        // always the same value: function attributes are shared by every host thread that drives this device
        if (cudaFuncSetAttribute(fns[M - 1], cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) {
          cudaGetLastError();
          break;
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fns[M - 1], 256, smem) != cudaSuccess || nb < 1) { cudaGetLastError(); break; }
        const int resident = ctx->sm_count * std::min(nb, max_ctas);
        const int grid = (ntiles + M - 1) / M;
        if (grid > resident) continue;
        coop_grid = (unsigned)grid; coop_M = M; coop_cap = cap; coop_smem = smem; coop_fn = fns[M - 1];
        if (getenv("TCS_DEBUG"))
          fprintf(stderr, "[tcs] chunk plan: n=%d Q=%d tiles=%d cap=%d M=%d grid=%d smem=%zu occupancy=%d\n", n, Qw, ntiles, cap, M, grid, smem, nb);
        TRY(coop_part.alloc(ctx, 2 * (size_t)grid * Kw));
        S.ring = S.tiled = false;
        S.plan_stale = true;
        return TCS_OK;
      }
    }
    return spmm_plan(ctx, S);
  };
  {
    const unsigned sleep_ns = (unsigned)env_int("TCS_COOP_SLEEP", 0);
    CK(cudaMemcpyToSymbolAsync(c_coop_sleep, &sleep_ns, 4, 0, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  // coarse levels of the preconditioner (amg.cuh); with them the solve runs on the per-kernel path
  amg::Hierarchy AH;
  TRY(amg::setup(ctx, S, Kp0, AH));
  if (AH.on) {
    coop_disabled = true;
    // a few dozen iterations in all: re-planning the SpMM for a narrower working set costs more than it saves
    if (!getenv("TCS_COMPACT")) compact_enabled = false;
  }
  TRY(plan_iteration());
  kernel_at_start = coop_grid ? 4 : S.ring ? 3 : S.tiled ? 2 : 1;
  const float4* B4 = (const float4*)S.B.p;
  float4* P4 = (float4*)S.P.p;
  float4* R4 = (float4*)S.R.p;
  float4* AP4 = (float4*)S.AP.p;
  LAUNCH(ctx, k_pcg_init, L.grid, L.block, L.smem, n, Qw, B4, S.diag.p, invd.p, R4, P4, Xw.p, (float4*)Xd.p, d_rz, d_bb);
  if (AH.on) {
    TRY(amg::apply(ctx, AH, Qw, R4, d_rz));
    LAUNCH(ctx, amg::k_prolong, div_up((size_t)n * Qw, 256), 256, 0, n, Qw, 1.0f, amg::fine_agg(AH), amg::result(AH), P4);
  }
  CK(cudaMemcpyAsync(d_rz0, d_rz, (size_t)Kp0 * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  if (nseg) {
    double* d_bb_seg = seg_rr.p + (size_t)nseg * Kp0;
    for (int sg = 0; sg < nseg; ++sg) {
      const int r0 = seg->row_begin[sg], nr = seg->row_begin[sg + 1] - r0;
      if (nr > 0)
        LAUNCH(ctx, k_col_sumsq, std::min(L.grid, div_up(nr, 256 / Qw)), L.block, L.smem, r0, nr, Qw, B4, d_bb_seg + (size_t)sg * Kp0);
    }
    CK(cudaMemcpyAsync(h_seg_bb.data(), d_bb_seg, (size_t)nseg * Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  std::vector<double> h((size_t)kScalRows * Kp0);
  // sampled SpMM timing: one event pair per launch for the first launches (full width)
  const int kMaxSamples = 64;
  std::vector<cudaEvent_t> evs;
  if (opts.time_spmm) {
    evs.resize(2 * kMaxSamples);
    for (auto& e : evs) CK(cudaEventCreate(&e));
  }
  int samples = 0;
  converged = 0;
  int it = 0;
  // TCS_DEBUG: wall-clock per phase (we synchronise at every check, so host time brackets device time)
  auto phase_t0 = std::chrono::steady_clock::now();
  int phase_it0 = 0;
  auto phase_report = [&](const char* what) {
    if (!getenv("TCS_DEBUG")) return;
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - phase_t0).count();
    fprintf(stderr, "[tcs] %s: iterations %d..%d at width %d, kernel %d: %.1f ms = %.1f us / iteration\n", what, phase_it0, it,
            Kw, S.ring ? 3 : S.tiled ? 2 : 1, ms, it > phase_it0 ? 1e3 * ms / (it - phase_it0) : 0.0);
    phase_t0 = std::chrono::steady_clock::now();
    phase_it0 = it;
  };
  // ---- re-pack the working set to the columns that are not done (newKw = their count rounded up to 4)
  auto compact = [&](int newKw, int active_cols) -> int {
    const size_t nkw = (size_t)n * Kw;
    // ---- drop the converged columns
    phase_report("compaction");
    if (!S.X.p) {
      TRY(S.X.alloc(ctx, nk));
      CK(cudaMemsetAsync(S.X.p, 0, nk * 8, ctx->stream));
    }
    std::vector<int> cols(2 * (size_t)Kp0, -1), newmap;  // [0, Kp0): scatter targets; [Kp0, 2 Kp0): repack sources
    for (int kw = 0; kw < Kw; ++kw) {
      cols[kw] = wmap[kw];
      if (wmap[kw] >= 0 && !done[wmap[kw]]) { cols[Kp0 + newmap.size()] = kw; newmap.push_back(wmap[kw]); }
    }
    newmap.resize(newKw, -1);
    CK(cudaMemcpyAsync(d_cols.p, cols.data(), 2 * (size_t)Kp0 * 4, cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(ctx, k_scatter_columns, div_up(nkw, 256), 256, 0, (size_t)n, Kw, Kp0, d_cols.p, Xw.p, S.X.p);
    const size_t nk2 = (size_t)n * newKw;
    DBuf<float> P2, R2, B2, Xd2, AP2;
    DBuf<double> X2;
    TRY(P2.alloc(ctx, nk2)); TRY(R2.alloc(ctx, nk2)); TRY(B2.alloc(ctx, nk2)); TRY(Xd2.alloc(ctx, nk2));
    TRY(AP2.alloc(ctx, nk2)); TRY(X2.alloc(ctx, nk2));
    const int* src = d_cols.p + Kp0;
    LAUNCH(ctx, k_repack_columns<float>, div_up(nk2, 256), 256, 0, (size_t)n, Kw, newKw, src, S.P.p, P2.p);
    LAUNCH(ctx, k_repack_columns<float>, div_up(nk2, 256), 256, 0, (size_t)n, Kw, newKw, src, S.R.p, R2.p);
    LAUNCH(ctx, k_repack_columns<float>, div_up(nk2, 256), 256, 0, (size_t)n, Kw, newKw, src,
           Bw.p ? Bw.p : S.B.p, B2.p);
    LAUNCH(ctx, k_repack_columns<double>, div_up(nk2, 256), 256, 0, (size_t)n, Kw, newKw, src, Xw.p, X2.p);
    CK(cudaMemsetAsync(Xd2.p, 0, nk2 * 4, ctx->stream));
    // per-column scalars: same permutation on the host (we are at a synchronisation point anyway)
    CK(cudaMemcpyAsync(h.data(), scal.p, (size_t)kScalRows * Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));  // also: `cols` was read by the copy above
    std::vector<double> h2((size_t)kScalRows * Kp0, 0.0);
    for (int r8 = 0; r8 < kScalRows; ++r8)
      for (int k2 = 0; k2 < newKw; ++k2)
        if (cols[Kp0 + k2] >= 0) h2[(size_t)r8 * Kp0 + k2] = h[(size_t)r8 * Kp0 + cols[Kp0 + k2]];
    CK(cudaMemcpyAsync(scal.p, h2.data(), (size_t)kScalRows * Kp0 * 8, cudaMemcpyHostToDevice, ctx->stream));
    for (int k = 0; k < Kp0; ++k) h_frozen[k] = 0;
    CK(cudaMemcpyAsync(frozen.p, h_frozen, (size_t)Kp0 * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));  // h2 is a local vector
    S.P.swap(P2); S.R.swap(R2); S.AP.swap(AP2); Bw.swap(B2); Xw.swap(X2); Xd.swap(Xd2);
    Kw = newKw; Qw = Kw / 4; wmap = newmap;
    S.Kp = Kw;
    compacted = true;
    L = pcg_launch_shape(ctx, n, Qw);
    TRY(plan_iteration());
    B4 = (const float4*)Bw.p; P4 = (float4*)S.P.p; R4 = (float4*)S.R.p; AP4 = (float4*)S.AP.p;
    if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] it=%d: %d columns left, working width %d\n", it, active_cols, Kw);
    phase_report("re-plan");
    return TCS_OK;
  };
  if (opts.column_count > 0 && (opts.column_begin > 0 || opts.column_begin + opts.column_count < K)) {
    // label-column shard: the other columns count as done from the start
    int mine = 0;
    for (int k = 0; k < K; ++k) {
      if (k < opts.column_begin || k >= opts.column_begin + opts.column_count) done[k] = 1;
      else ++mine;
    }
    if (mine == 0) return tcs_fail(ctx, TCS_ERR_INPUT, "column shard [%d, %d) selects none of the %d colours", opts.column_begin,
                                   opts.column_begin + opts.column_count, K);
    const int newKw = std::max(4, (mine + 3) & ~3);
    if (newKw < Kw) {
      TRY(compact(newKw, mine));
    } else {
      for (int kw = 0; kw < Kw; ++kw)
        if (wmap[kw] >= 0 && done[wmap[kw]]) h_frozen[kw] = 1;
      CK(cudaMemcpyAsync(frozen.p, h_frozen, (size_t)Kp0 * 4, cudaMemcpyHostToDevice, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
    }
  }
  // Every check costs a pass over the float64 solution and a float64 SpMM (8 % of the iterations' time on C4 at one
  // check per 50 iterations).  While the slowest column is still far from the tolerance - more than a factor
  // TCS_CHECK_FAR (default 30; 0 = off) above it, i.e. well over a thousand Jacobi-PCG iterations away on these
  // systems - the next interval is twice check_every.
  const double far_factor = (double)env_int("TCS_CHECK_FAR", 30);
  int next_chunk = opts.check_every;
  // With the hierarchy the solve takes tens of iterations, not thousands, and float32 p / r / Ap put a floor under
  // the true residual of ONE recurrence (eps32 |A| |x| sqrt(n) / |b|: 2e-5 .. 2e-4 on these systems, measured) - the
  // recurrence residual keeps falling below it, ever more slowly, while the true one does not move.  So the solve is
  // classic mixed-precision iterative refinement: a STAGE runs PCG until the recurrence residual of every column
  // has dropped by `amg_stage` (1e-3) relative to the stage's start (or to the tolerance); then x is folded into
  // float64, the float64 true residual replaces r, and the next stage starts a fresh direction sequence from it.
  // Two stages reach 1e-6; a restart costs ~6 iterations (CPU experiment, scripts/amg_lab.py).  Between stage ends
  // the host only probes the recurrence norms (a 1 KB read every `amg_probe` iterations): nothing disturbs CG.
  const int amg_probe = std::max(1, env_int("TCS_AMG_PROBE", 4));
  const double amg_stage = amg::env_double("TCS_AMG_STAGE", 1e-3);
  const int amg_restart = env_int("TCS_AMG_RESTART", 1);
  const int amg_x64 = env_int("TCS_AMG_X64", 0);
  std::vector<double> stage_start((size_t)Kp0, 1.0);
  if (AH.on) next_chunk = amg_probe;
  bool replaced = true;    // the last check replaced r: the next step uses the exact line search (it > 0)
  while (it < opts.max_iterations) {
    const int chunk = std::min(next_chunk, opts.max_iterations - it);
    if (coop_grid) {
      CK(cudaMemsetAsync(coop_bar.p, 0, 4, ctx->stream));
      int a_n = n, a_Q = Qw, a_chunk = chunk, a_it = it, a_stride = Kp0, a_cap = coop_cap;
      const uint32_t* a_rowptr = S.rowptr.p;
      const int* a_col = S.col.p;
      const float *a_val = S.val.p, *a_diag = S.gdiag.p, *a_invd = invd.p;
      const float4* a_R = R4;
      float4* a_Xd = (float4*)Xd.p;
      const double *a_rz0 = d_rz0, *a_pr = it > 0 ? d_pr : nullptr;
      const int* a_frozen = frozen.p;
      double a_done = done_factor;
      unsigned* a_bar = coop_bar.p;
      long long* a_dbg = coop_dbg.p;
      double* a_part = coop_part.p;
      void* args[] = {&a_n, &a_Q, &a_chunk, &a_it, &a_stride, &a_cap, &a_rowptr, &a_col, &a_val, &a_diag, &a_invd, &P4, &a_R,
                      &a_Xd, &d_pAp, &d_rz, &a_rz0, &a_pr, &a_frozen, &a_done, &a_bar, &a_part, &a_dbg};
      const cudaError_t lerr = cudaLaunchCooperativeKernel(coop_fn, dim3(coop_grid), dim3(256), args, coop_smem, ctx->stream);
      if (lerr == cudaErrorCooperativeLaunchTooLarge || lerr == cudaErrorLaunchOutOfResources) {
        // the CTAs cannot all be resident right now: this solve continues with one launch per kernel
        cudaGetLastError();
        if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] cooperative launch refused (%s): per-kernel path\n", cudaGetErrorString(lerr));
        // (the refused kernel touched nothing: the accumulators are in the state the per-kernel path expects)
        coop_grid = 0;
        coop_disabled = true;
        TRY(spmm_plan(ctx, S));
        continue;
      }
      CK(lerr);
      ctx->launches++;
      if (coop_dbg.p && it == 0) {
        long long hd[16];
        CK(cudaMemcpyAsync(hd, coop_dbg.p, sizeof hd, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        for (int b = 0; b < 2; ++b)
          fprintf(stderr, "[tcs] k_pcg_chunk<%d> grid %u %s CTA, cycles per iteration: spmm %lld reduce %lld barrier %lld | update %lld reduce %lld barrier %lld | dir %lld barrier %lld\n",
                  coop_M, coop_grid, b ? "last" : "first", hd[8 * b] / chunk, hd[8 * b + 1] / chunk, hd[8 * b + 2] / chunk, hd[8 * b + 3] / chunk,
                  hd[8 * b + 4] / chunk, hd[8 * b + 5] / chunk, hd[8 * b + 6] / chunk, hd[8 * b + 7] / chunk);
      }
      it += chunk;
      spmm_launches += chunk;
    } else
    for (int j = 0; j < chunk; ++j, ++it) {
      const int s0 = it % 3, s1 = (it + 1) % 3, s2 = (it + 2) % 3;
      const double* pr = (j == 0 && it > 0 && replaced) ? d_pr : nullptr;  // first step after a residual replacement
      const bool timed = opts.time_spmm && samples < kMaxSamples && !dropped_columns;
      if (timed) CK(cudaEventRecord(evs[2 * samples], ctx->stream));
      launch_spmm(ctx, S, P4, AP4, d_pAp + s0 * Kp0);
      if (timed) { CK(cudaEventRecord(evs[2 * samples + 1], ctx->stream)); ++samples; }
      ++spmm_launches;
      LAUNCH_PDL(ctx, n, k_pcg_update, L.grid_update, L.block, L.smem, n, Qw, AP4, R4, invd.p, d_rz + s0 * Kp0, d_pAp + s0 * Kp0,
             d_rz + s1 * Kp0, frozen.p, d_rz0, done_factor, pr, AH.on ? d_rr3 + s1 * Kp0 : (double*)nullptr);
      if (AH.on) TRY(amg::apply(ctx, AH, Qw, R4, d_rz + s1 * Kp0));
      LAUNCH_PDL(ctx, n, k_pcg_dir, L.grid_dir, L.block, 0, n, Qw, R4, P4, (float4*)Xd.p, invd.p, d_rz + s0 * Kp0,
             d_pAp + s0 * Kp0, d_rz + s1 * Kp0, d_pAp + s1 * Kp0, d_rz + s2 * Kp0,
             AH.on ? d_rr3 + s2 * Kp0 : (double*)nullptr, frozen.p, d_rz0, done_factor, pr,
             AH.on ? amg::fine_agg(AH) : (const uint32_t*)nullptr, AH.on ? amg::result(AH) : (const float4*)nullptr,
             AH.on ? amg_restart : 0, (AH.on && amg_x64) ? Xw.p : (double*)nullptr);
    }
    const size_t nkw = (size_t)n * Kw;
    bool replace = true;
    if (AH.on) {
      // probe: recurrence ||r||^2 of the iteration just queued
      CK(cudaMemcpyAsync(ctx->h_pin, d_rr3 + (it % 3) * Kp0, (size_t)Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync((char*)ctx->h_pin + 512, d_bb, (size_t)Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      const double* prr = (const double*)ctx->h_pin;
      const double* pbb = (const double*)((char*)ctx->h_pin + 512);
      bool stage_open = false;
      for (int kw = 0; kw < Kw; ++kw)
        if (wmap[kw] >= 0 && !done[wmap[kw]] && pbb[kw] > 0.0) {
          const double v = std::sqrt(prr[kw] / pbb[kw]);
          if (v == v && v > std::max(0.7 * opts.tolerance * seg_scale[kw], amg_stage * stage_start[kw])) stage_open = true;
        }
      replaced = false;
      if (stage_open && it < opts.max_iterations) continue;  // keep iterating, undisturbed
    }
    if (replace) {
      // fold the float32 increment into the float64 solution; the float64 true residual replaces the recurrence
      // residual, and rz of the coming iteration is recomputed from it
      LAUNCH(ctx, k_flush_x, div_up(nkw, 256), 256, 0, nkw, Xw.p, Xd.p);
      CK(cudaMemsetAsync(d_rr, 0, (size_t)Kp0 * 8, ctx->stream));
      CK(cudaMemsetAsync(d_rz + (it % 3) * Kp0, 0, (size_t)Kp0 * 8, ctx->stream));
      CK(cudaMemsetAsync(d_pr, 0, (size_t)Kp0 * 8, ctx->stream));
      if (nseg) {
        CK(cudaMemsetAsync(seg_rr.p, 0, (size_t)nseg * Kp0 * 8, ctx->stream));
        for (int sg = 0; sg < nseg; ++sg) {
          const int r0 = seg->row_begin[sg], nr = seg->row_begin[sg + 1] - r0;
          if (nr > 0)
            LAUNCH(ctx, k_true_residual, std::min(L.grid, div_up(nr, 256 / Qw)), L.block, L.smem, r0, nr, Qw, S.rowptr.p, S.col.p,
                   S.val.p, S.diag.p, B4, Xw.p, P4, R4, seg_rr.p + (size_t)sg * Kp0, d_rz + (it % 3) * Kp0, d_pr);
        }
      } else {
        LAUNCH(ctx, k_true_residual, L.grid, L.block, L.smem, 0, n, Qw, S.rowptr.p, S.col.p, S.val.p, S.diag.p, B4, Xw.p,
               P4, R4, d_rr, d_rz + (it % 3) * Kp0, d_pr);
      }
      if (AH.on) TRY(amg::apply(ctx, AH, Qw, R4, d_rz + (it % 3) * Kp0));
      CK(cudaMemcpyAsync(ctx->h_pin, d_rr, (size_t)2 * Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
      if (nseg) CK(cudaMemcpyAsync(h_seg_rr.data(), seg_rr.p, (size_t)nseg * Kp0 * 8, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      memcpy(h.data(), ctx->h_pin, (size_t)2 * Kp0 * 8);
      if (nseg)   // aggregate norms of the batch: what the recurrence probes see
        for (int kw = 0; kw < Kw; ++kw) {
          double srr = 0.0, sbb = 0.0;
          for (int sg = 0; sg < nseg; ++sg) { srr += h_seg_rr[(size_t)sg * Kp0 + kw]; sbb += h_seg_bb[(size_t)sg * Kp0 + kw]; }
          h[kw] = srr; h[Kp0 + kw] = sbb;
        }
      replaced = true;
    }
    int active_cols = 0;
    for (int kw = 0; kw < Kw; ++kw) {
      const int k = wmap[kw];
      if (k < 0) continue;
      const double rr = h[kw], bb = h[Kp0 + kw];
      const double agg = bb > 0.0 ? std::sqrt(rr / bb) : 0.0;
      rel[k] = agg;
      if (nseg) {   // a column is done when the worst SCENE is: the probes see the aggregate, scaled by agg / worst
        double worst = 0.0;
        for (int sg = 0; sg < nseg; ++sg) {
          const double b2 = h_seg_bb[(size_t)sg * Kp0 + kw], r2 = h_seg_rr[(size_t)sg * Kp0 + kw];
          if (b2 > 0.0) worst = std::max(worst, std::sqrt(r2 / b2));
        }
        rel[k] = worst;
        seg_scale[kw] = (worst > 0.0 && agg > 0.0) ? std::min(1.0, agg / worst) : 1.0;
      }
      if (rel[k] <= opts.tolerance) { done[k] = 1; h_frozen[kw] = 1; }
      if (!done[k]) ++active_cols;
      stage_start[kw] = agg;
    }
    if (getenv("TCS_DEBUG")) {
      fprintf(stderr, "[tcs] it=%d rel:", it);
      for (int k = 0; k < K; ++k) fprintf(stderr, " %.2e", rel[k]);
      fprintf(stderr, "\n");
    }
    if (active_cols == 0) { converged = 1; phase_report("done"); break; }
    {
      double worst = 0.0;
      for (int kw = 0; kw < Kw; ++kw)
        if (wmap[kw] >= 0 && !done[wmap[kw]]) worst = std::max(worst, rel[wmap[kw]]);
      next_chunk = (far_factor > 0.0 && worst > far_factor * opts.tolerance) ? 2 * opts.check_every : opts.check_every;
      if (AH.on) next_chunk = amg_probe;
    }
    const int newKw = std::max(4, (active_cols + 3) & ~3);
    if (compact_enabled && newKw < Kw && it < opts.max_iterations) {
      TRY(compact(newKw, active_cols));
      dropped_columns = true;
    } else {
      CK(cudaMemcpyAsync(frozen.p, h_frozen, (size_t)Kp0 * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
  }
  if (opts.column_count > 0)
    for (int k = 0; k < K; ++k)
      if (k < opts.column_begin || k >= opts.column_begin + opts.column_count) rel[k] = 0.0;  // not this shard's
  // the float64 solution of all columns, n x Kp0
  if (!compacted) {
    S.X.swap(Xw);
  } else {
    std::vector<int> cols(2 * (size_t)Kp0, -1);
    for (int kw = 0; kw < Kw; ++kw) cols[kw] = wmap[kw];
    CK(cudaMemcpyAsync(d_cols.p, cols.data(), 2 * (size_t)Kp0 * 4, cudaMemcpyHostToDevice, ctx->stream));
    const size_t nkw = (size_t)n * Kw;
    LAUNCH(ctx, k_scatter_columns, div_up(nkw, 256), 256, 0, (size_t)n, Kw, Kp0, d_cols.p, Xw.p, S.X.p);
    CK(cudaStreamSynchronize(ctx->stream));
    // the SpMM plan and P belong to the narrow working set: the measurement hooks re-plan at full width
    S.Kp = Kp0;
    S.ring = S.tiled = false;
    S.plan_stale = true;
    S.P.release();
  }
  iterations = it;
  for (int sidx = 0; sidx < samples; ++sidx) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, evs[2 * sidx], evs[2 * sidx + 1]));
    spmm_sum_ms += ms;
    ++spmm_timed;
  }
  for (auto& e : evs) cudaEventDestroy(e);
  // the work vectors are only needed again by tcs_bench_spmm, which re-allocates them
  S.R.release(); S.AP.release();
  res.iterations = iterations;
  res.converged = converged;
  for (int k = 0; k < K; ++k) res.rel[k] = rel[k];
  res.spmm_launches = spmm_launches;
  res.kernel = kernel_at_start;  // the kernel planned for the full label block
  res.spmm_ms = spmm_timed ? spmm_sum_ms / spmm_timed * spmm_launches : 0.f;
  return TCS_OK;
}

// assignColors + getPotentials for `nrows` rows of a solved system (X: their nrows x Kp float64 solution, row_vertex:
// their vertices): labels into g->colours, best potentials into g->best_val, near-tie count into info
int tcs_impl_argmax(tcs_graph* g, const uint32_t* row_vertex, int nrows, int K, int Kp, const double* X,
                    const tcs_solver_opts& opts, float* d_pot, tcs_solve_info* info) {
  tcs_ctx* ctx = g->ctx;
  TRY(g->best_val.alloc(ctx, g->V));
  if (g->V) LAUNCH(ctx, k_best_init, div_up(g->V, 256), 256, 0, g->V, g->colours.p, g->best_val.p);
  DBuf<unsigned long long> ties;
  TRY(ties.alloc(ctx, 1));
  CK(cudaMemsetAsync(ties.p, 0, 8, ctx->stream));
  int k0 = 0, k1 = K;
  if (opts.column_count > 0) { k0 = std::max(0, opts.column_begin); k1 = std::min(K, opts.column_begin + opts.column_count); }
  if (nrows > 0)
    LAUNCH(ctx, k_argmax, div_up(nrows, 256), 256, 0, nrows, K, Kp, k0, k1, X, row_vertex, g->colours.p, d_pot, g->V, ties.p,
           g->best_val.p);
  if (info) {
    CK(cudaMemcpyAsync(ctx->h_pin, ties.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    info->near_ties = (long long)*(unsigned long long*)ctx->h_pin;
  }
  return TCS_OK;
}

int tcs_impl_random_walker(tcs_graph* g, int K, const tcs_solver_opts* opts_in, tcs_solve_info* info) {
  tcs_ctx* ctx = g->ctx;
  tcs_solver_opts opts;
  tcs_solver_opts_default(&opts);
  if (opts_in) opts = *opts_in;
  if (opts.check_every < 1) opts.check_every = 50;
  if (K < 0 || K > TCS_MAX_LABELS) return tcs_fail(ctx, TCS_ERR_INPUT, "number of labels %d outside 0..%d", K, TCS_MAX_LABELS);
  const size_t V = g->V;
  if (info) {
    float keep[12];
    memcpy(keep, info->stage_ms, sizeof keep);
    memset(info, 0, sizeof *info);
    memcpy(info->stage_ms, keep, sizeof keep);
    info->converged = 1;
    info->ncolours = K;
    info->ncomponents = g->ncomp;
  }
  if (V == 0 || K == 0) return TCS_OK;
  cudaEvent_t e_asm0 = ctx->ev[8], e_asm1 = ctx->ev[9], e_solve1 = ctx->ev[10], e_arg1 = ctx->ev[11];
  CK(cudaEventRecord(e_asm0, ctx->stream));
  TRY(tcs_impl_assemble(g, K));
  CK(cudaEventRecord(e_asm1, ctx->stream));
  tcs_system& S = g->sys;
  const int n = S.n;
  if (info) { info->rows = n; info->nnz = S.nnz_off + n; }

  float* d_pot = nullptr;
  if (opts.store_potentials) {
    TRY(g->potentials.alloc(ctx, V * (size_t)K));
    g->pot_K = K;
    CK(cudaMemsetAsync(g->potentials.p, 0, V * (size_t)K * 4, ctx->stream));
    LAUNCH(ctx, k_seed_potentials, div_up(V, 256), 256, 0, V, g->colours.p, g->potentials.p);
    d_pot = g->potentials.p;
  } else {
    g->potentials.release();
    g->pot_K = 0;
  }

  tcs_pcg_result res;
  if (n == 0) {   // every vertex is a seed or isolated: the labels are the seeds
    TRY(g->best_val.alloc(ctx, V));
    LAUNCH(ctx, k_best_init, div_up(V, 256), 256, 0, V, g->colours.p, g->best_val.p);
  }
  if (n > 0) {
    TRY(tcs_impl_pcg_solve(ctx, S, K, opts, nullptr, res));
    CK(cudaEventRecord(e_solve1, ctx->stream));
    TRY(tcs_impl_argmax(g, S.row_vertex.p, n, K, S.Kp, S.X.p, opts, d_pot, info));
    CK(cudaEventRecord(e_arg1, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (info) {
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, e_asm0, e_asm1)); info->stage_ms[8] = ms;
      CK(cudaEventElapsedTime(&ms, e_asm1, e_solve1)); info->stage_ms[9] = ms; info->solve_ms = ms;
      CK(cudaEventElapsedTime(&ms, e_solve1, e_arg1)); info->stage_ms[10] = ms;
    }
  }
  if (info) {
    info->iterations = res.iterations;
    info->converged = res.converged;
    for (int k = 0; k < K; ++k) info->rel_residual[k] = res.rel[k];
    info->spmm_launches = res.spmm_launches;
    info->spmm_kernel = res.kernel;
    info->spmm_ms = res.spmm_ms;
  }
  CK(cudaGetLastError());
  if (!res.converged) return tcs_fail(ctx, TCS_ERR_NOT_CONVERGED, "PCG did not reach %.1e in %d iterations", opts.tolerance, res.iterations);
  return TCS_OK;
}

extern "C" int tcs_random_walker(tcs_graph* g, uint32_t* colours, int ncolours, const tcs_solver_opts* opts,
                                 float* potentials, tcs_solve_info* info) {
  if (!g || (g->V && !colours)) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  CK(cudaSetDevice(ctx->device));
  const size_t V = g->V;
  for (size_t v = 0; v < V; ++v)
    if (colours[v] > (uint32_t)ncolours) return tcs_fail(ctx, TCS_ERR_INPUT, "colour %u of vertex %zu exceeds ncolours", colours[v], v);
  TRY(g->colours.alloc(ctx, V));
  if (V) CK(cudaMemcpyAsync(g->colours.p, colours, 4 * V, cudaMemcpyHostToDevice, ctx->stream));
  tcs_solver_opts o;
  tcs_solver_opts_default(&o);
  if (opts) o = *opts;
  if (potentials) o.store_potentials = 1;
  int st = tcs_impl_random_walker(g, ncolours, &o, info);
  if (st != TCS_OK && st != TCS_ERR_NOT_CONVERGED) return st;
  if (V) CK(cudaMemcpyAsync(colours, g->colours.p, 4 * V, cudaMemcpyDeviceToHost, ctx->stream));
  if (potentials && V && ncolours) CK(cudaMemcpyAsync(potentials, g->potentials.p, 4 * V * (size_t)ncolours, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return st;
}

namespace {
// the measurement hooks work on the full label block: after a solve that ended on a compacted column set, allocate
// a full-width P again and plan the SpMM for it
int ensure_full_plan(tcs_ctx* ctx, tcs_system& S) {
  if (!S.plan_stale) return TCS_OK;
  const size_t nk = (size_t)S.n * S.Kp;
  TRY(S.P.alloc(ctx, nk)); TRY(S.AP.alloc(ctx, nk));
  CK(cudaMemsetAsync(S.P.p, 0, nk * 4, ctx->stream));
  TRY(spmm_plan(ctx, S));
  return TCS_OK;
}
}  // namespace

// bench hook: the SpMM alone on the system of the last solve, `reps` launches between two events
extern "C" int tcs_bench_spmm(tcs_graph* g, int reps, float* avg_ms, long long* algorithmic_bytes) {
  if (!g || reps < 1) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  tcs_system& S = g->sys;
  if (S.n == 0 || (!S.P.p && !S.plan_stale)) return tcs_fail(ctx, TCS_ERR_STATE, "no assembled system (run the random walker first)");
  CK(cudaSetDevice(ctx->device));
  TRY(ensure_full_plan(ctx, S));
  const int n = S.n, Kp = S.Kp, Q = Kp / 4;
  TRY(S.AP.alloc(ctx, (size_t)n * Kp));
  DBuf<double> acc;
  TRY(acc.alloc(ctx, Kp));
  CK(cudaMemsetAsync(acc.p, 0, (size_t)Kp * 8, ctx->stream));
  const PcgLaunch L = pcg_launch_shape(ctx, n, Q);
  for (int i = 0; i < 3; ++i) launch_spmm(ctx, S, (const float4*)S.P.p, (float4*)S.AP.p, acc.p);
  CK(cudaEventRecord(ctx->ev[12], ctx->stream));
  for (int i = 0; i < reps; ++i) launch_spmm(ctx, S, (const float4*)S.P.p, (float4*)S.AP.p, acc.p);
  CK(cudaEventRecord(ctx->ev[13], ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, ctx->ev[12], ctx->ev[13]));
  if (avg_ms) *avg_ms = ms / reps;
  // B_spmm = 8 nnz + 4 (n + 1) + 4 n K + 4 n K, nnz counting the diagonal, K = label columns
  if (algorithmic_bytes)
    *algorithmic_bytes = 8ll * (S.nnz_off + n) + 4ll * (n + 1) + 8ll * (long long)n * S.K;
  S.AP.release();
  return TCS_OK;
}

namespace {
__global__ void k_fill_random(size_t count, unsigned seed, float* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= count) return;
  unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull + seed;  // splitmix64
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  out[i] = (float)((double)(z >> 11) * (1.0 / 9007199254740992.0) * 2.0 - 1.0);
}
__global__ void k_max_diff(size_t count, const float* __restrict__ a, const float* __restrict__ b,
                           unsigned long long* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= count) return;
  // non-negative doubles order like their bit patterns
  atomicMax(&out[0], (unsigned long long)__double_as_longlong(fabs((double)a[i] - (double)b[i])));
  atomicMax(&out[1], (unsigned long long)__double_as_longlong(fabs((double)a[i])));
}
}  // namespace

extern "C" int tcs_check_spmm(tcs_graph* g, unsigned seed, double* max_abs_diff, double* max_abs_ref,
                              double* max_rel_dot_diff, int* kernel) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  tcs_system& S = g->sys;
  if (S.n == 0 || (!S.P.p && !S.plan_stale)) return tcs_fail(ctx, TCS_ERR_STATE, "no assembled system (run the random walker first)");
  CK(cudaSetDevice(ctx->device));
  TRY(ensure_full_plan(ctx, S));
  const int n = S.n, Kp = S.Kp, Q = Kp / 4;
  const size_t nk = (size_t)n * Kp;
  DBuf<float> ap1, ap2;
  DBuf<double> dots;
  DBuf<unsigned long long> mx;
  TRY(ap1.alloc(ctx, nk)); TRY(ap2.alloc(ctx, nk)); TRY(dots.alloc(ctx, 2 * (size_t)Kp)); TRY(mx.alloc(ctx, 2));
  CK(cudaMemsetAsync(dots.p, 0, 2 * (size_t)Kp * 8, ctx->stream));
  CK(cudaMemsetAsync(mx.p, 0, 16, ctx->stream));
  CK(cudaMemsetAsync(ap2.p, 0xFF, nk * 4, ctx->stream));
  LAUNCH(ctx, k_fill_random, div_up(nk, 256), 256, 0, nk, seed, S.P.p);
  const PcgLaunch L = pcg_launch_shape(ctx, n, Q);
  LAUNCH(ctx, k_spmm, L.grid, L.block, L.smem, n, Q, S.rowptr.p, S.col.p, S.val.p, S.gdiag.p, (const float4*)S.P.p,
         (float4*)ap1.p, dots.p);
  launch_spmm(ctx, S, (const float4*)S.P.p, (float4*)ap2.p, dots.p + Kp);
  LAUNCH(ctx, k_max_diff, div_up(nk, 256), 256, 0, nk, ap1.p, ap2.p, mx.p);
  std::vector<double> h(2 * Kp);
  unsigned long long hm[2];
  CK(cudaMemcpyAsync(h.data(), dots.p, 2 * (size_t)Kp * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(hm, mx.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  double rel = 0.0;
  for (int k = 0; k < S.K; ++k) rel = std::max(rel, std::fabs(h[k] - h[Kp + k]) / std::max(1e-300, std::fabs(h[k])));
  double d0, d1;
  memcpy(&d0, &hm[0], 8); memcpy(&d1, &hm[1], 8);
  if (max_abs_diff) *max_abs_diff = d0;
  if (max_abs_ref) *max_abs_ref = d1;
  if (max_rel_dot_diff) *max_rel_dot_diff = rel;
  if (kernel) *kernel = S.ring ? 3 : S.tiled ? 2 : 1;
  CK(cudaGetLastError());
  return TCS_OK;
}

extern "C" int tcs_graph_download_best(tcs_graph* g, float* best_potential, uint32_t* best_colour) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (g->V == 0) return TCS_OK;
  if (!g->best_val.p || !g->colours.p) return tcs_fail(ctx, TCS_ERR_STATE, "no solve on this graph yet");
  CK(cudaSetDevice(ctx->device));
  if (best_potential) CK(cudaMemcpyAsync(best_potential, g->best_val.p, 4 * g->V, cudaMemcpyDeviceToHost, ctx->stream));
  if (best_colour) CK(cudaMemcpyAsync(best_colour, g->colours.p, 4 * g->V, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}
