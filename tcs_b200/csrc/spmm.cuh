// tcs_b200/csrc/spmm.cuh - the tiled CG SpMM (included by pcg.cu).
//
// v1 (k_spmm in pcg.cu) is bound by L1 wavefronts: every gathered 64-byte P row and every per-thread col/val load is
// its own 128-byte-line wavefront (profiles/r01_spmm_v1_ncu_summary.md).  In Morton vertex order 90 % of a row's
// neighbours lie within +-256 rows, and a tile of 256 consecutive rows touches only ~1.6 distinct P rows per row, so:
//
//   * one persistent CTA per half SM walks tiles of R consecutive rows;
//   * per tile, one elected thread issues TMA bulk copies (cp.async.bulk -> UBLKCP) of the tile's contiguous CSR
//     segment (col, val) and of the P window [r0 - H, r1 + H) into shared memory, completion on an mbarrier;
//   * neighbours outside the window ("far", ~7 % of the entries but only ~70 distinct rows per tile) are resolved
//     once per solve: per tile a de-duplicated list of far rows, and far entries of the CSR carry -1 - slot;
//     the kernel fetches the tile's far rows with cp.async into a far buffer behind the window;
//   * the Q lanes of a row read {offset, weight} pairs as shared-memory broadcasts (LDS.64) and gather neighbour
//     rows with LDS.128 - the inner loop is LDS.64, IADD, LDS.128, 4 FFMA, no branch, no global access;
//   * AP rows leave as fully coalesced 128-bit stores, p.Ap dots are accumulated in fp64 per thread and reduced
//     with warp shuffles once per CTA.
#pragma once

#include <cstdint>

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on the mbarrier (bytes multiple of 16, 16-byte aligned)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}


// ---- one-time preprocessing of the far entries (per solve) ----------------------------------------------------
__global__ void k_far_flags(int n, int R, int H, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                            const float* __restrict__ val, int2* __restrict__ cv, uint32_t* __restrict__ flags) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n) return;
  const int r0 = (row / R) * R, r1 = min(n, r0 + R);
  const int w0 = max(0, r0 - H), w1 = min(n, r1 + H);
  for (uint32_t a = rowptr[row]; a < rowptr[row + 1]; ++a) {
    const int c = col[a];
    cv[a] = make_int2(c, __float_as_int(val[a]));
    flags[a] = (c < w0 || c >= w1) ? 1u : 0u;
  }
}
__global__ void k_far_scatter(int n, int R, const uint32_t* __restrict__ rowptr, const int* __restrict__ col,
                              const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos,
                              unsigned long long* __restrict__ keys, uint32_t* __restrict__ ent) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n) return;
  const unsigned long long tile = (unsigned long long)(row / R);
  for (uint32_t a = rowptr[row]; a < rowptr[row + 1]; ++a)
    if (flags[a]) {
      const uint32_t j = pos[a];
      keys[j] = tile << 32 | (uint32_t)col[a];
      ent[j] = a;
    }
}
__global__ void k_far_heads(const unsigned long long* __restrict__ keys, uint32_t nfar, uint32_t* __restrict__ head) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nfar) head[j] = (j == 0 || keys[j] != keys[j - 1]) ? 1u : 0u;
}
// tile_far_ptr[t] = index (among the distinct far rows) of the first far row of tile t
__global__ void k_far_tile_ptr(int ntiles, const unsigned long long* __restrict__ keys, uint32_t nfar,
                               const uint32_t* __restrict__ uidx, uint32_t* __restrict__ tile_far_ptr) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > ntiles) return;
  uint32_t lo = 0, hi = nfar;  // first j with tile(keys[j]) >= t
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    if ((int)(keys[mid] >> 32) < t) lo = mid + 1; else hi = mid;
  }
  tile_far_ptr[t] = uidx[lo];  // uidx[nfar] = number of distinct far rows
}
__global__ void k_far_assign(uint32_t nfar, const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ ent,
                             const uint32_t* __restrict__ head, const uint32_t* __restrict__ uidx,
                             const uint32_t* __restrict__ tile_far_ptr, int2* __restrict__ cv,
                             int* __restrict__ far_rows, uint32_t* __restrict__ max_far) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nfar) return;
  const unsigned long long k = keys[j];
  const uint32_t t = (uint32_t)(k >> 32);
  const uint32_t uid = uidx[j] + head[j] - 1u;
  if (head[j]) far_rows[uid] = (int)(uint32_t)k;
  const uint32_t slot = uid - tile_far_ptr[t];
  cv[ent[j]].x = -1 - (int)slot;
  if (head[j]) atomicMax(max_far, slot + 1u);
}

// ---- the kernel ---------------------------------------------------------------------------------------------------
// Per tile:
//  1. one thread issues TMA bulk copies (P window, {col,val} pairs, diagonal slice) -> mbarrier; meanwhile every
//     Q-lane group fetches far rows with cp.async; the next tile's rowptr slice is prefetched into registers
//  2. scan: column index -> byte offset of the neighbour row inside [window | far buffer] (in place)
//  3. branch-free gather loop out of shared memory, coalesced AP store, fp64 p.Ap partials
__global__ void __launch_bounds__(256, 2)
    k_spmm_tiled(int n, int Q, SpmmTile T, const uint32_t* __restrict__ rowptr, const int2* __restrict__ cv,
                 const int* __restrict__ far_rows, const uint32_t* __restrict__ tile_far_ptr,
                 const float* __restrict__ diag, const float4* __restrict__ P, float4* __restrict__ AP,
                 double* __restrict__ pAp) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  uint32_t* rp_s = (uint32_t*)(smem + T.off_rp);       // R + 3 entries: rowptr slice, far_ptr[t], far_ptr[t+1]
  const float* diag_s = (const float*)(smem + T.off_diag);
  int2* cv_s = (int2*)(smem + T.off_cv);
  unsigned char* P_s = smem + T.off_p;
  double* red = (double*)(smem + T.off_red);

  const int rows_per_pass = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_pass;
  const int ntiles = (n + T.R - 1) / T.R;
  const int rowbytes = 16 * Q;
  const int far_off = (T.R + 2 * T.H) * rowbytes;  // byte offset of the far buffer inside P_s
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // rowptr slice of the first tile
  {
    const int r0 = blockIdx.x * T.R;
    if (r0 < n) {
      const int rows = min(n, r0 + T.R) - r0;
      for (int i = threadIdx.x; i <= rows; i += blockDim.x) rp_s[i] = rowptr[r0 + i];
      if (threadIdx.x < 2) rp_s[T.R + 1 + threadIdx.x] = tile_far_ptr[blockIdx.x + threadIdx.x];
    }
  }
  __syncthreads();
  uint32_t phase = 0;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int r0 = tile * T.R, r1 = min(n, r0 + T.R);
    const int w0 = max(0, r0 - T.H), w1 = min(n, r1 + T.H);
    const uint32_t seg0 = rp_s[0], seg1 = rp_s[r1 - r0];
    const uint32_t fp0 = rp_s[T.R + 1], fp1 = rp_s[T.R + 2];
    const uint32_t a_lo = seg0 & ~1u, a_hi = (seg1 + 1u) & ~1u;
    if (threadIdx.x == 0) {
      const uint32_t pbytes = (uint32_t)(w1 - w0) * (uint32_t)rowbytes, sbytes = (a_hi - a_lo) * 8u;
      const uint32_t dbytes = (uint32_t)(((r1 - r0) * 4 + 15) & ~15);
      mbar_expect_tx(bar, pbytes + sbytes + dbytes);
      bulk_g2s(P_s, P + (size_t)w0 * Q, pbytes, bar);
      if (sbytes) bulk_g2s(cv_s, cv + a_lo, sbytes, bar);
      bulk_g2s((void*)diag_s, diag + r0, dbytes, bar);
    }
    // far rows of this tile: one row per Q-lane group and step, all in flight together
    if (active)
      for (uint32_t j = local; j < fp1 - fp0; j += rows_per_pass) {
        const int c = far_rows[fp0 + j];
        cp_async16(P_s + far_off + j * rowbytes + q * 16, &P[(size_t)c * Q + q]);
      }
    // prefetch the next tile's rowptr slice into registers (stored after the compute phase)
    const int nt = tile + gridDim.x;
    uint32_t nrp0 = 0, nrp1 = 0;
    if (nt < ntiles) {
      const int nr0 = nt * T.R, nrows = min(n, nr0 + T.R) - nr0;
      if ((int)threadIdx.x <= nrows) nrp0 = rowptr[nr0 + threadIdx.x];
      if ((int)threadIdx.x + 256 <= nrows) nrp1 = rowptr[nr0 + threadIdx.x + 256];
      if (threadIdx.x >= 254) nrp1 = tile_far_ptr[nt + (threadIdx.x - 254)];
    }
    mbar_wait(bar, phase);
    phase ^= 1u;
    // scan: column index -> byte offset inside P_s
    for (uint32_t a = seg0 - a_lo + threadIdx.x; a < seg1 - a_lo; a += blockDim.x) {
      const int c = cv_s[a].x;
      cv_s[a].x = c >= 0 ? (c - w0) * rowbytes : far_off + (-1 - c) * rowbytes;
    }
    cp_async_wait_all();
    __syncthreads();
    if (active) {
      const unsigned char* Pq = P_s + q * 16;
      for (int lr = local; lr < r1 - r0; lr += rows_per_pass) {
        uint32_t a = rp_s[lr] - a_lo;
        const uint32_t a1 = rp_s[lr + 1] - a_lo;
        const float4 p = *(const float4*)(Pq + (r0 + lr - w0) * rowbytes);
        const float d = diag_s[lr];
        const F4p pp = f4_pack(p);
        F4p yy = f4_pack(make_float4(d * p.x, d * p.y, d * p.z, d * p.w));
#pragma unroll 4
        for (; a < a1; ++a) {
          const int2 e = cv_s[a];
          const float w = __int_as_float(e.y);
          const float4 t = *(const float4*)(Pq + e.x);
          f4_diff_fma(yy, w, t, pp);
        }
        const float4 y = f4_unpack(yy);
        AP[(size_t)(r0 + lr) * Q + q] = y;
        acc[0] += (double)p.x * (double)y.x; acc[1] += (double)p.y * (double)y.y;
        acc[2] += (double)p.z * (double)y.z; acc[3] += (double)p.w * (double)y.w;
      }
    }
    __syncthreads();  // all reads of this tile's buffers are done
    if (nt < ntiles) {
      const int nr0 = nt * T.R, nrows = min(n, nr0 + T.R) - nr0;
      if ((int)threadIdx.x <= nrows) rp_s[threadIdx.x] = nrp0;
      if ((int)threadIdx.x + 256 <= nrows) rp_s[threadIdx.x + 256] = nrp1;
      if (threadIdx.x >= 254) rp_s[T.R + 1 + (threadIdx.x - 254)] = nrp1;
    }
    __syncthreads();
  }
  reduce_columns(acc, q, Q, active, red, pAp);
}

}  // namespace
