// tcs_b200/csrc/solve.cu - connected components, seed look-up, Laplacian assembly (CSR on device), the
// multi-right-hand-side Jacobi-preconditioned conjugate gradient, argmax labelling.
// Replaces createSubgraphsFromConnectedComponents (graph/impl/common.hpp:134-149), the seed painting of
// RandomWalkerSegmentation::segment (impl/random_walker_segmentation.hpp:144-152) and
// detail::RandomWalker (impl/random_walker.hpp:95-243; its Eigen::SimplicialCholesky solve becomes PCG).
#include <cfloat>
#include <cmath>
#include <algorithm>

#include <mutex>

#include "common.cuh"

namespace {

// ---- connected components: lock-free union-find, larger root hooked under the smaller one -----------------
// => the root of every component is its smallest vertex id, which is what boost::connected_components'
// numbering (ids in order of the smallest member, appendix A.7) needs.
__device__ __forceinline__ uint32_t uf_find(uint32_t* parent, uint32_t v) {
  uint32_t p = parent[v];
  while (p != v) {
    uint32_t gp = parent[p];
    if (gp != p) parent[v] = gp;  // path halving (benign race: only ever shortens towards an ancestor)
    v = p;
    p = gp;
  }
  return v;
}

__global__ void k_iota(uint32_t* p, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = (uint32_t)i;
}

__global__ void k_cc_hook(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                          uint32_t* parent) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    if (u >= v) continue;  // every undirected edge once
    uint32_t ru = uf_find(parent, u), rv = uf_find(parent, (uint32_t)v);
    while (ru != rv) {
      if (ru < rv) { uint32_t t = ru; ru = rv; rv = t; }  // ru is the larger root
      uint32_t old = atomicCAS(&parent[ru], ru, rv);
      if (old == ru) break;
      ru = uf_find(parent, ru);
      rv = uf_find(parent, rv);
    }
  }
}

__global__ void k_cc_flatten(size_t V, uint32_t* parent, uint32_t* __restrict__ is_root) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t r = (uint32_t)v;
  while (true) { uint32_t p = parent[r]; if (p == r) break; r = p; }
  parent[v] = r;  // racing writers all write roots; readers follow to the same root
  is_root[v] = (r == (uint32_t)v) ? 1u : 0u;
}

__global__ void k_cc_assign(size_t V, const uint32_t* __restrict__ parent, const uint32_t* __restrict__ rank,
                            int32_t* __restrict__ comp) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t r = parent[v];
  while (parent[r] != r) r = parent[r];
  comp[v] = (int32_t)rank[r];
}

// ---- seeds ---------------------------------------------------------------------------------------------------
// exact 1-NN of every seed among the vertices (pcl::getApproximateIndices -> KdTreeFLANN, appendix A.6),
// squared L2 accumulated x, y, z; ties -> lowest vertex id.  One block per seed.
__global__ void k_seed_nn(size_t V, const float* __restrict__ x, const float* __restrict__ y,
                          const float* __restrict__ z, const float* __restrict__ seeds, int32_t* __restrict__ out) {
  const int s = blockIdx.x;
  const float sx = seeds[3 * s], sy = seeds[3 * s + 1], sz = seeds[3 * s + 2];
  float best = __int_as_float(0x7F800000);
  uint32_t bi = 0xFFFFFFFFu;
  for (size_t v = threadIdx.x; v < V; v += blockDim.x) {
    float dx = sx - x[v], dy = sy - y[v], dz = sz - z[v];
    float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    if (d < best) { best = d; bi = (uint32_t)v; }
  }
  __shared__ float sb[32];
  __shared__ uint32_t si[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    uint32_t oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
    if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
      if (sb[w] < best || (sb[w] == best && si[w] < bi)) { best = sb[w]; bi = si[w]; }
    out[s] = (int32_t)bi;
  }
}

// Large graphs: the vertices are read ONCE.  A block keeps kSeedTile vertices in registers (12 per thread) and walks the
// seeds (staged in shared memory, 256 at a time); per (tile, seed) the block's best (distance, vertex) goes to
// part[tile * ns + seed], k_seed_nn_merge takes the minimum over the tiles.  Same distance arithmetic and tie rule
// (lowest vertex id) as k_seed_nn: a thread's vertices ascend, every merge compares (distance, id) - identical
// results.  (One block per seed re-read the 34 MB of C4 vertices 160 times: 2.6 ms; this is 0.2 ms.)
constexpr int kSeedPerThread = 12;
constexpr int kSeedTile = 256 * kSeedPerThread;
__global__ void __launch_bounds__(256) k_seed_nn_tiled(size_t V, int ns, const float* __restrict__ x, const float* __restrict__ y,
                                                       const float* __restrict__ z, const float* __restrict__ seeds,
                                                       float* __restrict__ part_d, uint32_t* __restrict__ part_i) {
  __shared__ float seed_s[3 * 256];
  __shared__ float wd[256 * 8];
  __shared__ uint32_t wi[256 * 8];
  const size_t v0 = (size_t)blockIdx.x * kSeedTile;
  float vx[kSeedPerThread], vy[kSeedPerThread], vz[kSeedPerThread];
#pragma unroll
  for (int j = 0; j < kSeedPerThread; ++j) {
    const size_t v = v0 + threadIdx.x + 256 * (size_t)j;
    // padding lanes hold NaN: their distance never compares smaller
    vx[j] = v < V ? x[v] : __int_as_float(0x7FC00000);
    vy[j] = v < V ? y[v] : 0.f;
    vz[j] = v < V ? z[v] : 0.f;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int s0 = 0; s0 < ns; s0 += 256) {
    const int nb = min(256, ns - s0);
    __syncthreads();   // the previous batch's partials are read
    for (int i = threadIdx.x; i < 3 * nb; i += 256) seed_s[i] = seeds[3 * (size_t)s0 + i];
    __syncthreads();
    for (int s = 0; s < nb; ++s) {
      const float sx = seed_s[3 * s], sy = seed_s[3 * s + 1], sz = seed_s[3 * s + 2];
      float best = __int_as_float(0x7F800000);
      uint32_t bi = 0xFFFFFFFFu;
#pragma unroll
      for (int j = 0; j < kSeedPerThread; ++j) {
        const float dx = sx - vx[j], dy = sy - vy[j], dz = sz - vz[j];
        const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        if (d < best) { best = d; bi = (uint32_t)(v0 + threadIdx.x + 256 * (size_t)j); }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
        const uint32_t oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
        if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
      }
      if (lane == 0) { wd[s * 8 + warp] = best; wi[s * 8 + warp] = bi; }
    }
    __syncthreads();
    for (int s = threadIdx.x; s < nb; s += 256) {
      float best = wd[s * 8];
      uint32_t bi = wi[s * 8];
      for (int w = 1; w < 8; ++w)
        if (wd[s * 8 + w] < best || (wd[s * 8 + w] == best && wi[s * 8 + w] < bi)) { best = wd[s * 8 + w]; bi = wi[s * 8 + w]; }
      part_d[(size_t)blockIdx.x * ns + s0 + s] = best;
      part_i[(size_t)blockIdx.x * ns + s0 + s] = bi;
    }
  }
}
// one warp per seed: minimum (distance, vertex) over the tiles
__global__ void __launch_bounds__(256) k_seed_nn_merge(int ns, int tiles, const float* __restrict__ part_d,
                                                       const uint32_t* __restrict__ part_i, int32_t* __restrict__ out) {
  const int s = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (s >= ns) return;
  float best = __int_as_float(0x7F800000);
  uint32_t bi = 0xFFFFFFFFu;
  for (int t = lane; t < tiles; t += 32) {
    const float d = part_d[(size_t)t * ns + s];
    const uint32_t i = part_i[(size_t)t * ns + s];
    if (d < best || (d == best && i < bi)) { best = d; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    const uint32_t oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
    if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if (lane == 0) out[s] = (int32_t)bi;
}

// ---- allocation-free variant of the seed look-up (tcs_nearest_vertices_noalloc) ----------------------------------
// Scratch of the variant: a file-scope device array (every GPU that loads the library holds these 276 KB until exit).
//   seeds   : 3 floats per seed (staging of the caller's host array)
//   nearest : 1 int32 per seed (result, copied back to the caller)
//   part_d / part_i : per (seed, chunk of the vertex range) the smallest distance and its vertex
// Every slot a call reads is written by that call first, so nothing has to start at zero; the array keeps the last
// call's contents, and two calls that overlap on the GPU (different streams / contexts) race on it.
constexpr int kNoallocChunks = 32;
struct NoallocScratch {
  float seeds[3 * TCS_NOALLOC_MAX_SEEDS];
  int32_t nearest[TCS_NOALLOC_MAX_SEEDS];
  float part_d[TCS_NOALLOC_MAX_SEEDS * kNoallocChunks];
  uint32_t part_i[TCS_NOALLOC_MAX_SEEDS * kNoallocChunks];
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoallocScratch)];

// grid (chunks, seeds): block (c, s) scans chunk c of the vertices for seed s; same distance arithmetic and tie rule
// (lowest vertex id) as k_seed_nn, so the two-stage result is identical to the one-block-per-seed kernel's
__global__ void __launch_bounds__(256) k_seed_nn_partial(size_t V, const float* __restrict__ x, const float* __restrict__ y,
                                                         const float* __restrict__ z, const float* __restrict__ seeds,
                                                         float* __restrict__ part_d, uint32_t* __restrict__ part_i) {
  const int s = blockIdx.y, c = blockIdx.x, chunks = gridDim.x;
  const size_t per = (V + chunks - 1) / chunks;
  const size_t v0 = (size_t)c * per, v1 = v0 + per < V ? v0 + per : V;
  const float sx = seeds[3 * s], sy = seeds[3 * s + 1], sz = seeds[3 * s + 2];
  float best = __int_as_float(0x7F800000);
  uint32_t bi = 0xFFFFFFFFu;
  for (size_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
    float dx = sx - x[v], dy = sy - y[v], dz = sz - z[v];
    float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    if (d < best) { best = d; bi = (uint32_t)v; }
  }
  __shared__ float sb[8];
  __shared__ uint32_t si[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    uint32_t oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
    if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
      if (sb[w] < best || (sb[w] == best && si[w] < bi)) { best = sb[w]; bi = si[w]; }
    part_d[(size_t)s * chunks + c] = best;   // written for every (seed, chunk), empty chunks included
    part_i[(size_t)s * chunks + c] = bi;
  }
}
__global__ void k_seed_nn_final(int ns, int chunks, const float* __restrict__ part_d, const uint32_t* __restrict__ part_i,
                                int32_t* __restrict__ out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  float best = __int_as_float(0x7F800000);
  uint32_t bi = 0xFFFFFFFFu;
  for (int c = 0; c < chunks; ++c) {
    const float d = part_d[(size_t)s * chunks + c];
    const uint32_t i = part_i[(size_t)s * chunks + c];
    if (d < best || (d == best && i < bi)) { best = d; bi = i; }
  }
  out[s] = (int32_t)bi;
}

// colours[nearest[s]] = colour[s] in seed order: later seeds overwrite (impl/random_walker_segmentation.hpp:151-152)
__global__ void k_paint_seeds(const int32_t* __restrict__ nearest, const uint32_t* __restrict__ colour, int ns,
                              uint32_t* __restrict__ colours) {
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int s = 0; s < ns; ++s)
      if (nearest[s] >= 0) colours[nearest[s]] = colour[s];
}

// ---- Laplacian assembly (RandomWalker::computeVertexDegrees + buildLinearSystem, impl/random_walker.hpp:103-184)
// degree = sum of incident weights in edge order; rows = unlabelled vertices with |degree| >= FLT_EPSILON
__global__ void k_degree_rows(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ eid,
                              const float* __restrict__ ew, const uint32_t* __restrict__ colours,
                              float* __restrict__ degree, uint32_t* __restrict__ isrow) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  float d = 0.0f;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) d = __fadd_rn(d, ew[eid[a]]);
  degree[v] = d;
  isrow[v] = (colours[v] == 0 && fabsf(d) >= FLT_EPSILON) ? 1u : 0u;
}

__global__ void k_row_counts(size_t V, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                             const uint32_t* __restrict__ isrow, const uint32_t* __restrict__ row_scan,
                             uint32_t* __restrict__ counts, int* __restrict__ row_of) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  if (!isrow[v]) { row_of[v] = -1; return; }
  uint32_t r = row_scan[v];
  row_of[v] = (int)r;
  uint32_t c = 0;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    if (u != (uint32_t)v && isrow[u]) ++c;
  }
  counts[r] = c;
}

// One thread per row: off-diagonal entries -w towards unlabelled rows, B(row, colour-1) += w towards seeds
// (sequential float adds in out-edge order, like Eigen's setFromTriplets over the reference's triplet order).
__global__ void k_fill_system(size_t V, int Kp, const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                              const uint32_t* __restrict__ eid, const float* __restrict__ ew,
                              const uint32_t* __restrict__ colours, const float* __restrict__ degree,
                              const int* __restrict__ row_of, const uint32_t* __restrict__ srowptr,
                              int* __restrict__ scol, float* __restrict__ sval, float* __restrict__ diag,
                              float* __restrict__ B, uint32_t* __restrict__ row_vertex) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  int r = row_of[v];
  if (r < 0) return;
  row_vertex[r] = (uint32_t)v;
  diag[r] = degree[v];
  uint32_t o = srowptr[r];
  float* b = B + (size_t)r * Kp;
  for (uint32_t a = rowptr[v]; a < rowptr[v + 1]; ++a) {
    uint32_t u = col[a];
    float w = ew[eid[a]];
    uint32_t cu = colours[u];
    if (cu) b[cu - 1] = __fadd_rn(b[cu - 1], w);
    else if (u != (uint32_t)v && row_of[u] >= 0) { scol[o] = row_of[u]; sval[o] = -w; ++o; }
  }
}

}  // namespace

// ---- C ABI: components, nearest vertices ----------------------------------------------------------------------
int tcs_impl_components(tcs_graph* g) {
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V;
  g->ncomp = 0;
  if (V == 0) return TCS_OK;
  DBuf<uint32_t> parent, is_root, rank;
  TRY(parent.alloc(ctx, V)); TRY(is_root.alloc(ctx, V + 1)); TRY(rank.alloc(ctx, V + 1));
  TRY(g->comp.alloc(ctx, V));
  LAUNCH(ctx, k_iota, div_up(V, 256), 256, 0, parent.p, V);
  LAUNCH(ctx, k_cc_hook, div_up(V, 128), 128, 0, V, g->rowptr.p, g->col.p, parent.p);
  LAUNCH(ctx, k_cc_flatten, div_up(V, 256), 256, 0, V, parent.p, is_root.p);
  TRY(tcs_impl_scan_u32(ctx, is_root.p, rank.p, V));
  LAUNCH(ctx, k_cc_assign, div_up(V, 256), 256, 0, V, parent.p, rank.p, g->comp.p);
  uint32_t nc = 0;
  TRY(tcs_impl_read_u32(ctx, rank.p + V, &nc));
  g->ncomp = (int)nc;
  return TCS_OK;
}

extern "C" int tcs_components(tcs_graph* g, int* ncomponents, int32_t* component) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  CK(cudaSetDevice(ctx->device));
  TRY(tcs_impl_components(g));
  if (ncomponents) *ncomponents = g->ncomp;
  if (component && g->V) {
    CK(cudaMemcpyAsync(component, g->comp.p, 4 * g->V, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return TCS_OK;
}

int tcs_impl_nearest_vertices(tcs_graph* g, const float* seed_xyz, size_t ns, DBuf<int32_t>& d_nearest) {
  tcs_ctx* ctx = g->ctx;
  TRY(d_nearest.alloc(ctx, ns));
  if (ns == 0) return TCS_OK;
  DBuf<float> d_seeds;
  TRY(d_seeds.alloc(ctx, 3 * ns));
  CK(cudaMemcpyAsync(d_seeds.p, seed_xyz, 12 * ns, cudaMemcpyHostToDevice, ctx->stream));
  if (g->V >= (size_t)16 * kSeedTile && ns >= 8) {
    const int tiles = (int)div_up(g->V, (size_t)kSeedTile);
    DBuf<float> part_d;
    DBuf<uint32_t> part_i;
    TRY(part_d.alloc(ctx, (size_t)tiles * ns)); TRY(part_i.alloc(ctx, (size_t)tiles * ns));
    LAUNCH(ctx, k_seed_nn_tiled, (unsigned)tiles, 256, 0, g->V, (int)ns, g->x.p, g->y.p, g->z.p, d_seeds.p, part_d.p, part_i.p);
    LAUNCH(ctx, k_seed_nn_merge, div_up(ns, 8), 256, 0, (int)ns, tiles, part_d.p, part_i.p, d_nearest.p);
  } else {
    LAUNCH(ctx, k_seed_nn, (unsigned)ns, 512, 0, g->V, g->x.p, g->y.p, g->z.p, d_seeds.p, d_nearest.p);
  }
  CK(cudaGetLastError());
  return TCS_OK;
}

extern "C" int tcs_nearest_vertices(tcs_graph* g, const float* seed_xyz, size_t ns, int32_t* out) {
  if (!g || (ns && (!seed_xyz || !out))) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  CK(cudaSetDevice(ctx->device));
  DBuf<int32_t> d;
  TRY(tcs_impl_nearest_vertices(g, seed_xyz, ns, d));
  if (ns) CK(cudaMemcpyAsync(out, d.p, 4 * ns, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

// Device address of the scratch array, resolved once per device with that device current (in host code the array's
// name is only a stand-in for the symbol, not an address a kernel could use).
static int noalloc_scratch_ptr(tcs_ctx* ctx, NoallocScratch** out) {
  static std::mutex mu;
  static NoallocScratch* ptrs[64] = {};
  if (ctx->device < 0 || ctx->device >= 64) return tcs_fail(ctx, TCS_ERR_INPUT, "device index %d outside 0..63", ctx->device);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[ctx->device]) {
    void* p = nullptr;
    CK(cudaSetDevice(ctx->device));
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return tcs_fail(ctx, TCS_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[ctx->device] = (NoallocScratch*)p;
  }
  *out = ptrs[ctx->device];
  return TCS_OK;
}

extern "C" int tcs_nearest_vertices_noalloc(tcs_graph* g, const float* seed_xyz, size_t ns, int32_t* out) {
  if (!g || (ns && (!seed_xyz || !out))) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (ns > (size_t)TCS_NOALLOC_MAX_SEEDS)
    return tcs_fail(ctx, TCS_ERR_INPUT, "tcs_nearest_vertices_noalloc: %zu seeds exceed the fixed scratch of TCS_NOALLOC_MAX_SEEDS = %d "
                    "(use tcs_nearest_vertices)", ns, TCS_NOALLOC_MAX_SEEDS);
  CK(cudaSetDevice(ctx->device));
  NoallocScratch* sc = nullptr;
  TRY(noalloc_scratch_ptr(ctx, &sc));   // ns = 0: set-up only (resolves the address; call once outside a capture)
  if (ns == 0) return TCS_OK;
  if (g->V == 0) {
    for (size_t i = 0; i < ns; ++i) out[i] = -1;
    return TCS_OK;
  }
  CK(cudaMemcpyAsync(sc->seeds, seed_xyz, 12 * ns, cudaMemcpyHostToDevice, ctx->stream));
  LAUNCH(ctx, k_seed_nn_partial, dim3(kNoallocChunks, (unsigned)ns), 256, 0, g->V, g->x.p, g->y.p, g->z.p, sc->seeds,
         sc->part_d, sc->part_i);
  LAUNCH(ctx, k_seed_nn_final, div_up(ns, 128), 128, 0, (int)ns, kNoallocChunks, sc->part_d, sc->part_i, sc->nearest);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, sc->nearest, 4 * ns, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

int tcs_impl_paint_seeds(tcs_graph* g, const DBuf<int32_t>& nearest, const uint32_t* h_colour, size_t ns) {
  tcs_ctx* ctx = g->ctx;
  TRY(g->colours.alloc(ctx, g->V));
  if (g->V) CK(cudaMemsetAsync(g->colours.p, 0, 4 * g->V, ctx->stream));
  if (ns == 0 || g->V == 0) return TCS_OK;
  DBuf<uint32_t> d_col;
  TRY(d_col.alloc(ctx, ns));
  CK(cudaMemcpyAsync(d_col.p, h_colour, 4 * ns, cudaMemcpyHostToDevice, ctx->stream));
  LAUNCH(ctx, k_paint_seeds, 1, 1, 0, nearest.p, d_col.p, (int)ns, g->colours.p);
  // h_colour must stay valid until the copy has been consumed
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

// Assemble L (off-diagonal CSR + diagonal) and B on the device from g->colours (seeds) and g->ew.
int tcs_impl_assemble(tcs_graph* g, int K) {
  tcs_ctx* ctx = g->ctx;
  tcs_system& S = g->sys;
  const size_t V = g->V;
  S.K = K;
  S.Kp = std::max(4, (K + 3) & ~3);
  S.n = 0;
  S.nnz_off = 0;
  if (V == 0) return TCS_OK;
  DBuf<float> degree;
  DBuf<uint32_t> isrow, row_scan, counts;
  TRY(degree.alloc(ctx, V)); TRY(isrow.alloc(ctx, V + 1)); TRY(row_scan.alloc(ctx, V + 1));
  LAUNCH(ctx, k_degree_rows, div_up(V, 128), 128, 0, V, g->rowptr.p, g->eid.p, g->ew.p, g->colours.p, degree.p, isrow.p);
  TRY(tcs_impl_scan_u32(ctx, isrow.p, row_scan.p, V));
  uint32_t n = 0;
  TRY(tcs_impl_read_u32(ctx, row_scan.p + V, &n));
  S.n = (int)n;
  TRY(S.row_of.alloc(ctx, V));
  TRY(counts.alloc(ctx, (size_t)n + 1));
  TRY(S.rowptr.alloc(ctx, (size_t)n + 1 + 8));  // + 8: the SpMM kernels copy 16-byte aligned slices
  LAUNCH(ctx, k_row_counts, div_up(V, 128), 128, 0, V, g->rowptr.p, g->col.p, isrow.p, row_scan.p, counts.p, S.row_of.p);
  TRY(tcs_impl_scan_u32(ctx, counts.p, S.rowptr.p, n));
  uint32_t nnz = 0;
  TRY(tcs_impl_read_u32(ctx, S.rowptr.p + n, &nnz));
  S.nnz_off = nnz;
  TRY(S.col.alloc(ctx, nnz)); TRY(S.val.alloc(ctx, (size_t)nnz + 16)); TRY(S.diag.alloc(ctx, (size_t)n + 8));
  TRY(S.B.alloc(ctx, (size_t)n * S.Kp)); TRY(S.row_vertex.alloc(ctx, n));
  if (n) {
    CK(cudaMemsetAsync(S.B.p, 0, (size_t)n * S.Kp * 4, ctx->stream));
    LAUNCH(ctx, k_fill_system, div_up(V, 128), 128, 0, V, S.Kp, g->rowptr.p, g->col.p, g->eid.p, g->ew.p, g->colours.p,
           degree.p, S.row_of.p, S.rowptr.p, S.col.p, S.val.p, S.diag.p, S.B.p, S.row_vertex.p);
  }
  CK(cudaGetLastError());
  return TCS_OK;
}
