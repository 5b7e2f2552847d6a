// tcs_b200/csrc/knn.cu - NearestNeighborsGraphBuilder on the device.
//
// Replaces pcl::graph::NearestNeighborsGraphBuilder::compute, graph/impl/nearest_neighbors_graph_builder.hpp:49-119:
// vertices are the finite points in index order (:58-72), every vertex is joined to its num_neighbors nearest other
// points (KNN, :95-97) or to the nearest num_neighbors among the points closer than `radius` (RADIUS, :100-103), an edge
// is added once (`if (!boost::edge(i, nb).second) add_edge`, :105-107), point_to_vertex_map :112-116.
//
// The reference searches with a FLANN kd-tree (exact search).  Here: a uniform grid over the bounding box, cell size
// tuned so that an occupied cell holds about num_neighbors / 2 points, points counting-sorted by cell (radix sort of
// 63-bit cell keys), an open-addressing hash from cell key to the cell's point range, and one thread per query that
// visits the cells shell by shell (Chebyshev rings around its own cell) keeping the best num_neighbors candidates
// ordered by (d^2, vertex id).  A query stops as soon as no unvisited cell can hold a closer point (or one inside the
// radius); queries still short of neighbours after kMaxRing rings (isolated outliers) fall back to a scan of all
// points, so the result is the exact k-nearest set for any input.  Distances are FLANN's L2_Simple<float>:
// ((dx*dx) + dy*dy) + dz*dz in float32 (this file is compiled with -fmad=false).
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "common.cuh"

namespace {

constexpr int kMaxNeighbors = 64;  // capacity of the per-thread candidate list
constexpr int kMaxRing = 6;        // rings visited before a query falls back to the exhaustive scan

struct NnGrid {
  double mn[3];
  double h;
  int dim[3];
};

__device__ __forceinline__ uint32_t nn_f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
inline float nn_ord2f(uint32_t o) {
  const uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

__device__ __forceinline__ bool nn_finite(const float4& p) { return isfinite(p.x) && isfinite(p.y) && isfinite(p.z); }

// flag[i] = point i is finite (:58-66); bounding box of the finite points (ordered-uint atomics)
__global__ void k_nn_flag(const float4* __restrict__ cloud, size_t n, uint32_t* __restrict__ flag, uint32_t* __restrict__ mm) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
  if (i < n) {
    const float4 p = cloud[i];
    const bool f = nn_finite(p);
    flag[i] = f ? 1u : 0u;
    if (f) {
      const float c[3] = {p.x, p.y, p.z};
#pragma unroll
      for (int a = 0; a < 3; ++a) lo[a] = hi[a] = nn_f2ord(c[a]);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    lo[a] = __reduce_min_sync(0xFFFFFFFFu, lo[a]);
    hi[a] = __reduce_max_sync(0xFFFFFFFFu, hi[a]);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      if (lo[a] != 0xFFFFFFFFu) atomicMin(&mm[a], lo[a]);
      if (hi[a] != 0u) atomicMax(&mm[3 + a], hi[a]);
    }
  }
}

// vertex v = rank of the finite point (copyPointCloud(*input_, *indices_, *cloud), :71); point_to_vertex_map :112-116
__global__ void k_nn_vertices(const float4* __restrict__ cloud, size_t n, const uint32_t* __restrict__ flag,
                              const uint32_t* __restrict__ rank, int has_color, float* __restrict__ x, float* __restrict__ y,
                              float* __restrict__ z, uint32_t* __restrict__ rgba, uint32_t* __restrict__ p2v) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!flag[i]) { p2v[i] = TCS_NIL; return; }
  const uint32_t v = rank[i];
  const float4 p = cloud[i];
  x[v] = p.x; y[v] = p.y; z[v] = p.z;
  rgba[v] = has_color ? __float_as_uint(p.w) : 0u;
  p2v[i] = v;
}

__device__ __forceinline__ void nn_cell_of(const NnGrid& G, float px, float py, float pz, int c[3]) {
  const float p[3] = {px, py, pz};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    int ca = (int)floor(((double)p[a] - G.mn[a]) / G.h);
    c[a] = min(max(ca, 0), G.dim[a] - 1);
  }
}
__device__ __forceinline__ uint64_t nn_key(int cx, int cy, int cz) {
  return (uint64_t)cx | ((uint64_t)cy << 21) | ((uint64_t)cz << 42);
}

__global__ void k_nn_cell_keys(NnGrid G, size_t V, const float* __restrict__ x, const float* __restrict__ y,
                               const float* __restrict__ z, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
  const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  int c[3];
  nn_cell_of(G, x[v], y[v], z[v], c);
  key[v] = nn_key(c[0], c[1], c[2]);
  idx[v] = (uint32_t)v;
}

// head[t] = 1 where sorted position t starts a new cell
__global__ void k_nn_heads(const uint64_t* __restrict__ key, size_t V, uint32_t* __restrict__ head) {
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= V) return;
  head[t] = (t == 0 || key[t] != key[t - 1]) ? 1u : 0u;
}

__device__ __forceinline__ uint32_t nn_hash(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
  return (uint32_t)k;
}

// cell table: key -> first sorted position; cell_end via the next head.  Also the sorted point records {x, y, z, id}.
__global__ void k_nn_cells(const uint64_t* __restrict__ key, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ head,
                           const uint32_t* __restrict__ cell_of_pos, size_t V, const float* __restrict__ x,
                           const float* __restrict__ y, const float* __restrict__ z, unsigned long long* __restrict__ tkey,
                           uint32_t* __restrict__ tval, uint32_t mask, uint32_t* __restrict__ cell_start, float4* __restrict__ spts) {
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= V) return;
  const uint32_t v = idx[t];
  spts[t] = make_float4(x[v], y[v], z[v], __uint_as_float(v));
  if (!head[t]) return;
  const uint32_t c = cell_of_pos[t];  // exclusive scan of head = index of this cell
  cell_start[c] = (uint32_t)t;
  const uint64_t k = key[t];
  uint32_t s = nn_hash(k) & mask;
  while (true) {
    const unsigned long long old = atomicCAS(&tkey[s], 0xFFFFFFFFFFFFFFFFull, (unsigned long long)k);
    if (old == 0xFFFFFFFFFFFFFFFFull) { tval[s] = c; break; }
    s = (s + 1) & mask;
  }
}

__device__ __forceinline__ void nn_insert(float* bd, uint32_t* bi, int& cnt, int k, float d, uint32_t j) {
  if (cnt == k && !(d < bd[k - 1] || (d == bd[k - 1] && j < bi[k - 1]))) return;
  int pos = cnt < k ? cnt : k - 1;
  while (pos > 0 && (bd[pos - 1] > d || (bd[pos - 1] == d && bi[pos - 1] > j))) {
    bd[pos] = bd[pos - 1];
    bi[pos] = bi[pos - 1];
    --pos;
  }
  bd[pos] = d;
  bi[pos] = j;
  if (cnt < k) ++cnt;
}

// One thread per query (in cell order, so a warp's queries read the same cells).
__global__ void __launch_bounds__(128) k_nn_search(NnGrid G, size_t V, int k, int use_radius, float r2,
                                                   const float4* __restrict__ spts, const unsigned long long* __restrict__ tkey,
                                                   const uint32_t* __restrict__ tval, uint32_t mask,
                                                   const uint32_t* __restrict__ cell_start, uint32_t ncells,
                                                   uint32_t* __restrict__ nbr, uint32_t* __restrict__ cnt_out) {
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t >= V) return;
  const float4 me = spts[t];
  const uint32_t v = __float_as_uint(me.w);
  float bd[kMaxNeighbors];
  uint32_t bi[kMaxNeighbors];
  int cnt = 0;
  int c[3];
  nn_cell_of(G, me.x, me.y, me.z, c);
  bool done = false;
  for (int r = 0; r <= kMaxRing && !done; ++r) {
    for (int dz = -r; dz <= r; ++dz) {
      const int cz = c[2] + dz;
      if (cz < 0 || cz >= G.dim[2]) continue;
      for (int dy = -r; dy <= r; ++dy) {
        const int cy = c[1] + dy;
        if (cy < 0 || cy >= G.dim[1]) continue;
        const bool face = (dz == -r || dz == r || dy == -r || dy == r);
        const int step = (face || r == 0) ? 1 : 2 * r;  // interior rows of the shell: only dx = -r and dx = +r
        for (int dx = -r; dx <= r; dx += step) {
          const int cx = c[0] + dx;
          if (cx < 0 || cx >= G.dim[0]) continue;
          const uint64_t key = nn_key(cx, cy, cz);
          uint32_t s = nn_hash(key) & mask;
          uint32_t cell = TCS_NIL;
          while (true) {
            const unsigned long long kk = tkey[s];
            if (kk == (unsigned long long)key) { cell = tval[s]; break; }
            if (kk == 0xFFFFFFFFFFFFFFFFull) break;
            s = (s + 1) & mask;
          }
          if (cell == TCS_NIL) continue;
          const uint32_t b = cell_start[cell], e = cell + 1 < ncells ? cell_start[cell + 1] : (uint32_t)V;
          for (uint32_t a = b; a < e; ++a) {
            const float4 o = spts[a];
            const uint32_t j = __float_as_uint(o.w);
            if (j == v) continue;
            const float ex = me.x - o.x, ey = me.y - o.y, ez = me.z - o.z;
            const float d = ex * ex + ey * ey + ez * ez;
            if (use_radius && !(d < r2)) continue;
            nn_insert(bd, bi, cnt, k, d, j);
          }
        }
      }
    }
    // every unvisited point is at least m away: m = distance from the query to the boundary of the visited cube
    double m = 1e300;
    bool covers = true;
    const float p[3] = {me.x, me.y, me.z};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const int lo = c[a] - r, hi = c[a] + r;
      if (lo > 0) { m = fmin(m, (double)p[a] - (G.mn[a] + (double)lo * G.h)); covers = false; }
      if (hi < G.dim[a] - 1) { m = fmin(m, (G.mn[a] + (double)(hi + 1) * G.h) - (double)p[a]); covers = false; }
    }
    if (covers) { done = true; break; }
    m = fmax(m, 0.0) * (1.0 - 1e-4);  // float32 rounding of d^2 and of the cell assignment
    const double m2 = m * m;
    if (cnt == k && (double)bd[k - 1] < m2) done = true;
    if (use_radius && m2 >= (double)r2) done = true;
  }
  if (!done) {
    // isolated query: exhaustive scan
    cnt = 0;
    for (size_t a = 0; a < V; ++a) {
      const float4 o = spts[a];
      const uint32_t j = __float_as_uint(o.w);
      if (j == v) continue;
      const float ex = me.x - o.x, ey = me.y - o.y, ez = me.z - o.z;
      const float d = ex * ex + ey * ey + ez * ez;
      if (use_radius && !(d < r2)) continue;
      nn_insert(bd, bi, cnt, k, d, j);
    }
  }
  for (int j = 0; j < cnt; ++j) nbr[(size_t)v * k + j] = bi[j];
  cnt_out[v] = (uint32_t)cnt;
}

// The reference's loop adds (i, nb) unless the edge exists already.  With the lists known up front: the pair {i, nb}
// can only have been added earlier by vertex nb < i listing i, so (i, nb) is kept iff i < nb or i is not in nb's list.
__global__ void k_nn_keep(size_t V, int k, const uint32_t* __restrict__ nbr, const uint32_t* __restrict__ cnt,
                          unsigned long long* __restrict__ keep, uint32_t* __restrict__ nkeep) {
  const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  unsigned long long m = 0;
  const uint32_t c = cnt[v];
  for (uint32_t j = 0; j < c; ++j) {
    const uint32_t nb = nbr[v * k + j];
    bool kp = v < nb;
    if (!kp) {
      kp = true;
      const uint32_t cn = cnt[nb];
      for (uint32_t i = 0; i < cn; ++i)
        if (nbr[(size_t)nb * k + i] == (uint32_t)v) { kp = false; break; }
    }
    if (kp) m |= 1ull << j;
  }
  keep[v] = m;
  nkeep[v] = (uint32_t)__popcll(m);
}

__global__ void k_nn_emit(size_t V, int k, const uint32_t* __restrict__ nbr, const unsigned long long* __restrict__ keep,
                          const uint32_t* __restrict__ off, uint32_t* __restrict__ esrc, uint32_t* __restrict__ etgt) {
  const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  unsigned long long m = keep[v];
  uint32_t o = off[v];
  while (m) {
    const int j = __ffsll((long long)m) - 1;
    m &= m - 1;
    esrc[o] = (uint32_t)v;
    etgt[o] = nbr[v * k + j];
    ++o;
  }
}

}  // namespace

extern "C" int tcs_build_nn_graph(tcs_ctx* ctx, int num_neighbors, double radius, int use_radius, tcs_graph** out) {
  if (!ctx || !out) return TCS_ERR_INPUT;
  *out = nullptr;
  CK(cudaSetDevice(ctx->device));
  const size_t n = ctx->n_points;
  if (n && !ctx->d_cloud) return tcs_fail(ctx, TCS_ERR_STATE, "no input cloud");
  if (num_neighbors < 0 || num_neighbors > kMaxNeighbors)
    return tcs_fail(ctx, TCS_ERR_INPUT, "number of neighbours %d outside 0..%d", num_neighbors, kMaxNeighbors);
  if (use_radius && !(radius > 0.0)) return tcs_fail(ctx, TCS_ERR_INPUT, "search radius must be positive");
  if (n >= 0x7FFFFFFFu) return tcs_fail(ctx, TCS_ERR_INPUT, "cloud too large for 32-bit ids");
  tcs_graph* g = new tcs_graph;
  g->ctx = ctx;
  g->N = n;
  g->has_color = ctx->has_color;
  g->from_builder = true;
  auto body = [&]() -> int {
    TRY(g->p2v.alloc(ctx, n));
    TRY(g->rowptr.alloc(ctx, 1));
    CK(cudaMemsetAsync(g->rowptr.p, 0, 4, ctx->stream));
    if (n == 0) return TCS_OK;  // empty cloud -> empty graph (:51-56)
    DBuf<uint32_t> flag, rank;
    TRY(flag.alloc(ctx, n + 1)); TRY(rank.alloc(ctx, n + 1));
    uint32_t* mm = (uint32_t*)ctx->d_scal;
    {
      const uint32_t init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
      memcpy(ctx->h_pin, init, sizeof init);
      CK(cudaMemcpyAsync(mm, ctx->h_pin, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
    }
    LAUNCH(ctx, k_nn_flag, div_up(n, 256), 256, 0, ctx->d_cloud, n, flag.p, mm);
    TRY(tcs_impl_scan_u32(ctx, flag.p, rank.p, n));
    uint32_t V = 0;
    TRY(tcs_impl_read_u32(ctx, rank.p + n, &V));
    TRY(tcs_impl_alloc_vertices(g, V));
    LAUNCH(ctx, k_nn_vertices, div_up(n, 256), 256, 0, ctx->d_cloud, n, flag.p, rank.p, ctx->has_color ? 1 : 0, g->x.p, g->y.p,
           g->z.p, g->rgba.p, g->p2v.p);
    const int k = num_neighbors;
    if (V == 0 || k == 0) {
      TRY(g->rowptr.alloc(ctx, (size_t)V + 1));
      CK(cudaMemsetAsync(g->rowptr.p, 0, ((size_t)V + 1) * 4, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      return TCS_OK;
    }
    uint32_t hmm[6];
    CK(cudaMemcpyAsync(ctx->h_pin, mm, sizeof hmm, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    memcpy(hmm, ctx->h_pin, sizeof hmm);
    NnGrid G;
    double ext = 0.0;
    for (int a = 0; a < 3; ++a) {
      G.mn[a] = (double)nn_ord2f(hmm[a]);
      ext = std::max(ext, (double)nn_ord2f(hmm[3 + a]) - G.mn[a]);
    }
    // cell size: aim at `target` points per occupied cell (surface-like data: occupancy grows with h^2)
    const double target = std::max(2.0, 0.5 * k);
    const double hmin = ext > 0.0 ? ext / 1048575.0 : 1.0;
    double h = ext > 0.0 ? std::max(hmin, ext / std::sqrt(std::max(1.0, (double)V / target))) : 1.0;
    DBuf<uint64_t> key0, key1;
    DBuf<uint32_t> idx0, idx1, head, cellpos;
    TRY(key0.alloc(ctx, V)); TRY(key1.alloc(ctx, V)); TRY(idx0.alloc(ctx, V)); TRY(idx1.alloc(ctx, V));
    TRY(head.alloc(ctx, (size_t)V + 1)); TRY(cellpos.alloc(ctx, (size_t)V + 1));
    uint32_t ncells = 0;
    for (int attempt = 0; attempt < 5; ++attempt) {
      G.h = h;
      for (int a = 0; a < 3; ++a) G.dim[a] = (int)std::floor(((double)nn_ord2f(hmm[3 + a]) - G.mn[a]) / h) + 1;
      LAUNCH(ctx, k_nn_cell_keys, div_up(V, 256), 256, 0, G, (size_t)V, g->x.p, g->y.p, g->z.p, key0.p, idx0.p);
      TRY(tcs_impl_sort_pairs_u64(ctx, key0.p, key1.p, idx0.p, idx1.p, V, 63));
      LAUNCH(ctx, k_nn_heads, div_up(V, 256), 256, 0, key1.p, (size_t)V, head.p);
      TRY(tcs_impl_scan_u32(ctx, head.p, cellpos.p, V));
      TRY(tcs_impl_read_u32(ctx, cellpos.p + V, &ncells));
      const double occ = (double)V / (double)std::max(1u, ncells);
      if (getenv("TCS_DEBUG")) fprintf(stderr, "[tcs] nn grid: h=%.5g dims=%dx%dx%d cells=%u occupancy=%.2f\n", h, G.dim[0], G.dim[1], G.dim[2], ncells, occ);
      if ((occ >= 0.5 * target && occ <= 2.0 * target) || attempt == 4) break;
      double f = std::sqrt(target / occ);
      f = std::min(8.0, std::max(0.125, f));
      const double hn = std::max(hmin, h * f);
      if (hn == h || (ncells == 1 && f > 1.0)) break;  // cannot refine further / everything already in one cell
      h = hn;
    }
    uint32_t tsize = 64;
    while (tsize < 2 * (size_t)ncells) tsize <<= 1;
    DBuf<unsigned long long> tkey;
    DBuf<uint32_t> tval, cell_start;
    DBuf<float4> spts;
    TRY(tkey.alloc(ctx, tsize)); TRY(tval.alloc(ctx, tsize)); TRY(cell_start.alloc(ctx, (size_t)ncells + 1));
    TRY(spts.alloc(ctx, V));
    CK(cudaMemsetAsync(tkey.p, 0xFF, (size_t)tsize * 8, ctx->stream));
    LAUNCH(ctx, k_nn_cells, div_up(V, 256), 256, 0, key1.p, idx1.p, head.p, cellpos.p, (size_t)V, g->x.p, g->y.p, g->z.p, tkey.p,
           tval.p, tsize - 1, cell_start.p, spts.p);
    DBuf<uint32_t> nbr, cnt, nkeep, off;
    DBuf<unsigned long long> keep;
    TRY(nbr.alloc(ctx, (size_t)V * k)); TRY(cnt.alloc(ctx, V)); TRY(nkeep.alloc(ctx, (size_t)V + 1));
    TRY(off.alloc(ctx, (size_t)V + 1)); TRY(keep.alloc(ctx, V));
    const float r2 = (float)(radius * radius);
    LAUNCH(ctx, k_nn_search, div_up(V, 128), 128, 0, G, (size_t)V, k, use_radius ? 1 : 0, r2, spts.p, tkey.p, tval.p, tsize - 1,
           cell_start.p, ncells, nbr.p, cnt.p);
    LAUNCH(ctx, k_nn_keep, div_up(V, 256), 256, 0, (size_t)V, k, nbr.p, cnt.p, keep.p, nkeep.p);
    TRY(tcs_impl_scan_u32(ctx, nkeep.p, off.p, V));
    uint32_t E = 0;
    TRY(tcs_impl_read_u32(ctx, off.p + V, &E));
    g->E = E;
    TRY(g->esrc.alloc(ctx, E)); TRY(g->etgt.alloc(ctx, E)); TRY(g->ew.alloc(ctx, E));
    if (E) {
      CK(cudaMemsetAsync(g->ew.p, 0, (size_t)E * 4, ctx->stream));
      LAUNCH(ctx, k_nn_emit, div_up(V, 256), 256, 0, (size_t)V, k, nbr.p, keep.p, off.p, g->esrc.p, g->etgt.p);
    }
    TRY(tcs_impl_build_out_edges(g));
    CK(cudaStreamSynchronize(ctx->stream));
    return TCS_OK;
  };
  const int st = body();
  if (st != TCS_OK) { delete g; return st; }
  *out = g;
  return TCS_OK;
}
