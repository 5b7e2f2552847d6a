// tcs_b200/csrc/build.cu - voxel-grid graph builder on the device.
// Replaces pcl::graph::VoxelGridGraphBuilder::compute (graph/impl/voxel_grid_graph_builder.hpp:69-143):
//   octree insertion + leaf walk  ->  Morton keys, radix sort, segmented reduce (one thread per voxel walks its
//                                      points in ascending index order => bit-exact centroids)
//   27 unordered_map look-ups     ->  open-addressing hash of packed voxel keys, 26-neighbour probe, CSR fill
// Compiled with -fmad=false: every float operation below is the reference's IEEE operation, no contraction.
#include <cfloat>
#include <cmath>
#include <algorithm>
#include <cub/cub.cuh>

#include "common.cuh"

int tcs_impl_sort_pairs_u32(tcs_ctx* ctx, uint32_t* key_in, uint32_t* key_out, uint32_t* val_in, uint32_t* val_out,
                            size_t n, int end_bit) {
  size_t bytes = 0;
  CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, key_in, key_out, val_in, val_out, (int)n, 0, end_bit, ctx->stream));
  DBuf<char> tmp;
  TRY(tmp.alloc(ctx, bytes));
  CK(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, key_in, key_out, val_in, val_out, (int)n, 0, end_bit, ctx->stream));
  ctx->launches += 1 + (end_bit + 7) / 8;
  return TCS_OK;
}

int tcs_impl_sort_pairs_u64(tcs_ctx* ctx, uint64_t* key_in, uint64_t* key_out, uint32_t* val_in, uint32_t* val_out,
                            size_t n, int end_bit) {
  size_t bytes = 0;
  CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, key_in, key_out, val_in, val_out, (int)n, 0, end_bit, ctx->stream));
  DBuf<char> tmp;
  TRY(tmp.alloc(ctx, bytes));
  CK(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, key_in, key_out, val_in, val_out, (int)n, 0, end_bit, ctx->stream));
  ctx->launches += 1 + (end_bit + 7) / 8;
  return TCS_OK;
}

namespace {

// voxel_grid_graph_builder.hpp:79-87: x /= z, y /= z, z = log z.  std::log(float) is DEFINED as the float64
// logarithm rounded once to float32 in both the oracle and here.
__device__ __forceinline__ bool transform_point(const float4 p, float& tx, float& ty, float& tz) {
  tx = p.x / p.z;
  ty = p.y / p.z;
  tz = (float)log((double)p.z);
  return isfinite(tx) && isfinite(ty) && isfinite(tz);
}

__device__ __forceinline__ uint32_t f2ord(float f) {
  uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
inline float ord2f(uint32_t o) {
  uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
  float f;
  memcpy(&f, &b, 4);
  return f;
}

// mm[0..2] = min (ordered-uint), mm[3..5] = max, mm[6] = finite count
__global__ void k_transform_minmax(const float4* __restrict__ cloud, size_t n, uint32_t* __restrict__ mm) {
  uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
  uint32_t cnt = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float t[3];
    if (!transform_point(cloud[i], t[0], t[1], t[2])) continue;
    ++cnt;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      uint32_t o = f2ord(t[a]);
      lo[a] = min(lo[a], o);
      hi[a] = max(hi[a], o);
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = min(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], s));
      hi[a] = max(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], s));
    }
    cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, s);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(&mm[a], lo[a]);
      atomicMax(&mm[3 + a], hi[a]);
    }
    if (cnt) atomicAdd(&mm[6], cnt);
  }
}

// OctreePointCloud::adoptBoundingBoxToPoint (appendix A.2) grows the tree while points are inserted one by one:
// a point's key is taken with the box that was current at ITS insertion, and every later doubling pushes the old
// root one level down (the keys inserted so far gain a top bit on the axes that grew downwards).  The host finds the
// few growth events (k_first_outside) and hands them over as epochs: points [estart[e], estart[e+1]) use emn[e],
// then OR eor[e] (the top bits of all later doublings).
#define TCS_MAX_EPOCHS 24
struct BoxParams {
  double mn[3];
  double res;
  int depth;
  int desc;
  int nepoch;  // 1 = no growth (mn[] alone is used)
  uint32_t estart[TCS_MAX_EPOCHS];
  double emn[TCS_MAX_EPOCHS][3];
  uint32_t eor[TCS_MAX_EPOCHS][3];
};

// lowest index >= from of a finite point outside [mn, mx)  (the next point adoptBoundingBoxToPoint would grow for)
struct BoxD { double mn[3], mx[3]; };
__global__ void k_first_outside(const float4* __restrict__ cloud, size_t n, size_t from, BoxD b,
                                uint32_t* __restrict__ out) {
  uint32_t best = 0xFFFFFFFFu;
  for (size_t i = from + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float t[3];
    if (!transform_point(cloud[i], t[0], t[1], t[2])) continue;
    bool outside = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) outside = outside || (double)t[a] < b.mn[a] || (double)t[a] >= b.mx[a];
    if (outside) { best = (uint32_t)i; break; }  // indices ascend along the grid-stride loop
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, s));
  if ((threadIdx.x & 31) == 0 && best != 0xFFFFFFFFu) atomicMin(out, best);
}

// transformed coordinates of the point *idx (device-side so that they are the same floats k_sort_keys sees)
__global__ void k_transformed_point_guarded(const float4* __restrict__ cloud, const uint32_t* __restrict__ idx,
                                            float* __restrict__ out) {
  float t[3] = {0.f, 0.f, 0.f};
  if (*idx != 0xFFFFFFFFu) transform_point(cloud[*idx], t[0], t[1], t[2]);
  out[0] = t[0]; out[1] = t[1]; out[2] = t[2];
}

__device__ __forceinline__ uint64_t spread3(uint32_t v) {  // 21 bits -> every third bit
  uint64_t x = v & 0x1FFFFFu;
  x = (x | x << 32) & 0x1F00000000FFFFull;
  x = (x | x << 16) & 0x1F0000FF0000FFull;
  x = (x | x << 8) & 0x100F00F00F00F00Full;
  x = (x | x << 4) & 0x10C30C30C30C30C3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}
__device__ __forceinline__ uint32_t compact3(uint64_t x) {
  x &= 0x1249249249249249ull;
  x = (x ^ (x >> 2)) & 0x10C30C30C30C30C3ull;
  x = (x ^ (x >> 4)) & 0x100F00F00F00F00Full;
  x = (x ^ (x >> 8)) & 0x1F0000FF0000FFull;
  x = (x ^ (x >> 16)) & 0x1F00000000FFFFull;
  x = (x ^ (x >> 32)) & 0x1FFFFFull;
  return (uint32_t)x;
}

// OctreePointCloud::genOctreeKeyforPoint (appendix A.2) + child index x<<2|y<<1|z per level => x-major Morton.
// Sort key: leaves are visited in descending (PCL <= 1.8) Morton order, so sort ascending on the complement;
// non-finite points get the sentinel 1 << 3*depth and end up behind every voxel.
template <typename KeyT>
__global__ void k_sort_keys(const float4* __restrict__ cloud, size_t n, BoxParams bp, KeyT* __restrict__ keys,
                            uint32_t* __restrict__ idx) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float tx, ty, tz;
  const uint64_t sentinel = 1ull << (3 * bp.depth);
  uint64_t k = sentinel;
  if (transform_point(cloud[i], tx, ty, tz)) {
    uint32_t kx, ky, kz;
    if (bp.nepoch <= 1) {
      kx = (uint32_t)(((double)tx - bp.mn[0]) / bp.res);
      ky = (uint32_t)(((double)ty - bp.mn[1]) / bp.res);
      kz = (uint32_t)(((double)tz - bp.mn[2]) / bp.res);
    } else {
      int e = 0;
      while (e + 1 < bp.nepoch && i >= bp.estart[e + 1]) ++e;
      kx = (uint32_t)(((double)tx - bp.emn[e][0]) / bp.res) | bp.eor[e][0];
      ky = (uint32_t)(((double)ty - bp.emn[e][1]) / bp.res) | bp.eor[e][1];
      kz = (uint32_t)(((double)tz - bp.emn[e][2]) / bp.res) | bp.eor[e][2];
    }
    uint64_t m = spread3(kx) << 2 | spread3(ky) << 1 | spread3(kz);
    k = bp.desc ? (~m & (sentinel - 1)) : m;
  }
  keys[i] = (KeyT)k;
  idx[i] = (uint32_t)i;
}

template <typename KeyT>
__global__ void k_heads(const KeyT* __restrict__ keys, size_t nfinite, uint32_t* __restrict__ flag) {
  size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (j >= nfinite) return;
  flag[j] = (j == 0 || keys[j] != keys[j - 1]) ? 1u : 0u;
}

__global__ void k_voxel_starts(const uint32_t* __restrict__ flag, const uint32_t* __restrict__ scan, size_t nfinite,
                               uint32_t* __restrict__ vstart) {
  size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (j > nfinite) return;
  if (j == nfinite) { vstart[scan[nfinite]] = (uint32_t)nfinite; return; }
  if (flag[j]) vstart[scan[j]] = (uint32_t)j;
}

// pcl::CentroidPoint (appendix A.3): float sums in ascending point-index order, divided by (float)n
template <typename KeyT>
__global__ void k_voxel_reduce(const float4* __restrict__ cloud, const KeyT* __restrict__ keys,
                               const uint32_t* __restrict__ idx, const uint32_t* __restrict__ vstart, size_t V,
                               int depth, int desc, int has_color, float* __restrict__ ox, float* __restrict__ oy,
                               float* __restrict__ oz, uint32_t* __restrict__ orgba, uint32_t* __restrict__ okeys,
                               uint64_t* __restrict__ opacked, uint32_t* __restrict__ p2v) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t j0 = vstart[v], j1 = vstart[v + 1];
  float sx = 0.f, sy = 0.f, sz = 0.f, sr = 0.f, sg = 0.f, sb = 0.f, sa = 0.f;
  for (uint32_t j = j0; j < j1; ++j) {
    uint32_t i = idx[j];
    float4 p = cloud[i];
    sx += p.x; sy += p.y; sz += p.z;
    uint32_t c = __float_as_uint(p.w);
    sr += (float)((c >> 16) & 0xFF);
    sg += (float)((c >> 8) & 0xFF);
    sb += (float)(c & 0xFF);
    sa += (float)((c >> 24) & 0xFF);
    p2v[i] = (uint32_t)v;
  }
  float cnt = (float)(j1 - j0);
  ox[v] = sx / cnt; oy[v] = sy / cnt; oz[v] = sz / cnt;
  orgba[v] = has_color ? ((uint32_t)(sa / cnt) << 24 | (uint32_t)(sr / cnt) << 16 | (uint32_t)(sg / cnt) << 8 |
                          (uint32_t)(sb / cnt))
                       : 0u;
  uint64_t m = (uint64_t)keys[j0];
  if (desc) m = ~m & ((1ull << (3 * depth)) - 1);
  uint32_t kx = compact3(m >> 2), ky = compact3(m >> 1), kz = compact3(m);
  okeys[3 * v] = kx; okeys[3 * v + 1] = ky; okeys[3 * v + 2] = kz;
  opacked[v] = (uint64_t)kx << 42 | (uint64_t)ky << 21 | (uint64_t)kz;
}

}  // namespace

namespace {

// ---- hashed 26-neighbourhood search (voxel_grid_graph_builder.hpp:121-141) --------------------------------
#define HASH_EMPTY 0xFFFFFFFFFFFFFFFFull
__device__ __forceinline__ uint32_t hash64(uint64_t k) {  // murmur3 finaliser
  k ^= k >> 33; k *= 0xFF51AFD7ED558CCDull; k ^= k >> 33; k *= 0xC4CEB9FE1A85EC53ull; k ^= k >> 33;
  return (uint32_t)k;
}

__global__ void k_hash_insert(const uint64_t* __restrict__ packed, size_t V, unsigned long long* __restrict__ tkey,
                              uint32_t* __restrict__ tval, uint32_t mask) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint64_t k = packed[v];
  uint32_t s = hash64(k) & mask;
  for (;;) {
    unsigned long long prev = atomicCAS(&tkey[s], HASH_EMPTY, (unsigned long long)k);
    if (prev == HASH_EMPTY) { tval[s] = (uint32_t)v; return; }  // keys are unique: no prev == k case
    s = (s + 1) & mask;
  }
}

__device__ __forceinline__ uint32_t hash_find(const unsigned long long* __restrict__ tkey,
                                              const uint32_t* __restrict__ tval, uint32_t mask, uint64_t k) {
  uint32_t s = hash64(k) & mask;
  for (;;) {
    unsigned long long t = tkey[s];
    if (t == k) return tval[s];
    if (t == HASH_EMPTY) return TCS_NIL;
    s = (s + 1) & mask;
  }
}

// Direction index d = (dx+1)*9 + (dy+1)*3 + (dz+1): the reference's lexicographic (dx, dy, dz) loop order.
// nbr is d-major: nbr[d * V + v].  lower[v] = neighbours already visited when v is processed (id < v) -> the
// edges v adds itself; lowmask bit d = that neighbour is lower.
__global__ void k_probe(const uint32_t* __restrict__ keys, size_t V, const unsigned long long* __restrict__ tkey,
                        const uint32_t* __restrict__ tval, uint32_t mask, uint32_t* __restrict__ nbr,
                        uint32_t* __restrict__ lowcount, uint32_t* __restrict__ degree,
                        uint32_t* __restrict__ lowmask) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t kx = keys[3 * v], ky = keys[3 * v + 1], kz = keys[3 * v + 2];
  uint32_t nlow = 0, ndeg = 0, lm = 0;
#pragma unroll
  for (int d = 0; d < 27; ++d) {
    const int dx = d / 9 - 1, dy = (d / 3) % 3 - 1, dz = d % 3 - 1;
    uint32_t u = TCS_NIL;
    bool skip = (d == 13) || (dx < 0 && kx == 0) || (dy < 0 && ky == 0) || (dz < 0 && kz == 0);
    // depth 21: a component at 2^21 - 1 has no +1 neighbour inside the tree, and its 21-bit field would carry into the
    // next axis of the packed key (aliasing (kx, 2^21, kz) with (kx + 1, 0, kz))
    skip = skip || (dx > 0 && kx == 0x1FFFFFu) || (dy > 0 && ky == 0x1FFFFFu) || (dz > 0 && kz == 0x1FFFFFu);
    if (!skip) {
      uint64_t k = (uint64_t)(kx + dx) << 42 | (uint64_t)(ky + dy) << 21 | (uint64_t)(kz + dz);
      u = hash_find(tkey, tval, mask, k);
    }
    nbr[(size_t)d * V + v] = u;
    if (u != TCS_NIL) {
      ++ndeg;
      if (u < v) { ++nlow; lm |= 1u << d; }
    }
  }
  lowcount[v] = nlow;
  degree[v] = ndeg;
  lowmask[v] = lm;
}

// Edge list in insertion order (edge id = edge_off[v] + rank in (dx,dy,dz) order) and the out-edge lists:
// [edges v added itself, in (dx,dy,dz) order] ++ [edges added later by higher ids, ascending id] (appendix A.7).
__global__ void k_fill_edges(size_t V, const uint32_t* __restrict__ nbr, const uint32_t* __restrict__ edge_off,
                             const uint32_t* __restrict__ rowptr, const uint32_t* __restrict__ lowmask,
                             uint32_t* __restrict__ esrc, uint32_t* __restrict__ etgt, uint32_t* __restrict__ col,
                             uint32_t* __restrict__ eid) {
  size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  uint32_t e = edge_off[v], a = rowptr[v];
  uint32_t up_u[26], up_e[26];
  int nup = 0;
#pragma unroll 1
  for (int d = 0; d < 27; ++d) {
    uint32_t u = nbr[(size_t)d * V + v];
    if (u == TCS_NIL) continue;
    if (u < v) {
      esrc[e] = (uint32_t)v; etgt[e] = u;
      col[a] = u; eid[a] = e;
      ++e; ++a;
    } else {
      // edge added by u towards v: direction from u to v is 26 - d; its rank among u's own edges
      int dd = 26 - d;
      uint32_t r = __popc(lowmask[u] & ((1u << dd) - 1u));
      uint32_t ue = edge_off[u] + r;
      int p = nup++;
      while (p > 0 && up_u[p - 1] > u) { up_u[p] = up_u[p - 1]; up_e[p] = up_e[p - 1]; --p; }
      up_u[p] = u; up_e[p] = ue;
    }
  }
  for (int p = 0; p < nup; ++p) { col[a] = up_u[p]; eid[a] = up_e[p]; ++a; }
}

__global__ void k_fill_u32(uint32_t* p, size_t n, uint32_t v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// OctreePointCloud::getKeyBitSize (appendix A.2) - host, float64, identical code path to the CPU semantics
struct OctreeBox {
  double res, mn[3], mx[3];
  unsigned depth = 0;
  bool negative_oversize = false;
  void fit() {
    const float eps = FLT_EPSILON;
    unsigned mk[3];
    for (int a = 0; a < 3; ++a) mk[a] = static_cast<unsigned>((mx[a] - mn[a]) / res);
    unsigned max_voxels = std::max(std::max(std::max(mk[0], mk[1]), mk[2]), 2u);
    double l2 = std::log((double)max_voxels) / std::log(2.0);
    depth = std::max(std::min(64u, static_cast<unsigned>(std::ceil(l2 - eps))), 0u);
    double side = static_cast<double>(1ull << depth) * res - eps;
    for (int a = 0; a < 3; ++a) {
      double over = (side - (mx[a] - mn[a])) / 2.0;
      if (over < 0) negative_oversize = true;
      mn[a] -= over;
      mx[a] += over;
    }
  }
  // getKeyBitSize() on a tree that already holds leaves: depth from the current box, max = min + side
  void refit_nonempty() {
    const float eps = FLT_EPSILON;
    unsigned mk[3];
    for (int a = 0; a < 3; ++a) mk[a] = static_cast<unsigned>((mx[a] - mn[a]) / res);
    unsigned max_voxels = std::max(std::max(std::max(mk[0], mk[1]), mk[2]), 2u);
    double l2 = std::log((double)max_voxels) / std::log(2.0);
    depth = std::max(std::min(64u, static_cast<unsigned>(std::ceil(l2 - eps))), 0u);
    double side = static_cast<double>(1ull << depth) * res - eps;
    for (int a = 0; a < 3; ++a) mx[a] = mn[a] + side;
  }
};

template <typename KeyT>
int build_sorted(tcs_ctx* ctx, tcs_graph* g, const BoxParams& bp, size_t n, size_t nfinite, uint32_t* Vout,
                 DBuf<uint64_t>& packed) {
  DBuf<KeyT> k0, k1;
  DBuf<uint32_t> i0, i1;
  TRY(k0.alloc(ctx, n)); TRY(k1.alloc(ctx, n)); TRY(i0.alloc(ctx, n)); TRY(i1.alloc(ctx, n));
  LAUNCH(ctx, k_sort_keys<KeyT>, div_up(n, 256), 256, 0, ctx->d_cloud, n, bp, k0.p, i0.p);
  if (sizeof(KeyT) == 4)
    TRY(tcs_impl_sort_pairs_u32(ctx, (uint32_t*)k0.p, (uint32_t*)k1.p, i0.p, i1.p, n, 3 * bp.depth + 1));
  else
    TRY(tcs_impl_sort_pairs_u64(ctx, (uint64_t*)k0.p, (uint64_t*)k1.p, i0.p, i1.p, n, 3 * bp.depth + 1));
  k0.release(); i0.release();
  DBuf<uint32_t> flag, scan;
  TRY(flag.alloc(ctx, nfinite + 1)); TRY(scan.alloc(ctx, nfinite + 1));
  LAUNCH(ctx, k_heads<KeyT>, div_up(nfinite, 256), 256, 0, k1.p, nfinite, flag.p);
  TRY(tcs_impl_scan_u32(ctx, flag.p, scan.p, nfinite));
  uint32_t V = 0;
  TRY(tcs_impl_read_u32(ctx, scan.p + nfinite, &V));
  *Vout = V;
  DBuf<uint32_t> vstart;
  TRY(vstart.alloc(ctx, (size_t)V + 1));
  LAUNCH(ctx, k_voxel_starts, div_up(nfinite + 1, 256), 256, 0, flag.p, scan.p, nfinite, vstart.p);
  TRY(tcs_impl_alloc_vertices(g, V));
  TRY(g->keys.alloc(ctx, 3 * (size_t)V));
  TRY(packed.alloc(ctx, V));
  LAUNCH(ctx, k_voxel_reduce<KeyT>, div_up(V, 128), 128, 0, ctx->d_cloud, k1.p, i1.p, vstart.p, (size_t)V, bp.depth,
         bp.desc, ctx->has_color ? 1 : 0, g->x.p, g->y.p, g->z.p, g->rgba.p, g->keys.p, packed.p, g->p2v.p);
  CK(cudaGetLastError());
  return TCS_OK;
}

}  // namespace

extern "C" int tcs_build_voxel_graph(tcs_ctx* ctx, float resolution, int leaf_order_desc, tcs_graph** out) {
  if (!ctx || !out) return TCS_ERR_INPUT;
  *out = nullptr;
  CK(cudaSetDevice(ctx->device));
  const size_t n = ctx->n_points;
  if (n && !ctx->d_cloud) return tcs_fail(ctx, TCS_ERR_STATE, "no input cloud");
  if (n && !(resolution > 0.0f)) return tcs_fail(ctx, TCS_ERR_INPUT, "voxel resolution must be positive");
  tcs_graph* g = new tcs_graph;
  g->ctx = ctx;
  g->N = n;
  g->has_color = ctx->has_color;
  g->from_builder = true;
  auto fail = [&](int st) { delete g; return st; };
#define TRYG(call) do { int _s = (call); if (_s != TCS_OK) return fail(_s); } while (0)
#define CKG(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return fail(tcs_fail(ctx, TCS_ERR_CUDA, "%s:%d %s", __FILE__, __LINE__, cudaGetErrorString(_e))); } while (0)
  TRYG(g->p2v.alloc(ctx, n));
  // empty cloud -> empty graph, no error (voxel_grid_graph_builder.hpp:72-77)
  TRYG(g->rowptr.alloc(ctx, 1));
  CKG(cudaMemsetAsync(g->rowptr.p, 0, 4, ctx->stream));
  if (n == 0) { *out = g; return TCS_OK; }
  LAUNCH(ctx, k_fill_u32, div_up(n, 256), 256, 0, g->p2v.p, n, TCS_NIL);

  // :79-90 transform + getMinMax3D over finite points
  uint32_t* mm = (uint32_t*)ctx->d_scal;
  {
    uint32_t init[7] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u};
    memcpy(ctx->h_pin, init, sizeof init);
    CKG(cudaMemcpyAsync(mm, ctx->h_pin, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
  }
  unsigned grid = std::min<unsigned>(div_up(n, 256), (unsigned)ctx->sm_count * 8);
  LAUNCH(ctx, k_transform_minmax, grid, 256, 0, ctx->d_cloud, n, mm);
  uint32_t hmm[7];
  CKG(cudaMemcpyAsync(ctx->h_pin, mm, sizeof hmm, cudaMemcpyDeviceToHost, ctx->stream));
  CKG(cudaStreamSynchronize(ctx->stream));
  memcpy(hmm, ctx->h_pin, sizeof hmm);
  const size_t nfinite = hmm[6];
  if (nfinite == 0) { CKG(cudaStreamSynchronize(ctx->stream)); *out = g; return TCS_OK; }

  // :93-96 OctreePointCloud(resolution).defineBoundingBox(min, max)
  OctreeBox box;
  box.res = (double)resolution;
  for (int a = 0; a < 3; ++a) {
    double lo = (double)ord2f(hmm[a]), hi = (double)ord2f(hmm[3 + a]);
    box.mn[a] = std::min(lo, hi);
    box.mx[a] = std::max(lo, hi);
  }
  box.fit();
  BoxParams bp;
  bp.nepoch = 1;
  bp.estart[0] = 0;
  for (int a = 0; a < 3; ++a) { bp.emn[0][a] = box.mn[a]; bp.eor[0][a] = 0; }
  {
    // Points outside [min, max) make PCL grow the tree while inserting (adoptBoundingBoxToPoint); this happens when
    // max_voxels is an exact power of two (the fitted cube is then a hair smaller than the extent).  The extreme
    // coordinates tell exactly whether any point is outside; if so, replay the growth events in insertion order.
    bool outside = false;
    for (int a = 0; a < 3; ++a) {
      double lo = (double)ord2f(hmm[a]), hi = (double)ord2f(hmm[3 + a]);
      outside = outside || std::min(lo, hi) < box.mn[a] || std::max(lo, hi) >= box.mx[a];
    }
    uint32_t* d_idx = (uint32_t*)ctx->d_scal + 8;
    float* d_pt = (float*)ctx->d_scal + 9;
    size_t from = 0;
    while (outside && from < n) {
      CKG(cudaMemsetAsync(d_idx, 0xFF, 4, ctx->stream));
      BoxD bd;
      for (int a = 0; a < 3; ++a) { bd.mn[a] = box.mn[a]; bd.mx[a] = box.mx[a]; }
      unsigned gr = std::min<unsigned>(div_up(n - from, 256), (unsigned)ctx->sm_count * 8);
      LAUNCH(ctx, k_first_outside, gr, 256, 0, ctx->d_cloud, n, from, bd, d_idx);
      LAUNCH(ctx, k_transformed_point_guarded, 1, 1, 0, ctx->d_cloud, d_idx, d_pt);
      CKG(cudaMemcpyAsync(ctx->h_pin, d_idx, 16, cudaMemcpyDeviceToHost, ctx->stream));
      CKG(cudaStreamSynchronize(ctx->stream));
      uint32_t idx;
      float pf[3];
      memcpy(&idx, ctx->h_pin, 4);
      memcpy(pf, (char*)ctx->h_pin + 4, 12);
      if (idx == 0xFFFFFFFFu) break;
      const double p[3] = {(double)pf[0], (double)pf[1], (double)pf[2]};
      if (bp.nepoch >= TCS_MAX_EPOCHS)
        return fail(tcs_fail(ctx, TCS_ERR_DEPTH, "octree grew more than %d times", TCS_MAX_EPOCHS - 1));
      for (;;) {
        bool up[3], any = false;
        for (int a = 0; a < 3; ++a) {
          up[a] = p[a] >= box.mx[a];
          any = any || up[a] || p[a] < box.mn[a];
        }
        if (!any) break;
        if (box.depth >= 31) return fail(tcs_fail(ctx, TCS_ERR_DEPTH, "octree depth exceeds 31 while growing"));
        // the old root becomes child ((!upX)<<2 | (!upY)<<1 | !upZ) of the new root
        const uint32_t bit = 1u << box.depth;
        for (int e = 0; e < bp.nepoch; ++e)
          for (int a = 0; a < 3; ++a) if (!up[a]) bp.eor[e][a] |= bit;
        double side = static_cast<double>(1ull << box.depth) * box.res;
        for (int a = 0; a < 3; ++a) if (!up[a]) box.mn[a] -= side;
        box.depth++;
        side = static_cast<double>(1ull << box.depth) * box.res - FLT_EPSILON;
        for (int a = 0; a < 3; ++a) box.mx[a] = box.mn[a] + side;
        box.refit_nonempty();
      }
      const int e = bp.nepoch++;
      bp.estart[e] = idx;
      for (int a = 0; a < 3; ++a) { bp.emn[e][a] = box.mn[a]; bp.eor[e][a] = 0; }
      from = (size_t)idx + 1;
    }
  }
  if (box.depth > 21 || box.depth == 0) return fail(tcs_fail(ctx, TCS_ERR_DEPTH, "octree depth %u outside 1..21", box.depth));
  g->depth = (int)box.depth;
  for (int a = 0; a < 3; ++a) { bp.mn[a] = box.mn[a]; g->bbmin[a] = box.mn[a]; }
  bp.res = box.res;
  bp.depth = (int)box.depth;
  bp.desc = leaf_order_desc ? 1 : 0;

  uint32_t V = 0;
  DBuf<uint64_t> packed;
  if (3 * bp.depth + 1 <= 32) TRYG(build_sorted<uint32_t>(ctx, g, bp, n, nfinite, &V, packed));
  else TRYG(build_sorted<uint64_t>(ctx, g, bp, n, nfinite, &V, packed));

  // hash of voxel keys, 26-neighbour probe
  uint32_t tsize = 64;
  while (tsize < 2 * (size_t)V) tsize <<= 1;
  DBuf<unsigned long long> tkey;
  DBuf<uint32_t> tval, nbr, lowcount, degree, lowmask, edge_off;
  TRYG(tkey.alloc(ctx, tsize)); TRYG(tval.alloc(ctx, tsize));
  CKG(cudaMemsetAsync(tkey.p, 0xFF, (size_t)tsize * 8, ctx->stream));
  LAUNCH(ctx, k_hash_insert, div_up(V, 256), 256, 0, packed.p, (size_t)V, tkey.p, tval.p, tsize - 1);
  TRYG(nbr.alloc(ctx, 27 * (size_t)V));
  TRYG(lowcount.alloc(ctx, (size_t)V + 1)); TRYG(degree.alloc(ctx, (size_t)V + 1)); TRYG(lowmask.alloc(ctx, V));
  TRYG(edge_off.alloc(ctx, (size_t)V + 1));
  LAUNCH(ctx, k_probe, div_up(V, 128), 128, 0, g->keys.p, (size_t)V, tkey.p, tval.p, tsize - 1, nbr.p, lowcount.p,
         degree.p, lowmask.p);
  TRYG(g->rowptr.alloc(ctx, (size_t)V + 1));
  TRYG(tcs_impl_scan_u32(ctx, lowcount.p, edge_off.p, V));
  TRYG(tcs_impl_scan_u32(ctx, degree.p, g->rowptr.p, V));
  uint32_t E = 0;
  TRYG(tcs_impl_read_u32(ctx, edge_off.p + V, &E));
  g->E = E;
  TRYG(g->esrc.alloc(ctx, E)); TRYG(g->etgt.alloc(ctx, E)); TRYG(g->ew.alloc(ctx, E));
  TRYG(g->col.alloc(ctx, 2 * (size_t)E)); TRYG(g->eid.alloc(ctx, 2 * (size_t)E));
  if (E) CKG(cudaMemsetAsync(g->ew.p, 0, (size_t)E * 4, ctx->stream));
  LAUNCH(ctx, k_fill_edges, div_up(V, 128), 128, 0, (size_t)V, nbr.p, edge_off.p, g->rowptr.p, lowmask.p, g->esrc.p,
         g->etgt.p, g->col.p, g->eid.p);
  CKG(cudaGetLastError());
#undef TRYG
#undef CKG
  *out = g;
  return TCS_OK;
}
