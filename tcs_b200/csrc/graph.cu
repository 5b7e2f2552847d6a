// tcs_b200/csrc/graph.cu - device graph container: user-graph upload (setInputGraph path), out-edge (CSR)
// construction from an edge list, and the download accessors of the C ABI.
#include "common.cuh"

// ---- out-edge lists from an edge list --------------------------------------------------------------------
// boost::adjacency_list<vecS, vecS, undirectedS, ..., listS>: add_edge(u, v) appends to the out-edge vectors of
// u and of v, so every out-edge list is ordered by ascending edge id.
__global__ void k_count_incidence(const uint32_t* __restrict__ src, const uint32_t* __restrict__ tgt, size_t E,
                                  uint32_t* __restrict__ deg) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= E) return;
  atomicAdd(&deg[src[e]], 1u);
  atomicAdd(&deg[tgt[e]], 1u);
}

// One thread per vertex would need the incident edges; instead: stable fill by edge-id order using a per-vertex
// rank = number of incident edges with a smaller id.  Computed with a segmented approach: sort-free because we
// process edges in blocks in order?  Not parallel-safe.  We instead build (vertex, eid) pairs and radix sort by
// vertex (stable), which preserves ascending edge id inside every vertex.
__global__ void k_make_pairs(const uint32_t* __restrict__ src, const uint32_t* __restrict__ tgt, size_t E,
                             uint32_t* __restrict__ key, uint32_t* __restrict__ val) {
  size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (e >= E) return;
  // slot 2e = source side, 2e+1 = target side; a self loop (u,u) appears twice in u's list, like BGL's
  key[2 * e] = src[e];
  val[2 * e] = (uint32_t)(2 * e);
  key[2 * e + 1] = tgt[e];
  val[2 * e + 1] = (uint32_t)(2 * e + 1);
}

__global__ void k_pairs_to_csr(const uint32_t* __restrict__ sorted_val, const uint32_t* __restrict__ src,
                               const uint32_t* __restrict__ tgt, size_t E2, uint32_t* __restrict__ col,
                               uint32_t* __restrict__ eid) {
  size_t a = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (a >= E2) return;
  uint32_t s = sorted_val[a];
  uint32_t e = s >> 1;
  col[a] = (s & 1u) ? src[e] : tgt[e];
  eid[a] = e;
}

int tcs_impl_sort_pairs_u32(tcs_ctx* ctx, uint32_t* key_in, uint32_t* key_out, uint32_t* val_in, uint32_t* val_out,
                            size_t n, int end_bit);

int tcs_impl_build_out_edges(tcs_graph* g) {
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V, E = g->E;
  TRY(g->rowptr.alloc(ctx, V + 1));
  TRY(g->col.alloc(ctx, 2 * E));
  TRY(g->eid.alloc(ctx, 2 * E));
  DBuf<uint32_t> deg;
  TRY(deg.alloc(ctx, V + 1));
  CK(cudaMemsetAsync(deg.p, 0, (V + 1) * 4, ctx->stream));
  if (E) LAUNCH(ctx, k_count_incidence, div_up(E, 256), 256, 0, g->esrc.p, g->etgt.p, E, deg.p);
  TRY(tcs_impl_scan_u32(ctx, deg.p, g->rowptr.p, V));
  if (E) {
    DBuf<uint32_t> k0, k1, v0, v1;
    TRY(k0.alloc(ctx, 2 * E)); TRY(k1.alloc(ctx, 2 * E)); TRY(v0.alloc(ctx, 2 * E)); TRY(v1.alloc(ctx, 2 * E));
    LAUNCH(ctx, k_make_pairs, div_up(E, 256), 256, 0, g->esrc.p, g->etgt.p, E, k0.p, v0.p);
    int bits = 1;
    while (bits < 32 && (1ull << bits) < V) ++bits;
    TRY(tcs_impl_sort_pairs_u32(ctx, k0.p, k1.p, v0.p, v1.p, 2 * E, bits));
    LAUNCH(ctx, k_pairs_to_csr, div_up(2 * E, 256), 256, 0, v1.p, g->esrc.p, g->etgt.p, 2 * E, g->col.p, g->eid.p);
  }
  CK(cudaGetLastError());
  return TCS_OK;
}

// ---- user graph (setInputGraph) ----------------------------------------------------------------------------
__global__ void k_split3(const float* __restrict__ a, size_t n, float* __restrict__ x, float* __restrict__ y,
                         float* __restrict__ z) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  x[i] = a[3 * i]; y[i] = a[3 * i + 1]; z[i] = a[3 * i + 2];
}
__global__ void k_split4(const float* __restrict__ a, size_t n, float* __restrict__ x, float* __restrict__ y,
                         float* __restrict__ z, float* __restrict__ w) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  x[i] = a[4 * i]; y[i] = a[4 * i + 1]; z[i] = a[4 * i + 2]; w[i] = a[4 * i + 3];
}
__global__ void k_join3(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z,
                        size_t n, float* __restrict__ a) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  a[3 * i] = x[i]; a[3 * i + 1] = y[i]; a[3 * i + 2] = z[i];
}
__global__ void k_join4(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z,
                        const float* __restrict__ w, size_t n, float* __restrict__ a) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  a[4 * i] = x[i]; a[4 * i + 1] = y[i]; a[4 * i + 2] = z[i]; a[4 * i + 3] = w[i];
}

int tcs_impl_alloc_vertices(tcs_graph* g, size_t V) {
  tcs_ctx* ctx = g->ctx;
  g->V = V;
  DBuf<float>* f[] = {&g->x, &g->y, &g->z, &g->nx, &g->ny, &g->nz, &g->nw, &g->curv};
  for (auto* b : f) {
    TRY(b->alloc(ctx, V));
    if (V) CK(cudaMemsetAsync(b->p, 0, V * 4, ctx->stream));
  }
  TRY(g->rgba.alloc(ctx, V));
  if (V) CK(cudaMemsetAsync(g->rgba.p, 0, V * 4, ctx->stream));
  return TCS_OK;
}

static int upload(tcs_ctx* ctx, void* d, const void* h, size_t bytes) {
  if (bytes) CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return TCS_OK;
}
static int download(tcs_ctx* ctx, void* h, const void* d, size_t bytes) {
  if (bytes) CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_upload_xyz(tcs_graph* g, const float* xyz) {
  if (!g || !xyz) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  DBuf<float> t;
  TRY(t.alloc(ctx, 3 * g->V));
  TRY(upload(ctx, t.p, xyz, 12 * g->V));
  if (g->V) LAUNCH(ctx, k_split3, div_up(g->V, 256), 256, 0, t.p, g->V, g->x.p, g->y.p, g->z.p);
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_upload_normals(tcs_graph* g, const float* normals4, const float* curvature) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (normals4) {
    DBuf<float> t;
    TRY(t.alloc(ctx, 4 * g->V));
    TRY(upload(ctx, t.p, normals4, 16 * g->V));
    if (g->V) LAUNCH(ctx, k_split4, div_up(g->V, 256), 256, 0, t.p, g->V, g->nx.p, g->ny.p, g->nz.p, g->nw.p);
    CK(cudaStreamSynchronize(ctx->stream));
  }
  if (curvature) { TRY(upload(ctx, g->curv.p, curvature, 4 * g->V)); CK(cudaStreamSynchronize(ctx->stream)); }
  return TCS_OK;
}

extern "C" int tcs_graph_upload_weights(tcs_graph* g, const float* weight) {
  if (!g || !weight) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  TRY(upload(ctx, g->ew.p, weight, 4 * g->E));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_create(tcs_ctx* ctx, size_t V, const float* xyz, const uint32_t* rgba,
                                const float* normals4, const float* curvature, size_t E, const uint32_t* src,
                                const uint32_t* tgt, const float* weight, tcs_graph** out) {
  if (!ctx || !out) return TCS_ERR_INPUT;
  *out = nullptr;
  if (V >= 0x7FFFFFFFu || E >= 0x3FFFFFFFu) return tcs_fail(ctx, TCS_ERR_INPUT, "graph too large for 32-bit ids");
  if (E && (!src || !tgt)) return tcs_fail(ctx, TCS_ERR_INPUT, "edge arrays missing");
  for (size_t e = 0; e < E; ++e)
    if (src[e] >= V || tgt[e] >= V) return tcs_fail(ctx, TCS_ERR_INPUT, "edge %zu references a vertex >= V", e);
  CK(cudaSetDevice(ctx->device));
  tcs_graph* g = new tcs_graph;
  g->ctx = ctx;
  g->has_color = rgba != nullptr;
  int st = tcs_impl_alloc_vertices(g, V);
  if (st == TCS_OK && xyz && V) st = tcs_graph_upload_xyz(g, xyz);
  if (st == TCS_OK && rgba) st = upload(ctx, g->rgba.p, rgba, 4 * V);
  if (st == TCS_OK && (normals4 || curvature)) st = tcs_graph_upload_normals(g, normals4, curvature);
  g->E = E;
  if (st == TCS_OK) st = g->esrc.alloc(ctx, E);
  if (st == TCS_OK) st = g->etgt.alloc(ctx, E);
  if (st == TCS_OK) st = g->ew.alloc(ctx, E);
  if (st == TCS_OK) st = upload(ctx, g->esrc.p, src, 4 * E);
  if (st == TCS_OK) st = upload(ctx, g->etgt.p, tgt, 4 * E);
  if (st == TCS_OK) {
    if (weight) st = upload(ctx, g->ew.p, weight, 4 * E);
    else if (E && cudaMemsetAsync(g->ew.p, 0, 4 * E, ctx->stream) != cudaSuccess) st = TCS_ERR_CUDA;
  }
  if (st == TCS_OK) st = tcs_impl_build_out_edges(g);
  if (st == TCS_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) st = tcs_fail(ctx, TCS_ERR_CUDA, "sync failed");
  if (st != TCS_OK) { delete g; return st; }
  *out = g;
  return TCS_OK;
}

extern "C" void tcs_graph_destroy(tcs_graph* g) {
  if (!g) return;
  cudaSetDevice(g->ctx->device);
  delete g;
}

extern "C" int tcs_graph_sizes(const tcs_graph* g, size_t* V, size_t* E, size_t* N, int* depth) {
  if (!g) return TCS_ERR_INPUT;
  if (V) *V = g->V;
  if (E) *E = g->E;
  if (N) *N = g->N;
  if (depth) *depth = g->depth;
  return TCS_OK;
}

extern "C" int tcs_graph_bbox_min(const tcs_graph* g, double* m) {
  if (!g || !m) return TCS_ERR_INPUT;
  for (int a = 0; a < 3; ++a) m[a] = g->bbmin[a];
  return TCS_OK;
}

extern "C" int tcs_graph_download_vertices(tcs_graph* g, float* xyz, uint32_t* rgba, float* normals4,
                                           float* curvature) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  const size_t V = g->V;
  if (V == 0) return TCS_OK;
  DBuf<float> t;
  TRY(t.alloc(ctx, 4 * V));
  if (xyz) {
    LAUNCH(ctx, k_join3, div_up(V, 256), 256, 0, g->x.p, g->y.p, g->z.p, V, t.p);
    TRY(download(ctx, xyz, t.p, 12 * V));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  if (normals4) {
    LAUNCH(ctx, k_join4, div_up(V, 256), 256, 0, g->nx.p, g->ny.p, g->nz.p, g->nw.p, V, t.p);
    TRY(download(ctx, normals4, t.p, 16 * V));
  }
  if (rgba) TRY(download(ctx, rgba, g->rgba.p, 4 * V));
  if (curvature) TRY(download(ctx, curvature, g->curv.p, 4 * V));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_keys(tcs_graph* g, uint32_t* keys3) {
  if (!g || !keys3) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (g->keys.n < 3 * g->V) return tcs_fail(ctx, TCS_ERR_STATE, "graph has no voxel keys (not built by the voxel builder)");
  TRY(download(ctx, keys3, g->keys.p, 12 * g->V));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_edges(tcs_graph* g, uint32_t* src, uint32_t* tgt, float* weight) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (src) TRY(download(ctx, src, g->esrc.p, 4 * g->E));
  if (tgt) TRY(download(ctx, tgt, g->etgt.p, 4 * g->E));
  if (weight) TRY(download(ctx, weight, g->ew.p, 4 * g->E));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_csr(tcs_graph* g, uint32_t* rowptr, uint32_t* col, uint32_t* eid) {
  if (!g) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (rowptr) {
    if (g->rowptr.p) TRY(download(ctx, rowptr, g->rowptr.p, 4 * (g->V + 1)));
    else rowptr[0] = 0;
  }
  if (col) TRY(download(ctx, col, g->col.p, 8 * g->E));
  if (eid) TRY(download(ctx, eid, g->eid.p, 8 * g->E));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_point_to_vertex(tcs_graph* g, uint32_t* map) {
  if (!g || !map) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  TRY(download(ctx, map, g->p2v.p, 4 * g->N));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_colours(tcs_graph* g, uint32_t* colours) {
  if (!g || !colours) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (g->colours.n < g->V) return tcs_fail(ctx, TCS_ERR_STATE, "no segmentation result on this graph");
  TRY(download(ctx, colours, g->colours.p, 4 * g->V));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" int tcs_graph_download_potentials(tcs_graph* g, float* potentials) {
  if (!g || !potentials) return TCS_ERR_INPUT;
  tcs_ctx* ctx = g->ctx;
  if (!g->potentials.p) return tcs_fail(ctx, TCS_ERR_STATE, "potentials were not stored (store_potentials = 0)");
  TRY(download(ctx, potentials, g->potentials.p, 4 * g->V * (size_t)g->pot_K));
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

extern "C" const uint32_t* tcs_graph_d_colours(const tcs_graph* g) { return g ? g->colours.p : nullptr; }
extern "C" const int32_t* tcs_graph_d_point_cluster(const tcs_graph* g) { return g ? g->point_cluster.p : nullptr; }
extern "C" const float* tcs_graph_d_best_potential(const tcs_graph* g) { return g ? g->best_val.p : nullptr; }
