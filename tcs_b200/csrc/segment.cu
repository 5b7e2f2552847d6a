// tcs_b200/csrc/segment.cu - the whole class path in one call: RandomWalkerSegmentation::setSeeds + segment
// (impl/random_walker_segmentation.hpp:82-224) with preComputeGraph (:103-136), or the CLI pipeline
// (src/random_walker_segmentation_app.cpp:122-134 with the factory defaults) for TCS_PRESET_CLI.
#include <algorithm>
#include <map>
#include <set>

#include "common.cuh"

namespace {

// cluster assembly (:200-223): clusters[colour - 1] <- point, nil / unlabelled vertex -> last cluster (index K)
__global__ void k_point_cluster(size_t N, size_t V, int K, const uint32_t* __restrict__ p2v,
                                const uint32_t* __restrict__ colours, int32_t* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  uint32_t v = p2v[i];
  uint32_t c = (v < V) ? colours[v] : 0u;
  out[i] = c ? (int32_t)c - 1 : K;
}

}  // namespace

int tcs_impl_point_clusters(tcs_graph* g, int K) {
  tcs_ctx* ctx = g->ctx;
  TRY(g->point_cluster.alloc(ctx, g->N));
  if (g->N == 0) return TCS_OK;
  if (!g->p2v.p) return tcs_fail(ctx, TCS_ERR_STATE, "graph has no point-to-vertex map");
  LAUNCH(ctx, k_point_cluster, div_up(g->N, 256), 256, 0, g->N, g->V, K, g->p2v.p, g->colours.p, g->point_cluster.p);
  CK(cudaGetLastError());
  return TCS_OK;
}

namespace {

// setSeeds (:82-101): unique labels in ascending order -> colours 1..K
int seed_colours(tcs_ctx* ctx, const uint32_t* seed_label, size_t ns, std::vector<uint32_t>& seed_colour, int* K_out) {
  std::set<uint32_t> labels(seed_label, seed_label + ns);
  std::map<uint32_t, uint32_t> label_colour;
  for (uint32_t l : labels) { uint32_t c = (uint32_t)label_colour.size() + 1; label_colour[l] = c; }
  const int K = (int)labels.size();
  if (K > TCS_MAX_LABELS) return tcs_fail(ctx, TCS_ERR_INPUT, "more than %d distinct seed labels", TCS_MAX_LABELS);
  seed_colour.resize(ns);
  for (size_t i = 0; i < ns; ++i) seed_colour[i] = label_colour[seed_label[i]];
  *K_out = K;
  return TCS_OK;
}

// preComputeGraph (:103-136) or the CLI pipeline on the uploaded cloud, then the seed painting of segment (:144-152).
// Records the stage boundaries in ev[0..8] when ev != NULL.  *out owns the graph also on failure (caller destroys).
int prepare_scene(tcs_ctx* ctx, float resolution, int preset, int mode, const float* seed_xyz,
                  const std::vector<uint32_t>& seed_colour, size_t ns, cudaEvent_t* ev, tcs_graph** out) {
  tcs_graph* g = nullptr;
  if (ev) CK(cudaEventRecord(ev[0], ctx->stream));
  TRY(tcs_build_voxel_graph(ctx, resolution, 1, &g));
  *out = g;
  int st = TCS_OK;
  auto step = [&](int s) { if (st == TCS_OK) st = s; };
  auto mark = [&](int i) { if (ev) cudaEventRecord(ev[i], ctx->stream); };
  if (g->V == 0) return TCS_OK;
  mark(1);
  step(tcs_normals(g, 0, mode));
  mark(2);
  if (preset == TCS_PRESET_CLI) {
    step(tcs_smoothen(g, 0.012f, 0.0012f));
    mark(3);
    step(tcs_normals(g, 0, mode));
  } else {
    mark(3);
  }
  mark(4);
  step(tcs_signed_curvatures(g));
  mark(5);
  tcs_term terms[4];
  if (preset == TCS_PRESET_CLASS) {  // impl/random_walker_segmentation.hpp:124-129
    terms[0] = {TCS_TERM_XYZ, 3.0f, 1.0f, TCS_NORM_GLOBAL};
    terms[1] = {TCS_TERM_NORMAL, 0.01f, 0.0f, TCS_NORM_NONE};
    terms[2] = {TCS_TERM_CURVATURE, 0.0001f, 0.0f, TCS_NORM_NONE};
    terms[3] = {TCS_TERM_RGB, 3.0f, 1.0f, TCS_NORM_GLOBAL};
  } else {                           // factory/edge_weight_computer_factory.h:20-25,41-46
    terms[0] = {TCS_TERM_XYZ, 3.0f, 1.0f, TCS_NORM_GLOBAL};
    terms[1] = {TCS_TERM_NORMAL, 0.01f, 0.01f, TCS_NORM_NONE};
    terms[2] = {TCS_TERM_CURVATURE, 0.0001f, 0.01f, TCS_NORM_NONE};
    terms[3] = {TCS_TERM_RGB, 3.0f, 1.0f, TCS_NORM_LOCAL};
  }
  step(tcs_edge_weights(g, terms, 4, 1e-5f, TCS_SMALL_COERCE, mode));
  mark(6);
  if (st == TCS_OK) step(tcs_impl_components(g));
  mark(7);
  DBuf<int32_t> nearest;
  if (st == TCS_OK) step(tcs_impl_nearest_vertices(g, seed_xyz, ns, nearest));
  if (st == TCS_OK) step(tcs_impl_paint_seeds(g, nearest, seed_colour.data(), ns));
  mark(8);
  return st;
}

// block-diagonal merge of the per-scene systems (rows and entries of scene s shifted by its offsets)
__global__ void k_merge_rowptr(int n, uint32_t nnz_off, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int last) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n || (last && i == n)) out[i] = in[i] + nnz_off;
}
__global__ void k_merge_entries(size_t nnz, int row_off, const int* __restrict__ col, const float* __restrict__ val,
                                int* __restrict__ col_out, float* __restrict__ val_out) {
  size_t a = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (a < nnz) { col_out[a] = col[a] + row_off; val_out[a] = val[a]; }
}
__global__ void k_merge_rhs(size_t n, int Kin, int Kout, const float* __restrict__ in, float* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n * Kout) return;
  const size_t row = i / Kout;
  const int k = (int)(i - row * Kout);
  out[i] = k < Kin ? in[row * Kin + k] : 0.f;
}

}  // namespace

extern "C" int tcs_segment_cloud(tcs_ctx* ctx, float resolution, int preset, int mode, const float* seed_xyz,
                                 const uint32_t* seed_label, size_t ns, const tcs_solver_opts* opts,
                                 int32_t* point_cluster, uint32_t* vertex_colour, tcs_solve_info* info,
                                 tcs_graph** out) {
  if (!ctx || !out) return TCS_ERR_INPUT;
  *out = nullptr;
  if (ns && (!seed_xyz || !seed_label)) return tcs_fail(ctx, TCS_ERR_INPUT, "seed arrays missing");
  if (preset != TCS_PRESET_CLASS && preset != TCS_PRESET_CLI) return tcs_fail(ctx, TCS_ERR_INPUT, "unknown preset");
  CK(cudaSetDevice(ctx->device));
  std::vector<uint32_t> seed_colour;
  int K = 0;
  TRY(seed_colours(ctx, seed_label, ns, seed_colour, &K));
  cudaEvent_t* ev = ctx->ev + 16;  // stage boundaries (the solver uses ev[8..13] of the context)
  tcs_graph* g = nullptr;
  int st = prepare_scene(ctx, resolution, preset, mode, seed_xyz, seed_colour, ns, ev, &g);
  if (!g) return st;
  const bool ran = g->V > 0;
  if (ran) {
    if (st == TCS_OK) st = tcs_impl_random_walker(g, K, opts, info);
  } else {
    TRY(g->colours.alloc(ctx, 0));
    if (info) { memset(info, 0, sizeof *info); info->converged = 1; info->ncolours = K; }
  }
  if (st != TCS_OK && st != TCS_ERR_NOT_CONVERGED) { tcs_graph_destroy(g); return st; }
  cudaEventRecord(ev[9], ctx->stream);
  int st2 = tcs_impl_point_clusters(g, K);
  cudaEventRecord(ev[10], ctx->stream);
  if (st2 != TCS_OK) { tcs_graph_destroy(g); return st2; }
  cudaError_t ce = cudaStreamSynchronize(ctx->stream);
  if (ce != cudaSuccess) { tcs_graph_destroy(g); return tcs_fail(ctx, TCS_ERR_CUDA, "segment: %s", cudaGetErrorString(ce)); }
  if (info && ran) {
    // stage_ms: build, normals, smoothen, normals2, signs, weights, components, seeds, [8..10 by the solver], labels
    float ms;
    for (int i = 0; i < 8; ++i)
      if (cudaEventElapsedTime(&ms, ev[i], ev[i + 1]) == cudaSuccess) info->stage_ms[i] = ms;
    if (cudaEventElapsedTime(&ms, ev[9], ev[10]) == cudaSuccess) info->stage_ms[11] = ms;
  }
  if (point_cluster && g->N) {
    ce = cudaMemcpyAsync(point_cluster, g->point_cluster.p, 4 * g->N, cudaMemcpyDeviceToHost, ctx->stream);
    if (ce != cudaSuccess) { tcs_graph_destroy(g); return tcs_fail(ctx, TCS_ERR_CUDA, "segment: %s", cudaGetErrorString(ce)); }
  }
  if (vertex_colour && g->V) {
    ce = cudaMemcpyAsync(vertex_colour, g->colours.p, 4 * g->V, cudaMemcpyDeviceToHost, ctx->stream);
    if (ce != cudaSuccess) { tcs_graph_destroy(g); return tcs_fail(ctx, TCS_ERR_CUDA, "segment: %s", cudaGetErrorString(ce)); }
  }
  cudaStreamSynchronize(ctx->stream);
  *out = g;
  return st;
}

// ---- batched class path (BASELINE configs[4]: many small scenes) -------------------------------------------------
// One segment() per scene in the reference (impl/random_walker_segmentation.hpp:138-224).  Here the graphs are built
// scene by scene (each stage a handful of launches on a 50 k-vertex graph), but the random-walker systems of all
// scenes of the batch are solved TOGETHER as one block-diagonal system: scenes are independent, so the merged matrix is
// their direct sum and label column k of the batch is the union of the scenes' k-th columns.  A 640 x 480 scene alone
// is latency-bound (tens of launches per iteration on 47 k rows); sixteen of them run through the large-system kernels
// (ring SpMM, multilevel hierarchy) at the cost of one.  The float64 convergence test is applied per scene and column.
extern "C" int tcs_segment_batch(tcs_ctx* ctx, tcs_scene* scenes, int nscenes, float resolution, int preset, int mode,
                                 const tcs_solver_opts* opts_in, tcs_solve_info* info) {
  if (!ctx || nscenes < 0 || (nscenes && !scenes)) return TCS_ERR_INPUT;
  if (preset != TCS_PRESET_CLASS && preset != TCS_PRESET_CLI) return tcs_fail(ctx, TCS_ERR_INPUT, "unknown preset");
  CK(cudaSetDevice(ctx->device));
  tcs_solver_opts opts;
  tcs_solver_opts_default(&opts);
  if (opts_in) opts = *opts_in;
  opts.store_potentials = 0;
  opts.column_begin = opts.column_count = 0;
  if (opts.check_every < 1) opts.check_every = 50;
  if (info) { memset(info, 0, sizeof *info); info->converged = 1; }
  struct Guard {
    std::vector<tcs_graph*> g;
    ~Guard() { for (tcs_graph* h : g) if (h) tcs_graph_destroy(h); }
  } graphs;
  graphs.g.assign((size_t)nscenes, nullptr);
  std::vector<int> K((size_t)nscenes, 0);
  cudaEvent_t* ev = ctx->ev + 16;
  CK(cudaEventRecord(ev[0], ctx->stream));
  // ---- per scene: upload, graph, features, weights, components, seeds, system
  tcs_segments seg;
  seg.row_begin.assign(1, 0);
  std::vector<long long> nnz_begin(1, 0);
  int Kmax = 0;
  for (int s = 0; s < nscenes; ++s) {
    tcs_scene& sc = scenes[s];
    sc.status = TCS_OK; sc.V = sc.E = 0; sc.ncolours = 0;
    if (sc.ns && (!sc.seed_xyz || !sc.seed_label)) return tcs_fail(ctx, TCS_ERR_INPUT, "scene %d: seed arrays missing", s);
    std::vector<uint32_t> seed_colour;
    TRY(seed_colours(ctx, sc.seed_label, sc.ns, seed_colour, &K[s]));
    if (sc.d_xyzc) TRY(tcs_cloud_set_device(ctx, sc.d_xyzc, sc.n, sc.d_has_color));
    else TRY(tcs_cloud_upload(ctx, sc.xyz, sc.xyz_stride, sc.rgba, sc.rgba_stride, sc.n));
    int st = prepare_scene(ctx, resolution, preset, mode, sc.seed_xyz, seed_colour, sc.ns, nullptr, &graphs.g[s]);
    if (st != TCS_OK) return st;
    tcs_graph* g = graphs.g[s];
    sc.V = g->V; sc.E = g->E; sc.ncolours = K[s];
    if (g->V == 0) TRY(g->colours.alloc(ctx, 0));
    if (g->V > 0 && K[s] > 0) TRY(tcs_impl_assemble(g, K[s]));
    seg.row_begin.push_back(seg.row_begin.back() + g->sys.n);
    nnz_begin.push_back(nnz_begin.back() + g->sys.nnz_off);
    Kmax = std::max(Kmax, K[s]);
  }
  CK(cudaEventRecord(ev[1], ctx->stream));
  // ---- merge
  tcs_system M;
  M.n = seg.row_begin.back();
  M.K = Kmax;
  M.Kp = std::max(4, (Kmax + 3) & ~3);
  M.nnz_off = nnz_begin.back();
  if ((long long)M.n != (long long)seg.row_begin.back() || nnz_begin.back() > 0x7FFFFFFFll)
    return tcs_fail(ctx, TCS_ERR_INPUT, "batch too large for one system: split it");
  tcs_pcg_result res;
  if (M.n > 0) {
    TRY(M.rowptr.alloc(ctx, (size_t)M.n + 1 + 8));
    TRY(M.col.alloc(ctx, (size_t)M.nnz_off)); TRY(M.val.alloc(ctx, (size_t)M.nnz_off + 16)); TRY(M.diag.alloc(ctx, (size_t)M.n + 8));
    TRY(M.B.alloc(ctx, (size_t)M.n * M.Kp));
    int last_nonempty = -1;
    for (int s = 0; s < nscenes; ++s) if (graphs.g[s]->sys.n > 0) last_nonempty = s;
    for (int s = 0; s < nscenes; ++s) {
      tcs_system& S = graphs.g[s]->sys;
      if (S.n == 0) continue;
      const int r0 = seg.row_begin[s];
      const long long a0 = nnz_begin[s];
      LAUNCH(ctx, k_merge_rowptr, div_up((size_t)S.n + 1, 256), 256, 0, S.n, (uint32_t)a0, S.rowptr.p, M.rowptr.p + r0, s == last_nonempty ? 1 : 0);
      if (S.nnz_off)
        LAUNCH(ctx, k_merge_entries, div_up((size_t)S.nnz_off, 256), 256, 0, (size_t)S.nnz_off, r0, S.col.p, S.val.p, M.col.p + a0, M.val.p + a0);
      CK(cudaMemcpyAsync(M.diag.p + r0, S.diag.p, (size_t)S.n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
      LAUNCH(ctx, k_merge_rhs, div_up((size_t)S.n * M.Kp, 256), 256, 0, (size_t)S.n, S.Kp, M.Kp, S.B.p, M.B.p + (size_t)r0 * M.Kp);
      // the scene's own copy of the matrix is no longer needed
      S.col.release(); S.val.release(); S.B.release();
    }
    CK(cudaEventRecord(ev[2], ctx->stream));
    TRY(tcs_impl_pcg_solve(ctx, M, Kmax, opts, &seg, res));
  } else {
    CK(cudaEventRecord(ev[2], ctx->stream));
  }
  CK(cudaEventRecord(ev[3], ctx->stream));
  // ---- per scene: labels, point clusters, results to the host
  for (int s = 0; s < nscenes; ++s) {
    tcs_graph* g = graphs.g[s];
    tcs_scene& sc = scenes[s];
    if (g->V > 0 && K[s] > 0)
      TRY(tcs_impl_argmax(g, g->sys.row_vertex.p, g->sys.n, K[s], M.Kp, g->sys.n ? M.X.p + (size_t)seg.row_begin[s] * M.Kp : nullptr,
                          opts, nullptr, nullptr));
    TRY(tcs_impl_point_clusters(g, K[s]));
    if (sc.point_cluster && g->N) CK(cudaMemcpyAsync(sc.point_cluster, g->point_cluster.p, 4 * g->N, cudaMemcpyDeviceToHost, ctx->stream));
    if (sc.vertex_colour && g->V) CK(cudaMemcpyAsync(sc.vertex_colour, g->colours.p, 4 * g->V, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaEventRecord(ev[4], ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (info) {
    info->iterations = res.iterations;
    info->converged = res.converged;
    info->rows = M.n;
    info->nnz = M.nnz_off + M.n;
    info->ncolours = Kmax;
    for (int k = 0; k < Kmax; ++k) info->rel_residual[k] = res.rel[k];
    info->spmm_launches = res.spmm_launches;
    info->spmm_kernel = res.kernel;
    info->spmm_ms = res.spmm_ms;
    float ms = 0.f;
    // stage_ms of a batch: [0] graphs + features + systems of all scenes, [8] merge, [9] the joint solve, [10] labels
    if (cudaEventElapsedTime(&ms, ev[0], ev[1]) == cudaSuccess) info->stage_ms[0] = ms;
    if (cudaEventElapsedTime(&ms, ev[1], ev[2]) == cudaSuccess) info->stage_ms[8] = ms;
    if (cudaEventElapsedTime(&ms, ev[2], ev[3]) == cudaSuccess) { info->stage_ms[9] = ms; info->solve_ms = ms; }
    if (cudaEventElapsedTime(&ms, ev[3], ev[4]) == cudaSuccess) info->stage_ms[10] = ms;
  }
  if (!res.converged) return tcs_fail(ctx, TCS_ERR_NOT_CONVERGED, "PCG did not reach %.1e in %d iterations", opts.tolerance, res.iterations);
  return TCS_OK;
}
