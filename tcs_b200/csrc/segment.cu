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

extern "C" int tcs_segment_cloud(tcs_ctx* ctx, float resolution, int preset, int mode, const float* seed_xyz,
                                 const uint32_t* seed_label, size_t ns, const tcs_solver_opts* opts,
                                 int32_t* point_cluster, uint32_t* vertex_colour, tcs_solve_info* info,
                                 tcs_graph** out) {
  if (!ctx || !out) return TCS_ERR_INPUT;
  *out = nullptr;
  if (ns && (!seed_xyz || !seed_label)) return tcs_fail(ctx, TCS_ERR_INPUT, "seed arrays missing");
  if (preset != TCS_PRESET_CLASS && preset != TCS_PRESET_CLI) return tcs_fail(ctx, TCS_ERR_INPUT, "unknown preset");
  CK(cudaSetDevice(ctx->device));
  // setSeeds (:82-101): unique labels in ascending order -> colours 1..K
  std::set<uint32_t> labels(seed_label, seed_label + ns);
  std::map<uint32_t, uint32_t> label_colour;
  for (uint32_t l : labels) { uint32_t c = (uint32_t)label_colour.size() + 1; label_colour[l] = c; }
  const int K = (int)labels.size();
  if (K > TCS_MAX_LABELS) return tcs_fail(ctx, TCS_ERR_INPUT, "more than %d distinct seed labels", TCS_MAX_LABELS);
  std::vector<uint32_t> seed_colour(ns);
  for (size_t i = 0; i < ns; ++i) seed_colour[i] = label_colour[seed_label[i]];

  cudaEvent_t* ev = ctx->ev + 16;  // stage boundaries (the solver uses ev[8..13] of the context)
  tcs_graph* g = nullptr;
  CK(cudaEventRecord(ev[0], ctx->stream));
  TRY(tcs_build_voxel_graph(ctx, resolution, 1, &g));
  int st = TCS_OK;
  bool ran = false;
  auto step = [&](int s) { if (st == TCS_OK) st = s; };
  if (g->V > 0) {
    ran = true;
    cudaEventRecord(ev[1], ctx->stream);
    step(tcs_normals(g, 0, mode));
    cudaEventRecord(ev[2], ctx->stream);
    if (preset == TCS_PRESET_CLI) {
      step(tcs_smoothen(g, 0.012f, 0.0012f));
      cudaEventRecord(ev[3], ctx->stream);
      step(tcs_normals(g, 0, mode));
    } else {
      cudaEventRecord(ev[3], ctx->stream);
    }
    cudaEventRecord(ev[4], ctx->stream);
    step(tcs_signed_curvatures(g));
    cudaEventRecord(ev[5], ctx->stream);
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
    cudaEventRecord(ev[6], ctx->stream);
    if (st == TCS_OK) step(tcs_impl_components(g));
    cudaEventRecord(ev[7], ctx->stream);
    // segment (:144-152): seeds paint their nearest vertex
    DBuf<int32_t> nearest;
    if (st == TCS_OK) step(tcs_impl_nearest_vertices(g, seed_xyz, ns, nearest));
    if (st == TCS_OK) step(tcs_impl_paint_seeds(g, nearest, seed_colour.data(), ns));
    cudaEventRecord(ev[8], ctx->stream);
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
