/* include/tcs_b200.h - C ABI of libtcs_b200.so: the B200 (sm_100a) implementation of the random-walker
 * point-cloud segmentation hot path of taketwo/tcs.
 *
 * The reference has no FFI layer: its boundary is the C++ template API compiled into
 * librandom_walker_segmentation (src/random_walker_segmentation.cpp:44).  This header is the thin C layer
 * underneath our C++ classes of the same names (include/tcs/ *.h); each entry point cites the reference
 * function it replaces (file:line into taketwo/tcs @16cc7ab8).
 *
 * Conventions
 *  - every function returns int: TCS_OK (0) or a negative tcs_status; nothing throws across the ABI;
 *    tcs_last_error(ctx) gives the message of the last failure on that context.
 *  - the caller owns host buffers; the context owns all device memory.  Pointers named d_* are DEVICE
 *    pointers, all others are HOST pointers.  NULL output pointers mean "not wanted".
 *  - one context = one device + one CUDA stream; it is single-threaded.  Different contexts (one or more
 *    per GPU) may be used from different host threads concurrently.
 *  - all work is enqueued on the context's stream; functions that return host data synchronise it.
 *  - there is NO CPU fallback: without a CUDA device tcs_ctx_create fails with TCS_ERR_CUDA.
 */
#ifndef TCS_B200_H
#define TCS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tcs_ctx tcs_ctx;
typedef struct tcs_graph tcs_graph;

typedef enum tcs_status {
  TCS_OK = 0,
  TCS_ERR_CUDA = -1,       /* a CUDA runtime call failed / no device */
  TCS_ERR_INPUT = -2,      /* invalid argument (reference: ComputeFailedException, impl/random_walker_segmentation.hpp:111) */
  TCS_ERR_DEPTH = -3,      /* octree deeper than 21 levels (63-bit Morton key) */
  TCS_ERR_NOT_CONVERGED = -4, /* PCG hit max_iterations (reference: "Random walker segmentation failed", :194) */
  TCS_ERR_STATE = -5       /* call order violated (e.g. no cloud uploaded) */
} tcs_status;

/* edge-weight terms: include/graph/edge_weight_computer_terms.h:55-140 */
enum { TCS_TERM_XYZ = 0, TCS_TERM_NORMAL = 1, TCS_TERM_CURVATURE = 2, TCS_TERM_RGB = 3 };
/* EdgeWeightComputer::NormalizationType, include/graph/edge_weight_computer.h:210-218 */
enum { TCS_NORM_NONE = 0, TCS_NORM_GLOBAL = 1, TCS_NORM_LOCAL = 2 };
/* EdgeWeightComputer::SmallWeightPolicy, include/graph/edge_weight_computer.h:222-235 */
enum { TCS_SMALL_IGNORE = 0, TCS_SMALL_COERCE = 1, TCS_SMALL_REMOVE = 2 };
/* arithmetic mode of the order-sensitive float32 reductions (normal moments, global normaliser):
 *   PARITY: the reference's sequential float32 order (one thread walks the list) - bit-exact with the reference order
 *   FAST  : float32 products, sums accumulated in float64 in parallel and rounded once to float32 */
enum { TCS_MODE_PARITY = 0, TCS_MODE_FAST = 1 };
/* weight presets */
enum { TCS_PRESET_CLASS = 0, /* RandomWalkerSegmentation::preComputeGraph, impl/random_walker_segmentation.hpp:120-131 */
       TCS_PRESET_CLI = 1    /* app pipeline, src/random_walker_segmentation_app.cpp:129-134 + factory defaults */ };

/* EdgeWeightComputer::addTerm<TermT>(influence, convex_influence_multiplier, normalization),
 * include/graph/edge_weight_computer.h:259-283 */
typedef struct tcs_term {
  int kind;           /* TCS_TERM_* */
  float influence;    /* sigma */
  float convex_mult;  /* convex influence multiplier */
  int normalization;  /* TCS_NORM_* */
} tcs_term;

typedef struct tcs_solver_opts {
  double tolerance;      /* relative residual ||b - Lx|| / ||b|| per label column; default 1e-6 */
  int max_iterations;    /* default 200000 */
  int check_every;       /* iterations between float64 true-residual replacements / convergence checks; default 50
                            (doubled while the slowest column is still more than 30x above the tolerance) */
  int store_potentials;  /* keep the dense V x K potentials on the device (getPotentials) */
  int time_spmm;         /* bracket sampled SpMM launches with CUDA events (fills tcs_solve_info.spmm_ms) */
  /* label-column shard: solve only the colour columns [column_begin, column_begin +
   * column_count); column_count = 0 solves all.  All seeds stay boundary conditions; the other columns' potentials
   * are reported as 0, vertices are labelled by the best column of THIS shard, and tcs_graph_download_best gives the
   * (potential, colour) pairs a caller merges across shards (maximum potential, lowest colour on ties). */
  int column_begin;
  int column_count;
} tcs_solver_opts;

#define TCS_MAX_LABELS 64

typedef struct tcs_solve_info {
  int iterations;               /* PCG iterations executed */
  int converged;                /* 1 if every column reached the tolerance */
  int rows;                     /* n: unknowns (unlabelled vertices with non-zero degree) */
  long long nnz;                /* off-diagonal entries of L (both triangles) + n diagonal entries */
  int ncolours;                 /* K */
  int ncomponents;
  long long near_ties;          /* vertices whose two largest potentials differ by < 1e-6 */
  double rel_residual[TCS_MAX_LABELS]; /* float64 true residual per colour column */
  float spmm_ms;                /* SAMPLED: mean device time of the first <= 64 (batched solve: 32) SpMM launches of the
                                   solve (one CUDA event pair each, ctx stream) x spmm_launches; 0 unless time_spmm */
  int spmm_launches;
  float solve_ms;               /* device time of the whole PCG loop */
  float stage_ms[12];           /* build, normals, smoothen, normals2, signs, weights, components, seeds,
                                   assemble, solve, argmax, point labels (device time, CUDA events) */
  int spmm_kernel;              /* which SpMM ran: 1 = k_spmm (generic), 2 = k_spmm_tiled, 3 = k_spmm_ring,
                                   4 = k_pcg_chunk (small systems: whole iterations in one persistent launch) */
} tcs_solve_info;

/* ---- context ------------------------------------------------------------------------------------------- */
int tcs_ctx_create(int device, tcs_ctx** out);
/* use an existing CUDA stream (cudaStream_t passed as void*), e.g. the caller's current stream */
int tcs_ctx_create_on_stream(int device, void* cuda_stream, tcs_ctx** out);
void tcs_ctx_destroy(tcs_ctx* ctx);
const char* tcs_last_error(const tcs_ctx* ctx);
/* Page-locked ("pinned") host memory, usable with every device of the process; NULL on failure.  No counterpart in the
 * reference (a CPU library).  Host buffers passed to tcs_cloud_upload / tcs_segment_batch are copied at full PCIe speed,
 * asynchronously, only when they are pinned; pageable buffers work but go through the driver's staging copy. */
void* tcs_host_alloc(size_t bytes);
void tcs_host_free(void* p);
int tcs_ctx_synchronize(tcs_ctx* ctx);
void* tcs_ctx_stream(tcs_ctx* ctx);
/* number of kernels launched by this library on this context since creation (bench.py's gpu_launches) */
long long tcs_ctx_launch_count(const tcs_ctx* ctx);
const char* tcs_version(void);

/* ---- input cloud: RandomWalkerSegmentation::setInputCloud (random_walker_segmentation.h:170) ------------ */
/* xyz: 3 floats per point, consecutive points xyz_stride bytes apart (12 for packed, 32 for pcl::PointXYZRGB);
 * rgba: one uint32 per point rgba_stride bytes apart, or NULL for uncoloured clouds (pcl::PointXYZ). */
int tcs_cloud_upload(tcs_ctx* ctx, const float* xyz, size_t xyz_stride, const uint32_t* rgba, size_t rgba_stride,
                     size_t n);
/* cloud already resident on the device as n records {x, y, z, rgba-bits} of 16 bytes; borrowed, not copied */
int tcs_cloud_set_device(tcs_ctx* ctx, const void* d_xyzc, size_t n, int has_color);

/* ---- graph construction -------------------------------------------------------------------------------- */
/* VoxelGridGraphBuilder::compute, graph/impl/voxel_grid_graph_builder.hpp:69-143.
 * leaf_order_desc: 1 = PCL <= 1.8 leaf-iterator order (the reference's), 0 = ascending Morton order. */
int tcs_build_voxel_graph(tcs_ctx* ctx, float resolution, int leaf_order_desc, tcs_graph** out);
/* NearestNeighborsGraphBuilder::compute, graph/impl/nearest_neighbors_graph_builder.hpp:49-119:
 * vertices = the finite points in index order; every vertex is joined to its num_neighbors (<= 64) nearest other
 * points (use_radius = 0, useNearestKSearch) or to the nearest num_neighbors among the points with
 * d^2 < (float)(radius^2) (use_radius = 1, useRadiusSearch); neighbours ordered by (d^2, index); edges in the
 * reference's insertion order, each undirected edge once. */
int tcs_build_nn_graph(tcs_ctx* ctx, int num_neighbors, double radius, int use_radius, tcs_graph** out);
/* RandomWalkerSegmentation::setInputGraph (random_walker_segmentation.h:175): user graph, edges in insertion
 * order; rgba / normals4 / curvature / weight may be NULL. */
int tcs_graph_create(tcs_ctx* ctx, size_t V, const float* xyz, const uint32_t* rgba, const float* normals4,
                     const float* curvature, size_t E, const uint32_t* src, const uint32_t* tgt,
                     const float* weight, tcs_graph** out);
void tcs_graph_destroy(tcs_graph* g);
int tcs_graph_sizes(const tcs_graph* g, size_t* V, size_t* E, size_t* N, int* depth);
int tcs_graph_bbox_min(const tcs_graph* g, double* min3);
int tcs_graph_download_vertices(tcs_graph* g, float* xyz, uint32_t* rgba, float* normals4, float* curvature);
int tcs_graph_download_keys(tcs_graph* g, uint32_t* keys3);
int tcs_graph_download_edges(tcs_graph* g, uint32_t* src, uint32_t* tgt, float* weight);
int tcs_graph_download_csr(tcs_graph* g, uint32_t* rowptr, uint32_t* col, uint32_t* eid);
/* GraphBuilder::getPointToVertexMap (graph/graph_builder.h:96-102); nil = 0xFFFFFFFF */
int tcs_graph_download_point_to_vertex(tcs_graph* g, uint32_t* map);
int tcs_graph_upload_xyz(tcs_graph* g, const float* xyz);
int tcs_graph_upload_normals(tcs_graph* g, const float* normals4, const float* curvature);
int tcs_graph_upload_weights(tcs_graph* g, const float* weight);

/* ---- per-vertex / per-edge features ---------------------------------------------------------------------- */
/* computeNormalsAndCurvatures(g, neighborhood_1ring), graph/impl/common.hpp:53-103 */
int tcs_normals(tcs_graph* g, int one_ring, int mode);
/* smoothen(g, spatial_sigma, influence_sigma), graph/impl/common.hpp:241-279 */
int tcs_smoothen(tcs_graph* g, float spatial_sigma, float influence_sigma);
/* computeSignedCurvatures(g), graph/impl/common.hpp:105-132 */
int tcs_signed_curvatures(tcs_graph* g);
/* EdgeWeightComputer::compute, graph/impl/edge_weight_computer.hpp:110-188 */
int tcs_edge_weights(tcs_graph* g, const tcs_term* terms, int nterms, float threshold, int policy, int mode);
/* EdgeWeightComputer::setTermBalancingFunction (graph/edge_weight_computer.h:207,325) + compute: the function that turns
 * a (normalised) term distance d and the term's influence s into a factor of the edge weight.  The reference accepts
 * any boost::function<float (float, float)>; the device offers this closed set (all are 1 for s <= 0, like the
 * reference's default):  GAUSSIAN exp(-d/s) (the default, edge_weight_computer.h:333-337) | INVERSE 1/(1 + d/s) |
 * GAUSSIAN_SQUARED exp(-(d/s)^2) | LINEAR max(0, 1 - d/s). */
enum { TCS_BALANCE_GAUSSIAN = 0, TCS_BALANCE_INVERSE = 1, TCS_BALANCE_GAUSSIAN_SQUARED = 2, TCS_BALANCE_LINEAR = 3 };
int tcs_edge_weights_ex(tcs_graph* g, const tcs_term* terms, int nterms, float threshold, int policy, int mode,
                        int balancing);
/* createSubgraphsFromConnectedComponents, graph/impl/common.hpp:134-149: component id per vertex, ids in
 * order of the smallest member vertex.  component (host, V ints) may be NULL. */
int tcs_components(tcs_graph* g, int* ncomponents, int32_t* component);

/* ---- random walker ------------------------------------------------------------------------------------------ */
void tcs_solver_opts_default(tcs_solver_opts* o);
/* seed -> nearest vertex (pcl::getApproximateIndices, impl/random_walker_segmentation.hpp:149-152) */
int tcs_nearest_vertices(tcs_graph* g, const float* seed_xyz, size_t ns, int32_t* out);
/* Allocation-free variant of tcs_nearest_vertices.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results as tcs_nearest_vertices (bit-identical: same distance arithmetic,
 * ties -> lowest vertex id), but every device buffer it needs (seed staging, per-(seed, chunk) partial minima, the
 * result) lives in ONE fixed file-scope device array of 276 KB that every GPU loading the library holds until exit; it
 * calls no cudaMalloc / cudaFree / allocating library routine.  Limits and rules:
 *  - ns <= TCS_NOALLOC_MAX_SEEDS, otherwise TCS_ERR_INPUT with a message (use tcs_nearest_vertices for more);
 *  - all calls on one GPU share the array, which keeps the last call's contents (every slot a call reads is written
 *    by that call first); calls that overlap on the GPU - different contexts / streams - RACE: serialise them;
 *  - ns = 0 only resolves the array's device address (once per device): call it once before a graph capture. */
#define TCS_NOALLOC_MAX_SEEDS 1024
int tcs_nearest_vertices_noalloc(tcs_graph* g, const float* seed_xyz, size_t ns, int32_t* out);
/* randomWalker(graph, weights, colours[, potentials, colour->column]) (random_walker.h:95-137,
 * impl/random_walker.hpp:95-243) on all connected components at once.
 * colours: V uint32; in: seed colours 1..ncolours (0 = unlabelled); out: labels.
 * potentials: V x ncolours floats, column-major (column = colour - 1), or NULL. */
int tcs_random_walker(tcs_graph* g, uint32_t* colours, int ncolours, const tcs_solver_opts* opts,
                      float* potentials, tcs_solve_info* info);
/* Whole class path: setInputCloud (already uploaded) + setSeeds + segment
 * (impl/random_walker_segmentation.hpp:82-224), or the CLI pipeline for preset = TCS_PRESET_CLI.
 * point_cluster: n int32 = cluster index of every input point (0..K-1 by ascending seed label, K = unlabelled);
 * vertex_colour: V uint32.  Either may be NULL: results then stay on the device (tcs_graph_d_*).
 * The graph is returned in *out (caller destroys it). */
int tcs_segment_cloud(tcs_ctx* ctx, float resolution, int preset, int mode, const float* seed_xyz,
                      const uint32_t* seed_label, size_t ns, const tcs_solver_opts* opts, int32_t* point_cluster,
                      uint32_t* vertex_colour, tcs_solve_info* info, tcs_graph** out);
/* Batched class path (BASELINE configs[4]): `nscenes` independent clouds, one segment() each in
 * the reference (impl/random_walker_segmentation.hpp:138-224).  Graphs, features and weights are computed scene by
 * scene; the random-walker systems of all scenes are solved together as ONE block-diagonal system (scenes are
 * independent: label column k of the batch is the union of the scenes' k-th columns), with the float64 convergence
 * test applied per scene and column.  Results per scene equal those of tcs_segment_cloud on that scene up to the
 * solver tolerance.  info (may be NULL) describes the joint solve. */
typedef struct tcs_scene {
  const float* xyz; size_t xyz_stride;         /* in: as tcs_cloud_upload */
  const uint32_t* rgba; size_t rgba_stride;    /* in: NULL for uncoloured clouds */
  size_t n;
  const void* d_xyzc; int d_has_color;         /* in: alternative to xyz / rgba - the cloud already on the device as n
                                                  records {x, y, z, rgba bits} of 16 bytes (borrowed); used when non-NULL */
  const float* seed_xyz; const uint32_t* seed_label; size_t ns;   /* in: setSeeds */
  int32_t* point_cluster;                      /* out: n int32 (may be NULL) */
  uint32_t* vertex_colour;                     /* out: capacity n >= V uint32 (may be NULL) */
  size_t V, E;                                 /* out */
  int ncolours;                                /* out: K of this scene */
  int status;                                  /* out */
} tcs_scene;
int tcs_segment_batch(tcs_ctx* ctx, tcs_scene* scenes, int nscenes, float resolution, int preset, int mode,
                      const tcs_solver_opts* opts, tcs_solve_info* info);
/* potentials kept by store_potentials (getPotentials, random_walker_segmentation.h:251): V x K column-major */
int tcs_graph_download_potentials(tcs_graph* g, float* potentials);
int tcs_graph_download_colours(tcs_graph* g, uint32_t* colours);
/* per vertex: the largest potential among the solved columns and its colour (seeds: 1 and their colour; unlabelled
 * vertices: 0 and 0) - what a label-column shard contributes to the merged labelling */
int tcs_graph_download_best(tcs_graph* g, float* best_potential, uint32_t* best_colour);
/* device-resident results (valid until the graph is destroyed) */
const uint32_t* tcs_graph_d_colours(const tcs_graph* g);
const int32_t* tcs_graph_d_point_cluster(const tcs_graph* g);
/* V floats: the device copy of tcs_graph_download_best's best_potential (its colours are tcs_graph_d_colours) - what a
 * label-column shard hands to the all-gather without staging through the host.  Written on the context stream:
 * tcs_ctx_synchronize before another stream or library reads it. */
const float* tcs_graph_d_best_potential(const tcs_graph* g);

/* ---- measurement hooks (bench.py): the CG SpMM kernel alone on the assembled system of the last solve ----- */
/* launches the SpMM `reps` times on the context stream between two CUDA events; returns average ms per launch */
int tcs_bench_spmm(tcs_graph* g, int reps, float* avg_ms, long long* algorithmic_bytes);
/* test hook: applies the generic SpMM kernel (k_spmm) and the kernel planned for this system (tiled / ring) to the
 * same pseudo-random dense block and returns max |difference|, max |reference| over the n x K outputs and over the
 * per-column p.Ap dots (relative), plus the planned kernel's id (tcs_solve_info.spmm_kernel) */
int tcs_check_spmm(tcs_graph* g, unsigned seed, double* max_abs_diff, double* max_abs_ref, double* max_rel_dot_diff,
                   int* kernel);

#ifdef __cplusplus
}
#endif
#endif /* TCS_B200_H */
