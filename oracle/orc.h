/* oracle/orc.h - CPU ORACLE for the tcs random-walker segmentation hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product (tcs_b200/, include/) never does.
 *
 * It is a single-threaded, order-exact float32 restatement of the reference's algorithms (citations are
 * file:line into taketwo/tcs @16cc7ab8) plus the third-party semantics (PCL 1.7.2 / Boost.Graph / Eigen 3.2)
 * those call, restated from their published behaviour because none of the three
 * libraries is available in this image.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or stored outputs for this path
 *  and cannot be compiled here (needs PCL, Boost.Graph, Eigen).  The oracle is
 * pinned only by first-principles known-answer tests (tests/test_oracle_kat.py) and by a scipy float64
 * cross-check of the linear solve.
 *
 * Build: g++ -O2 -ffp-contract=off (no FMA contraction: the reference is x86-64 SSE2 code).
 */
#ifndef TCS_ORACLE_ORC_H
#define TCS_ORACLE_ORC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_graph orc_graph;

/* term kinds: include/graph/edge_weight_computer_terms.h:55-140 */
enum { ORC_TERM_XYZ = 0, ORC_TERM_NORMAL = 1, ORC_TERM_CURVATURE = 2, ORC_TERM_RGB = 3 };
/* include/graph/edge_weight_computer.h:210-218 */
enum { ORC_NORM_NONE = 0, ORC_NORM_GLOBAL = 1, ORC_NORM_LOCAL = 2 };
/* include/graph/edge_weight_computer.h:222-235 */
enum { ORC_SMALL_IGNORE = 0, ORC_SMALL_COERCE = 1, ORC_SMALL_REMOVE = 2 };
/* arithmetic mode: 0 = reference order (sequential float32 sums); 1 = "fast" definition shared with the
 * GPU fast mode (order-free sums accumulated in float64, then rounded to float32) */
enum { ORC_MODE_PARITY = 0, ORC_MODE_FAST = 1 };
/* linear solvers for the random walker */
enum { ORC_SOLVE_LDLT_F32 = 0, ORC_SOLVE_LDLT_F64 = 1, ORC_SOLVE_PCG_F64 = 2 };

typedef struct orc_term {
  int kind;            /* ORC_TERM_* */
  float influence;     /* sigma */
  float convex_mult;   /* convex influence multiplier */
  int normalization;   /* ORC_NORM_* */
} orc_term;

/* status codes */
enum { ORC_OK = 0, ORC_ERR_INPUT = -1, ORC_ERR_DEPTH = -2, ORC_ERR_SOLVE = -3 };

/* ---- graph construction ------------------------------------------------------------------------- */
/* graph/impl/voxel_grid_graph_builder.hpp:69-143.  xyz: n*3 floats, rgba: n uint32 or NULL.
 * leaf_order_desc: 1 = PCL<=1.8 leaf iterator order (descending child index), 0 = ascending. */
orc_graph* orc_build_voxel_graph(const float* xyz, const uint32_t* rgba, size_t n, float resolution,
                                 int leaf_order_desc, int* status);
/* explicit graph (setInputGraph path): edges in insertion order; normals4/curvature/rgba/weights may be NULL */
orc_graph* orc_graph_from_edges(size_t V, const float* xyz, const uint32_t* rgba, const float* normals4,
                                const float* curvature, size_t E, const uint32_t* src, const uint32_t* tgt,
                                const float* weight);
/* graph/impl/nearest_neighbors_graph_builder.hpp:49-119: vertices = finite points in index order; edges to the
 * num_neighbors nearest other points (use_radius = 0) or to the nearest num_neighbors among those with
 * d^2 < (float)(radius^2) (use_radius = 1); neighbours ordered by (d^2, index), brute force. */
orc_graph* orc_build_nn_graph(const float* xyz, const uint32_t* rgba, size_t n, int num_neighbors, double radius,
                              int use_radius, int* status);
void orc_graph_free(orc_graph* g);

size_t orc_num_vertices(const orc_graph* g);
size_t orc_num_edges(const orc_graph* g);
size_t orc_num_points(const orc_graph* g);
int orc_depth(const orc_graph* g);
void orc_get_bbox_min(const orc_graph* g, double* min3);
void orc_get_vertices(const orc_graph* g, float* xyz, uint32_t* rgba, float* normals4, float* curvature);
void orc_get_keys(const orc_graph* g, uint32_t* keys3);
void orc_get_edges(const orc_graph* g, uint32_t* src, uint32_t* tgt, float* weight);
void orc_get_csr(const orc_graph* g, uint32_t* rowptr, uint32_t* col, uint32_t* eid);
void orc_get_point_to_vertex(const orc_graph* g, uint32_t* map); /* nil = 0xFFFFFFFF */
void orc_set_xyz(orc_graph* g, const float* xyz);
void orc_set_normals(orc_graph* g, const float* normals4, const float* curvature);
void orc_set_weights(orc_graph* g, const float* w);

/* ---- per-vertex / per-edge features ---------------------------------------------------------------- */
void orc_normals(orc_graph* g, int one_ring, int mode);               /* graph/impl/common.hpp:53-103 */
void orc_smoothen(orc_graph* g, float spatial_sigma, float influence_sigma); /* common.hpp:241-279 */
void orc_signed_curvatures(orc_graph* g);                             /* common.hpp:105-132 */
int orc_edge_weights(orc_graph* g, const orc_term* terms, int nterms, float threshold, int policy,
                     int mode);
/* the same with EdgeWeightComputer::setTermBalancingFunction: 0 Gaussian (default), 1 inverse, 2 squared Gaussian, 3 linear */
int orc_edge_weights_ex(orc_graph* g, const orc_term* terms, int nterms, float threshold, int policy, int mode, int balancing);                                       /* edge_weight_computer.hpp:110-188 */
int orc_components(const orc_graph* g, int32_t* component);           /* common.hpp:134-149 */

/* ---- random walker --------------------------------------------------------------------------------- */
/* nearest vertex per seed (impl/random_walker_segmentation.hpp:149-152) */
void orc_nearest_vertices(const orc_graph* g, const float* seed_xyz, size_t ns, int32_t* out);
/* impl/random_walker.hpp:95-243 applied per connected component (impl/random_walker_segmentation.hpp:165-196).
 * colours: V uint32, in = seeds (0 = unlabelled), out = labels.  potentials: V*ncolours floats, column-major
 * (column = colour-1), or NULL.  residuals: ncolours doubles (max over components of ||L x - b|| / ||b||
 * evaluated in float64) or NULL.  Returns ORC_OK or ORC_ERR_SOLVE. */
int orc_random_walker(orc_graph* g, uint32_t* colours, int ncolours, int solver, float* potentials,
                      double* residuals);
/* whole class path: setInputCloud + setSeeds + segment (impl/random_walker_segmentation.hpp:82-224).
 * preset: 0 = class preset (:120-131), 1 = CLI preset (src/random_walker_segmentation_app.cpp:129-134 with
 * factory/edge_weight_computer_factory.h:41-46 defaults).  point_cluster: n int32 = index of the cluster each
 * input point lands in (0..K, K = "unlabelled" cluster, clusters ordered by ascending seed label). */
orc_graph* orc_segment_cloud(const float* xyz, const uint32_t* rgba, size_t n, float resolution, int preset,
                             int mode, const float* seed_xyz, const uint32_t* seed_label, size_t ns,
                             int solver, int32_t* point_cluster, uint32_t* vertex_colour_out /* V, may be NULL */,
                             int* ncolours_out, double* stage_ms /* 10 doubles or NULL */, int* status);

/* evaluation: src/segmentation_evaluation.cpp:139-165 + src/hungarian/hungarian.cpp:8-54.
 * returns number of mislabelled points */
long orc_segmentation_error(const uint32_t* gt_label, const uint32_t* seg_label, size_t n);

#ifdef __cplusplus
}
#endif
#endif
