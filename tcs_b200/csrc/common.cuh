// tcs_b200/csrc/common.cuh - internal structures of libtcs_b200.so (context, device graph, helpers).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/tcs_b200.h"

#define TCS_NIL 0xFFFFFFFFu

struct tcs_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int sm_count = 148;
  long long launches = 0;
  std::string err;
  // input cloud: n records {x, y, z, rgba bits}
  float4* d_cloud = nullptr;
  bool own_cloud = false;
  size_t cloud_cap = 0;
  size_t n_points = 0;
  bool has_color = false;
  // scratch for host<->device scalars (pinned)
  void* h_pin = nullptr;  // 4 KB pinned
  void* d_scal = nullptr; // 4 KB device
  // staging buffers for uploads
  void* d_stage = nullptr;
  size_t stage_cap = 0;
  cudaEvent_t ev[32];
  int pdl = 1;   // programmatic dependent launch for the kernels of a PCG iteration (TCS_PDL=0 switches it off)
  long long pdl_rows = 65536;   // ... for launches of at most this many rows (TCS_PDL_ROWS)
};

int tcs_fail(tcs_ctx* ctx, int code, const char* fmt, ...);

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t _e = (call);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return tcs_fail(ctx, TCS_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
  } while (0)

#define TRY(call)            \
  do {                       \
    int _s = (call);         \
    if (_s != TCS_OK) return _s; \
  } while (0)

// stream-ordered device buffer
template <typename T>
struct DBuf {
  T* p = nullptr;
  size_t n = 0;
  tcs_ctx* ctx = nullptr;
  DBuf() {}
  DBuf(const DBuf&) = delete;
  DBuf& operator=(const DBuf&) = delete;
  ~DBuf() { release(); }
  int alloc(tcs_ctx* c, size_t count) {
    release();
    ctx = c;
    n = count;
    if (count == 0) return TCS_OK;
    cudaError_t e = cudaMallocAsync((void**)&p, count * sizeof(T), c->stream);
    if (e != cudaSuccess) {
      p = nullptr;
      return tcs_fail(c, TCS_ERR_CUDA, "cudaMallocAsync(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
    }
    return TCS_OK;
  }
  void release() {
    if (p && ctx) cudaFreeAsync(p, ctx->stream);
    p = nullptr;
    n = 0;
  }
  void swap(DBuf& o) {
    std::swap(p, o.p);
    std::swap(n, o.n);
    std::swap(ctx, o.ctx);
  }
};

// tile shape of k_spmm_tiled (spmm.cuh)
struct SpmmTile {
  int R;        // rows per tile (multiple of 256 / Q)
  int H;        // halo rows staged on either side
  int cap;      // capacity of the staged CSR segment in entries (even)
  int farcap;   // rows of the far buffer
  // shared-memory byte offsets
  int off_rp, off_diag, off_cv, off_p, off_red, total;
};

// shape of k_spmm_ring (spmm_ring.cuh): contiguous tile range per CTA, sliding ring window of P, S-stage TMA pipeline
struct SpmmRing {
  int R;          // rows per tile = consumer warps x G
  int G;          // rows per warp step = 32 / Qp
  int Qp;         // lanes per row group (= Q, or 4 for 48-byte rows padded to a 64-byte pitch)
  int fill_async; // ring rows arrive by 16-byte cp.async instead of TMA bulk copies (needed when pitch != 16 Q)
  int pitch;      // bytes between ring rows in shared memory (16 Q, or 64 when padded)
  int H;          // halo rows on either side of a tile that are guaranteed to be in the ring
  int S;          // pipeline stages (tiles in flight)
  int ring_rows;  // power of two
  int dup;        // second copy of the ring, shifted by dup_shift bytes: conflict-free gathers (see ring_pair_choice)
  int dup_shift;  // 64 for 64-byte rows (the other half of the bank line), 48 for 48-byte rows
  int cap;        // {offset, weight} entries per stage (multiple of 8)
  int farcap;     // far rows per stage
  int nctas;
  // shared-memory byte offsets (stage-relative for rp/diag/cv/far)
  int off_bar, off_red, off_ids, ids_bytes, off_stage0, stage_bytes, off_meta, off_cv, off_far, off_ring, ring_bytes, total;
};

// assembled random-walker system (kept for getPotentials / tcs_bench_spmm)
struct tcs_system {
  int n = 0;          // rows
  int K = 0;          // label columns
  int Kp = 0;         // padded column count used by the kernels (multiple of 4)
  long long nnz_off = 0;
  DBuf<uint32_t> rowptr;  // n+1
  DBuf<int> col;          // nnz_off (row indices)
  DBuf<float> val;        // nnz_off (-w)
  DBuf<float> diag;       // n (degree)
  DBuf<float> gdiag;      // n: diagonal excess d_i - sum_j w_ij (what the difference-form SpMM kernels multiply p_i by)
  DBuf<float> B;          // n x Kp row-major
  DBuf<uint32_t> row_vertex;  // n
  DBuf<int> row_of;       // V (-1 for non-rows)
  DBuf<double> X;         // n x Kp fp64
  DBuf<float> P, R, AP;   // n x Kp
  // k_spmm_tiled plan (spmm.cuh)
  bool tiled = false;
  SpmmTile tile;
  DBuf<int2> cv;                  // {col or -1 - far slot, weight bits} per off-diagonal entry
  DBuf<int> far_rows;             // per tile: de-duplicated rows outside the tile's window
  DBuf<uint32_t> tile_far_ptr;    // ntiles + 1 offsets into far_rows
  // k_spmm_ring plan (spmm_ring.cuh); shares cv / far_rows / tile_far_ptr, cv.x = shared-memory byte offset
  bool ring = false;
  bool plan_stale = false;        // the last solve ended on a compacted column set: hooks re-plan at full width
  SpmmRing rg;
  DBuf<uint4> ring_desc;          // per tile {first entry, end entry, first far row, end far row}
  DBuf<int> ring_cta_ptr;         // nctas + 1 tile offsets: the contiguous tile range of every CTA
  DBuf<long long> ring_clk;       // per-CTA cycles of a balancing launch
  DBuf<unsigned char> ring_meta;  // per tile: {entries}, slice table, row order, diagonal (ring_meta_bytes(R) each)
  DBuf<unsigned char> ring_blob;  // per tile: 16-bit shared-memory offsets then float weights, in consumption order
  long long ring_entries = 0;     // sliced entries incl. padding
};

struct tcs_graph {
  tcs_ctx* ctx = nullptr;
  size_t V = 0, E = 0, N = 0;
  bool has_color = false;
  bool from_builder = false;
  int depth = 0;
  double bbmin[3] = {0, 0, 0};
  // vertex bundle (PointXYZRGBNormal / PointNormal fields), SoA
  DBuf<float> x, y, z, nx, ny, nz, nw, curv;
  DBuf<uint32_t> rgba;
  DBuf<uint32_t> keys;  // 3 per vertex (builder graphs)
  // global edge list in insertion order
  DBuf<uint32_t> esrc, etgt;
  DBuf<float> ew;
  // out-edge lists (insertion order == ascending edge id)
  DBuf<uint32_t> rowptr, col, eid;
  DBuf<uint32_t> p2v;  // N
  DBuf<int32_t> comp;  // V
  int ncomp = -1;
  // results
  DBuf<uint32_t> colours;        // V
  DBuf<int32_t> point_cluster;   // N
  DBuf<float> best_val;          // V: largest potential among the solved columns (seeds: 1)
  DBuf<float> potentials;        // V x K column-major (store_potentials)
  int pot_K = 0;
  tcs_system sys;
};

static inline unsigned div_up(size_t a, unsigned b) { return (unsigned)((a + b - 1) / b); }

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                        \
  do {                                                                     \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);       \
    (ctx)->launches++;                                                     \
  } while (0)

// ---- packed float pairs: sm_100 executes two IEEE float operations per issue slot (FADD2 / FFMA2, PTX add/sub/fma
// .f32x2) - the same roundings as the scalar instructions, half the issue slots.  The SpMM kernels are issue-bound
// next to their shared-memory / L1 traffic (profiles/r02_spmm_ring_ncu_summary.md), so their inner loops use these.
struct F4p { unsigned long long lo, hi; };   // {x, y}, {z, w}
__device__ __forceinline__ unsigned long long f2_pack(float a, float b) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ F4p f4_pack(float4 v) { F4p r; r.lo = f2_pack(v.x, v.y); r.hi = f2_pack(v.z, v.w); return r; }
__device__ __forceinline__ float4 f4_unpack(const F4p& a) {
  float4 v;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(v.x), "=f"(v.y) : "l"(a.lo));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(v.z), "=f"(v.w) : "l"(a.hi));
  return v;
}
// acc += w * (v - p), per component: sub then fma, exactly what `acc.x += w * (v.x - p.x)` compiles to
__device__ __forceinline__ void f4_diff_fma(F4p& acc, float w, float4 v, const F4p& p) {
  const unsigned long long ww = f2_pack(w, w);
  unsigned long long d0, d1;
  asm("sub.f32x2 %0, %1, %2;" : "=l"(d0) : "l"(f2_pack(v.x, v.y)), "l"(p.lo));
  asm("sub.f32x2 %0, %1, %2;" : "=l"(d1) : "l"(f2_pack(v.z, v.w)), "l"(p.hi));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(acc.lo) : "l"(ww), "l"(d0), "l"(acc.lo));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(acc.hi) : "l"(ww), "l"(d1), "l"(acc.hi));
}

// ---- programmatic dependent launch (sm_90+).  An iteration of the solve is ~50 dependent kernels, many of them
// microseconds long: with plain stream order every boundary costs a full launch + drain gap.  Kernels launched with
// LAUNCH_PDL may be scheduled as soon as the kernel before them has let its dependents go (pdl_trigger, called right
// after its own pdl_wait) and SM resources are free; they then block in pdl_wait until that kernel has COMPLETED and
// its writes are visible - the data dependency is unchanged, the launch latency is hidden.  Every kernel that is ever
// launched with LAUNCH_PDL starts with pdl_wait() (a no-op under a plain launch).
// Measured (C4 scene / one 640x480 scene, solve time): every launch of the iteration programmatic 127.8 -> 132.7 ms /
// 7.15 -> 5.65 ms - it pays where the kernels are launch-bound and costs where they are bandwidth-bound (dependents
// scheduled early take SM resources from the kernel that still runs), hence the size limit.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// `rows`: size of the launch; only small kernels take the programmatic edge (tcs_ctx::pdl_rows)
#define LAUNCH_PDL(ctx, rows, kernel, grid, block, smem, ...)                                \
  do {                                                                                       \
    cudaLaunchConfig_t _cfg = {};                                                            \
    _cfg.gridDim = dim3(grid); _cfg.blockDim = dim3(block);                                  \
    _cfg.dynamicSmemBytes = (size_t)(smem); _cfg.stream = (ctx)->stream;                     \
    cudaLaunchAttribute _at[1];                                                              \
    _at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                          \
    _at[0].val.programmaticStreamSerializationAllowed = ((ctx)->pdl && (long long)(rows) <= (ctx)->pdl_rows) ? 1 : 0; \
    _cfg.attrs = _at; _cfg.numAttrs = 1;                                                     \
    cudaLaunchKernelEx(&_cfg, kernel, __VA_ARGS__);                                          \
    (ctx)->launches++;                                                                       \
  } while (0)

// ---- stage entry points shared between translation units ---------------------------------------------------
int tcs_impl_scan_u32(tcs_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, size_t n);  // exclusive sum, d_out[n] = total (n+1 outputs)
int tcs_impl_read_u32(tcs_ctx* ctx, const uint32_t* d_ptr, uint32_t* out);             // synchronising scalar read
int tcs_impl_build_out_edges(tcs_graph* g);                                              // CSR from an edge list (user graphs / REMOVE policy)
int tcs_impl_alloc_vertices(tcs_graph* g, size_t V);
// stable LSD radix sorts (CUB) on bits [0, end_bit)
int tcs_impl_sort_pairs_u32(tcs_ctx* ctx, uint32_t* key_in, uint32_t* key_out, uint32_t* val_in, uint32_t* val_out,
                            size_t n, int end_bit);
int tcs_impl_sort_pairs_u64(tcs_ctx* ctx, uint64_t* key_in, uint64_t* key_out, uint32_t* val_in, uint32_t* val_out,
                            size_t n, int end_bit);

// block-diagonal system of several scenes: rows [row_begin[s], row_begin[s + 1]) belong to scene s
struct tcs_segments {
  std::vector<int> row_begin;
  int count() const { return row_begin.empty() ? 0 : (int)row_begin.size() - 1; }
};
struct tcs_pcg_result {
  int iterations = 0, converged = 1, kernel = 0, spmm_launches = 0;
  float spmm_ms = 0.f;
  double rel[TCS_MAX_LABELS] = {0};
};

// solve.cu / pcg.cu
int tcs_impl_components(tcs_graph* g);
int tcs_impl_nearest_vertices(tcs_graph* g, const float* seed_xyz, size_t ns, DBuf<int32_t>& d_nearest);
int tcs_impl_paint_seeds(tcs_graph* g, const DBuf<int32_t>& nearest, const uint32_t* h_colour, size_t ns);
int tcs_impl_assemble(tcs_graph* g, int K);
// PCG + argmax on g->colours (device, seeds in / labels out); fills info (may be NULL)
int tcs_impl_random_walker(tcs_graph* g, int K, const tcs_solver_opts* opts, tcs_solve_info* info);
int tcs_impl_point_clusters(tcs_graph* g, int K);
int tcs_impl_pcg_solve(tcs_ctx* ctx, tcs_system& S, int K, const tcs_solver_opts& opts, const tcs_segments* seg,
                       tcs_pcg_result& res);
int tcs_impl_argmax(tcs_graph* g, const uint32_t* row_vertex, int nrows, int K, int Kp, const double* X,
                    const tcs_solver_opts& opts, float* d_pot, tcs_solve_info* info);
