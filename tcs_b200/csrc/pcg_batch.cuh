// tcs_b200/csrc/pcg_batch.cuh - batched PCG: many independent scenes in ONE set of kernel launches (included by pcg.cu).
//
// tcs_segment_batch merges the random-walker systems of a batch of scenes into one block-diagonal matrix.  Running
// plain PCG on the direct sum is correct but slow: one step length per column is shared by all scenes, so the Krylov
// polynomial has to fit the UNION of the scenes' spectra (measured: 64 iterations for one 640 x 480 scene, 520 for
// sixteen).  Here every (scene, column) pair keeps its OWN alpha / beta / convergence state - sixteen independent CG
// recurrences advanced by the same launches:
//   * matrix, vectors and the multilevel hierarchy are those of the merged system (aggregates never span scenes: there
//     is no edge between two scenes), so SpMM, cycle and vector kernels run at large-system efficiency;
//   * every kernel that forms or uses a scalar runs on a (blocks, scene) grid: blockIdx.y is the scene, its rows are
//     contiguous, the dot products land in [scene][column] slots and alpha / beta are read once per block;
//   * r.z is evaluated as sum_i r_i (r_i / d_i + e[agg_i]) on the fine level (= r.D^-1 r + (Z^T r).e without needing
//     the scenes' row ranges on the coarse level);
//   * staged mixed-precision refinement as in the single-scene solve, per (scene, column): a stage ends when every
//     pair has dropped by 1e-3 (or to the tolerance); the float64 true residual then replaces r and every recurrence
//     restarts from it (z, p = z).
#pragma once

namespace batch {

// out[s][c] += sum over the rows of scene s of a[row][c] * b[row][c]
__global__ void __launch_bounds__(256) k_dot(int Q, int Kp, const int* __restrict__ row_begin, const float4* __restrict__ A,
                                             const float4* __restrict__ B, double* __restrict__ out) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  extern __shared__ double sm[];
  const int s = blockIdx.y, r0 = row_begin[s], r1 = row_begin[s + 1];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double acc[4] = {0, 0, 0, 0};
  if (active)
    for (int row = r0 + blockIdx.x * rows_per_block + local; row < r1; row += gridDim.x * rows_per_block) {
      const float4 a = A[(size_t)row * Q + q], b = B[(size_t)row * Q + q];
      acc[0] += (double)a.x * b.x; acc[1] += (double)a.y * b.y; acc[2] += (double)a.z * b.z; acc[3] += (double)a.w * b.w;
    }
  reduce_columns(acc, q, Q, active, sm, out + (size_t)s * Kp, true);
}

// rz[s][c] += sum r (r / d + e[agg]) ;  rr[s][c] += sum r^2      (e / agg NULL: Jacobi alone)
__global__ void __launch_bounds__(256) k_dot_rz(int Q, int Kp, const int* __restrict__ row_begin, const float4* __restrict__ R,
                                                const float* __restrict__ invd, const uint32_t* __restrict__ agg,
                                                const float4* __restrict__ e, double* __restrict__ rz, double* __restrict__ rr) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  extern __shared__ double sm[];
  const int s = blockIdx.y, r0 = row_begin[s], r1 = row_begin[s + 1];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  const bool active = local < rows_per_block;
  double a1[4] = {0, 0, 0, 0}, a2[4] = {0, 0, 0, 0};
  if (active)
    for (int row = r0 + blockIdx.x * rows_per_block + local; row < r1; row += gridDim.x * rows_per_block) {
      const float4 r = R[(size_t)row * Q + q];
      const float d = invd[row];
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (agg) c = e[(size_t)agg[row] * Q + q];
      a1[0] += (double)r.x * ((double)r.x * d + c.x); a1[1] += (double)r.y * ((double)r.y * d + c.y);
      a1[2] += (double)r.z * ((double)r.z * d + c.z); a1[3] += (double)r.w * ((double)r.w * d + c.w);
      a2[0] += (double)r.x * r.x; a2[1] += (double)r.y * r.y; a2[2] += (double)r.z * r.z; a2[3] += (double)r.w * r.w;
    }
  reduce_columns(a1, q, Q, active, sm, rz + (size_t)s * Kp, true);
  __syncthreads();
  reduce_columns(a2, q, Q, active, sm, rr + (size_t)s * Kp, true);
}

// alpha[s][c] = rz / pAp for the live pairs (0 freezes a pair); clears the accumulators of the next dots
__global__ void k_alpha(int count, const double* __restrict__ rz, const double* __restrict__ pAp, const double* __restrict__ rz0,
                        const int* __restrict__ frozen, double done_factor, float* __restrict__ al, double* __restrict__ clear_a,
                        double* __restrict__ clear_b) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const double den = pAp[i], num = rz[i];
  const bool live = den > 0.0 && num > done_factor * rz0[i] && !frozen[i];
  const double ratio = live ? num / den : 0.0;
  al[i] = ratio < 1e30 ? (float)ratio : 0.0f;
  clear_a[i] = 0.0;
  clear_b[i] = 0.0;
}

// beta[s][c] = rz_new / rz ; clears the p.Ap accumulator of the next iteration
__global__ void k_beta(int count, const double* __restrict__ rz, const double* __restrict__ rz_new, const double* __restrict__ rz0,
                       const int* __restrict__ frozen, double done_factor, float* __restrict__ be, double* __restrict__ clear_a) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const double den = rz[i], num = rz_new[i];
  const bool live = den > done_factor * rz0[i] && den > 0.0 && !frozen[i];
  const double ratio = live ? num / den : 0.0;
  be[i] = ratio < 1e30 ? (float)ratio : 0.0f;
  clear_a[i] = 0.0;
}

// r -= alpha[s] Ap
__global__ void __launch_bounds__(256) k_update(int Q, int Kp, const int* __restrict__ row_begin, const float* __restrict__ al,
                                                const float4* __restrict__ AP, float4* __restrict__ R) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const int s = blockIdx.y, r0 = row_begin[s], r1 = row_begin[s + 1];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  if (local >= rows_per_block) return;
  const float4 a = *(const float4*)(al + (size_t)s * Kp + 4 * q);
  for (int row = r0 + blockIdx.x * rows_per_block + local; row < r1; row += gridDim.x * rows_per_block) {
    const size_t i = (size_t)row * Q + q;
    const float4 ap = AP[i];
    float4 r = R[i];
    // a frozen pair (alpha = 0) is left untouched: its Ap may hold inf or nan after float32 underflow
    if (a.x != 0.f) r.x -= a.x * ap.x;
    if (a.y != 0.f) r.y -= a.y * ap.y;
    if (a.z != 0.f) r.z -= a.z * ap.z;
    if (a.w != 0.f) r.w -= a.w * ap.w;
    R[i] = r;
  }
}

// x += alpha[s] p (float64) ; p = r / d + e[agg] + beta[s] p
__global__ void __launch_bounds__(256) k_dir(int Q, int Kp, const int* __restrict__ row_begin, const float* __restrict__ al,
                                             const float* __restrict__ be, const float4* __restrict__ R, float4* __restrict__ P,
                                             double* __restrict__ X, const float* __restrict__ invd,
                                             const uint32_t* __restrict__ agg, const float4* __restrict__ e) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  const int s = blockIdx.y, r0 = row_begin[s], r1 = row_begin[s + 1];
  const int rows_per_block = blockDim.x / Q;
  const int local = threadIdx.x / Q, q = threadIdx.x - local * Q;
  if (local >= rows_per_block) return;
  const float4 a = *(const float4*)(al + (size_t)s * Kp + 4 * q);
  const float4 b = *(const float4*)(be + (size_t)s * Kp + 4 * q);
  for (int row = r0 + blockIdx.x * rows_per_block + local; row < r1; row += gridDim.x * rows_per_block) {
    const size_t i = (size_t)row * Q + q;
    const float4 r = R[i];
    float4 p = P[i];
    const float d = invd[row];
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (agg) c = e[(size_t)agg[row] * Q + q];
    if (a.x != 0.f || a.y != 0.f || a.z != 0.f || a.w != 0.f) {
      double2* xp = (double2*)(X + 4 * i);
      double2 x0 = xp[0], x1 = xp[1];
      if (a.x != 0.f) x0.x += (double)a.x * p.x;
      if (a.y != 0.f) x0.y += (double)a.y * p.y;
      if (a.z != 0.f) x1.x += (double)a.z * p.z;
      if (a.w != 0.f) x1.y += (double)a.w * p.w;
      xp[0] = x0; xp[1] = x1;
    }
    p.x = r.x * d + c.x + (b.x != 0.f ? b.x * p.x : 0.f); p.y = r.y * d + c.y + (b.y != 0.f ? b.y * p.y : 0.f);
    p.z = r.z * d + c.z + (b.z != 0.f ? b.z * p.z : 0.f); p.w = r.w * d + c.w + (b.w != 0.f ? b.w * p.w : 0.f);
    P[i] = p;
  }
}

__global__ void k_inv_diag(int n, const float* __restrict__ diag, float* __restrict__ invd) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) invd[i] = 1.0f / diag[i];
}

}  // namespace batch

namespace {

// Batched solve of the block-diagonal system S (scenes = row ranges of seg).  Leaves res.kernel = 0 and returns TCS_OK
// without touching anything when the batch cannot run here (no hierarchy): the caller then falls back to the joint solve.
int pcg_solve_batched(tcs_ctx* ctx, tcs_system& S, int K, const tcs_solver_opts& opts, const tcs_segments& seg,
                      tcs_pcg_result& res, bool* ran) {
  *ran = false;
  const int n = S.n, Kp = S.Kp, Q = Kp / 4, ns = seg.count();
  const bool dbg = getenv("TCS_DEBUG") != nullptr;
  amg::Hierarchy AH;
  TRY(amg::setup(ctx, S, Kp, AH));
  if (!AH.on) return TCS_OK;
  *ran = true;
  const size_t nk = (size_t)n * Kp;
  const int count = ns * Kp;
  DBuf<double> X;
  DBuf<float> invd;
  DBuf<int> d_rb, frozen;
  TRY(X.alloc(ctx, nk)); TRY(S.P.alloc(ctx, nk)); TRY(S.R.alloc(ctx, nk)); TRY(S.AP.alloc(ctx, nk)); TRY(invd.alloc(ctx, n));
  TRY(d_rb.alloc(ctx, (size_t)ns + 1)); TRY(frozen.alloc(ctx, count));
  CK(cudaMemcpyAsync(d_rb.p, seg.row_begin.data(), ((size_t)ns + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemsetAsync(X.p, 0, nk * 8, ctx->stream));
  CK(cudaMemsetAsync(S.P.p, 0, nk * 4, ctx->stream));
  CK(cudaMemcpyAsync(S.R.p, S.B.p, nk * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  LAUNCH(ctx, batch::k_inv_diag, div_up(n, 256), 256, 0, n, S.diag.p, invd.p);
  TRY(spmm_plan(ctx, S));   // (after P / AP exist: the ring plan balances its tile ranges with trial launches)
  res.kernel = S.ring ? 3 : S.tiled ? 2 : 1;
  // scalars per (scene, column): pAp, rz (two generations), rr, bb, rz0 ; alpha, beta ; a dummy for the fused dots
  DBuf<double> scal;
  DBuf<float> fscal;
  TRY(scal.alloc(ctx, 7 * (size_t)count)); TRY(fscal.alloc(ctx, 2 * (size_t)count));
  CK(cudaMemsetAsync(scal.p, 0, 7 * (size_t)count * 8, ctx->stream));
  CK(cudaMemsetAsync(fscal.p, 0, 2 * (size_t)count * 4, ctx->stream));
  CK(cudaMemsetAsync(frozen.p, 0, (size_t)count * 4, ctx->stream));
  double *d_pAp = scal.p, *d_rz[2] = {scal.p + count, scal.p + 2 * count}, *d_rr = scal.p + 3 * count,
         *d_bb = scal.p + 4 * count, *d_rz0 = scal.p + 5 * count, *d_dummy = scal.p + 6 * count;
  float *d_al = fscal.p, *d_be = fscal.p + count;
  const float4 *B4 = (const float4*)S.B.p;
  float4 *P4 = (float4*)S.P.p, *R4 = (float4*)S.R.p, *AP4 = (float4*)S.AP.p;
  const uint32_t* agg = amg::fine_agg(AH);
  const float4* e1 = amg::result(AH);
  // grid of the per-scene kernels: about 8 CTAs per SM in total
  int max_rows = 0;
  for (int s = 0; s < ns; ++s) max_rows = std::max(max_rows, seg.row_begin[s + 1] - seg.row_begin[s]);
  const int rows_per_block = 256 / Q;
  const unsigned gx = (unsigned)std::max(1, std::min((max_rows + rows_per_block - 1) / rows_per_block, (8 * ctx->sm_count + ns - 1) / ns));
  const dim3 grid(gx, (unsigned)ns);
  const size_t dsm = (size_t)Kp * sizeof(double);
  const double done_factor = 1e-8 * opts.tolerance * opts.tolerance;
  LAUNCH(ctx, batch::k_dot, grid, 256, dsm, Q, Kp, d_rb.p, B4, B4, d_bb);
  std::vector<double> h_bb(count), h_rr(count), stage_start(count, 1.0);
  std::vector<char> done(count, 0);
  std::vector<int> h_frozen(count, 0);
  // z = D^-1 r + Z V(Z^T r), r.z, p = z: the start of every refinement stage
  int cur = 0;
  auto restart = [&]() -> int {
    TRY(amg::apply(ctx, AH, Q, R4, d_dummy));
    CK(cudaMemsetAsync(d_rz[cur], 0, (size_t)count * 8, ctx->stream));
    CK(cudaMemsetAsync(d_rr, 0, (size_t)count * 8, ctx->stream));
    CK(cudaMemsetAsync(d_pAp, 0, (size_t)count * 8, ctx->stream));
    CK(cudaMemsetAsync(fscal.p, 0, 2 * (size_t)count * 4, ctx->stream));   // alpha = beta = 0: p = z, x untouched
    LAUNCH(ctx, batch::k_dot_rz, grid, 256, dsm, Q, Kp, d_rb.p, R4, invd.p, agg, e1, d_rz[cur], d_rr);
    LAUNCH(ctx, batch::k_dir, grid, 256, 0, Q, Kp, d_rb.p, d_al, d_be, R4, P4, X.p, invd.p, agg, e1);
    return TCS_OK;
  };
  TRY(restart());
  CK(cudaMemcpyAsync(d_rz0, d_rz[cur], (size_t)count * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  CK(cudaMemcpyAsync(h_bb.data(), d_bb, (size_t)count * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  // pairs without a right-hand side (a scene with fewer labels than the batch's maximum) are done from the start
  int open_pairs = 0;
  for (int i = 0; i < count; ++i) {
    if (!(h_bb[i] > 0.0)) { done[i] = 1; h_frozen[i] = 1; }
    else ++open_pairs;
  }
  CK(cudaMemcpyAsync(frozen.p, h_frozen.data(), (size_t)count * 4, cudaMemcpyHostToDevice, ctx->stream));
  const int probe = std::max(1, env_int("TCS_AMG_PROBE", 4));
  const double stage_factor = amg::env_double("TCS_AMG_STAGE", 1e-3);
  int it = 0, converged = open_pairs == 0 ? 1 : 0, spmm_launches = 0;
  std::vector<double> rel(count, 0.0);
  const PcgLaunch L = pcg_launch_shape(ctx, n, Q);
  DBuf<double> seg_rr;
  TRY(seg_rr.alloc(ctx, (size_t)count));
  // sampled SpMM timing (bench.py's roofline): one event pair around each of the first launches
  const int kMaxSamples = 32;
  std::vector<cudaEvent_t> evs;
  int samples = 0;
  if (opts.time_spmm) {
    evs.resize(2 * kMaxSamples);
    for (auto& e : evs) CK(cudaEventCreate(&e));
  }
  // One iteration = ~55 launches, all with the same arguments every time except the ping-pong of the two r.z
  // accumulators.  A chunk of `probe` iterations is therefore captured ONCE per solve into a CUDA graph and replayed:
  // one cudaGraphLaunch per chunk instead of ~220 kernel launches (a single host process drives at most ~430 k
  // launches/s - with 24 worker threads on 8 GPUs that was the limit of tcs::BatchSegmenter - and dependent small
  // kernels start back to back on the device).  TCS_GRAPH=0 or SpMM timing (events cannot be read back from a
  // capture) select the launch-by-launch path.
  auto enqueue_iteration = [&](int c, bool timed) -> int {
    if (timed) CK(cudaEventRecord(evs[2 * samples], ctx->stream));
    launch_spmm(ctx, S, P4, AP4, d_dummy);
    if (timed) { CK(cudaEventRecord(evs[2 * samples + 1], ctx->stream)); ++samples; }
    LAUNCH_PDL(ctx, n, batch::k_dot, grid, 256, dsm, Q, Kp, d_rb.p, (const float4*)P4, (const float4*)AP4, d_pAp);
    LAUNCH_PDL(ctx, count, batch::k_alpha, div_up(count, 128), 128, 0, count, d_rz[c], d_pAp, d_rz0, frozen.p, done_factor, d_al,
           d_rz[c ^ 1], d_rr);
    LAUNCH_PDL(ctx, n, batch::k_update, grid, 256, 0, Q, Kp, d_rb.p, d_al, (const float4*)AP4, R4);
    TRY(amg::apply(ctx, AH, Q, R4, d_dummy));
    LAUNCH_PDL(ctx, n, batch::k_dot_rz, grid, 256, dsm, Q, Kp, d_rb.p, R4, invd.p, agg, e1, d_rz[c ^ 1], d_rr);
    LAUNCH_PDL(ctx, count, batch::k_beta, div_up(count, 128), 128, 0, count, d_rz[c], d_rz[c ^ 1], d_rz0, frozen.p, done_factor, d_be,
           d_pAp);
    LAUNCH_PDL(ctx, n, batch::k_dir, grid, 256, 0, Q, Kp, d_rb.p, d_al, d_be, R4, P4, X.p, invd.p, agg, e1);
    return TCS_OK;
  };
  struct ChunkGraph {
    cudaGraphExec_t exec[2] = {nullptr, nullptr};   // by the parity of `cur` at the start of the chunk
    long long nodes = 0;
    ~ChunkGraph() { for (auto e : exec) if (e) cudaGraphExecDestroy(e); }
  } cg;
  const bool use_graph = !opts.time_spmm && env_int("TCS_GRAPH", 1) != 0;
  while (!converged && it < opts.max_iterations) {
    const int chunk = std::min(probe, opts.max_iterations - it);
    if (use_graph && chunk == probe) {
      if (!cg.exec[cur]) {
        const long long before = ctx->launches;
        CK(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
        int st = TCS_OK, c = cur;
        for (int j = 0; j < chunk && st == TCS_OK; ++j, c ^= 1) st = enqueue_iteration(c, false);
        cudaGraph_t graph = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
        cg.nodes = ctx->launches - before;
        ctx->launches = before;                       // nothing ran yet: replays are counted below
        if (st != TCS_OK) { if (graph) cudaGraphDestroy(graph); return st; }
        CK(ce);
        const cudaError_t ie = cudaGraphInstantiate(&cg.exec[cur], graph, 0);
        cudaGraphDestroy(graph);
        CK(ie);
      }
      CK(cudaGraphLaunch(cg.exec[cur], ctx->stream));
      ctx->launches += cg.nodes;
      it += chunk;
      spmm_launches += chunk;
      if (chunk & 1) cur ^= 1;
    } else
    for (int j = 0; j < chunk; ++j, ++it) {
      TRY(enqueue_iteration(cur, opts.time_spmm && samples < kMaxSamples));
      ++spmm_launches;
      cur ^= 1;
    }
    // probe: recurrence ||r||^2 per (scene, column)
    CK(cudaMemcpyAsync(h_rr.data(), d_rr, (size_t)count * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    bool stage_open = false;
    for (int i = 0; i < count && !stage_open; ++i)
      if (!done[i]) {
        const double v = std::sqrt(h_rr[i] / h_bb[i]);
        if (v == v && v > std::max(0.7 * opts.tolerance, stage_factor * stage_start[i])) stage_open = true;
      }
    if (stage_open && it < opts.max_iterations) continue;
    // stage end: float64 true residual per scene replaces r; every recurrence restarts from it
    CK(cudaMemsetAsync(seg_rr.p, 0, (size_t)count * 8, ctx->stream));
    for (int s = 0; s < ns; ++s) {
      const int r0 = seg.row_begin[s], nr = seg.row_begin[s + 1] - r0;
      if (nr > 0)
        LAUNCH(ctx, k_true_residual, std::min(L.grid, div_up(nr, 256 / Q)), L.block, L.smem, r0, nr, Q, S.rowptr.p, S.col.p,
               S.val.p, S.diag.p, B4, X.p, (const float4*)P4, R4, seg_rr.p + (size_t)s * Kp, d_dummy, d_dummy);
    }
    CK(cudaMemcpyAsync(h_rr.data(), seg_rr.p, (size_t)count * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    open_pairs = 0;
    double worst = 0.0;
    for (int i = 0; i < count; ++i) {
      if (!(h_bb[i] > 0.0)) continue;
      rel[i] = std::sqrt(h_rr[i] / h_bb[i]);
      if (rel[i] <= opts.tolerance) { done[i] = 1; h_frozen[i] = 1; }
      if (!done[i]) { ++open_pairs; worst = std::max(worst, rel[i]); }
      stage_start[i] = rel[i];
    }
    if (dbg) fprintf(stderr, "[tcs] batched pcg: it=%d, %d of %d (scene, column) pairs open, worst %.2e\n", it, open_pairs, count, worst);
    if (open_pairs == 0) { converged = 1; break; }
    CK(cudaMemcpyAsync(frozen.p, h_frozen.data(), (size_t)count * 4, cudaMemcpyHostToDevice, ctx->stream));
    TRY(restart());
  }
  S.X.swap(X);
  S.R.release(); S.AP.release();
  if (samples) {
    CK(cudaStreamSynchronize(ctx->stream));
    float sum = 0.f;
    for (int i = 0; i < samples; ++i) {
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, evs[2 * i], evs[2 * i + 1]));
      sum += ms;
    }
    res.spmm_ms = sum / samples * spmm_launches;
  }
  for (auto& e : evs) cudaEventDestroy(e);
  res.iterations = it;
  res.converged = converged;
  res.spmm_launches = spmm_launches;
  for (int k = 0; k < K; ++k) {
    double w = 0.0;
    for (int s = 0; s < ns; ++s) w = std::max(w, rel[(size_t)s * Kp + k]);
    res.rel[k] = w;
  }
  return TCS_OK;
}

}  // namespace
