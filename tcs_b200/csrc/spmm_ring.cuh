// tcs_b200/csrc/spmm_ring.cuh - the CG SpMM, third generation (included by pcg.cu after spmm.cuh).
//
// k_spmm_tiled is bound by the shared-memory pipe (profiles/r01_spmm_tiled_ncu_summary.md): 9.5 wavefronts per warp
// step of 8 entries (bank conflicts on the 64-byte row gathers, a scan pass that rewrites column indices into offsets)
// and a load phase that only the second resident CTA hides.  This kernel keeps the tiling idea and changes the
// schedule and the data layout:
//
//   * one persistent CTA per SM owns a CONTIGUOUS range of tiles (ranges balanced at plan time with the kernel's own
//     per-CTA cycle counts); the window of P it needs slides, so P lives in a shared-memory ring addressed by
//     (row & (ring_rows - 1)): every P row enters the SM once instead of once per overlapping window;
//   * a producer group (4 warps) runs S - 1 tiles ahead of the 16 consumer warps.  Per tile it issues a handful of TMA
//     bulk copies (cp.async.bulk, completion on the stage's `full` mbarrier) - ONE record with the tile's slice table,
//     row order and diagonal, ONE blob with its entries, the new ring rows - spread over the four warps because a
//     UBLKCP issue costs ~200 cycles of a single thread, and fetches the tile's far rows with 16-byte cp.async;
//     consumers release a stage through its `empty` mbarrier - no __syncthreads in the steady state;
//   * entries are stored per tile in the order the warps consume them (see k_ring_perm): rows sorted by length, the G
//     rows of a warp step-major, so a warp's loads of {offset, weight} are contiguous (no bank conflicts) and all its
//     lanes run the same number of steps; the shared-memory offset of every gathered row is computed ONCE per solve
//     and stored as a 16-bit index in 16-byte units: 6 bytes per entry from HBM instead of 8; the inner loop is
//     LDS.U16 offset, LDS.32 weight, LDS.128 gather, 4 FFMA;
//   * 64-byte rows (K = 13..16) keep a second copy of the ring shifted by 64 bytes: the two row groups of a quarter
//     warp always read opposite halves of the 128-byte bank line, so the gather is conflict-free by construction.
//
// Ring invariant: ring_rows >= S * R + 2 * H.  The producer overwrites the slots of rows that are older than the
// window of the oldest tile still in flight only after that tile's `empty` barrier completed.
#pragma once

namespace {

constexpr int kRingWarps = 16;                      // consumer warps
constexpr int kRingProducers = 128;                 // producer threads (4 warps)
constexpr int kRingThreads = 32 * kRingWarps + kRingProducers;
constexpr int kRingMetaHdr = 16 + 4 * kRingWarps;   // per-tile record: {entries, 0, 0, 0}, slice table, then perm, diag

__host__ __device__ inline int ring_meta_bytes(int R) { return kRingMetaHdr + 2 * R + 4 * R; }

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// adds to the barrier's pending transaction bytes without arriving
__device__ __forceinline__ void mbar_expect_only(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

// Row order and entry layout inside a tile.  Rows are sorted by length (descending) so that the G rows a warp walks
// together have (nearly) the same number of entries; chunks of G sorted rows are rotated over the warps by the tile
// index so every warp sees long and short rows over time.  A warp's G rows form a SLICE stored step-major:
// entry (step k, row group g) sits at slice_start + k * G + g, rows shorter than the slice are padded with zero
// weights.  Writes the head of the tile's meta record:
//   u32 entries (incl. padding, rounded up to 8), 3 x u32 zero,
//   u32 slice[warp] = slice start inside the tile (low 16 bits) | steps (high 16 bits),
//   u16 perm[position] = local row index,  f32 diag[local row]
// One CTA of R threads per tile.
__global__ void k_ring_perm(int n, int R, int G, int sort_rows, const uint32_t* __restrict__ rowptr,
                            const float* __restrict__ diag, unsigned char* __restrict__ meta,
                            uint32_t* __restrict__ tile_entries) {
  extern __shared__ int len_s[];  // [R] by local row, [R] by position
  int* len_pos = len_s + R;
  const int t = blockIdx.x, i = threadIdx.x, row = t * R + i;
  unsigned char* m = meta + (size_t)t * ring_meta_bytes(R);
  uint32_t* hdr = (uint32_t*)m;
  uint16_t* perm = (uint16_t*)(m + kRingMetaHdr);
  float* dg = (float*)(m + kRingMetaHdr + 2 * R);
  const int len = row < n ? (int)(rowptr[row + 1] - rowptr[row]) : 0;
  len_s[i] = len;
  dg[i] = row < n ? diag[row] : 0.f;
  __syncthreads();
  int rank = i;
  const int nw = R / G;
  if (sort_rows) {
    rank = 0;
    for (int k = 0; k < R; ++k) {
      const int lk = len_s[k];
      rank += (lk > len || (lk == len && k < i)) ? 1 : 0;
    }
    rank = ((rank / G + t) % nw) * G + rank % G;
  }
  perm[rank] = (uint16_t)i;
  len_pos[rank] = len;
  __syncthreads();
  if (i == 0) {
    uint32_t start = 0;
    for (int w = 0; w < nw; ++w) {
      int L = 0;
      for (int g = 0; g < G; ++g) L = max(L, len_pos[w * G + g]);
      hdr[4 + w] = start | ((uint32_t)L << 16);
      start += (uint32_t)(L * G);
    }
    const uint32_t e = (start + 7u) & ~7u;
    hdr[0] = e; hdr[1] = 0u; hdr[2] = 0u; hdr[3] = 0u;
    tile_entries[t] = e;
  }
}

// Shared-memory placement of one gathered row: the two candidate byte offsets (ring copy 1 / copy 2; equal for a far
// row or when the ring is not duplicated).
struct RingCand { int o0, o1; };
// 16-byte units of the 128-byte bank line a Q-unit row at byte offset `off` covers: [u, u + Q) mod 8
__device__ __forceinline__ bool ring_rows_collide(int off_a, int off_b, int Q) {
  if (off_a == off_b) return false;  // same words for the same lane index: a broadcast, not a conflict
  const int d = ((off_b >> 4) - (off_a >> 4)) & 7;
  return d < Q || 8 - d < Q;
}
// The two row groups of a quarter warp (even / odd position) gather one row each in the same LDS.128: pick the ring
// copies so that the two rows do not share a bank.  64-byte rows: copy 2 is shifted by 64 bytes, the rows end up in
// opposite halves of the line.  48-byte rows: copy 2 is shifted by 48 bytes; for every distance of the two row starts
// one of the four combinations is collision-free.  Returns the combination as (copy of even) | (copy of odd) << 1.
__device__ __forceinline__ int ring_pair_choice(RingCand a, RingCand b, int Q) {
  if (!ring_rows_collide(a.o0, b.o0, Q)) return 0;
  if (!ring_rows_collide(a.o0, b.o1, Q)) return 2;
  if (!ring_rows_collide(a.o1, b.o0, Q)) return 1;
  if (!ring_rows_collide(a.o1, b.o1, Q)) return 3;
  return 0;
}

// Fills the tile blobs: per tile `entries` 16-bit offsets (shared-memory offset of the gathered row in 16-byte units,
// 227 KB = 14 528 units) followed by `entries` float weights, at byte 6 * tile_start[t].  cv holds
// {col | -1 - far slot, weight bits} in CSR order.  One thread per (tile, position); with a duplicated ring a thread
// also walks the entries of its quarter-warp partner (position ^ 1) to choose the copies of the pair jointly.
__global__ void k_ring_fill(int n, int Q, SpmmRing T, int ntiles, const uint32_t* __restrict__ rowptr,
                            const unsigned char* __restrict__ meta, const uint32_t* __restrict__ tile_start,
                            const int2* __restrict__ cv, unsigned char* __restrict__ blob) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= ntiles * T.R) return;
  const int t = idx / T.R, pos = idx - t * T.R;
  const int ws = pos / T.G, g = pos - ws * T.G;
  const unsigned char* m = meta + (size_t)t * ring_meta_bytes(T.R);
  const uint32_t entries = ((const uint32_t*)m)[0];
  const uint32_t sl = ((const uint32_t*)m)[4 + ws];
  const uint16_t* perm = (const uint16_t*)(m + kRingMetaHdr);
  unsigned char* b = blob + (size_t)tile_start[t] * 6;
  uint16_t* off16 = (uint16_t*)b + (sl & 0xFFFFu) + g;
  float* wout = (float*)(b + (size_t)entries * 2) + (sl & 0xFFFFu) + g;
  const int L = (int)(sl >> 16);
  const int stage = t % T.S;  // a tile's stage is its global index mod S: independent of the CTA ranges
  const int rowbytes = 16 * Q, pitch = T.pitch, mask = T.ring_rows - 1;
  const int far_base = T.off_stage0 + stage * T.stage_bytes + T.off_far;
  const int ring2 = T.off_ring + T.ring_bytes + T.dup_shift;
  const bool pairs = T.dup && T.Qp == 4;   // two row groups per quarter warp
  // candidate placements of entry k of a row (padding: the row's own ring slot with a zero weight)
  auto cand = [&](int row, uint32_t a, uint32_t a1, int k, float* w) -> RingCand {
    RingCand c;
    int col = row;
    *w = 0.f;
    if (row >= n) { c.o0 = c.o1 = T.off_ring; return c; }
    if (a + (uint32_t)k < a1) {
      const int2 e = cv[a + k];
      *w = __int_as_float(e.y);
      col = e.x;
      if (col < 0) { c.o0 = c.o1 = far_base + (-1 - col) * rowbytes; return c; }
    }
    c.o0 = T.off_ring + (col & mask) * pitch;
    c.o1 = T.dup ? ring2 + (col & mask) * pitch : c.o0;
    return c;
  };
  const int row = t * T.R + perm[pos];
  const int prow = t * T.R + perm[pos ^ 1];
  uint32_t a = 0, a1 = 0, pa = 0, pa1 = 0;
  if (row < n) { a = rowptr[row]; a1 = rowptr[row + 1]; }
  if (pairs && prow < n) { pa = rowptr[prow]; pa1 = rowptr[prow + 1]; }
  for (int k = 0; k < L; ++k) {
    float w, pw;
    const RingCand mine = cand(row, a, a1, k, &w);
    int off = mine.o0;
    if (pairs) {
      const RingCand other = cand(prow, pa, pa1, k, &pw);
      const int choice = (pos & 1) ? ring_pair_choice(other, mine, Q) : ring_pair_choice(mine, other, Q);
      const int my_copy = (pos & 1) ? (choice >> 1) : (choice & 1);
      off = my_copy ? mine.o1 : mine.o0;
    }
    off16[(size_t)k * T.G] = (uint16_t)(off >> 4);
    wout[(size_t)k * T.G] = w;
  }
}

// per-tile descriptor {first entry, entries, first far row, end far row}; zero-padded behind the last tile
__global__ void k_ring_desc(int ntiles, const uint32_t* __restrict__ tile_start,
                            const uint32_t* __restrict__ tile_far_ptr, uint4* __restrict__ desc) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles + 4) return;
  uint4 d = make_uint4(0u, 0u, 0u, 0u);
  if (t < ntiles) d = make_uint4(tile_start[t], tile_start[t + 1] - tile_start[t], tile_far_ptr[t], tile_far_ptr[t + 1]);
  desc[t] = d;
}

__global__ void __launch_bounds__(kRingThreads, 1)
    k_spmm_ring(int n, int Q, SpmmRing T, const unsigned char* __restrict__ meta, const unsigned char* __restrict__ blob,
                const int* __restrict__ far_rows, const uint4* __restrict__ desc, const int* __restrict__ cta_ptr,
                const float4* __restrict__ P, float4* __restrict__ AP, double* __restrict__ pAp,
                long long* __restrict__ cta_clk) {
  pdl_wait();      // programmatic dependent launch: the kernel before this one has completed
  pdl_trigger();   // ... and the one after it may be scheduled
  extern __shared__ __align__(128) unsigned char smem[];
  const long long clk0 = clock64();
  uint64_t* full = (uint64_t*)(smem + T.off_bar);
  uint64_t* empty = full + 8;
  uint64_t* idbar = empty + 8;  // [2]: far-row id lists
  double* red = (double*)(smem + T.off_red);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t0 = cta_ptr[blockIdx.x], t1 = cta_ptr[blockIdx.x + 1];  // contiguous tile range, balanced by work
  const int rowbytes = 16 * Q, pitch = T.pitch, mask = T.ring_rows - 1;
  const int metabytes = ring_meta_bytes(T.R);
  if (threadIdx.x == 0) {
    for (int s = 0; s < T.S; ++s) {
      mbar_init(&full[s], 1 + kRingProducers);  // the expect_tx arrive + one cp.async arrive per producer thread
      mbar_init(&empty[s], kRingWarps);
    }
    mbar_init(&idbar[0], 1);
    mbar_init(&idbar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // Qp lanes per row group (Qp = Q, or 4 for the padded 48-byte rows: lane 3 of every group idles)
  const int g = lane / T.Qp, q = lane - g * T.Qp;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  bool active = false;
  int s = t0 % T.S;   // stage of tile t = t mod S; jm = j mod S and kq = (j / S) & 1 count its uses inside this CTA
  int jm = 0;
  uint32_t kq = 0;

  if (warp >= kRingWarps) {
    // ---------------------------------------------------------------- producer group (4 warps)
    // Lane 0 of each producer warp issues one kind of TMA bulk copy (a UBLKCP issue is ~200 cycles of one thread):
    //   warp 0: the tile's meta record and entry blob     warp 1: new ring rows
    //   warp 2: new ring rows, second copy                warp 3: far-row id list of the NEXT tile
    // and all 128 threads fetch the tile's far rows with 16-byte cp.async (UBLKCP takes uniform operands: one bulk
    // copy per far row would serialise the warp), joining the stage barrier through cp.async.mbarrier.arrive.noinc.
    // Nothing sits behind a dependent global load: tile descriptors are prefetched two tiles ahead into registers.
    const int pid = threadIdx.x - 32 * kRingWarps;
    const int pw = pid >> 5;
    unsigned char* ring = smem + T.off_ring + (pw == 2 ? T.ring_bytes + T.dup_shift : 0);
    // u / Q = umulhi(u, qmagic) for u < 2^29 and Q >= 2 (Q = 1 would need 2^32: handled apart)
    const uint32_t qmagic = Q > 1 ? (uint32_t)((0x100000000ull + (unsigned)Q - 1u) / (unsigned)Q) : 0u;
    int hi_prev = 0;
    uint4 d0 = desc[t0], d1 = desc[t0 + 1];
    if (pid == 96) {  // id list of the first tile
      const uint32_t f0 = d0.z & ~3u, f1 = (d0.w + 3u) & ~3u;
      mbar_expect_tx(&idbar[0], (f1 - f0) * 4u);
      if (f1 > f0) bulk_g2s(smem + T.off_ids, far_rows + f0, (f1 - f0) * 4u, &idbar[0]);
    }
    for (int t = t0, j = 0; t < t1; ++t, ++j) {
      const uint4 d2 = desc[t + 2];
      // all producer threads are done with the id buffer the next bulk copy overwrites
      asm volatile("bar.sync 1, %0;" ::"n"(kRingProducers) : "memory");
      if (pid == 96 && t + 1 < t1) {
        const uint32_t f0 = d1.z & ~3u, f1 = (d1.w + 3u) & ~3u;
        uint64_t* ib = &idbar[(j + 1) & 1];
        mbar_expect_tx(ib, (f1 - f0) * 4u);
        if (f1 > f0) bulk_g2s(smem + T.off_ids + ((j + 1) & 1) * T.ids_bytes, far_rows + f0, (f1 - f0) * 4u, ib);
      }
      if (j >= T.S) mbar_wait(&empty[s], kq ^ 1u);
      unsigned char* st = smem + T.off_stage0 + s * T.stage_bytes;
      if (pid == 0) {
        mbar_expect_tx(&full[s], (uint32_t)metabytes + d0.y * 6u);
        bulk_g2s(st + T.off_meta, meta + (size_t)t * metabytes, (uint32_t)metabytes, &full[s]);
        if (d0.y) bulk_g2s(st + T.off_cv, blob + (size_t)d0.x * 6, d0.y * 6u, &full[s]);
      } else if (!T.fill_async && (pid == 32 || (pid == 64 && T.dup))) {
        const int r0 = t * T.R, r1 = min(n, r0 + T.R);
        const int lo = j == 0 ? max(0, r0 - T.H) : hi_prev;
        const int hi = max(lo, min(n, r1 + T.H));
        if (hi > lo) mbar_expect_only(&full[s], (uint32_t)(hi - lo) * (uint32_t)rowbytes);
        for (int x = lo; x < hi;) {
          const int slot = x & mask;
          const int cnt = min(hi - x, T.ring_rows - slot);
          bulk_g2s(ring + (size_t)slot * rowbytes, P + (size_t)x * Q, (uint32_t)cnt * rowbytes, &full[s]);
          x += cnt;
        }
        hi_prev = hi;
      }
      if (T.fill_async) {
        // padded ring rows (48-byte rows at a 64-byte pitch): a bulk copy cannot pad, so all producer threads fill
        // the ring with 16-byte cp.async, both copies
        const int r0 = t * T.R, r1 = min(n, r0 + T.R);
        const int lo = j == 0 ? max(0, r0 - T.H) : hi_prev;
        const int hi = max(lo, min(n, r1 + T.H));
        unsigned char* rg0 = smem + T.off_ring;
        const uint32_t nrc = (uint32_t)(hi - lo) * (uint32_t)Q;
        for (uint32_t u = pid; u < nrc; u += kRingProducers) {
          const uint32_t i = (Q > 1 ? __umulhi(u, qmagic) : u), ch = u - i * (uint32_t)Q;
          const int x = lo + (int)i;
          const float4* src = &P[(size_t)x * Q + ch];
          unsigned char* dst = rg0 + (size_t)(x & mask) * pitch + ch * 16u;
          cp_async16(dst, src);
          if (T.dup) cp_async16(dst + T.ring_bytes + T.dup_shift, src);
        }
        hi_prev = hi;
      }
      // far rows of this tile
      mbar_wait(&idbar[j & 1], (uint32_t)(j >> 1) & 1u);
      const int* ids = (const int*)(smem + T.off_ids + (j & 1) * T.ids_bytes) + (d0.z & 3u);
      const uint32_t nch = (d0.w - d0.z) * (uint32_t)Q;
      for (uint32_t u = pid; u < nch; u += kRingProducers) {
        const uint32_t i = (Q > 1 ? __umulhi(u, qmagic) : u), ch = u - i * (uint32_t)Q;
        cp_async16(st + T.off_far + u * 16u, &P[(size_t)ids[i] * Q + ch]);
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[s])) : "memory");
      d0 = d1; d1 = d2;
      if (++s == T.S) s = 0;
      if (++jm == T.S) { jm = 0; kq ^= 1u; }
    }
  } else {
    // ---------------------------------------------------------------- consumer warps: one row per Q-lane group
    active = g < T.G && q < Q;
    const int pos = warp * T.G + g;
    const unsigned char* Pq = smem + q * 16;
    const int own_ring2 = T.off_ring + T.ring_bytes + T.dup_shift;
    const bool own_by_parity = T.dup && T.dup_shift == 64;   // 64-byte rows: the own-row read follows the position parity
    const int G = T.G;
    for (int t = t0; t < t1; ++t) {
      mbar_wait(&full[s], kq);
      const unsigned char* st = smem + T.off_stage0 + s * T.stage_bytes;
      if (active) {
        const uint32_t* hdr = (const uint32_t*)(st + T.off_meta);
        const uint32_t entries = hdr[0], sl = hdr[4 + warp];
        const int lr = (int)((const uint16_t*)(st + T.off_meta + kRingMetaHdr))[pos];
        const int row = t * T.R + lr;
        const int L = (int)(sl >> 16);
        const uint16_t* off_s = (const uint16_t*)(st + T.off_cv) + (sl & 0xFFFFu) + g;
        const float* w_s = (const float*)(st + T.off_cv + entries * 2u) + (sl & 0xFFFFu) + g;
        const int own = (own_by_parity && ((row ^ pos) & 1)) ? own_ring2 : T.off_ring;
        const float4 p = *(const float4*)(Pq + own + (row & mask) * pitch);
        const float d = ((const float*)(st + T.off_meta + kRingMetaHdr + 2 * T.R))[lr];
        const F4p pp = f4_pack(p);
        F4p yy = f4_pack(make_float4(d * p.x, d * p.y, d * p.z, d * p.w));
#pragma unroll 4
        for (int k = 0; k < L; ++k) {
          const uint32_t off = off_s[k * G];
          const float w = w_s[k * G];
          const float4 v = *(const float4*)(Pq + (off << 4));
          f4_diff_fma(yy, w, v, pp);   // y += w * (v - p): two FADD2 + two FFMA2
        }
        const float4 y = f4_unpack(yy);
        if (row < n) {
          AP[(size_t)row * Q + q] = y;
          acc[0] += (double)p.x * (double)y.x; acc[1] += (double)p.y * (double)y.y;
          acc[2] += (double)p.z * (double)y.z; acc[3] += (double)p.w * (double)y.w;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      if (++s == T.S) s = 0;
      if (++jm == T.S) { jm = 0; kq ^= 1u; }
    }
  }
  reduce_columns(acc, q < Q ? q : 0, Q, active, red, pAp);
  if (cta_clk && threadIdx.x == 0) cta_clk[blockIdx.x] = clock64() - clk0;  // feeds the plan's range balancing
}

}  // namespace
