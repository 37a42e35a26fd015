// Per-cell ("stock-month") kernels, exact-fp32 FFMA path.
//
// One tile = up to 128 consecutive stocks of ONE month.  A persistent CTA of 256 threads
// walks tiles with a static stride; per tile it
//   * pulls the [ncell, F] fp32 characteristic slab with one 1-D bulk copy (TMA engine) into a
//     staging buffer, transposes it to unit-major smem, and immediately re-arms the copy for
//     its next tile so the load overlaps the math;
//   * runs the SDF MLP  z1 = W1x.x + a_t, h1 = relu, z2 = W2.h1 + b2, h2 = relu, u = W3.h2 + b3
//     (reference: src/model.py:208-219, 261-268; the tile-over-N concat of :253-255 is the
//     per-month bias a_t) as register-tiled fp32 GEMMs, weights resident in smem;
//   * runs the moment net  q = tanh(Wqx.x + c_t)  (src/model.py:132-161, :514-516).
// The forward kernel emits u and q (everything that sums over stocks or months happens in step_reduce.cu);
// the backward kernel recomputes the activations instead of reading saved ones and produces per-CTA
// weight-gradient partials plus per-tile da_t / dc_t partials.
#pragma once
#include "dlap_common.cuh"

namespace dlap {

struct CellParams {
  // panel (caller-owned, read-only)
  const float* x;            // [T,N,F]
  const float* R;            // [T,N]      backward kernels only
  const uint8_t* mask;       // [T,N]      torch.bool bytes
  int T, N, F;
  long long x_bytes;         // T*N*F*4, to keep the bulk copy inside the tensor
  // parameters (device, reference layouts)
  const float* W1; int ldw1; // fc_layers.0.weight [64, ldw1], characteristics are columns 0..F-1
  const float* W2;           // fc_layers.3.weight [64,64]
  const float* b2;           // [64]
  const float* W3;           // output_proj.weight [1,64]
  const float* b3;           // [1]
  const float* Wq; int ldwq; int wq_off;  // moment output_proj.weight [8, ldwq]; chars start at wq_off
  // per-month vectors
  const float* a_t;          // [T,64]  W1m.h_t + b1
  const float* c_t;          // [T,8]   Wqm.macro_t + bq
  // forward outputs
  float* u;                  // [T,N]
  float* q;                  // [8,T,N]  tanh outputs of the moment net   (null: skip the moment net)
  // backward inputs
  const float* row_stats;    // [T,RS_W]
  const double* dS2;         // [T,2]  dS_t for unit cotangents of L_c and of L_u (step_reduce.cu, phase 4)
  const float* cot[2];       // device scalars dL/dL_c, dL/dL_u: dS_t = cot[0] dS2[t][0] + cot[1] dS2[t][1]
  int ncot;                  // 2
  const float* dpr_cot;      // [T] cotangent of portfolio_returns, or null
  const float* G;            // [N,GS]  unit loss coefficients per stock: G_0..G_7 (times the L_c cotangent), G_u
  const float* scalars;      // [SC_W]  (n_bar)
  int weighted;              // weighted_loss flag
  // residual-loss regulariser (src/model.py:435-483; null = off): its gradient reaches u as
  //   du_ti += a_t (R_ti - Rbar_t) + res_b[t] (u_ti - mu_t);  a_t rides in dpr_cot, the second term needs the
  //   forward's u (left-packed for the tensor-core kernels, dense for fp32) and one coefficient per month
  const float* res_b;        // [T]
  const float* u_fwd;        // [T,N]
  // backward outputs
  float* wpart;              // [grid, WP_STRIDE] per-CTA weight-gradient partials
  float* dadc_part;          // [T,tpm,72]  da_t (64) then dc_t (8)
  int tpm;                   // tiles per month
  // dropout after each ReLU of the SDF net (src/model.py:213), train mode only
  uint32_t drop_thr;         // drop when a 16-bit uniform < drop_thr (0 = off)
  float drop_scale;          // 1 / (1 - drop_thr/65536)
  uint32_t seed_lo, seed_hi; // per-forward seed; the backward pass regenerates the same mask
  const uint32_t* seed_ptr;  // non-null: the 64-bit seed lives in device memory (a step replayed from a CUDA graph)
  long long* timeline;       // diagnostics (dlap_debug_timeline): per-tile clock64 stamps of CTA 0, or null
  // valid-cell index (dlap_panel_index); tensor-core kernels only, null = not built
  const int* index;
  // X images of every tile of the index (dlap_panel_image; cell_tc.cuh), or null: gather the rows of x instead
  const uint8_t* ximg;
  // non-null: a_t / c_t are still being written by a concurrent dlap_macro_fwd; months [0, *months_ready) are
  // complete (tensor-core forward only); month t of this (possibly month-sharded) panel is month months_base + t there
  const int* months_ready;
  int months_base;
  int months_stride;         // rows of a_t / c_t between consecutive months of this panel (1; n for a rank that owns
                             // every n-th month of the series)
  int cyc_sh;                // log2 of the block size of the block-cyclic tile order used with months_ready
};

// dL/dP_t of month t from the two unit sums and the loss cotangents (src/model.py:367 backwards):
// dS_t = gc dS2[t][0] + gu dS2[t][1] (+ the cotangent of portfolio_returns).  Written as a loop over P.ncot = 2
// terms on purpose: straight-line code lets ptxas hoist these once-a-month loads to the top of the per-tile path,
// where the tensor-core backward kernel has no register to spare (measured: 40 bytes of spills, +18 us).
__device__ __forceinline__ float month_dP(const CellParams& P, int t) {
  double dS = 0.0;
#pragma unroll 1
  for (int k = 0; k < P.ncot; ++k) dS += (double)__ldg(P.cot[k]) * P.dS2[2 * (size_t)t + k];
  if (P.dpr_cot) dS += (double)__ldg(P.dpr_cot + t);
  const float* rs = P.row_stats + (size_t)t * RS_W;
  return (float)(dS * (P.weighted ? (double)P.scalars[SC_NBAR] / (double)rs[RS_NT] : 1.0));
}

// ---- valid-cell index ----------------------------------------------------------------------
// The reference pushes every cell through both nets and multiplies by the mask afterwards
// (src/model.py:261-279, :376-382); a masked cell therefore contributes exactly 0 to every loss, weight
// and gradient.  dlap_panel_index lists, per month, the stocks with mask = 1 in ascending order and cuts
// each list into tiles of <= 128 cells; the tensor-core kernels walk those tiles only.
//   int32 index[]:  [0] total tiles | [4, 4+T] tile_start (exclusive prefix, [T] = total) | cnt [T] |
//                   vidx [T, N] (16-B aligned): row t holds cnt[t] stock numbers, rest undefined |
//                   rank [T, ceil(N/32)]: unmasked stocks of month t before stock 32 w |
//                   tmonth [total tiles <= T ceil(N/128)]: month of every tile
// Per-cell intermediates of the tensor-core path (u, g) are stored "left-packed": the value of the j-th listed
// stock of month t sits at [t*N + j].  A per-month consumer reads a contiguous run; a per-stock consumer finds
// its position as rank[t][i / 32] + popc(ballot(mask) below its lane).
__host__ __device__ inline int idx_off_ts() { return 4; }
__host__ __device__ inline int idx_off_cnt(int T) { return 4 + T + 1; }
__host__ __device__ inline long long idx_off_vidx(int T) { return (4 + 2 * T + 1 + 3) & ~3; }
__host__ __device__ inline long long idx_off_rank(int T, int N) { return idx_off_vidx(T) + (long long)T * N; }
__host__ __device__ inline long long idx_off_tmonth(int T, int N) {
  return (idx_off_rank(T, N) + (long long)T * ((N + 31) / 32) + 3) & ~3ll;
}
__host__ __device__ inline long long idx_ints(int T, int N) {
  return idx_off_tmonth(T, N) + (((long long)T * ((N + TILE - 1) / TILE) + 3) & ~3ll);
}

__device__ __forceinline__ const int* vidx_row(const CellParams& P, int t) {
  return P.index + idx_off_vidx(P.T) + (size_t)t * P.N;
}

// Walks the tile list in ascending order: month of a tile, its position in the month's list.
struct MonthCursor {
  int t, first, next, cnt;   // month, its first tile, first tile of the following month, valid cells
  // first use: `tile` < total tiles
  __device__ __forceinline__ void init(const CellParams& P, int tile) {
    const int* ts = P.index + idx_off_ts();
    t = __ldg(P.index + idx_off_tmonth(P.T, P.N) + tile);
    first = __ldg(ts + t); next = __ldg(ts + t + 1);
    cnt = __ldg(P.index + idx_off_cnt(P.T) + t);
  }
  // move forward to `tile` (>= the current one, < total tiles): the next month as a rule, a table look-up when
  // the tile lies further ahead (empty months; block-cyclic tile order, TileMap)
  __device__ __forceinline__ void seek(const CellParams& P, int tile) {
    if (tile < next) return;
    const int* ts = P.index + idx_off_ts();
    ++t; first = next; next = __ldg(ts + t + 1);
    if (tile >= next) {
      t = __ldg(P.index + idx_off_tmonth(P.T, P.N) + tile);
      first = __ldg(ts + t); next = __ldg(ts + t + 1);
    }
    cnt = __ldg(P.index + idx_off_cnt(P.T) + t);
    // consume the load inside this (warp-uniform, once-a-month) branch: its first consumer would otherwise be an
    // instruction on the per-tile path, whose scoreboard wait then also catches every other load in flight
    cnt = __shfl_sync(0xffffffffu, cnt, 0);
  }
  __device__ __forceinline__ int j0(int tile) const { return (tile - first) * TILE; }
  __device__ __forceinline__ int ncell(int tile) const { return min(TILE, cnt - (tile - first) * TILE); }
};

// Which tiles a persistent CTA works on, in its own order: local tile i -> tile(i).
//   contiguous (backward kernel): CTA c owns [c * chunk, (c + 1) * chunk), so a month's sums stay in one CTA;
//   block-cyclic (forward kernel): blocks of 2^sh tiles dealt round-robin to the CTAs, so that every CTA walks
//   the months from the first to the last -- the order a concurrent LSTM recurrence delivers a_t in.
struct TileMap {
  int base, sh, jump, count;
  __device__ __forceinline__ int tile(int i) const { return base + i + (i >> sh) * jump; }
  static __device__ __forceinline__ TileMap contiguous(int ntiles, int cta, int ncta) {
    const int chunk = (ntiles + ncta - 1) / ncta;
    const int lo = min(ntiles, cta * chunk);
    return TileMap{lo, 31, 0, min(ntiles, lo + chunk) - lo};
  }
  static __device__ __forceinline__ TileMap cyclic(int ntiles, int cta, int ncta, int sh) {
    const int B = 1 << sh, nblk = (ntiles + B - 1) >> sh;
    const int mine = cta < nblk ? (nblk - cta + ncta - 1) / ncta : 0;      // blocks cta, cta + ncta, ...
    int n = mine << sh;
    if (mine > 0 && cta + (mine - 1) * ncta == nblk - 1) n -= (nblk << sh) - ntiles;   // the last block may be short
    return TileMap{cta << sh, sh, (ncta - 1) << sh, n};
  }
};

// spin until month t of a concurrently produced a_t / c_t is complete (CellParams::months_ready); `seen` caches
// the last value read.  A producer that never shows up (it could not be scheduled beside this kernel) turns
// into a trap after ~4 s instead of a hang.
__device__ __forceinline__ void wait_month(const int* ready, int t, int& seen) {
  if (t < seen) return;
  unsigned long long t0 = 0;
  for (;;) {
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(ready) : "memory");
    if (t < seen) return;
    __nanosleep(200);
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > 4000000000ull) __trap();
  }
}

// Gather of characteristic rows by one warp.  Lane l (< ROWS) knows the stock number of row l of the batch
// (n_lane; a cell past the end of the tile repeats the tile's last row, so only real, finite panel rows are
// ever fetched).  Row j is fetched by ONE coalesced warp-wide load -- lane p takes columns 2p, 2p+1 -- so ROWS
// independent loads are in flight per warp; scattered 184-B rows need that memory-level parallelism (per-row
// TMA bulk copies and cp.async both serialise: measured 78 and 190 cycles per row and warp).
// The loads are unconditional on clamped addresses (lanes past the row re-read its last element; whoever
// consumes the registers ignores them): a predicated load followed by a select makes ptxas wait for each load
// before issuing the next one.  Addresses are a 64-bit base plus a 32-bit byte offset (one IMAD.WIDE).
// A row lands in ONE 64-bit register per lane (low word = column 2p, high word = column 2p+1): 64-bit scalars
// live in aligned register pairs.  WIDE (8-byte loads: F even, x 8-B aligned) is a template parameter so that
// the row registers have a single definition inside the caller's loop: with two definitions (a prologue, or
// an if/else over the load width) ptxas gives them different homes and copies every loaded value after the
// loads -- one full-latency wait per load, inside the issue loop.
template <int ROWS, bool WIDE>
__device__ __forceinline__ void gather_load(const CellParams& P, int t, int n_lane, int lane, uint64_t (&r)[ROWS]) {
  const char* xt = reinterpret_cast<const char*>(P.x + (size_t)t * P.N * P.F);
  const uint32_t pitch = (uint32_t)P.F * 4u;                 // N * F * 4 < 2^32 (checked by dlap_make_plan)
  if (WIDE) {
    const char* base = xt + 8 * min(lane, (P.F >> 1) - 1);
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      const uint32_t nr = (uint32_t)__shfl_sync(0xffffffffu, n_lane, j);
      r[j] = __ldg(reinterpret_cast<const unsigned long long*>(base + nr * pitch));
    }
  } else {
    const char* b0 = xt + 4 * min(2 * lane, P.F - 1);
    const char* b1 = xt + 4 * min(2 * lane + 1, P.F - 1);
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      const uint32_t nr = (uint32_t)__shfl_sync(0xffffffffu, n_lane, j);
      const uint32_t lo = __ldg(reinterpret_cast<const unsigned int*>(b0 + nr * pitch));
      const uint32_t hi = __ldg(reinterpret_cast<const unsigned int*>(b1 + nr * pitch));
      r[j] = ((uint64_t)hi << 32) | lo;
    }
  }
}
__device__ __forceinline__ bool gather_wide(const CellParams& P) {
  return ((P.F & 1) == 0) && ((reinterpret_cast<uintptr_t>(P.x) & 7) == 0);
}
// L2 prefetch of this lane's own row (stock n of month t), issued two batches ahead of its gather_load so that
// the loads find the row in L2 instead of paying the HBM latency of a scattered 184-B read
__device__ __forceinline__ void prefetch_row(const CellParams& P, int t, int n) {
  const char* a = reinterpret_cast<const char*>(P.x + ((size_t)t * P.N + n) * P.F);
  const int bytes = P.F * 4;
  asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
  if (bytes > 64) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 64));
  if (bytes > 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 128));
  asm volatile("prefetch.global.L2 [%0];" ::"l"(a + bytes - 4));
}

// layout of one CTA's weight-gradient partial
__host__ __device__ inline int wp_off_w2() { return 0; }
__host__ __device__ inline int wp_off_w1() { return 64 * 64; }
__host__ __device__ inline int wp_off_wq(int Fr) { return 64 * 64 + 64 * Fr; }
__host__ __device__ inline int wp_off_w3(int Fr) { return 64 * 64 + 72 * Fr; }
__host__ __device__ inline int wp_off_b2(int Fr) { return 64 * 64 + 72 * Fr + 64; }
__host__ __device__ inline int wp_stride(int Fr) { return 64 * 64 + 72 * Fr + 128; }

struct SmemFwd {
  float* stage; float* xs; float* hs; float* W1s; float* W2s; float* Wqs;
  float* b2s; float* w3s; float* as; float* cs; float* us; float* sc; double* red; uint64_t* bar;
};

__host__ __device__ inline int round4(int v) { return (v + 3) & ~3; }

// Tensor-core kernels give every CTA a CONTIGUOUS range of `chunk` valid-cell tiles; per-month partial sums
// (da_t | dc_t) are then written once per (month, CTA) "segment": month t is touched by the consecutive CTAs
// tile_start[t] / chunk .. (tile_start[t+1] - 1) / chunk, never more than its tile count.
__host__ __device__ inline int tc_chunk(long long ntiles, int grid) { return (int)((ntiles + grid - 1) / grid); }

// bytes of dynamic smem for the forward kernel
__host__ __device__ inline size_t smem_fwd_bytes(int F) {
  const int Fr = round4(F);
  size_t fl = (size_t)(TILE * F + 8) + (size_t)Fr * TILE + 64 * TILE + (size_t)Fr * 64 + 64 * 64 +
              (size_t)Fr * 8 + 64 + 64 + 64 + 8 + TILE + 8;
  return fl * 4 + 64 * 8 + 16 + 16;
}

__device__ __forceinline__ void carve_fwd(unsigned char* base, int F, SmemFwd& s) {
  const int Fr = round4(F);
  float* p = reinterpret_cast<float*>(base);
  s.stage = p; p += TILE * F + 8;
  p = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(p) + 15) & ~uintptr_t(15));
  s.xs = p; p += Fr * TILE;
  s.hs = p; p += 64 * TILE;
  s.W1s = p; p += Fr * 64;
  s.W2s = p; p += 64 * 64;
  s.Wqs = p; p += Fr * 8;
  s.b2s = p; p += 64;
  s.w3s = p; p += 64;
  s.as = p; p += 64;
  s.cs = p; p += 8;
  s.us = p; p += TILE;
  s.sc = p; p += 8;
  p = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(p) + 7) & ~uintptr_t(7));
  s.red = reinterpret_cast<double*>(p); p += 64 * 2;
  s.bar = reinterpret_cast<uint64_t*>(p);
}

// ---- tile streaming ----------------------------------------------------------------------
struct TileLoad {
  const float* src;   // first float of the tile in global memory
  int ncell;          // valid cells
  int shift;          // floats between the 16-B aligned copy start and src (0..3)
  uint32_t bytes;     // bulk-copy size (multiple of 16), 0 => cooperative fallback
};

__device__ __forceinline__ TileLoad plan_tile(const CellParams& P, int t, int n0) {
  TileLoad L;
  L.ncell = min(TILE, P.N - n0);
  L.src = P.x + ((size_t)t * P.N + n0) * P.F;
  const uintptr_t a = reinterpret_cast<uintptr_t>(L.src);
  const uintptr_t a0 = a & ~uintptr_t(15);
  const uintptr_t e1 = (a + (size_t)L.ncell * P.F * 4 + 15) & ~uintptr_t(15);
  const uintptr_t xend = reinterpret_cast<uintptr_t>(P.x) + (uintptr_t)P.x_bytes;
  L.shift = (int)((a - a0) >> 2);
  L.bytes = (e1 <= xend && (a & 3) == 0) ? (uint32_t)(e1 - a0) : 0u;
  return L;
}

// Starts the copy of a tile into `stage`.  Bulk path: one elected thread arms the mbarrier and
// issues the copy.  Fallback (copy would run past the tensor): every thread copies floats.
__device__ __forceinline__ void issue_tile(const TileLoad& L, int F, float* stage, uint64_t* bar) {
  if (L.bytes) {
    if (threadIdx.x == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar, L.bytes);
      bulk_g2s(stage, L.src - L.shift, L.bytes, bar);
    }
  } else {
    const int n = L.ncell * F;
    for (int i = threadIdx.x; i < n; i += NT) stage[L.shift + i] = __ldg(L.src + i);
  }
}

// stage (cell-major, as in HBM) -> xs (swizzled unit-major [Fr][128]); zero-fills missing
// cells/rows
__device__ __forceinline__ void transpose_tile(const TileLoad& L, int F, const float* stage,
                                               float* xs) {
  const int Fr = round4(F);
  const int cell = threadIdx.x & (TILE - 1);
  const bool live = cell < L.ncell;
  const float* row = stage + L.shift + cell * F;
  for (int k = threadIdx.x >> 7; k < Fr; k += NT / TILE)
    xs[act_idx(k, cell)] = (live && k < F) ? row[k] : 0.f;
}

// ---- weights -> smem ---------------------------------------------------------------------
__device__ __forceinline__ void load_weights_fwd(const CellParams& P, SmemFwd& s) {
  const int F = P.F, Fr = round4(F);
  for (int i = threadIdx.x; i < Fr * 64; i += NT) {
    const int k = i >> 6, j = i & 63;
    s.W1s[i] = (k < F) ? __ldg(P.W1 + (size_t)j * P.ldw1 + k) : 0.f;
  }
  for (int i = threadIdx.x; i < 64 * 64; i += NT) {
    const int k = i >> 6, j = i & 63;
    s.W2s[i] = __ldg(P.W2 + j * 64 + k);
  }
  for (int i = threadIdx.x; i < Fr * 8; i += NT) {
    const int k = i >> 3, q = i & 7;
    s.Wqs[i] = (k < F && P.Wq) ? __ldg(P.Wq + (size_t)q * P.ldwq + P.wq_off + k) : 0.f;
  }
  if (threadIdx.x < 64) {
    s.b2s[threadIdx.x] = __ldg(P.b2 + threadIdx.x);
    s.w3s[threadIdx.x] = __ldg(P.W3 + threadIdx.x);
  }
}

// ---- register-tiled GEMM pieces ------------------------------------------------------------
// thread (ty = tid>>3, tx = tid&7) owns cells 4*ty..4*ty+3 and units 4*tx+i, 32+4*tx+i.
__device__ __forceinline__ int unit_of(int tx, int o) { return (o < 4) ? 4 * tx + o : 28 + 4 * tx + o; }

#define DLAP_FMA4x8(acc, av, w0, w1)                                   \
  {                                                                    \
    const float a_[4] = {av.x, av.y, av.z, av.w};                      \
    _Pragma("unroll") for (int c_ = 0; c_ < 4; ++c_) {                 \
      acc[c_][0] = fmaf(a_[c_], w0.x, acc[c_][0]);                     \
      acc[c_][1] = fmaf(a_[c_], w0.y, acc[c_][1]);                     \
      acc[c_][2] = fmaf(a_[c_], w0.z, acc[c_][2]);                     \
      acc[c_][3] = fmaf(a_[c_], w0.w, acc[c_][3]);                     \
      acc[c_][4] = fmaf(a_[c_], w1.x, acc[c_][4]);                     \
      acc[c_][5] = fmaf(a_[c_], w1.y, acc[c_][5]);                     \
      acc[c_][6] = fmaf(a_[c_], w1.z, acc[c_][6]);                     \
      acc[c_][7] = fmaf(a_[c_], w1.w, acc[c_][7]);                     \
    }                                                                  \
  }

// z1 = a_t + W1x.x  -> acc ; W1s [k][64]
__device__ __forceinline__ void layer1(const SmemFwd& s, int F, int ty, int tx, float acc[4][8]) {
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const float b = s.as[unit_of(tx, o)];
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[c][o] = b;
  }
  const float4* xs4 = reinterpret_cast<const float4*>(s.xs);
  const float4* w4 = reinterpret_cast<const float4*>(s.W1s);
#pragma unroll 2
  for (int k = 0; k < F; ++k) {
    const float4 xv = xs4[act_idx4(k, ty)];
    const float4 w0 = w4[k * 16 + tx], w1 = w4[k * 16 + 8 + tx];
    DLAP_FMA4x8(acc, xv, w0, w1);
  }
}

// ---- dropout: counter-based, keyed on (seed, stock-month, layer, unit) ----------------------------
__device__ __forceinline__ uint32_t drop_seed_lo(const CellParams& P) { return P.seed_ptr ? __ldg(P.seed_ptr) : P.seed_lo; }
__device__ __forceinline__ uint32_t drop_seed_hi(const CellParams& P) { return P.seed_ptr ? __ldg(P.seed_ptr + 1) : P.seed_hi; }
__device__ __forceinline__ uint32_t mix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
// acc <- relu(acc) * (0 | scale); one 32-bit hash decides two neighbouring units of one cell
__device__ __forceinline__ void relu_dropout(const CellParams& P, uint32_t cell0, int layer, int tx,
                                             float acc[4][8]) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const uint32_t key = mix32((cell0 + c) ^ drop_seed_lo(P)) ^ drop_seed_hi(P);
#pragma unroll
    for (int pr = 0; pr < 4; ++pr) {
      const int j0 = unit_of(tx, 2 * pr);
      const uint32_t r = mix32(key + (uint32_t)(layer * 32 + (j0 >> 1)) * 0x9E3779B9u);
      acc[c][2 * pr] = ((r & 0xffffu) >= P.drop_thr) ? fmaxf(acc[c][2 * pr], 0.f) * P.drop_scale : 0.f;
      acc[c][2 * pr + 1] = ((r >> 16) >= P.drop_thr) ? fmaxf(acc[c][2 * pr + 1], 0.f) * P.drop_scale : 0.f;
    }
  }
}

// relu(acc) -> hs (swizzled unit-major)
__device__ __forceinline__ void store_relu(float* hs, int ty, int tx, float acc[4][8]) {
  float4* hs4 = reinterpret_cast<float4*>(hs);
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const int j = unit_of(tx, o);
    float4 v;
    v.x = fmaxf(acc[0][o], 0.f); v.y = fmaxf(acc[1][o], 0.f);
    v.z = fmaxf(acc[2][o], 0.f); v.w = fmaxf(acc[3][o], 0.f);
    hs4[act_idx4(j, ty)] = v;
  }
}

// z2 = b2 + W2.h1 -> acc
__device__ __forceinline__ void layer2(const SmemFwd& s, int ty, int tx, float acc[4][8]) {
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const float b = s.b2s[unit_of(tx, o)];
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[c][o] = b;
  }
  const float4* hs4 = reinterpret_cast<const float4*>(s.hs);
  const float4* w4 = reinterpret_cast<const float4*>(s.W2s);
#pragma unroll 4
  for (int k = 0; k < 64; ++k) {
    const float4 hv = hs4[act_idx4(k, ty)];
    const float4 w0 = w4[k * 16 + tx], w1 = w4[k * 16 + 8 + tx];
    DLAP_FMA4x8(acc, hv, w0, w1);
  }
}

// moment pre-activations: thread (cell = tid&127, half = tid>>7) does moments 4*half..4*half+3
__device__ __forceinline__ void moment_pre(const SmemFwd& s, int F, float qp[4]) {
  const int cell = threadIdx.x & (TILE - 1), half = threadIdx.x >> 7;
#pragma unroll
  for (int i = 0; i < 4; ++i) qp[i] = s.cs[4 * half + i];
  const float4* wq4 = reinterpret_cast<const float4*>(s.Wqs);
#pragma unroll 2
  for (int k = 0; k < F; ++k) {
    const float xv = s.xs[act_idx(k, cell)];
    const float4 w = wq4[k * 2 + half];
    qp[0] = fmaf(xv, w.x, qp[0]); qp[1] = fmaf(xv, w.y, qp[1]);
    qp[2] = fmaf(xv, w.z, qp[2]); qp[3] = fmaf(xv, w.w, qp[3]);
  }
}

}  // namespace dlap
