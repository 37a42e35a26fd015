// Forward per-cell kernel on the 5th-gen tensor cores (tcgen05 + TMEM), precision "bf16x3".
// Same contract as cell_fwd_f32_kernel (cell_fwd.cu); see cell_tc.cuh for the arithmetic.
//
// The kernel walks the valid-cell tiles of dlap_panel_index (cell_f32.cuh): a tile is <= 128 unmasked
// stocks of one month, gathered row by row.
//
// h1 never touches shared memory: the epilogue writes relu(z1), split into bf16 hi/lo, back into the TMEM
// columns of its own z1 accumulator, and layer 2 takes its A operand from there (tcgen05.mma with A in TMEM).
//
// The LSTM beside it.  With FUSED the LAST CTA of the grid does not work on cells: it runs the macro path
// (macro_role.cuh: input projection, the serial recurrence, a_t / c_t rows) and publishes the months as they
// complete; the other CTAs wait for a month's row before they touch the month's first tile and walk the tiles
// in month order.  Producer and consumers are CTAs of ONE grid that is launched co-resident (cooperative
// launch, one CTA per SM), so the wait cannot outlive the producer's start.  The progress word is put back to
// zero by the last CTA that leaves the kernel.
//
// Persistent CTA, one per SM, 16 warps with fixed roles; three tiles are in flight:
//   warp 1      MMA issuer (one lane): layer 1 of tile i is issued before layer 2 of tile i-1
//   warps 0, 2-7  gather + convert (warp 2 also owns the TMEM allocation): batches of 32 rows, round-robin over
//               the seven warps, are fetched with coalesced loads (one row per load, a column pair per lane),
//               split into bf16 hi/lo and stored straight into the X image; the warp that handles the first
//               batch of a tile also writes the per-month bias column of W1
//   warps 8-11  epilogue of even tiles   } z1 -> relu -> split -> H1 image ; tanh(q) ;
//   warps 12-15 epilogue of odd tiles    } z2 -> relu -> .W3 -> u
#include "cell_tc.cuh"
#include "macro_role.cuh"

namespace dlap {
using namespace tcl;

namespace {
constexpr int NT_TC = 512;
constexpr int BR = 32;                                   // rows per gather batch (64 registers in flight per lane)
constexpr int NB = TILE / BR;                            // batches per tile
constexpr int NGW = 7;                                   // gather warps: 0, 2 (also TMEM owner), 3, 4..7
constexpr int NXS = 3;                                   // slots of X image and W1 image
constexpr int NDA = 4;                                   // layer-1 accumulator (then h1) slots in TMEM
struct SmemTc {
  uint8_t* base;
  __device__ uint8_t* xh(int sl) const { return base + sl * 2 * X_IMG; }
  __device__ uint8_t* xl(int sl) const { return xh(sl) + X_IMG; }
  __device__ uint8_t* w1h(int sl) const { return base + NXS * 2 * X_IMG + sl * 2 * W1_IMG; }
  __device__ uint8_t* w1l(int sl) const { return w1h(sl) + W1_IMG; }
  __device__ uint8_t* w2h() const { return base + NXS * 2 * (X_IMG + W1_IMG); }
  __device__ uint8_t* w2l() const { return w2h() + W2_IMG; }
  __device__ float* b2s() const { return reinterpret_cast<float*>(w2l() + W2_IMG); }
  __device__ float* w3s() const { return b2s() + 64; }
  __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(w3s() + 64); }
  __device__ uint32_t* tmem_slot() const { return reinterpret_cast<uint32_t*>(bars() + 24); }
};
enum { F_XOP_FULL = 0, F_XOP_EMPTY = 3, F_D1_FULL = 6, F_H1_FULL = 10,
       F_D2_FULL = 14, F_D2_EMPTY = 16 };

constexpr size_t SMEM_TC_BYTES = (size_t)NXS * 2 * (X_IMG + W1_IMG) + 2 * W2_IMG +
                                 128 * 4 + 24 * 8 + 16 + 256;

// TMEM columns: four layer-1 accumulators (layer 1 runs three tiles ahead of layer 2, and each epilogue group
// has two tiles in flight), two layer-2 ones.
// After E1 the first 64 columns of a layer-1 slot hold h1 as packed bf16 pairs, in two blocks of 32 units:
// columns [32b, 32b+16) = hi of units 32b .. 32b+31, [32b+16, 32b+32) = lo  (each block is rewritten in place
// by the thread that just read it).  K step ks (units 16ks ..) of layer 2 reads 8 columns of each.
constexpr uint32_t COL_A = 0, COL_A_STRIDE = 96, COL_B0 = 384, COL_B1 = 448, TMEM_COLS = 512;
__device__ __forceinline__ uint32_t h1_col_hi(int ks) { return (uint32_t)((ks >> 1) * 32 + (ks & 1) * 8); }
__device__ __forceinline__ uint32_t h1_col_lo(int ks) { return h1_col_hi(ks) + 16; }
}  // namespace

// diagnostics: stamp slot k of local tile i (CTA 0 only, one thread per role)
#define DLAP_STAMP(i, k)                                                                    \
  do {                                                                                      \
    if (P.timeline && blockIdx.x == 0 && (i) < 256) P.timeline[(i) * 16 + (k)] = clock64(); \
  } while (0)

// every CTA's last act when the macro role runs in this grid: the last one out zeroes both words of `sync`
// (sync[0]: months published so far, sync[1]: CTAs that have left)
__device__ __forceinline__ void fused_exit(int* sync) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(sync + 1, 1) == (int)gridDim.x - 1) {
      atomicExch(sync, 0);
      atomicExch(sync + 1, 0);
    }
  }
}

template <bool MOM, bool DROP, bool FUSED>
__global__ void __launch_bounds__(NT_TC, 1) cell_fwd_tc_kernel(const CellParams P, const MacroParams MP) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  static_assert(MT == NT_TC, "the macro role is written for the per-cell kernel's CTA size");
  if (FUSED && blockIdx.x == gridDim.x - 1) {
    macro_fwd_role<4>(MP, reinterpret_cast<float*>(smem_raw));
    fused_exit(MP.months_ready);
    return;
  }
  const int ncta = FUSED ? (int)gridDim.x - 1 : (int)gridDim.x;     // CTAs that work on cells
  SmemTc s;
  s.base = tc::smem_align128(smem_raw);
  uint64_t* bar = s.bars();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = __ldg(P.index);                       // tiles of valid cells (dlap_panel_index)
  // A contiguous tile range per CTA as a rule.  When the LSTM recurrence runs beside this kernel (months_ready),
  // blocks of four tiles are dealt round-robin instead: every CTA then needs the months in the order they are made.
  const TileMap tm = P.months_ready ? TileMap::cyclic(ntiles, blockIdx.x, ncta, P.cyc_sh)
                                    : TileMap::contiguous(ntiles, blockIdx.x, ncta);
  const int my_tiles = tm.count;

  // ---- one-time set-up ----
  for (int sl = 0; sl < NXS; ++sl) fill_w1_images(P, s.w1h(sl), s.w1l(sl), tid, NT_TC);
  fill_w2_images(P, s.w2h(), s.w2l(), tid, NT_TC);
  for (int sl = 0; sl < NXS; ++sl) init_x_images(s.xh(sl), s.xl(sl), tid, NT_TC);
  if (tid < 64) { s.b2s()[tid] = __ldg(P.b2 + tid); s.w3s()[tid] = __ldg(P.W3 + tid); }
  if (tid == 0) {
    for (int i = 0; i < NXS; ++i) { mbar_init(bar + F_XOP_FULL + i, NB); mbar_init(bar + F_XOP_EMPTY + i, 1); }
    for (int i = 0; i < NDA; ++i) { mbar_init(bar + F_D1_FULL + i, 1); mbar_init(bar + F_H1_FULL + i, 4); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar + F_D2_FULL + i, 1); mbar_init(bar + F_D2_EMPTY + i, 4); }
    fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s.tmem_slot(), TMEM_COLS);
  fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = *s.tmem_slot();

  if (warp == 1) {
    // ===== MMA issuer (all lanes wait, lane 0 issues) =====
    {
      const uint32_t id1 = tc::idesc_bf16(128, N1, 0, 0), id2 = tc::idesc_bf16(128, 64, 0, 0);
      const uint64_t w2h = tc::smem_desc(smem_u32(s.w2h()), 64 * 16, 128), w2l = tc::smem_desc(smem_u32(s.w2l()), 64 * 16, 128);
      // Layer 1 runs three tiles ahead of layer 2 (four accumulator slots, three X / W1 slots): L1(0) L1(1) L1(2)
      // | L2(0) L1(3) | L2(1) L1(4) | ...  An epilogue group starts E1 of its next tile while layer 2 of its
      // current one runs, and a late gather delays only layer 1, never layer 2.
      auto issue_l1 = [&](int i) {
        const int xs = i % NXS, use = i / NXS;
        mbar_wait(bar + F_XOP_FULL + xs, use & 1);
        tc::fence_after_sync();
        if (lane == 0) DLAP_STAMP(i, 8);
        const uint64_t ah = tc::smem_desc(smem_u32(s.xh(xs)), X_LBO, 128), al = tc::smem_desc(smem_u32(s.xl(xs)), X_LBO, 128);
        const uint64_t bh = tc::smem_desc(smem_u32(s.w1h(xs)), N1 * 16, 128), bl = tc::smem_desc(smem_u32(s.w1l(xs)), N1 * 16, 128);
        const uint32_t d = tbase + COL_A + (i % NDA) * COL_A_STRIDE;
        if (tc::elect_one()) {
#pragma unroll
          for (int ks = 0; ks < KG1 / 2; ++ks)
            umma3(d, tc::desc_advance(ah, ks * 2 * X_LBO), tc::desc_advance(al, ks * 2 * X_LBO),
                  tc::desc_advance(bh, ks * 2 * N1 * 16), tc::desc_advance(bl, ks * 2 * N1 * 16), id1, ks == 0);
          tc::umma_commit(bar + F_XOP_EMPTY + xs);
          tc::umma_commit(bar + F_D1_FULL + i % NDA);
        }
        __syncwarp();
        if (lane == 0) DLAP_STAMP(i, 9);
      };
      for (int j = 0; j < 3 && j < my_tiles; ++j) issue_l1(j);
      for (int j = 0; j < my_tiles; ++j) {           // layer 2 of tile j, then layer 1 of tile j + 3
        const int sl = j & 1, use = j >> 1;
        mbar_wait(bar + F_H1_FULL + j % NDA, (j / NDA) & 1);
        mbar_wait(bar + F_D2_EMPTY + sl, (use & 1) ^ 1);
        tc::fence_after_sync();
        if (lane == 0) DLAP_STAMP(j, 10);
        const uint32_t a = tbase + COL_A + (j % NDA) * COL_A_STRIDE;     // h1 of tile j, in its z1 slot
        const uint32_t d = tbase + (sl ? COL_B1 : COL_B0);
        if (tc::elect_one()) {
#pragma unroll
          for (int ks = 0; ks < KG2 / 2; ++ks) {
            const uint64_t bh = tc::desc_advance(w2h, ks * 2 * 64 * 16), bl = tc::desc_advance(w2l, ks * 2 * 64 * 16);
            tc::umma_bf16_ts(d, a + h1_col_hi(ks), bh, id2, ks == 0 ? 0u : 1u);
            tc::umma_bf16_ts(d, a + h1_col_hi(ks), bl, id2, 1u);
            tc::umma_bf16_ts(d, a + h1_col_lo(ks), bh, id2, 1u);
          }
          tc::umma_commit(bar + F_D2_FULL + sl);
        }
        __syncwarp();
        if (lane == 0) DLAP_STAMP(j, 11);
        if (j + 3 < my_tiles) issue_l1(j + 3);
      }
    }
  } else if (warp < 8) {
    // ===== gather + convert (cell_tc.cuh); stamps 0-3 of timeline row q / NGW: loop top, slot free, converted,
    // loads issued
    const int gw = warp == 0 ? 0 : warp - 1;
    long long* tl = (P.timeline && blockIdx.x == 0 && warp == 4 && lane == 0) ? P.timeline : nullptr;
    if (P.ximg) {                   // the panel's X images exist: one warp feeds the slots with bulk copies
      if (warp == 0)
        image_role<NB, NXS, MOM>(P, lane, tm, s.xh(0), s.w1h(0) + (KG1 - 1) * N1 * 16, 2 * W1_IMG, W1_IMG,
                                 bar + F_XOP_FULL, bar + F_XOP_EMPTY);
    } else if (gather_wide(P))
      gather_role<BR, NGW, NXS, true, MOM>(P, gw, lane, tm, s.xh(0), s.w1h(0) + (KG1 - 1) * N1 * 16,
                                         2 * W1_IMG, W1_IMG, bar + F_XOP_FULL,
                                      bar + F_XOP_EMPTY, tl, 0, true);
    else
      gather_role<BR, NGW, NXS, false, MOM>(P, gw, lane, tm, s.xh(0), s.w1h(0) + (KG1 - 1) * N1 * 16,
                                          2 * W1_IMG, W1_IMG, bar + F_XOP_FULL,
                                       bar + F_XOP_EMPTY, tl, 0, true);
  } else if (warp >= 8) {
    // ===== epilogue: group e handles tiles with (i & 1) == e ; thread = TMEM lane = cell =====
    const int e = (warp - 8) >> 2, q4 = warp & 3, cell = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    const float b3 = __ldg(P.b3);
    MonthCursor mc;
    if (e < my_tiles) mc.init(P, tm.tile(e));
    int n_nxt = -1, t_nxt = 0, pos_nxt = 0;
    bool live_nxt = false;
    auto load_row = [&](int tile) {
      mc.seek(P, tile);
      t_nxt = mc.t;
      pos_nxt = mc.j0(tile) + cell;               // position in the month's list: where u, q of this cell go
      live_nxt = cell < mc.ncell(tile);
      if (DROP) n_nxt = live_nxt ? __ldg(vidx_row(P, mc.t) + pos_nxt) : -1;     // the dropout key is the DENSE cell number
    };
    if (e < my_tiles) load_row(tm.tile(e));
    // per-tile state that E1 hands to E2 (a group has two tiles in flight: E1 of its next tile runs while layer 2
    // of the current one is on the tensor core)
    struct TileState { size_t pi; uint32_t gi; bool live; };
    // E1 of local tile i: h1 = relu(z1) -> TMEM, and the moment outputs
    auto do_e1 = [&](int i) {
      TileState S;
      const int t = t_nxt, n = n_nxt;
      S.pi = (size_t)t * P.N + pos_nxt;
      const uint32_t dA = tbase + lane_base + COL_A + (i % NDA) * COL_A_STRIDE;
      S.live = live_nxt;
      S.gi = (uint32_t)((size_t)t * P.N + (DROP && S.live ? n : 0));
      if (i + 2 < my_tiles) load_row(tm.tile(i + 2));
      const bool stamp = q4 == 0 && lane == 0;
      mbar_wait(bar + F_D1_FULL + i % NDA, (i / NDA) & 1);
      tc::fence_after_sync();
      if (stamp) DLAP_STAMP(i, 12);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tc::tmem_ld32(dA + c * 32, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) v[k] = fmaxf(v[k], 0.f);
        if (DROP) tc_dropout<32>(P, S.gi, 0, c * 32, v);
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) tc::split2(v[2 * k], v[2 * k + 1], hi[k], lo[k]);
        tc::tmem_st16(dA + c * 32, hi);              // over the z1 columns this thread has just read
        tc::tmem_st16(dA + c * 32 + 16, lo);
      }
      float q[8];
      if (MOM) {
        tc::tmem_ld8(dA + 64, q);
        tc::tmem_wait_ld();
      }
      tc::tmem_wait_st();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + F_H1_FULL + i % NDA);
      if (stamp) DLAP_STAMP(i, 13);
      if (MOM && S.live) {
        const size_t plane = (size_t)P.T * P.N;
#pragma unroll
        for (int k = 0; k < 8; ++k) P.q[(size_t)k * plane + S.pi] = tc::tanh_fast(q[k]);
      }
      return S;
    };
    TileState cur{}, nxt{};
    if (e < my_tiles) cur = do_e1(e);
    for (int i = e; i < my_tiles; i += 2) {
      const int tile = tm.tile(i), sl = i & 1, use = i >> 1;
      const bool stamp = q4 == 0 && lane == 0;
      const uint32_t dB = tbase + lane_base + (sl ? COL_B1 : COL_B0);
      if (i + 2 < my_tiles) nxt = do_e1(i + 2);      // while layer 2 of tile i runs
      const bool live = cur.live;
      const size_t pi = cur.pi;
      const uint32_t gi = cur.gi;
      // -- after layer 2: u = b3 + W3 . relu(z2 + b2)
      mbar_wait(bar + F_D2_FULL + sl, use & 1);
      tc::fence_after_sync();
      if (stamp) DLAP_STAMP(i, 14);
      float up[4] = {b3, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tc::tmem_ld32(dB + c * 32, v);
        const float4* b4 = reinterpret_cast<const float4*>(s.b2s() + c * 32);
        const float4* w4 = reinterpret_cast<const float4*>(s.w3s() + c * 32);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 bb = b4[k];
          v[4 * k] = fmaxf(v[4 * k] + bb.x, 0.f);         v[4 * k + 1] = fmaxf(v[4 * k + 1] + bb.y, 0.f);
          v[4 * k + 2] = fmaxf(v[4 * k + 2] + bb.z, 0.f); v[4 * k + 3] = fmaxf(v[4 * k + 3] + bb.w, 0.f);
        }
        if (DROP) tc_dropout<32>(P, gi, 1, c * 32, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 ww = w4[k];
          up[0] = fmaf(v[4 * k], ww.x, up[0]);     up[1] = fmaf(v[4 * k + 1], ww.y, up[1]);
          up[2] = fmaf(v[4 * k + 2], ww.z, up[2]); up[3] = fmaf(v[4 * k + 3], ww.w, up[3]);
        }
      }
      const float u = (up[0] + up[1]) + (up[2] + up[3]);
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + F_D2_EMPTY + sl);
      if (live) P.u[pi] = u;
      if (stamp) DLAP_STAMP(i, 15);
      cur = nxt;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tbase, TMEM_COLS);
  if (FUSED) fused_exit(MP.months_ready);
}

template <bool MOM, bool DROP, bool FUSED>
static cudaError_t launch_fwd_tc_variant(const CellParams& P, const MacroParams& MP, int grid, bool cooperative,
                                         cudaStream_t st) {
  size_t smem = SMEM_TC_BYTES;
  if (FUSED) {
    const size_t need = macro_fwd_smem_bytes(MP.T, MP.M, 4) + 128;
    if (need > smem) smem = need;
  }
  auto kern = cell_fwd_tc_kernel<MOM, DROP, FUSED>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  if (cooperative) {
    void* args[] = {(void*)&P, (void*)&MP};
    return cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(NT_TC), args, smem, st);
  }
  kern<<<grid, NT_TC, smem, st>>>(P, MP);
  return cudaGetLastError();
}

static long long* g_timeline_fwd = nullptr;
void set_fwd_tc_timeline(long long* p) { g_timeline_fwd = p; }

// grid: CTAs that work on cells.  MP0 non-null: one more CTA runs the macro path beside them (H = 4 only; the caller
// has put MP0->months_ready into P0.months_ready as well), and the grid is launched co-resident when `cooperative`.
cudaError_t launch_cell_fwd_tc(const CellParams& P0, const MacroParams* MP0, int grid, bool cooperative,
                               cudaStream_t st) {
  CellParams P = P0;
  P.timeline = g_timeline_fwd;
  const bool drop = P.drop_thr != 0, mom = P.q != nullptr;
  if (MP0) {
    const MacroParams& MP = *MP0;
    if (mom) return drop ? launch_fwd_tc_variant<true, true, true>(P, MP, grid + 1, cooperative, st)
                         : launch_fwd_tc_variant<true, false, true>(P, MP, grid + 1, cooperative, st);
    return drop ? launch_fwd_tc_variant<false, true, true>(P, MP, grid + 1, cooperative, st)
                : launch_fwd_tc_variant<false, false, true>(P, MP, grid + 1, cooperative, st);
  }
  MacroParams MP{};
  if (mom) return drop ? launch_fwd_tc_variant<true, true, false>(P, MP, grid, false, st)
                       : launch_fwd_tc_variant<true, false, false>(P, MP, grid, false, st);
  return drop ? launch_fwd_tc_variant<false, true, false>(P, MP, grid, false, st)
              : launch_fwd_tc_variant<false, false, false>(P, MP, grid, false, st);
}

}  // namespace dlap
