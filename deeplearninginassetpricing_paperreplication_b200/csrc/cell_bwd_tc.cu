// Backward per-cell kernel on the tensor cores (tcgen05 + TMEM), precision "bf16x3", with activation
// recompute.  Same mathematics as cell_bwd_f32_kernel (cell_bwd.cu).
//
// Per tile of 128 stock-months, on split bf16 operands (3 MMAs per fp32 product):
//   L1    D_A[128,80] = X.[W1x;Wqx]^T (+ a_t|c_t through the ones column)      -> E1: h1, relu mask, dpre
//   L2    D_B[128,64] = H1.W2^T                                                -> E2: dz2, dW3, db2
//   WG2   D_W2[64,64] += DZ2^T.H1       (dW2)
//   DG    D_C[128,64] = DZ2.W2                                                 -> E3: dz1, da_t
//   WG1   D_W1[64,72] += X^T.[DZ1|DPRE]  (dW1x^T and dWq^T in one MMA: dz1 and dpre are adjacent images)
// All activation tiles live in shared memory as [group of 8 units][cell][8] bf16 images, so the same
// bytes serve as K-major operand (forward / dgrad) and MN-major operand (weight gradients).
//
// Tiles are the valid-cell tiles of dlap_panel_index (<= 128 unmasked stocks of one month, gathered row
// by row); masked cells contribute exactly 0 to every gradient and are never visited.
//
// Per-month sums.  X carries a ones column (k = 46), so row 46 of the dW1 accumulator is sum_cells [dz1 | dpre]
// = da_t | dc_t of the tiles accumulated so far: nothing has to be summed over cells by the CUDA cores.  A CTA
// owns a CONTIGUOUS range of tiles, so a month's tiles are consecutive; when the month changes, the epilogue
// threads that own that TMEM lane wait for the month's last dW1 MMA, copy the row out as this CTA's partial
// of the month and write zeros back.  dW3 and db2 (sums over all tiles of the CTA) are reduced over the 32
// cells of a warp tile by tile with shuffles and kept as one value per lane.
//
// Scheduling.  A CTA owns a CONTIGUOUS range of tiles, so the tiles of one month are consecutive: the
// per-month sums (da_t, dc_t) and the per-CTA sums (dW3, db2) are accumulated in registers per cell and
// reduced across cells only when the month changes / at the end.  H1 is double buffered and dz1 is
// written over the (dead) H1 of its own tile, so the epilogue of tile i+1's layer 1 runs while the
// tensor core works on dW2/dgrad of tile i, and layer 2 of tile i+1 runs under E3 of tile i.
//
// 24 warps (80 registers per thread): 1 MMA issuer; 0 (also TMEM owner), 2, 3, 20-23 gather + convert (batches
// of 16 rows are fetched with coalesced loads, one row per load and a column pair per lane, split into bf16
// hi/lo and stored straight into the X image; batches go round-robin over the seven warps); 4-19 epilogue (four warps per TMEM lane quarter; each thread handles 16 of the 64 hidden units of its cell).
#include "cell_tc.cuh"

namespace dlap {
using namespace tcl;

namespace {
constexpr int NT_B = 768;
constexpr int NEPI = 16;                               // epilogue warps
constexpr int BR = 16;                                   // rows per gather batch (32 registers in flight per lane)
constexpr int NB = TILE / BR;                            // batches per tile
constexpr int NGW = 7;                                   // gather warps (0, 2, 3, 20..23)
constexpr int NXS = 3;                                   // X image slots
constexpr uint32_t G5_BLK = N1 * 16;                     // one k-group of a W1 image: [N1][8] bf16
constexpr uint32_t W1_STATIC = (KG1 - 1) * G5_BLK;       // k-groups 0..4
constexpr uint32_t DPS_IMG = TILE * 16;                // one group: [cell][8 moments]

struct SmemB {
  uint8_t* base;
  // X images: NXS slots, so the gather warps run a full tile ahead of the last reader (dW1) of a slot
  __device__ uint8_t* xh(int sl) const { return base + sl * 2 * X_IMG; }
  __device__ uint8_t* xl(int sl) const { return xh(sl) + X_IMG; }
  // W1 image: k-groups 0..4 once (static); k-group 5, which carries the per-month bias column, once per X slot.
  // The MMA descriptor of the last K step reaches its second group through the leading-dimension offset.
  __device__ uint8_t* w1h() const { return base + NXS * 2 * X_IMG; }
  __device__ uint8_t* w1l() const { return w1h() + W1_STATIC; }
  __device__ uint8_t* g5h(int sl) const { return w1l() + W1_STATIC + sl * G5_BLK; }
  __device__ uint8_t* g5l(int sl) const { return g5h(sl) + NXS * G5_BLK; }
  __device__ uint8_t* w2h() const { return base + NXS * 2 * X_IMG + 2 * W1_STATIC + 2 * NXS * G5_BLK; }
  __device__ uint8_t* w2l() const { return w2h() + W2_IMG; }
  // per slot: [H hi (8 groups) | DPRE hi (1 group) | H lo | DPRE lo]: dz1 (written over h1) and dpre are
  // adjacent, so one MN-major B operand of 9 groups (N = 72) feeds dW1x and dWq in the same MMA
  __device__ uint8_t* hh(int sl) const { return w2l() + W2_IMG + sl * 2 * (H_IMG + DPS_IMG); }
  __device__ uint8_t* dph(int sl) const { return hh(sl) + H_IMG; }
  __device__ uint8_t* hl(int sl) const { return dph(sl) + DPS_IMG; }
  __device__ uint8_t* dpl(int sl) const { return hl(sl) + H_IMG; }
  __device__ uint8_t* dzh() const { return w2l() + W2_IMG + 4 * (H_IMG + DPS_IMG); }
  __device__ uint8_t* dzl() const { return dzh() + H_IMG; }
  __device__ float* b2s() const { return reinterpret_cast<float*>(dzl() + H_IMG); }
  __device__ float* w3s() const { return b2s() + 64; }
  __device__ float* red() const { return w3s() + 64; }                       // [16 warps][16][2] + [4][8]
  __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(red() + NEPI * 32 + 32); }
  __device__ uint32_t* tmem_slot() const { return reinterpret_cast<uint32_t*>(bars() + 24); }
};
constexpr size_t SMEM_B_BYTES = (size_t)NXS * 2 * X_IMG + 2 * W1_STATIC + 2 * NXS * G5_BLK + 2 * W2_IMG + 6 * H_IMG + 4 * DPS_IMG +
                                128 * 4 + (NEPI * 32 + 32) * 4 + 24 * 8 + 16 + 128;

enum { B_XOP_FULL = 0, B_XOP_EMPTY = 3, B_D1_FULL = 6, B_H1_FULL = 8,
       B_W1_DONE = 10, B_D2_FULL = 12, B_DZ2_FULL = 13, B_D3_FULL = 14, B_DZ1_FULL = 15, B_ALL_DONE = 16 };

constexpr uint32_t C_A0 = 0, C_A1 = 96, C_B = 192, C_C = 256, C_W2 = 320, C_W1 = 384, TMEM_COLS = 512;   // C_W1: 72 cols

// Sum over the 32 lanes of 16 per-lane values: every lane L returns sum_lanes v[L & 15].
__device__ __forceinline__ float lane_transpose_sum16(float (&v)[16], int lane) {
#pragma unroll
  for (int s = 8; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 512;" ::: "memory"); }
}  // namespace

// diagnostics: stamp slot k of local tile i (CTA 0 only, one thread per role)
#define DLAP_STAMP(i, k)                                                                    \
  do {                                                                                      \
    if (P.timeline && blockIdx.x == 0 && (i) < 256) P.timeline[(i) * 16 + (k)] = clock64(); \
  } while (0)

// RES: the residual-loss term of du (CellParams::res_b), a variant of its own so that the kernel everybody else runs
// keeps its register allocation (80 registers, no spills)
template <bool SDF, bool MOM, bool DROP, bool RES = false>
__global__ void __launch_bounds__(NT_B, 1) cell_bwd_tc_kernel(const CellParams P) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  SmemB s;
  s.base = tc::smem_align128(smem_raw);
  uint64_t* bar = s.bars();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = __ldg(P.index);                       // tiles of valid cells (dlap_panel_index)
  const int chunk = tc_chunk(ntiles, gridDim.x);
  const int tile_lo = min(ntiles, (int)blockIdx.x * chunk);
  const int my_tiles = min(ntiles, tile_lo + chunk) - tile_lo;
  const int Fr = round4(P.F);

  // ---- one-time set-up ----
  if (tid == 0) DLAP_STAMP(255, 0);                          // diagnostics: kernel entry
  const long long t_entry = (P.timeline && tid == 0) ? clock64() : 0;
  fill_w1_groups(P, s.w1h(), s.w1l(), 0, KG1 - 1, tid, NT_B);
  for (int sl = 0; sl < NXS; ++sl) fill_w1_groups(P, s.g5h(sl), s.g5l(sl), KG1 - 1, 1, tid, NT_B);
  fill_w2_images(P, s.w2h(), s.w2l(), tid, NT_B);
  for (int sl = 0; sl < NXS; ++sl) init_x_images(s.xh(sl), s.xl(sl), tid, NT_B);
  for (int sl = 0; sl < 2; ++sl)
    for (int i = tid; i < (int)DPS_IMG / 4; i += NT_B) {
      reinterpret_cast<uint32_t*>(s.dph(sl))[i] = 0u;
      reinterpret_cast<uint32_t*>(s.dpl(sl))[i] = 0u;
    }
  if (tid < 64) { s.b2s()[tid] = __ldg(P.b2 + tid); s.w3s()[tid] = __ldg(P.W3 + tid); }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar + B_D1_FULL + i, 1);     mbar_init(bar + B_H1_FULL + i, NEPI);
      mbar_init(bar + B_W1_DONE + i, 1);
    }
    for (int i = 0; i < NXS; ++i) { mbar_init(bar + B_XOP_FULL + i, NB); mbar_init(bar + B_XOP_EMPTY + i, 1); }
    mbar_init(bar + B_D2_FULL, 1);  mbar_init(bar + B_DZ2_FULL, NEPI);
    mbar_init(bar + B_D3_FULL, 1);  mbar_init(bar + B_DZ1_FULL, NEPI);
    mbar_init(bar + B_ALL_DONE, 1);
    fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(s.tmem_slot(), TMEM_COLS);
  fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = *s.tmem_slot();
  if (tid == 0) DLAP_STAMP(255, 1);                          // set-up done

  if (warp == 1) {
    // ===== MMA issuer (all lanes wait, lane 0 issues) =====
    const uint32_t id_l1 = tc::idesc_bf16(128, N1, 0, 0), id_l2 = tc::idesc_bf16(128, 64, 0, 0);
    const uint32_t id_dg = tc::idesc_bf16(128, 64, 0, 1);                 // A K-major, B = W2 MN-major
    const uint32_t id_w2 = tc::idesc_bf16(64, 64, 1, 1);
    const uint32_t id_w1q = tc::idesc_bf16(64, 72, 1, 1);                 // A = X (chars), B = [dz1 | dpre]
    const uint32_t KM_A = TILE * 16;                                      // K-major activation: LBO
    const uint64_t dzh_k = tc::smem_desc(smem_u32(s.dzh()), KM_A, 128), dzl_k = tc::smem_desc(smem_u32(s.dzl()), KM_A, 128);
    const uint64_t dzh_m = tc::smem_desc(smem_u32(s.dzh()), 128, KM_A), dzl_m = tc::smem_desc(smem_u32(s.dzl()), 128, KM_A);
    const uint64_t w2h_k = tc::smem_desc(smem_u32(s.w2h()), 64 * 16, 128), w2l_k = tc::smem_desc(smem_u32(s.w2l()), 64 * 16, 128);
    const uint64_t w2h_m = tc::smem_desc(smem_u32(s.w2h()), 128, 64 * 16), w2l_m = tc::smem_desc(smem_u32(s.w2l()), 128, 64 * 16);

    auto issue_l1 = [&](int i) {
      const int sl = i & 1, xs = i % NXS, xuse = i / NXS;
      mbar_wait(bar + B_XOP_FULL + xs, xuse & 1);
      tc::fence_after_sync();
      if (tc::elect_one()) {
        const uint64_t ah = tc::smem_desc(smem_u32(s.xh(xs)), X_LBO, 128), al = tc::smem_desc(smem_u32(s.xl(xs)), X_LBO, 128);
        const uint64_t bh = tc::smem_desc(smem_u32(s.w1h()), G5_BLK, 128), bl = tc::smem_desc(smem_u32(s.w1l()), G5_BLK, 128);
        // last K step: k-group 4 is static, k-group 5 (bias column of this slot) sits g5 - g4 bytes further
        const uint32_t g4h = smem_u32(s.w1h()) + (KG1 - 2) * G5_BLK, g4l = smem_u32(s.w1l()) + (KG1 - 2) * G5_BLK;
        const uint64_t bh5 = tc::smem_desc(g4h, smem_u32(s.g5h(xs)) - g4h, 128);
        const uint64_t bl5 = tc::smem_desc(g4l, smem_u32(s.g5l(xs)) - g4l, 128);
        const uint32_t d = tbase + (sl ? C_A1 : C_A0);
#pragma unroll
        for (int ks = 0; ks < KG1 / 2; ++ks) {
          const bool last = ks == KG1 / 2 - 1;
          umma3(d, tc::desc_advance(ah, ks * 2 * X_LBO), tc::desc_advance(al, ks * 2 * X_LBO),
                last ? bh5 : tc::desc_advance(bh, ks * 2 * G5_BLK), last ? bl5 : tc::desc_advance(bl, ks * 2 * G5_BLK),
                id_l1, ks == 0);
        }
        tc::umma_commit(bar + B_D1_FULL + sl);
      }
      __syncwarp();
    };
    // layer 2 + moment weight gradient of tile i (after its E1)
    auto issue_l2 = [&](int i) {
      const int sl = i & 1, use = i >> 1;
      mbar_wait(bar + B_H1_FULL + sl, use & 1);
      tc::fence_after_sync();
      if (tc::elect_one()) {
        if (SDF) {
          const uint64_t hh_k = tc::smem_desc(smem_u32(s.hh(sl)), KM_A, 128), hl_k = tc::smem_desc(smem_u32(s.hl(sl)), KM_A, 128);
#pragma unroll
          for (int ks = 0; ks < KG2 / 2; ++ks)
            umma3(tbase + C_B, tc::desc_advance(hh_k, ks * 2 * KM_A), tc::desc_advance(hl_k, ks * 2 * KM_A),
                  tc::desc_advance(w2h_k, ks * 2 * 64 * 16), tc::desc_advance(w2l_k, ks * 2 * 64 * 16), id_l2, ks == 0);
          tc::umma_commit(bar + B_D2_FULL);
        }
        if (MOM && !SDF) {   // moment-only step: nothing else will read X(i); dWq^T = X^T.[.. | dpre] right away
          const uint64_t xh_m = tc::smem_desc(smem_u32(s.xh(i % NXS)), 128, X_LBO), xl_m = tc::smem_desc(smem_u32(s.xl(i % NXS)), 128, X_LBO);
          const uint64_t bh_m = tc::smem_desc(smem_u32(s.hh(sl)), 128, KM_A), bl_m = tc::smem_desc(smem_u32(s.hl(sl)), 128, KM_A);
#pragma unroll
          for (int ks = 0; ks < TILE / 16; ++ks)
            umma3(tbase + C_W1, tc::desc_advance(xh_m, ks * 256), tc::desc_advance(xl_m, ks * 256),
                  tc::desc_advance(bh_m, ks * 256), tc::desc_advance(bl_m, ks * 256), id_w1q, i == 0 && ks == 0);
        }
      }
      __syncwarp();
    };

    if (my_tiles > 0) { issue_l1(0); issue_l2(0); }
    if (my_tiles > 1) issue_l1(1);
    for (int i = 0; i < my_tiles; ++i) {
      const int sl = i & 1, par = i & 1;
      const bool first = (i == 0);
      const uint64_t xh_m = tc::smem_desc(smem_u32(s.xh(i % NXS)), 128, X_LBO), xl_m = tc::smem_desc(smem_u32(s.xl(i % NXS)), 128, X_LBO);
      const uint64_t hh_m = tc::smem_desc(smem_u32(s.hh(sl)), 128, KM_A), hl_m = tc::smem_desc(smem_u32(s.hl(sl)), 128, KM_A);
      if (SDF) {
        // -- after E2(i): dW2 and dgrad
        mbar_wait(bar + B_DZ2_FULL, par);
        tc::fence_after_sync();
        if (lane == 0) DLAP_STAMP(i, 8);
        if (tc::elect_one()) {
#pragma unroll
          for (int ks = 0; ks < TILE / 16; ++ks)
            umma3(tbase + C_W2, tc::desc_advance(dzh_m, ks * 256), tc::desc_advance(dzl_m, ks * 256),
                  tc::desc_advance(hh_m, ks * 256), tc::desc_advance(hl_m, ks * 256), id_w2, first && ks == 0);
#pragma unroll
          for (int ks = 0; ks < KG2 / 2; ++ks)
            umma3(tbase + C_C, tc::desc_advance(dzh_k, ks * 2 * KM_A), tc::desc_advance(dzl_k, ks * 2 * KM_A),
                  tc::desc_advance(w2h_m, ks * 256), tc::desc_advance(w2l_m, ks * 256), id_dg, ks == 0);
          tc::umma_commit(bar + B_D3_FULL);
        }
        __syncwarp();
        if (lane == 0) DLAP_STAMP(i, 9);
      }
      if (!SDF) {              // moment-only: the one reader of X(i) / H(i), dWq(i), was issued by issue_l2(i); the
        if (tc::elect_one()) { // epilogue may wait for it (month change) before it lets issue_l2(i + 1) go
          tc::umma_commit(bar + B_W1_DONE + sl);
          tc::umma_commit(bar + B_XOP_EMPTY + i % NXS);
        }
        __syncwarp();
      }
      if (i + 1 < my_tiles) issue_l2(i + 1);            // runs under E3(i)
      if (lane == 0) DLAP_STAMP(i, 10);
      if (SDF) {
        // -- after E3(i): [dW1x | dWq]^T = X^T . [dz1 | dpre]; dz1 sits in the H buffer of this tile, dpre next to it
        mbar_wait(bar + B_DZ1_FULL, par);
        tc::fence_after_sync();
        if (lane == 0) DLAP_STAMP(i, 11);
        if (tc::elect_one()) {
#pragma unroll
          for (int ks = 0; ks < TILE / 16; ++ks)
            umma3(tbase + C_W1, tc::desc_advance(xh_m, ks * 256), tc::desc_advance(xl_m, ks * 256),
                  tc::desc_advance(hh_m, ks * 256), tc::desc_advance(hl_m, ks * 256), id_w1q, first && ks == 0);
        }
        __syncwarp();
        if (tc::elect_one()) {                                 // every reader of X(i) / H(i) has been issued
          tc::umma_commit(bar + B_W1_DONE + sl);
          tc::umma_commit(bar + B_XOP_EMPTY + i % NXS);
        }
        __syncwarp();
      }
      if (i + 2 < my_tiles) issue_l1(i + 2);
      if (lane == 0) DLAP_STAMP(i, 12);
    }
    if (tc::elect_one()) tc::umma_commit(bar + B_ALL_DONE);
    __syncwarp();
  } else if (warp == 0 || warp == 2 || warp == 3 || warp >= 20) {
    // ===== gather + convert (cell_tc.cuh); stamps 4-7 of row q / NGW: loop top, slot free, converted, loads issued
    const int gw = warp == 0 ? 0 : (warp < 4 ? warp - 1 : warp - 17);
    long long* tl = (P.timeline && blockIdx.x == 0 && gw == 1 && lane == 0) ? P.timeline + 256 * 16 * 2 : nullptr;
    if (P.ximg) {                   // the panel's X images exist (dlap_panel_image): one warp, one bulk copy per tile
      if (warp == 0)
        image_role<NB, NXS, true>(P, lane, TileMap{tile_lo, 31, 0, my_tiles}, s.xh(0), s.g5h(0), G5_BLK, NXS * G5_BLK,
                                  bar + B_XOP_FULL, bar + B_XOP_EMPTY);
    } else if (gather_wide(P))
      gather_role<BR, NGW, NXS, true, true>(P, gw, lane, TileMap{tile_lo, 31, 0, my_tiles}, s.xh(0), s.g5h(0), G5_BLK, NXS * G5_BLK, bar + B_XOP_FULL,
                                       bar + B_XOP_EMPTY, tl, 0, true);
    else
      gather_role<BR, NGW, NXS, false, true>(P, gw, lane, TileMap{tile_lo, 31, 0, my_tiles}, s.xh(0), s.g5h(0), G5_BLK, NXS * G5_BLK, bar + B_XOP_FULL,
                                        bar + B_XOP_EMPTY, tl, 0, true);
  } else if (warp >= 4 && warp < 20) {
    // ===== epilogue: thread = (cell, 16 of the hidden units) =====
    const int ew = warp - 4, q4 = warp & 3, hf = ew >> 2, cell = q4 * 32 + lane, U0 = hf * 16;
    const bool mom_warp = MOM && hf == 3;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    // dW3 / db2 of unit U0 + (lane & 15), summed over this warp's 32 cells tile by tile (the sixteen per-cell
    // running sums each would not fit the register budget: they spilled, and E2 crawled through local memory)
    float w3l = 0.f, b2l = 0.f;

    float du_cur = 0.f, du_nxt = 0.f;
    uint32_t m1_cur = 0u, m1_nxt = 0u;

    // Per-cell inputs of the next E1 tile are fetched one phase early and its stock number one tile before
    // that (HBM latency off the critical path); the per-month scalars (dP_t, Rbar_t, S_t) are recomputed
    // only when the month changes.
    MonthCursor mc;
    int nn_n = -1;                                       // stock of this thread's cell in the next tile to prefetch (month: mc.t)
    int pf_t = 0;                                        // month of the prefetched tile
    uint32_t pf_gi = 0u;
    bool pf_live = false;
    float pf_r = 0.f, pf_g[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) pf_g[k] = 0.f;
    int ms_month = -1;
    float ms_dP = 0.f, ms_rbar = 0.f, ms_S = 0.f;
    float pf_u = 0.f, ms_rb = 0.f, ms_mu = 0.f;          // RES only
    auto load_n = [&](int i) {
      const int tile = tile_lo + i;
      mc.seek(P, tile);
      nn_n = (cell < mc.ncell(tile)) ? __ldg(vidx_row(P, mc.t) + mc.j0(tile) + cell) : -1;
    };
    auto prefetch = [&](int i) {                         // mc / nn_n describe tile i
      const int t = mc.t, n = nn_n;
      pf_t = t; pf_live = n >= 0;
      const size_t gi = (size_t)t * P.N + (pf_live ? n : 0);
      pf_gi = (uint32_t)gi;
      pf_r = pf_live ? __ldg(P.R + gi) : 0.f;            // consumed one phase later: no stall here
      if (RES) pf_u = pf_live ? __ldg(P.u_fwd + (size_t)t * P.N + mc.j0(tile_lo + i) + cell) : 0.f;   // left-packed u
      if (mom_warp) {                                    // the stock's eight coefficients: one 32-byte piece of its G row
        const float4* gr = reinterpret_cast<const float4*>(P.G + (size_t)(pf_live ? n : 0) * GS);
        const float4 ga = __ldg(gr), gb = __ldg(gr + 1);
        pf_g[0] = ga.x; pf_g[1] = ga.y; pf_g[2] = ga.z; pf_g[3] = ga.w;
        pf_g[4] = gb.x; pf_g[5] = gb.y; pf_g[6] = gb.z; pf_g[7] = gb.w;
        if (!pf_live) {
#pragma unroll
          for (int k = 0; k < 8; ++k) pf_g[k] = 0.f;
        }
      }
      if (t != ms_month) {
        ms_month = t;
        const float* rs = P.row_stats + (size_t)t * RS_W;
        ms_dP = month_dP(P, t);
        ms_rbar = rs[RS_RBAR];
        // S_t only ever multiplies the moment-side cotangent: carry dL/dL_c with it (G holds UNIT coefficients)
        ms_S = rs[RS_S] * __ldg(P.cot[0]);
        if (RES) { ms_rb = __ldg(P.res_b + t); ms_mu = rs[RS_MU]; }
      }
      if (i + 1 < my_tiles) load_n(i + 1);
    };
    int e1_t = 0;
    uint32_t gi_cur = 0u, gi_nxt = 0u;                   // dense cell number (dropout key) of tile i / i+1

    const float dscale = DROP ? P.drop_scale : 1.f;
    // E1 of local tile i: h1 -> H[sl] image, relu mask, this cell's du ; moment warps: dpre
    auto do_e1 = [&](int i) {
      const int sl = i & 1, use = i >> 1;
      const int t = pf_t;
      e1_t = pf_t; gi_nxt = pf_gi;
      const float pf_m = pf_live ? 1.f : 0.f;            // every listed cell is unmasked
      du_nxt = pf_m * ms_dP * (pf_r - ms_rbar);
      if (RES) du_nxt = fmaf(pf_m * ms_rb, pf_u - ms_mu, du_nxt);
      const float cq = pf_r * pf_m * ms_S;
      float gk[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) gk[k] = pf_g[k];
      const uint32_t dA = tbase + lane_base + (sl ? C_A1 : C_A0);
      mbar_wait(bar + B_W1_DONE + sl, (use & 1) ^ 1);   // dW1x of tile i-2 no longer reads H[sl]
      mbar_wait(bar + B_D1_FULL + sl, use & 1);
      tc::fence_after_sync();
      m1_nxt = 0u;
      if (SDF) {
        float v[16];
        tc::tmem_ld16(dA + U0, v);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = fmaxf(v[k], 0.f);
        if (DROP) tc_dropout<16>(P, gi_nxt, 0, U0, v);                // the image holds h1 AFTER dropout
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (v[k] > 0.f) m1_nxt |= 1u << k;                          // kept and z1 > 0
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint4 hi, lo;
          tc::split8(v + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.hh(sl) + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.hl(sl) + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
      }
      if (mom_warp) {
        float q[8];
        tc::tmem_ld8(dA + 64, q);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float qq = tc::tanh_fast(q[k]);
          q[k] = gk[k] * cq * (1.f - qq * qq);           // dpre_k
        }
        uint4 hi, lo;
        tc::split8(q, hi, lo);
        *reinterpret_cast<uint4*>(s.dph(sl) + cell * 16) = hi;
        *reinterpret_cast<uint4*>(s.dpl(sl) + cell * 16) = lo;
      }
      tc::fence_before_sync();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + B_H1_FULL + sl);
    };

    // The finished month's da_t | dc_t partial of this CTA: row 46 of the dW1 accumulator (M = 64 layout: TMEM
    // lane 32*(46/16) + 46%16 = 78, i.e. lane 14 of the third lane quarter), read and reset to zero.  i_last: the
    // month's last tile; its dW1 MMA must be complete, and the next one cannot start before these threads have
    // passed E3 / E1 of the following tile, so the accumulator is quiet in between.
    auto flush_month = [&](int t, int i_last) {
      if (q4 != 2) return;
      mbar_wait(bar + B_W1_DONE + (i_last & 1), (i_last >> 1) & 1);
      tc::fence_after_sync();
      // CTAs that touch month t are consecutive, starting with the owner of the month's first tile
      const int seg = (int)blockIdx.x - __ldg(P.index + idx_off_ts() + t) / chunk;
      float* out = P.dadc_part + ((size_t)t * P.tpm + seg) * 72;
      const uint32_t row = tbase + lane_base + C_W1;
      {
        float v[16];
        tc::tmem_ld16(row + U0, v);
        tc::tmem_wait_ld();
        if (lane == 14) {
#pragma unroll
          for (int k = 0; k < 16; ++k) { out[U0 + k] = SDF ? v[k] : 0.f; v[k] = 0.f; }
        }
        tc::tmem_st8(row + U0, reinterpret_cast<const uint32_t*>(v));
        tc::tmem_st8(row + U0 + 8, reinterpret_cast<const uint32_t*>(v) + 8);
      }
      if (hf == 3) {
        float q[8];
        tc::tmem_ld8(row + 64, q);
        tc::tmem_wait_ld();
        if (lane == 14) {
#pragma unroll
          for (int k = 0; k < 8; ++k) { out[64 + k] = MOM ? q[k] : 0.f; q[k] = 0.f; }
        }
        tc::tmem_st8(row + 64, reinterpret_cast<const uint32_t*>(q));
      }
      tc::tmem_wait_st();
      tc::fence_before_sync();
    };

    if (my_tiles > 0) {
      mc.init(P, tile_lo);
      load_n(0);
      prefetch(0);
      do_e1(0);
    }
    for (int i = 0; i < my_tiles; ++i) {
      const int sl = i & 1, par = i & 1;
      const int t = e1_t;
      du_cur = du_nxt; m1_cur = m1_nxt; gi_cur = gi_nxt;
      const bool stamp = ew == 0 && lane == 0;
      if (stamp) DLAP_STAMP(i, 0);
      if (i + 1 < my_tiles) prefetch(i + 1);
      if (SDF) {
        // ---- E2: dz2 = du W3 [z2 > 0] -> DZ image ; dW3 += du h2 ; db2 += dz2
        mbar_wait(bar + B_D2_FULL, par);
        tc::fence_after_sync();
        if (stamp) DLAP_STAMP(i, 1);
        float v[16], cw[16];
        tc::tmem_ld16(tbase + lane_base + C_B + U0, v);
        tc::tmem_wait_ld();
        const float4* b4 = reinterpret_cast<const float4*>(s.b2s() + U0);
        const float4* w4 = reinterpret_cast<const float4*>(s.w3s() + U0);
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 bb = b4[k4];
          const float b_[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) v[4 * k4 + e] = fmaxf(v[4 * k4 + e] + b_[e], 0.f);
        }
        if (DROP) tc_dropout<16>(P, gi_cur, 1, U0, v);                // h2 after dropout: > 0 iff kept and z2 > 0
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 ww = w4[k4];
          const float w_[4] = {ww.x, ww.y, ww.z, ww.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = 4 * k4 + e;
            const float h2 = v[k];
            cw[k] = du_cur * h2;
            v[k] = (h2 > 0.f) ? du_cur * w_[e] * dscale : 0.f;
          }
        }
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint4 hi, lo;
          tc::split8(v + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.dzh() + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.dzl() + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
        tc::fence_before_sync();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar + B_DZ2_FULL);
        if (stamp) DLAP_STAMP(i, 2);
        // off the critical path (the tensor core is busy with dW2 / dgrad now): this tile's dW3, db2 contributions
        w3l += lane_transpose_sum16(cw, lane);
        b2l += lane_transpose_sum16(v, lane);
      }
      if (!SDF) {              // moment-only: dWq(i + 1) follows E1(i + 1) directly, so a finished month goes out first
        const int tn = (i + 1 < my_tiles) ? pf_t : -1;
        if (tn != t) flush_month(t, i);
      }
      if (i + 1 < my_tiles) do_e1(i + 1);               // overlaps dW2 / dgrad of tile i on the tensor core
      if (stamp) DLAP_STAMP(i, 3);
      if (SDF) {
        // ---- E3: dz1 = (dz2 . W2) [z1 > 0] -> H[sl] image (h1 of this tile is dead) ; da_t
        mbar_wait(bar + B_D3_FULL, par);
        tc::fence_after_sync();
        if (stamp) DLAP_STAMP(i, 4);
        float v[16];
        tc::tmem_ld16(tbase + lane_base + C_C + U0, v);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          v[k] = ((m1_cur >> k) & 1u) ? v[k] * dscale : 0.f;
        }
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint4 hi, lo;
          tc::split8(v + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.hh(sl) + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.hl(sl) + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
        tc::fence_before_sync();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar + B_DZ1_FULL);
        if (stamp) DLAP_STAMP(i, 5);
      }
      if (SDF) {
        const int t_next = (i + 1 < my_tiles) ? e1_t : -1;   // do_e1(i + 1) has run
        if (t_next != t) flush_month(t, i);
      }
    }

    // ---- per-CTA partials: accumulators live in TMEM (M = 64: row r sits in lane 32*(r/16) + r%16)
    mbar_wait(bar + B_ALL_DONE, 0);
    tc::fence_after_sync();
    float* wp = P.wpart + (size_t)blockIdx.x * wp_stride(Fr);
    const int row = q4 * 16 + lane;                  // valid for lane < 16
    const bool have = my_tiles > 0;
    {
      float v[16];
      if (SDF) { tc::tmem_ld16(tbase + lane_base + C_W2 + U0, v); tc::tmem_wait_ld(); }
      if (lane < 16)
        for (int k = 0; k < 16; ++k) wp[wp_off_w2() + row * 64 + U0 + k] = (SDF && have) ? v[k] : 0.f;
      // [dW1x | dWq]^T: row = characteristic f (valid for f < F), columns 0..63 = hidden unit, 64..71 = moment
      if (SDF) { tc::tmem_ld16(tbase + lane_base + C_W1 + U0, v); tc::tmem_wait_ld(); }
      if (lane < 16 && row < Fr)
        for (int k = 0; k < 16; ++k) wp[wp_off_w1() + (U0 + k) * Fr + row] = (SDF && have && row < P.F) ? v[k] : 0.f;
      if (hf == 3) {
        float q[8];
        if (MOM) { tc::tmem_ld8(tbase + lane_base + C_W1 + 64, q); tc::tmem_wait_ld(); }
        if (lane < 16 && row < Fr)
          for (int k = 0; k < 8; ++k) wp[wp_off_wq(Fr) + k * Fr + row] = (MOM && have && row < P.F) ? q[k] : 0.f;
      }
    }
    // dW3 / db2: reduce over this warp's cells, then over the four lane quarters in a fixed order
    float* red = s.red();
    const float a = w3l, b = b2l;
    if (lane < 16) { red[(ew * 16 + lane) * 2] = a; red[(ew * 16 + lane) * 2 + 1] = b; }
    epi_barrier();
    const int et = tid - 128;
    if (et < 64) {
      const int h = et >> 4, k = et & 15;
      float sa = 0.f, sb = 0.f;
      for (int q = 0; q < 4; ++q) { sa += red[((h * 4 + q) * 16 + k) * 2]; sb += red[((h * 4 + q) * 16 + k) * 2 + 1]; }
      wp[wp_off_w3(Fr) + et] = SDF ? sa : 0.f;
      wp[wp_off_b2(Fr) + et] = SDF ? sb : 0.f;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (tid == 0) DLAP_STAMP(255, 2);                          // every role done
  if (P.timeline && tid == 0 && blockIdx.x < 2048)           // per-CTA active cycles and tile count
    P.timeline[2 * 4096 + 2048 + blockIdx.x] = ((clock64() - t_entry) << 8) | (long long)(my_tiles & 255);
  if (warp == 0) tc::tmem_dealloc(tbase, TMEM_COLS);
}

template <bool SDF, bool MOM, bool DROP, bool RES = false>
static cudaError_t launch_bwd_tc_variant(const CellParams& P, int grid, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(cell_bwd_tc_kernel<SDF, MOM, DROP, RES>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_B_BYTES);
  if (e != cudaSuccess) return e;
  cell_bwd_tc_kernel<SDF, MOM, DROP, RES><<<grid, NT_B, SMEM_B_BYTES, st>>>(P);
  return cudaGetLastError();
}

static long long* g_timeline = nullptr;
void set_bwd_tc_timeline(long long* p) { g_timeline = p; }

cudaError_t launch_cell_bwd_tc(const CellParams& P0, int grid, bool sdf, bool mom, cudaStream_t st) {
  CellParams P = P0;
  P.timeline = g_timeline;
  const bool drop = P.drop_thr != 0 && sdf;
  if (sdf && P.res_b) {            // the residual-loss term only ever reaches the SDF net
    if (mom) return drop ? launch_bwd_tc_variant<true, true, true, true>(P, grid, st)
                         : launch_bwd_tc_variant<true, true, false, true>(P, grid, st);
    return drop ? launch_bwd_tc_variant<true, false, true, true>(P, grid, st)
                : launch_bwd_tc_variant<true, false, false, true>(P, grid, st);
  }
  if (sdf && mom) return drop ? launch_bwd_tc_variant<true, true, true>(P, grid, st)
                              : launch_bwd_tc_variant<true, true, false>(P, grid, st);
  if (sdf) return drop ? launch_bwd_tc_variant<true, false, true>(P, grid, st)
                       : launch_bwd_tc_variant<true, false, false>(P, grid, st);
  return launch_bwd_tc_variant<false, true, false>(P, grid, st);
}

}  // namespace dlap
