// Tensor-core variant of the per-cell kernels (precision "bf16x3").
//
// Every fp32 contraction D = A.B^T of the SDF MLP / moment net is issued as three tcgen05 BF16 MMAs
// on split operands  a = a_hi + a_lo  (|a - a_hi - a_lo| <= 2^-18 |a|):
//     D = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi            (fp32 accumulation in TMEM)
// which keeps ~fp32 accuracy (measured against the oracle in tests/) at 3x the MMA count.
//
// Shared-memory operand images use the no-swizzle canonical layout [group of 8][row][8] (tc_sm100.cuh).
//   X  [6][128+1][8] characteristics of the tile; k = 46 holds 1.0 (the "ones column"), k = 47 = 0.  The group
//                    stride is padded by one row (2064 B) so that a warp holding one row as a column pair per
//                    lane stores it without bank conflicts (rows_to_image)
//   W1 [6][80][8]    rows 0..63 = W1[:, :F], rows 64..71 = Wq[:, M:], rows 72..79 = 0;
//                    column k = 46 carries the per-month bias a_t | c_t, so both first-layer biases
//                    are added by the MMA itself
//   H1 [8][128][8]   relu(z1) written by the epilogue warps, A operand of layer 2
//   W2 [8][64][8]
#pragma once
#include "cell_f32.cuh"
#include "tc_sm100.cuh"

namespace dlap {
namespace tcl {

constexpr int KG1 = 6;                  // k-groups of layer 1 (K = 48)
constexpr int N1 = 80;                  // layer-1 MMA N: 64 hidden + 8 moments + 8 pad
constexpr int KG2 = 8;                  // k-groups of layer 2 (K = 64)
constexpr uint32_t X_LBO = TILE * 16 + 16;       // byte stride between the k-groups of an X image
constexpr uint32_t X_IMG = KG1 * X_LBO;          // 12384 B per hi or lo image
constexpr uint32_t W1_IMG = KG1 * N1 * 16;       //  7680 B
constexpr uint32_t H_IMG = KG2 * TILE * 16;      // 16384 B
constexpr uint32_t W2_IMG = KG2 * 64 * 16;       //  8192 B
constexpr int ONES_K = 46;              // index of the ones column inside X (needs F <= 46)

// X image pair before the first tile: zeros, and 1.0 in column k = 46 of the hi image.  The gather warps only
// ever overwrite columns < F, so the ones column and the zero padding survive every tile.
__device__ __forceinline__ void init_x_images(uint8_t* hi, uint8_t* lo, int tid, int nthr) {
  for (uint32_t i = tid; i < X_IMG / 4; i += nthr) {
    const uint32_t byte = 4 * i, g = byte / X_LBO, w = byte % X_LBO;
    const bool ones = g == (uint32_t)(ONES_K >> 3) && w < TILE * 16 && (w & 15) == (uint32_t)(ONES_K & 7) * 2;
    reinterpret_cast<uint32_t*>(hi)[i] = ones ? 0x00003F80u : 0u;     // bf16 (1.0, 0.0)
    reinterpret_cast<uint32_t*>(lo)[i] = 0u;
  }
}

// ROWS rows of a tile (cells cell0 .. cell0+ROWS-1), held as one 64-bit register per lane (lane p = columns
// 2p, 2p+1; gather_load), -> the bf16 hi/lo X image pair.  Lane p writes 4 bytes per row and image at
// [group p/4][cell][(p%4)*4]; with the padded group stride the lanes of a warp hit distinct banks.
template <int ROWS>
__device__ __forceinline__ void rows_to_image(const uint64_t (&r)[ROWS], uint8_t* xh, uint8_t* xl, int cell0, int lane,
                                              int F) {
  const bool odd_tail = 2 * lane + 1 >= F;       // last pair of an odd F: column 2p+1 does not exist
  if (2 * lane < F) {
    const uint32_t off = (uint32_t)(lane >> 2) * X_LBO + (uint32_t)cell0 * 16 + (uint32_t)(lane & 3) * 4;
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      uint32_t h, l;
      const float x = __uint_as_float((uint32_t)r[j]), y = __uint_as_float((uint32_t)(r[j] >> 32));
      tc::split2(x, odd_tail ? 0.f : y, h, l);
      *reinterpret_cast<uint32_t*>(xh + off + j * 16) = h;
      *reinterpret_cast<uint32_t*>(xl + off + j * 16) = l;
    }
  }
}

// ---- gather + convert role (forward and backward kernels) -----------------------------------------------
// The tiles of a CTA (tm: local tile i -> tile) are cut into batches of BR rows; batch q = (local tile q / NB, rows BR * (q % NB) ..) goes to
// gather warp q % NGW.  A batch is fetched with coalesced loads (gather_load), waits in registers, and is
// written into the X image of its tile's slot (rows_to_image) once the slot is free; the warp that handles
// the first batch of a tile also writes the per-month bias column (k = 46) of the slot's W1 image:
// a_t (units 0..63) | c_t (moments) | 0.  x0: hi X image of slot 0 (slots are 2 images apart); g5h0: the hi
// [N1][8] block of k-group 5 of slot 0's W1 image, g5_stride bytes between slots, g5_lo bytes to the lo block.
//
// Loop order matters.  ptxas tracks loads with six scoreboards per warp, so a wait placed after the row loads of
// the next batch shares a scoreboard with them and stalls for the whole gather: every consumer of a long-latency
// load (month cursor, bias, stock numbers) sits BEFORE the row loads are issued, and the stock-number load for
// a later batch is issued AFTER them, once the old value is dead, so that it lands in the loop-carried register
// itself (a register copy at the loop end would wait for it).  Stock numbers run two batches ahead of the row
// loads: those rows are prefetched to L2 in the meantime.  The loop starts one step early (q < 0: loads only)
// so that the row registers have a single definition.
template <int BR, int NGW, int NSLOT, bool WIDE, bool MOM>
__device__ __forceinline__ void gather_role(const CellParams& P, int gw, int lane, const TileMap tm,
                                            uint8_t* x0, uint8_t* g5h0, uint32_t g5_stride, uint32_t g5_lo,
                                            uint64_t* xop_full, uint64_t* xop_empty,
                                            long long* tl /* stamps or null */, int tl_slot, bool tl_per_batch) {
  constexpr int NB = TILE / BR;
  const int nbatch = tm.count * NB;
  MonthCursor mc;
  if (tm.count > 0) mc.init(P, tm.tile(0));
  int t_seek = 0;
  // where lane l's stock number of batch q lives (cells past the end of the tile repeat its last row); moves the
  // month cursor, whose rare loads are consumed right here
  auto n_addr = [&](int q) {
    const int tile = tm.tile(q / NB), cell = (q % NB) * BR + (lane & (BR - 1));
    mc.seek(P, tile);
    t_seek = mc.t;
    return vidx_row(P, mc.t) + mc.j0(tile) + min(cell, mc.ncell(tile) - 1);
  };
  // entries lane, lane + 32, lane + 64 of the bias column.  The recurrence that produces a_t may still be running
  // (months_ready): wait for the month, and read past L1 (the lines can change during this kernel).
  int months_seen = 0;
  auto load_bias = [&](int t, float (&b)[3]) {
    const size_t row = (size_t)t * P.months_stride;
    if (P.months_ready) wait_month(P.months_ready, P.months_base + (int)row, months_seen);
    b[0] = __ldcg(P.a_t + row * 64 + lane);
    b[1] = __ldcg(P.a_t + row * 64 + 32 + lane);
    b[2] = (MOM && lane < 8) ? __ldcg(P.c_t + row * 8 + lane) : 0.f;
  };
  uint64_t rows[BR];
#pragma unroll
  for (int j = 0; j < BR; ++j) rows[j] = 0ull;
  float bias[3] = {0.f, 0.f, 0.f};
  int n_nxt = 0, t_nxt = 0;                        // this warp's next batch: rows loaded next
  int n_pf = 0, t_pf = 0;                          // the batch after it: rows prefetched to L2, loaded one batch later
  if (gw < nbatch) { n_nxt = __ldg(n_addr(gw)); t_nxt = t_seek; }
  if (gw + NGW < nbatch) { n_pf = __ldg(n_addr(gw + NGW)); t_pf = t_seek; }
  for (int q = gw - NGW; q < nbatch; q += NGW) {
    long long* st = (tl && q >= 0) ? tl + (size_t)(tl_per_batch ? q / NGW : q / NB) * 16 + tl_slot : nullptr;
    if (st && (tl_per_batch ? q / NGW : q / NB) >= 256) st = nullptr;
    if (q >= 0) {
      const int i = q / NB, b = q % NB, sl = i % NSLOT, use = i / NSLOT;
      uint8_t* xh = x0 + sl * 2 * X_IMG;
      uint8_t* g5h = g5h0 + sl * g5_stride;
      if (st) st[0] = clock64();
      __nv_bfloat16 bh[3], bl[3];
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        bh[e] = __float2bfloat16_rn(bias[e]);
        bl[e] = __float2bfloat16_rn(bias[e] - __bfloat162float(bh[e]));
      }
      mbar_wait(xop_empty + sl, (use & 1) ^ 1);
      if (st) st[1] = clock64();
      rows_to_image<BR>(rows, xh, xh + X_IMG, b * BR, lane, P.F);        // waits for the row loads
      if (b == 0) {
#pragma unroll
        for (int e = 0; e < 3; ++e) {
          const int unit = lane + 32 * e;
          if (unit < N1) {
            const int off = unit * 16 + (ONES_K - 8 * (KG1 - 1)) * 2;
            *reinterpret_cast<__nv_bfloat16*>(g5h + off) = bh[e];
            *reinterpret_cast<__nv_bfloat16*>(g5h + g5_lo + off) = bl[e];
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(xop_full + sl);
      if (st) st[2] = clock64();
    }
    if (q + NGW < nbatch) {
      const int* pn = (q + 3 * NGW < nbatch) ? n_addr(q + 3 * NGW) : nullptr;
      const int n2 = n_pf, t2 = t_pf;
      if ((q + NGW) % NB == 0) load_bias(t_nxt, bias);
      gather_load<BR, WIDE>(P, t_nxt, n_nxt, lane, rows);
      if (q + 2 * NGW < nbatch) prefetch_row(P, t2, n2);
      n_nxt = n2; t_nxt = t2;
      if (pn) { n_pf = __ldg(pn); t_pf = t_seek; }
    }
    if (st) st[3] = clock64();
  }
}

// ---- the same role when the panel's X images already exist in HBM (CellParams::ximg, dlap_panel_image) -----------
// The characteristics do not change from one step to the next (full-batch training re-reads the same panel every
// epoch, src/train.py:45-103), so gathering 184-byte rows and splitting them into bf16 hi/lo pairs can be done once
// per panel instead of twice per step: dlap_panel_image stores, for every tile of the valid-cell index, exactly the
// bytes the gather warps would have put into shared memory (hi image, then lo image: 2 * X_IMG contiguous bytes).
// ONE warp then feeds the tensor core: per tile one 1-D bulk copy (TMA engine) of 24 KB into the free slot,
// completing on the slot's `full` barrier, and the per-month bias column of the slot's W1 image as before.  The
// barrier still counts NB arrivals per tile (the gather role's batches): lane 0 makes all of them.
template <int NB, int NSLOT, bool MOM>
__device__ __forceinline__ void image_role(const CellParams& P, int lane, const TileMap tm, uint8_t* x0, uint8_t* g5h0,
                                           uint32_t g5_stride, uint32_t g5_lo, uint64_t* xop_full, uint64_t* xop_empty) {
  static_assert(NB >= 2, "the arrivals that follow the bias column publish it");
  MonthCursor mc;
  if (tm.count > 0) mc.init(P, tm.tile(0));
  int months_seen = 0;
  for (int i = 0; i < tm.count; ++i) {
    const int tile = tm.tile(i), sl = i % NSLOT, use = i / NSLOT;
    mc.seek(P, tile);
    mbar_wait(xop_empty + sl, (use & 1) ^ 1);
    if (lane == 0) {
      mbar_expect_tx(xop_full + sl, 2 * X_IMG);                         // arrival 1 of NB, and the bytes to come
      bulk_g2s_stream(x0 + (size_t)sl * 2 * X_IMG, P.ximg + (size_t)tile * (2 * X_IMG), 2 * X_IMG, xop_full + sl);
    }
    // bias column a_t | c_t | 0 of this tile's month (the recurrence that makes a_t may still be running)
    const size_t row = (size_t)mc.t * P.months_stride;
    if (P.months_ready) wait_month(P.months_ready, P.months_base + (int)row, months_seen);
    const float b[3] = {__ldcg(P.a_t + row * 64 + lane), __ldcg(P.a_t + row * 64 + 32 + lane),
                        (MOM && lane < 8) ? __ldcg(P.c_t + row * 8 + lane) : 0.f};
    uint8_t* g5h = g5h0 + sl * g5_stride;
#pragma unroll
    for (int e = 0; e < 3; ++e) {
      const int unit = lane + 32 * e;
      if (unit < N1) {
        const __nv_bfloat16 bh = __float2bfloat16_rn(b[e]);
        const __nv_bfloat16 bl = __float2bfloat16_rn(b[e] - __bfloat162float(bh));
        const int off = unit * 16 + (ONES_K - 8 * (KG1 - 1)) * 2;
        *reinterpret_cast<__nv_bfloat16*>(g5h + off) = bh;
        *reinterpret_cast<__nv_bfloat16*>(g5h + g5_lo + off) = bl;
      }
    }
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
#pragma unroll
      for (int k = 1; k < NB; ++k) mbar_arrive(xop_full + sl);
    }
  }
}

// fill k-groups g0 .. g0+ng-1 of a W1 image pair (hi, lo: contiguous [ng][N1][8] blocks) from global
// parameters; the bias column k = 46 is left zero (the gather warps write it per tile)
__device__ __forceinline__ void fill_w1_groups(const CellParams& P, uint8_t* hi, uint8_t* lo, int g0, int ng, int tid,
                                               int nthr) {
  for (int i = tid; i < ng * N1 * 4; i += nthr) {       // one bf16x2 pair per step
    const int g = g0 + i / (N1 * 4), n = (i / 4) % N1, pr = i % 4;
    const int k0 = g * 8 + pr * 2;
    float v[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int k = k0 + e;
      float w = 0.f;
      if (k < P.F) {
        if (n < 64) w = __ldg(P.W1 + (size_t)n * P.ldw1 + k);
        else if (n < 72 && P.Wq) w = __ldg(P.Wq + (size_t)(n - 64) * P.ldwq + P.wq_off + k);
      }
      v[e] = w;
    }
    uint32_t h, l;
    tc::split2(v[0], v[1], h, l);
    reinterpret_cast<uint32_t*>(hi)[i] = h;
    reinterpret_cast<uint32_t*>(lo)[i] = l;
  }
}
// W2 image: element (row n = out unit j, k = in unit) = W2[j][k]
__device__ __forceinline__ void fill_w1_images(const CellParams& P, uint8_t* hi, uint8_t* lo, int tid, int nthr) {
  fill_w1_groups(P, hi, lo, 0, KG1, tid, nthr);
}
__device__ __forceinline__ void fill_w2_images(const CellParams& P, uint8_t* hi, uint8_t* lo, int tid, int nthr) {
  for (int i = tid; i < KG2 * 64 * 4; i += nthr) {
    const int g = i / (64 * 4), n = (i / 4) % 64, pr = i % 4;
    const int k0 = g * 8 + pr * 2;
    uint32_t h, l;
    tc::split2(__ldg(P.W2 + n * 64 + k0), __ldg(P.W2 + n * 64 + k0 + 1), h, l);
    reinterpret_cast<uint32_t*>(hi)[i] = h;
    reinterpret_cast<uint32_t*>(lo)[i] = l;
  }
}

// Dropout on N consecutive hidden units U0.. of one stock-month (N even): v[k] <- v[k] * (0 | scale).
// Same counter-based generator and keying as relu_dropout() of the fp32 kernels (cell_f32.cuh), so the
// two precisions drop exactly the same units for a given seed, forward and backward.
template <int N>
__device__ __forceinline__ void tc_dropout(const CellParams& P, uint32_t cell_global, int layer, int U0, float* v) {
  const uint32_t key = mix32(cell_global ^ drop_seed_lo(P)) ^ drop_seed_hi(P);
#pragma unroll
  for (int k = 0; k < N; k += 2) {
    const uint32_t r = mix32(key + (uint32_t)(layer * 32 + ((U0 + k) >> 1)) * 0x9E3779B9u);
    v[k] = ((r & 0xffffu) >= P.drop_thr) ? v[k] * P.drop_scale : 0.f;
    v[k + 1] = ((r >> 16) >= P.drop_thr) ? v[k + 1] * P.drop_scale : 0.f;
  }
}

// three-term split product for one K = 16 step
__device__ __forceinline__ void umma3(uint32_t d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl,
                                      uint32_t idesc, bool first) {
  tc::umma_bf16(d, ah, bh, idesc, first ? 0u : 1u);
  tc::umma_bf16(d, ah, bl, idesc, 1u);
  tc::umma_bf16(d, al, bh, idesc, 1u);
}

}  // namespace tcl
}  // namespace dlap
