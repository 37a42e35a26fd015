// Backward per-cell kernel on the tensor cores (tcgen05 + TMEM), precision "bf16x3", with activation
// recompute.  Same contract as cell_bwd_f32_kernel (cell_bwd.cu) except that the per-tile da_t | dc_t
// partials are written per epilogue warp quarter (4 slots per tile).
//
// Per tile of 128 stock-months, on split bf16 operands (3 MMAs per fp32 product):
//   L1    D_A[128,80] = X.[W1x;Wqx]^T (+ a_t|c_t through the ones column)      -> E1: h1, relu mask, dpre
//   L2    D_B[128,64] = H1.W2^T                                                -> E2: dz2, dW3, db2
//   MQ    D_Q[64,8]  += X^T.DPRE        (moment-net weight gradient, accumulated over the CTA's tiles)
//   WG2   D_W2[64,64] += DZ2^T.H1       (dW2)
//   DG    D_C[128,64] = DZ2.W2                                                 -> E3: dz1, da_t
//   WG1   D_W1[64,48] += DZ1^T.X        (dW1x)
// All activation tiles live in shared memory as [group of 8 units][cell][8] bf16 images, so the same
// bytes serve as K-major operand (forward / dgrad) and MN-major operand (weight gradients).
//
// 14 warps: 0 TMA producer + TMEM owner, 1 MMA issuer, 2-5 converter, 6-13 epilogue (two warps per
// TMEM lane quarter; each handles 32 of the 64 hidden units of its 32 cells).
#include "cell_tc.cuh"

namespace dlap {
using namespace tcl;

namespace {
constexpr int NT_B = 448;
constexpr int STAGE_BYTES = ((TILE * 46 + 8) * 4 + 127) & ~127;
constexpr uint32_t DPS_IMG = TILE * 16;            // one group: [cell][8 moments]

struct SmemB {
  uint8_t* base;
  __device__ float* stage(int i) const { return reinterpret_cast<float*>(base + i * STAGE_BYTES); }
  __device__ uint8_t* xh(int sl) const { return base + 2 * STAGE_BYTES + sl * 2 * X_IMG; }
  __device__ uint8_t* xl(int sl) const { return xh(sl) + X_IMG; }
  __device__ uint8_t* w1h(int sl) const { return base + 2 * STAGE_BYTES + 4 * X_IMG + sl * 2 * W1_IMG; }
  __device__ uint8_t* w1l(int sl) const { return w1h(sl) + W1_IMG; }
  __device__ uint8_t* w2h() const { return base + 2 * STAGE_BYTES + 4 * X_IMG + 4 * W1_IMG; }
  __device__ uint8_t* w2l() const { return w2h() + W2_IMG; }
  __device__ uint8_t* hh() const { return w2l() + W2_IMG; }
  __device__ uint8_t* hl() const { return hh() + H_IMG; }
  __device__ uint8_t* dzh() const { return hl() + H_IMG; }
  __device__ uint8_t* dzl() const { return dzh() + H_IMG; }
  __device__ uint8_t* dph() const { return dzl() + H_IMG; }
  __device__ uint8_t* dpl() const { return dph() + DPS_IMG; }
  __device__ float* b2s() const { return reinterpret_cast<float*>(dpl() + DPS_IMG); }
  __device__ float* w3s() const { return b2s() + 64; }
  __device__ float* red() const { return w3s() + 64; }                       // [8 warps][32][2]
  __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(red() + 8 * 32 * 2); }
  __device__ uint32_t* tmem_slot() const { return reinterpret_cast<uint32_t*>(bars() + 24); }
};
constexpr size_t SMEM_B_BYTES = 2 * (size_t)STAGE_BYTES + 4 * X_IMG + 4 * W1_IMG + 2 * W2_IMG + 4 * H_IMG + 2 * DPS_IMG +
                                128 * 4 + 8 * 32 * 2 * 4 + 24 * 8 + 16 + 128;

// barrier indices
enum { B_STAGE_FULL = 0, B_STAGE_EMPTY = 2, B_XOP_FULL = 4, B_XOP_EMPTY = 6, B_D1_FULL = 8, B_H1_FULL = 10,
       B_D2_FULL = 11, B_DZ2_FULL = 12, B_D3_FULL = 13, B_DZ1_FULL = 14, B_W1_DONE = 15, B_ALL_DONE = 16 };

// TMEM columns
constexpr uint32_t C_A0 = 0, C_A1 = 96, C_B = 192, C_C = 256, C_W2 = 320, C_W1 = 384, C_Q = 448, TMEM_COLS = 512;

// column sums over the 32 lanes of a warp of 32 per-lane values: lane L ends with sum_lanes v[L]
__device__ __forceinline__ float lane_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}
}  // namespace

template <bool SDF, bool MOM>
__global__ void __launch_bounds__(NT_B, 1) cell_bwd_tc_kernel(const CellParams P) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  SmemB s;
  s.base = tc::smem_align128(smem_raw);
  uint64_t* bar = s.bars();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = P.T * P.tpm;
  const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int Fr = round4(P.F);

  // ---- one-time set-up ----
  for (int sl = 0; sl < 2; ++sl) fill_w1_images(P, s.w1h(sl), s.w1l(sl), tid, NT_B);
  fill_w2_images(P, s.w2h(), s.w2l(), tid, NT_B);
  for (int i = tid; i < 2 * (int)DPS_IMG / 4; i += NT_B) reinterpret_cast<uint32_t*>(s.dph())[i] = 0u;
  if (tid < 64) { s.b2s()[tid] = __ldg(P.b2 + tid); s.w3s()[tid] = __ldg(P.W3 + tid); }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar + B_STAGE_FULL + i, 1);  mbar_init(bar + B_STAGE_EMPTY + i, 4);
      mbar_init(bar + B_XOP_FULL + i, 4);    mbar_init(bar + B_XOP_EMPTY + i, 1);
      mbar_init(bar + B_D1_FULL + i, 1);
    }
    mbar_init(bar + B_H1_FULL, 8);  mbar_init(bar + B_D2_FULL, 1);  mbar_init(bar + B_DZ2_FULL, 8);
    mbar_init(bar + B_D3_FULL, 1);  mbar_init(bar + B_DZ1_FULL, 8); mbar_init(bar + B_W1_DONE, 1);
    mbar_init(bar + B_ALL_DONE, 1);
    fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(s.tmem_slot(), TMEM_COLS);
  fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = *s.tmem_slot();

  if (warp == 0) {
    // ===== producer =====
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = blockIdx.x + i * gridDim.x, st = i & 1, use = i >> 1;
      mbar_wait(bar + B_STAGE_EMPTY + st, (use & 1) ^ 1);
      const TileLoad L = plan_tile(P, tile / P.tpm, (tile % P.tpm) * TILE);
      __syncwarp();
      if (L.bytes) {
        if (lane == 0) {
          mbar_expect_tx(bar + B_STAGE_FULL + st, L.bytes);
          bulk_g2s(s.stage(st), L.src - L.shift, L.bytes, bar + B_STAGE_FULL + st);
        }
      } else {
        const int n = L.ncell * P.F;
        for (int j = lane; j < n; j += 32) s.stage(st)[L.shift + j] = __ldg(L.src + j);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar + B_STAGE_FULL + st);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (all lanes wait, lane 0 issues) =====
    const uint32_t id_l1 = tc::idesc_bf16(128, N1, 0, 0), id_l2 = tc::idesc_bf16(128, 64, 0, 0);
    const uint32_t id_dg = tc::idesc_bf16(128, 64, 0, 1);                 // A K-major, B = W2 MN-major
    const uint32_t id_w2 = tc::idesc_bf16(64, 64, 1, 1), id_w1 = tc::idesc_bf16(64, 48, 1, 1);
    const uint32_t id_q = tc::idesc_bf16(64, 8, 1, 1);
    const uint32_t KM_A = TILE * 16;                                      // K-major activation: LBO
    // K-major images
    const uint64_t hh_k = tc::smem_desc(smem_u32(s.hh()), KM_A, 128), hl_k = tc::smem_desc(smem_u32(s.hl()), KM_A, 128);
    const uint64_t dzh_k = tc::smem_desc(smem_u32(s.dzh()), KM_A, 128), dzl_k = tc::smem_desc(smem_u32(s.dzl()), KM_A, 128);
    const uint64_t w2h_k = tc::smem_desc(smem_u32(s.w2h()), 64 * 16, 128), w2l_k = tc::smem_desc(smem_u32(s.w2l()), 64 * 16, 128);
    // MN-major views of the same bytes (LBO = 128: next 8 K rows; SBO: next group of 8 MN)
    const uint64_t hh_m = tc::smem_desc(smem_u32(s.hh()), 128, KM_A), hl_m = tc::smem_desc(smem_u32(s.hl()), 128, KM_A);
    const uint64_t dzh_m = tc::smem_desc(smem_u32(s.dzh()), 128, KM_A), dzl_m = tc::smem_desc(smem_u32(s.dzl()), 128, KM_A);
    const uint64_t w2h_m = tc::smem_desc(smem_u32(s.w2h()), 128, 64 * 16), w2l_m = tc::smem_desc(smem_u32(s.w2l()), 128, 64 * 16);
    const uint64_t dph_m = tc::smem_desc(smem_u32(s.dph()), 128, KM_A), dpl_m = tc::smem_desc(smem_u32(s.dpl()), 128, KM_A);

    auto issue_l1 = [&](int i) {
      const int sl = i & 1, use = i >> 1;
      mbar_wait(bar + B_XOP_FULL + sl, use & 1);
      tc::fence_after_sync();
      if (lane == 0) {
        const uint64_t ah = tc::smem_desc(smem_u32(s.xh(sl)), KM_A, 128), al = tc::smem_desc(smem_u32(s.xl(sl)), KM_A, 128);
        const uint64_t bh = tc::smem_desc(smem_u32(s.w1h(sl)), N1 * 16, 128), bl = tc::smem_desc(smem_u32(s.w1l(sl)), N1 * 16, 128);
        const uint32_t d = tbase + (sl ? C_A1 : C_A0);
#pragma unroll
        for (int ks = 0; ks < KG1 / 2; ++ks)
          umma3(d, tc::desc_advance(ah, ks * 2 * KM_A), tc::desc_advance(al, ks * 2 * KM_A),
                tc::desc_advance(bh, ks * 2 * N1 * 16), tc::desc_advance(bl, ks * 2 * N1 * 16), id_l1, ks == 0);
        tc::umma_commit(bar + B_D1_FULL + sl);
      }
      __syncwarp();
    };

    if (my_tiles > 0) issue_l1(0);
    for (int i = 0; i < my_tiles; ++i) {
      const int sl = i & 1, par = i & 1;
      const bool first = (i == 0);
      const uint64_t xh_m = tc::smem_desc(smem_u32(s.xh(sl)), 128, KM_A), xl_m = tc::smem_desc(smem_u32(s.xl(sl)), 128, KM_A);
      // -- after E1: layer 2 and the moment-net weight gradient
      mbar_wait(bar + B_H1_FULL, par);
      tc::fence_after_sync();
      if (lane == 0) {
        if (SDF) {
#pragma unroll
          for (int ks = 0; ks < KG2 / 2; ++ks)
            umma3(tbase + C_B, tc::desc_advance(hh_k, ks * 2 * KM_A), tc::desc_advance(hl_k, ks * 2 * KM_A),
                  tc::desc_advance(w2h_k, ks * 2 * 64 * 16), tc::desc_advance(w2l_k, ks * 2 * 64 * 16), id_l2, ks == 0);
          tc::umma_commit(bar + B_D2_FULL);
        }
        if (MOM) {
#pragma unroll
          for (int ks = 0; ks < TILE / 16; ++ks)
            umma3(tbase + C_Q, tc::desc_advance(xh_m, ks * 256), tc::desc_advance(xl_m, ks * 256),
                  tc::desc_advance(dph_m, ks * 256), tc::desc_advance(dpl_m, ks * 256), id_q, first && ks == 0);
        }
      }
      __syncwarp();
      if (i + 1 < my_tiles) issue_l1(i + 1);          // overlaps the epilogues of tile i
      if (SDF) {
        // -- after E2: dW2 and dgrad
        mbar_wait(bar + B_DZ2_FULL, par);
        tc::fence_after_sync();
        if (lane == 0) {
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
        // -- after E3: dW1x
        mbar_wait(bar + B_DZ1_FULL, par);
        tc::fence_after_sync();
        if (lane == 0) {
#pragma unroll
          for (int ks = 0; ks < TILE / 16; ++ks)
            umma3(tbase + C_W1, tc::desc_advance(dzh_m, ks * 256), tc::desc_advance(dzl_m, ks * 256),
                  tc::desc_advance(xh_m, ks * 256), tc::desc_advance(xl_m, ks * 256), id_w1, first && ks == 0);
          tc::umma_commit(bar + B_W1_DONE);
        }
        __syncwarp();
      }
      if (lane == 0) tc::umma_commit(bar + B_XOP_EMPTY + sl);     // every reader of X(i) has completed
      __syncwarp();
    }
    if (lane == 0) tc::umma_commit(bar + B_ALL_DONE);
    __syncwarp();
  } else if (warp >= 2 && warp < 6) {
    // ===== converter: thread = cell =====
    const int cell = tid - 64;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = blockIdx.x + i * gridDim.x, st = i & 1, sl = i & 1, use = i >> 1;
      const int t = tile / P.tpm, n0 = (tile % P.tpm) * TILE;
      const TileLoad L = plan_tile(P, t, n0);
      mbar_wait(bar + B_STAGE_FULL + st, use & 1);
      mbar_wait(bar + B_XOP_EMPTY + sl, (use & 1) ^ 1);
      const float* row = s.stage(st) + L.shift + cell * P.F;
      const bool live = cell < L.ncell;
#pragma unroll
      for (int g = 0; g < KG1; ++g) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int k = g * 8 + e;
          v[e] = (live && k < P.F) ? row[k] : 0.f;
        }
        if (g == KG1 - 1) v[ONES_K - 8 * (KG1 - 1)] = 1.f;
        uint4 hi, lo;
        tc::split8(v, hi, lo);
        *reinterpret_cast<uint4*>(s.xh(sl) + (g * TILE + cell) * 16) = hi;
        *reinterpret_cast<uint4*>(s.xl(sl) + (g * TILE + cell) * 16) = lo;
      }
      if (cell < N1) {
        float b = 0.f;
        if (cell < 64) b = __ldg(P.a_t + (size_t)t * 64 + cell);
        else if (cell < 72) b = __ldg(P.c_t + (size_t)t * 8 + cell - 64);
        const __nv_bfloat16 bh = __float2bfloat16_rn(b);
        const __nv_bfloat16 bl = __float2bfloat16_rn(b - __bfloat162float(bh));
        const int off = ((KG1 - 1) * N1 + cell) * 16 + (ONES_K - 8 * (KG1 - 1)) * 2;
        *reinterpret_cast<__nv_bfloat16*>(s.w1h(sl) + off) = bh;
        *reinterpret_cast<__nv_bfloat16*>(s.w1l(sl) + off) = bl;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) { mbar_arrive(bar + B_XOP_FULL + sl); mbar_arrive(bar + B_STAGE_EMPTY + st); }
    }
  } else if (warp >= 6) {
    // ===== epilogue: thread = (cell, half of the hidden units) =====
    const int q4 = warp & 3, hf = (warp - 6) >> 2, cell = q4 * 32 + lane, U0 = hf * 32;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    float dw3_acc = 0.f, db2_acc = 0.f;            // unit U0 + lane, summed over this warp's cells
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = blockIdx.x + i * gridDim.x, sl = i & 1, use = i >> 1, par = i & 1;
      const int t = tile / P.tpm, jt = tile % P.tpm, n0 = jt * TILE;
      const int ncell = min(TILE, P.N - n0);
      const bool live = cell < ncell;
      const size_t gi = (size_t)t * P.N + n0 + cell;
      float* dadc = P.dadc_part + (((size_t)t * P.tpm + jt) * 4 + q4) * 72;
      // per-month scalars and this cell's du, cq
      float m = 0.f, r = 0.f;
      if (live) { m = P.mask[gi] ? 1.f : 0.f; r = __ldg(P.R + gi); }
      const float* rs = P.row_stats + (size_t)t * RS_W;
      double dS = 0.0;
      for (int k = 0; k < P.ds_splits; ++k) dS += P.dS_part[(size_t)t * P.ds_splits + k];
      if (P.dpr_cot) dS += (double)__ldg(P.dpr_cot + t);
      const float dP = (float)(dS * (P.weighted ? (double)P.scalars[SC_NBAR] / (double)rs[RS_NT] : 1.0));
      const float du = m * dP * (r - rs[RS_RBAR]);
      const float cq = r * m * rs[RS_S];
      const uint32_t dA = tbase + lane_base + (sl ? C_A1 : C_A0);

      // ---- E1: h1 = relu(z1) -> H1 image ; relu mask ; dpre -> DPS image ; dc_t
      mbar_wait(bar + B_D1_FULL + sl, use & 1);
      tc::fence_after_sync();
      uint32_t m1 = 0u;
      if (SDF) {
        float v[32];
        tc::tmem_ld32(dA + U0, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          if (v[k] > 0.f) m1 |= 1u << k;
          v[k] = fmaxf(v[k], 0.f);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 hi, lo;
          tc::split8(v + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.hh() + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.hl() + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
      }
      if (MOM && hf == 1) {
        float q[8];
        tc::tmem_ld8(dA + 64, q);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float qq = tc::tanh_fast(q[k]);
          const float gk = live ? __ldg(P.G + (size_t)k * P.N + n0 + cell) : 0.f;
          q[k] = gk * cq * (1.f - qq * qq);              // dpre_k
        }
        uint4 hi, lo;
        tc::split8(q, hi, lo);
        *reinterpret_cast<uint4*>(s.dph() + cell * 16) = hi;
        *reinterpret_cast<uint4*>(s.dpl() + cell * 16) = lo;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float sum = warp_sum(q[k]);
          if (lane == 0) dadc[64 + k] = sum;
        }
      }
      if (!MOM && hf == 1 && lane < 8) dadc[64 + lane] = 0.f;
      tc::fence_before_sync();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + B_H1_FULL);

      if (SDF) {
        // ---- E2: dz2 = du W3 [z2 > 0] -> DZ image ; dW3 += du h2 ; db2 += dz2
        mbar_wait(bar + B_W1_DONE, par ^ 1);            // dW1x MMAs of the previous tile released DZ
        mbar_wait(bar + B_D2_FULL, par);
        tc::fence_after_sync();
        float pw[32], dz[32];
        {
          float v[32];
          tc::tmem_ld32(tbase + lane_base + C_B + U0, v);
          const float4* b4 = reinterpret_cast<const float4*>(s.b2s() + U0);
          const float4* w4 = reinterpret_cast<const float4*>(s.w3s() + U0);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float4 bb = b4[k], ww = w4[k];
            const float b_[4] = {bb.x, bb.y, bb.z, bb.w}, w_[4] = {ww.x, ww.y, ww.z, ww.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float z = v[4 * k + e] + b_[e];
              pw[4 * k + e] = du * fmaxf(z, 0.f);
              dz[4 * k + e] = (z > 0.f) ? du * w_[e] : 0.f;
            }
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 hi, lo;
          tc::split8(dz + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.dzh() + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.dzl() + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
        tc::fence_before_sync();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar + B_DZ2_FULL);
        dw3_acc += lane_transpose_sum(pw, lane);
        db2_acc += lane_transpose_sum(dz, lane);

        // ---- E3: dz1 = (dz2 . W2) [z1 > 0] -> DZ image (over dz2) ; da_t
        mbar_wait(bar + B_D3_FULL, par);
        tc::fence_after_sync();
        tc::tmem_ld32(tbase + lane_base + C_C + U0, dz);
#pragma unroll
        for (int k = 0; k < 32; ++k) dz[k] = ((m1 >> k) & 1u) ? dz[k] : 0.f;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 hi, lo;
          tc::split8(dz + 8 * g, hi, lo);
          *reinterpret_cast<uint4*>(s.dzh() + (((U0 >> 3) + g) * TILE + cell) * 16) = hi;
          *reinterpret_cast<uint4*>(s.dzl() + (((U0 >> 3) + g) * TILE + cell) * 16) = lo;
        }
        tc::fence_before_sync();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar + B_DZ1_FULL);
        dadc[U0 + lane] = lane_transpose_sum(dz, lane);
      } else if (hf == 0) {
        dadc[lane] = 0.f; dadc[32 + lane] = 0.f;
      }
    }

    // ---- per-CTA partials: accumulators live in TMEM (M = 64: row r sits in lane 32*(r/16) + r%16)
    mbar_wait(bar + B_ALL_DONE, 0);
    tc::fence_after_sync();
    float* wp = P.wpart + (size_t)blockIdx.x * wp_stride(Fr);
    const int row = q4 * 16 + lane;                  // valid for lane < 16
    if (SDF) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {                  // dW2[j][k]: this half's 32 columns
        float v[16];
        tc::tmem_ld16(tbase + lane_base + C_W2 + U0 + c * 16, v);
        tc::tmem_wait_ld();
        if (lane < 16 && my_tiles > 0)
          for (int k = 0; k < 16; ++k) wp[wp_off_w2() + row * 64 + U0 + c * 16 + k] = v[k];
        else if (lane < 16)
          for (int k = 0; k < 16; ++k) wp[wp_off_w2() + row * 64 + U0 + c * 16 + k] = 0.f;
      }
      if (hf == 0) {                                 // dW1x[j][f], f < 48
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float v[16];
          tc::tmem_ld16(tbase + lane_base + C_W1 + c * 16, v);
          tc::tmem_wait_ld();
          if (lane < 16)
            for (int k = 0; k < 16; ++k) {
              const int f = c * 16 + k;
              if (f < Fr) wp[wp_off_w1() + row * Fr + f] = (my_tiles > 0 && f < P.F) ? v[k] : 0.f;
            }
        }
      }
    } else {
      for (int k = hf * 32 + lane; k < 64 * 64 / 4; k += 64) wp[wp_off_w2() + q4 * 1024 + k] = 0.f;
      if (hf == 0) for (int k = lane; k < 16 * Fr; k += 32) wp[wp_off_w1() + q4 * 16 * Fr + k] = 0.f;
    }
    if (hf == 1) {                                   // dWq[k][f] = D_Q[f][k]
      float v[8];
      if (MOM) { tc::tmem_ld8(tbase + lane_base + C_Q, v); tc::tmem_wait_ld(); }
      if (lane < 16 && row < Fr)
        for (int k = 0; k < 8; ++k) wp[wp_off_wq(Fr) + k * Fr + row] = (MOM && my_tiles > 0 && row < P.F) ? v[k] : 0.f;
    }
    // dW3 / db2: combine the four lane quarters in a fixed order
    float* red = s.red();
    red[((warp - 6) * 32 + lane) * 2] = dw3_acc;
    red[((warp - 6) * 32 + lane) * 2 + 1] = db2_acc;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (q4 == 0) {
      float a = 0.f, b = 0.f;
      for (int w = 0; w < 4; ++w) {
        const int ww = hf * 4 + ((w + 2) & 3);       // warps 6..9 are quarters 2,3,0,1
        a += red[(ww * 32 + lane) * 2];
        b += red[(ww * 32 + lane) * 2 + 1];
      }
      wp[wp_off_w3(Fr) + U0 + lane] = a;
      wp[wp_off_b2(Fr) + U0 + lane] = b;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tbase, TMEM_COLS);
}

cudaError_t launch_cell_bwd_tc(const CellParams& P, int grid, bool sdf, bool mom, cudaStream_t st) {
  cudaError_t e = cudaSuccess;
#define DLAP_LAUNCH_TC(S, M)                                                                              \
  {                                                                                                       \
    e = cudaFuncSetAttribute(cell_bwd_tc_kernel<S, M>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                             (int)SMEM_B_BYTES);                                                          \
    if (e != cudaSuccess) return e;                                                                       \
    cell_bwd_tc_kernel<S, M><<<grid, NT_B, SMEM_B_BYTES, st>>>(P);                                        \
  }
  if (sdf && mom) DLAP_LAUNCH_TC(true, true)
  else if (sdf) DLAP_LAUNCH_TC(true, false)
  else DLAP_LAUNCH_TC(false, true)
#undef DLAP_LAUNCH_TC
  return cudaGetLastError();
}

}  // namespace dlap
