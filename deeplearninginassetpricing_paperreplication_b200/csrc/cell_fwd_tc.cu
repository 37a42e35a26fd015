// Forward per-cell kernel on the 5th-gen tensor cores (tcgen05 + TMEM), precision "bf16x3".
// Same contract as cell_fwd_f32_kernel (cell_fwd.cu); see cell_tc.cuh for the arithmetic.
//
// Persistent CTA, one per SM, 16 warps with fixed roles; two tiles are in flight (slot = tile & 1):
//   warp 0      TMA producer: 1-D bulk copies of fp32 characteristic slabs into a 2-stage ring
//   warp 1      MMA issuer (one lane): layer 1 of tile i is issued before layer 2 of tile i-1
//   warp 2      TMEM allocator
//   warps 4-7   converter: fp32 slab -> bf16 hi/lo X image (+ ones column), per-month bias column of W1
//   warps 8-11  epilogue of even tiles   } z1 -> relu -> split -> H1 image ; tanh(q) ;
//   warps 12-15 epilogue of odd tiles    } z2 -> relu -> .W3 -> u ; g = q R m ; row sums
#include "cell_tc.cuh"

namespace dlap {
using namespace tcl;

namespace {
constexpr int NT_TC = 512;
constexpr int STAGE_FLOATS = TILE * 46 + 8;

struct SmemTc {
  float* stage[2];
  uint8_t *xh[2], *xl[2], *w1h[2], *w1l[2], *hh[2], *hl[2], *w2h, *w2l;
  float *b2s, *w3s;
  uint64_t *stage_full, *stage_empty, *xop_full, *xop_empty, *d1_full, *h1_full, *d2_full, *d2_empty;
  uint32_t* tmem_slot;
};

constexpr size_t SMEM_TC_BYTES = 2 * (size_t)STAGE_FLOATS * 4 + 4 * X_IMG + 4 * W1_IMG + 4 * H_IMG + 2 * W2_IMG +
                                 128 * 4 + 16 * 8 + 16 + 256;

__device__ __forceinline__ void carve(uint8_t* base, SmemTc& s) {
  uint8_t* p = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(base) + 127) & ~uintptr_t(127));
  for (int i = 0; i < 2; ++i) { s.stage[i] = reinterpret_cast<float*>(p); p += ((STAGE_FLOATS * 4 + 127) & ~127); }
  for (int i = 0; i < 2; ++i) { s.xh[i] = p; p += X_IMG; s.xl[i] = p; p += X_IMG; }
  for (int i = 0; i < 2; ++i) { s.w1h[i] = p; p += W1_IMG; s.w1l[i] = p; p += W1_IMG; }
  for (int i = 0; i < 2; ++i) { s.hh[i] = p; p += H_IMG; s.hl[i] = p; p += H_IMG; }
  s.w2h = p; p += W2_IMG; s.w2l = p; p += W2_IMG;
  s.b2s = reinterpret_cast<float*>(p); p += 64 * 4;
  s.w3s = reinterpret_cast<float*>(p); p += 64 * 4;
  uint64_t* b = reinterpret_cast<uint64_t*>(p);
  s.stage_full = b; s.stage_empty = b + 2; s.xop_full = b + 4; s.xop_empty = b + 6;
  s.d1_full = b + 8; s.h1_full = b + 10; s.d2_full = b + 12; s.d2_empty = b + 14;
  s.tmem_slot = reinterpret_cast<uint32_t*>(b + 16);
}

// TMEM columns
constexpr uint32_t COL_A0 = 0, COL_B0 = 96, COL_A1 = 160, COL_B1 = 256, TMEM_COLS = 512;
}  // namespace

template <bool MOM>
__global__ void __launch_bounds__(NT_TC, 1) cell_fwd_tc_kernel(const CellParams P) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  SmemTc s;
  carve(smem_raw, s);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = P.T * P.tpm;
  const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  // ---- one-time set-up ----
  for (int sl = 0; sl < 2; ++sl) fill_w1_images(P, s.w1h[sl], s.w1l[sl], tid, NT_TC);
  fill_w2_images(P, s.w2h, s.w2l, tid, NT_TC);
  if (tid < 64) { s.b2s[tid] = __ldg(P.b2 + tid); s.w3s[tid] = __ldg(P.W3 + tid); }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(s.stage_full + i, 1);  mbar_init(s.stage_empty + i, 4);
      mbar_init(s.xop_full + i, 4);    mbar_init(s.xop_empty + i, 1);
      mbar_init(s.d1_full + i, 1);     mbar_init(s.h1_full + i, 4);
      mbar_init(s.d2_full + i, 1);     mbar_init(s.d2_empty + i, 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s.tmem_slot, TMEM_COLS);
  fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = *s.tmem_slot;

  if (warp == 0) {
    // ===== producer =====
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = blockIdx.x + i * gridDim.x, st = i & 1, use = i >> 1;
      mbar_wait(s.stage_empty + st, (use & 1) ^ 1);
      const TileLoad L = plan_tile(P, tile / P.tpm, (tile % P.tpm) * TILE);
      __syncwarp();
      if (L.bytes) {
        if (lane == 0) {
          mbar_expect_tx(s.stage_full + st, L.bytes);
          bulk_g2s(s.stage[st], L.src - L.shift, L.bytes, s.stage_full + st);
        }
      } else {                                       // copy would run past the tensor: plain loads
        const int n = L.ncell * P.F;
        for (int j = lane; j < n; j += 32) s.stage[st][L.shift + j] = __ldg(L.src + j);
        __syncwarp();
        if (lane == 0) mbar_arrive(s.stage_full + st);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (all lanes wait, lane 0 issues) =====
    {
      const uint32_t id1 = tc::idesc_bf16(128, N1, 0, 0), id2 = tc::idesc_bf16(128, 64, 0, 0);
      const uint64_t w2h = tc::smem_desc(smem_u32(s.w2h), 64 * 16, 128), w2l = tc::smem_desc(smem_u32(s.w2l), 64 * 16, 128);
      for (int i = 0; i <= my_tiles; ++i) {
        if (i < my_tiles) {                          // layer 1 of tile i
          const int sl = i & 1, use = i >> 1;
          mbar_wait(s.xop_full + sl, use & 1);
          tc::fence_after_sync();
          const uint64_t ah = tc::smem_desc(smem_u32(s.xh[sl]), TILE * 16, 128), al = tc::smem_desc(smem_u32(s.xl[sl]), TILE * 16, 128);
          const uint64_t bh = tc::smem_desc(smem_u32(s.w1h[sl]), N1 * 16, 128), bl = tc::smem_desc(smem_u32(s.w1l[sl]), N1 * 16, 128);
          const uint32_t d = tbase + (sl ? COL_A1 : COL_A0);
          if (lane == 0) {
#pragma unroll
            for (int ks = 0; ks < KG1 / 2; ++ks)
              umma3(d, tc::desc_advance(ah, ks * 2 * TILE * 16), tc::desc_advance(al, ks * 2 * TILE * 16),
                    tc::desc_advance(bh, ks * 2 * N1 * 16), tc::desc_advance(bl, ks * 2 * N1 * 16), id1, ks == 0);
            tc::umma_commit(s.xop_empty + sl);
            tc::umma_commit(s.d1_full + sl);
          }
          __syncwarp();
        }
        if (i > 0) {                                 // layer 2 of tile i-1
          const int j = i - 1, sl = j & 1, use = j >> 1;
          mbar_wait(s.h1_full + sl, use & 1);
          mbar_wait(s.d2_empty + sl, (use & 1) ^ 1);
          tc::fence_after_sync();
          const uint64_t ah = tc::smem_desc(smem_u32(s.hh[sl]), TILE * 16, 128), al = tc::smem_desc(smem_u32(s.hl[sl]), TILE * 16, 128);
          const uint32_t d = tbase + (sl ? COL_B1 : COL_B0);
          if (lane == 0) {
#pragma unroll
            for (int ks = 0; ks < KG2 / 2; ++ks)
              umma3(d, tc::desc_advance(ah, ks * 2 * TILE * 16), tc::desc_advance(al, ks * 2 * TILE * 16),
                    tc::desc_advance(w2h, ks * 2 * 64 * 16), tc::desc_advance(w2l, ks * 2 * 64 * 16), id2, ks == 0);
            tc::umma_commit(s.d2_full + sl);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== converter: thread = cell =====
    const int cell = tid - 128;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = blockIdx.x + i * gridDim.x, st = i & 1, sl = i & 1, use = i >> 1;
      const int t = tile / P.tpm, n0 = (tile % P.tpm) * TILE;
      const TileLoad L = plan_tile(P, t, n0);
      mbar_wait(s.stage_full + st, use & 1);
      mbar_wait(s.xop_empty + sl, (use & 1) ^ 1);
      const float* row = s.stage[st] + L.shift + cell * P.F;
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
        *reinterpret_cast<uint4*>(s.xh[sl] + (g * TILE + cell) * 16) = hi;
        *reinterpret_cast<uint4*>(s.xl[sl] + (g * TILE + cell) * 16) = lo;
      }
      if (cell < N1) {                               // bias column k = 46 of this slot's W1 image
        float b = 0.f;
        if (cell < 64) b = __ldg(P.a_t + (size_t)t * 64 + cell);
        else if (cell < 72 && MOM) b = __ldg(P.c_t + (size_t)t * 8 + cell - 64);
        const __nv_bfloat16 bh = __float2bfloat16_rn(b);
        const __nv_bfloat16 bl = __float2bfloat16_rn(b - __bfloat162float(bh));
        const int off = ((KG1 - 1) * N1 + cell) * 16 + (ONES_K - 8 * (KG1 - 1)) * 2;
        *reinterpret_cast<__nv_bfloat16*>(s.w1h[sl] + off) = bh;
        *reinterpret_cast<__nv_bfloat16*>(s.w1l[sl] + off) = bl;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) { mbar_arrive(s.xop_full + sl); mbar_arrive(s.stage_empty + st); }
    }
  } else if (warp >= 8) {
    // ===== epilogue: group e handles tiles with (i & 1) == e ; thread = TMEM lane = cell =====
    const int e = (warp - 8) >> 2, q4 = warp & 3, cell = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    const float b3 = __ldg(P.b3);
    for (int i = e; i < my_tiles; i += 2) {
      const int tile = blockIdx.x + i * gridDim.x, sl = i & 1, use = i >> 1;
      const int t = tile / P.tpm, jt = tile % P.tpm, n0 = jt * TILE;
      const int ncell = min(TILE, P.N - n0);
      const uint32_t dA = tbase + lane_base + (sl ? COL_A1 : COL_A0), dB = tbase + lane_base + (sl ? COL_B1 : COL_B0);
      // -- after layer 1: h1 = relu(z1) -> bf16 hi/lo image
      mbar_wait(s.d1_full + sl, use & 1);
      tc::fence_after_sync();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[16];
        tc::tmem_ld16(dA + c * 16, v);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = fmaxf(v[k], 0.f);
        uint4 hi, lo;
        tc::split8(v, hi, lo);
        *reinterpret_cast<uint4*>(s.hh[sl] + ((2 * c) * TILE + cell) * 16) = hi;
        *reinterpret_cast<uint4*>(s.hl[sl] + ((2 * c) * TILE + cell) * 16) = lo;
        tc::split8(v + 8, hi, lo);
        *reinterpret_cast<uint4*>(s.hh[sl] + ((2 * c + 1) * TILE + cell) * 16) = hi;
        *reinterpret_cast<uint4*>(s.hl[sl] + ((2 * c + 1) * TILE + cell) * 16) = lo;
      }
      float q[8];
      if (MOM) {
        tc::tmem_ld8(dA + 64, q);
        tc::tmem_wait_ld();
      }
      tc::fence_before_sync();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(s.h1_full + sl);
      if (MOM) {
#pragma unroll
        for (int k = 0; k < 8; ++k) q[k] = tanhf(q[k]);
      }
      // -- after layer 2: u = b3 + W3 . relu(z2 + b2)
      mbar_wait(s.d2_full + sl, use & 1);
      tc::fence_after_sync();
      float u = b3;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[16];
        tc::tmem_ld16(dB + c * 16, v);
        tc::tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 16; ++k) u = fmaf(fmaxf(v[k] + s.b2s[c * 16 + k], 0.f), s.w3s[c * 16 + k], u);
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(s.d2_empty + sl);
      // -- outputs
      const bool live = cell < ncell;
      const size_t gi = (size_t)t * P.N + n0 + cell;
      float m = 0.f, r = 0.f;
      if (live) {
        m = P.mask[gi] ? 1.f : 0.f;
        r = P.R ? __ldg(P.R + gi) : 0.f;
        P.u[gi] = u;
        if (MOM) {
          const size_t plane = (size_t)P.T * P.N;
          const float rm = r * m;
#pragma unroll
          for (int k = 0; k < 8; ++k) P.g[(size_t)k * plane + gi] = q[k] * rm;
        }
      }
      double v0 = 0, v1 = 0, v2 = 0, v3 = 0;
      if (live) {
        const double mu_ = (double)m * (double)u;
        v0 = m; v1 = mu_; v2 = mu_ * (double)r; v3 = (double)m * (double)r;
      }
      v0 = warp_sum(v0); v1 = warp_sum(v1); v2 = warp_sum(v2); v3 = warp_sum(v3);
      if (lane == 0) {
        double* rp = P.row_part + (((size_t)t * P.tpm + jt) * 4 + q4) * 4;
        rp[0] = v0; rp[1] = v1; rp[2] = v2; rp[3] = v3;
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tbase, TMEM_COLS);
}

cudaError_t launch_cell_fwd_tc(const CellParams& P, int grid, cudaStream_t st) {
  cudaError_t e;
  if (P.g) {
    e = cudaFuncSetAttribute(cell_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC_BYTES);
    if (e != cudaSuccess) return e;
    cell_fwd_tc_kernel<true><<<grid, NT_TC, SMEM_TC_BYTES, st>>>(P);
  } else {
    e = cudaFuncSetAttribute(cell_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC_BYTES);
    if (e != cudaSuccess) return e;
    cell_fwd_tc_kernel<false><<<grid, NT_TC, SMEM_TC_BYTES, st>>>(P);
  }
  return cudaGetLastError();
}

}  // namespace dlap
