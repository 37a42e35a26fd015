// Forward per-cell kernel (exact fp32).  See cell_f32.cuh for the tile scheme.
// Replaces, for one [T,N,F] panel: src/model.py:251-271 (SDF MLP on every cell) and
// :514-521 + :132-161 (moment net).  Dense outputs: u [T,N], q [8,T,N].
#include "cell_f32.cuh"

namespace dlap {

template <bool MOM, bool DROP>
__global__ void __launch_bounds__(NT, 1) cell_fwd_f32_kernel(const CellParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmemFwd s;
  carve_fwd(smem_raw, P.F, s);
  const int F = P.F;
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
  const int ntiles = P.T * P.tpm;

  if (tid == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  load_weights_fwd(P, s);
  __syncthreads();

  int tile = blockIdx.x;
  uint32_t phase = 0;
  TileLoad L;
  if (tile < ntiles) {
    L = plan_tile(P, tile / P.tpm, (tile % P.tpm) * TILE);
    issue_tile(L, F, s.stage, s.bar);
  }
  for (; tile < ntiles; tile += gridDim.x) {
    const int t = tile / P.tpm, jt = tile % P.tpm, n0 = jt * TILE;
    if (L.bytes) { mbar_wait(s.bar, phase); phase ^= 1; } else { __syncthreads(); }
    transpose_tile(L, F, s.stage, s.xs);
    if (tid < 64) s.as[tid] = __ldg(P.a_t + (size_t)t * P.months_stride * 64 + tid);
    if (MOM && tid < 8) s.cs[tid] = __ldg(P.c_t + (size_t)t * P.months_stride * 8 + tid);
    __syncthreads();
    const int ncell = L.ncell;
    const int nxt = tile + gridDim.x;
    if (nxt < ntiles) {           // re-arm the copy: it lands while this tile is computed
      L = plan_tile(P, nxt / P.tpm, (nxt % P.tpm) * TILE);
      issue_tile(L, F, s.stage, s.bar);
    }

    float acc[4][8];
    const uint32_t cell0 = (uint32_t)((size_t)t * P.N + n0 + 4 * ty);
    layer1(s, F, ty, tx, acc);
    if (DROP) relu_dropout(P, cell0, 0, tx, acc);
    store_relu(s.hs, ty, tx, acc);
    float qp[4];
    if (MOM) moment_pre(s, F, qp);
    __syncthreads();
    layer2(s, ty, tx, acc);
    if (DROP) relu_dropout(P, cell0, 1, tx, acc);
    // u = b3 + W3.relu(z2): 8 units per thread, then across the 8 tx lanes
    float up[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int o = 0; o < 8; ++o) {
      const float w = s.w3s[unit_of(tx, o)];
#pragma unroll
      for (int c = 0; c < 4; ++c) up[c] = fmaf(fmaxf(acc[c][o], 0.f), w, up[c]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      up[c] += __shfl_xor_sync(0xffffffffu, up[c], 1);
      up[c] += __shfl_xor_sync(0xffffffffu, up[c], 2);
      up[c] += __shfl_xor_sync(0xffffffffu, up[c], 4);
    }
    if (tx == 0) {
      const float b3 = __ldg(P.b3);
#pragma unroll
      for (int c = 0; c < 4; ++c) s.us[4 * ty + c] = up[c] + b3;
    }
    __syncthreads();

    // epilogue: one thread per (cell, half)
    const int cell = tid & (TILE - 1), half = tid >> 7;
    const bool live = cell < ncell;
    const size_t gi = (size_t)t * P.N + n0 + cell;
    if (MOM && live) {
      const size_t plane = (size_t)P.T * P.N;
#pragma unroll
      for (int i = 0; i < 4; ++i) P.q[(size_t)(4 * half + i) * plane + gi] = tanhf(qp[i]);
    }
    if (half == 0 && live) P.u[gi] = s.us[cell];
    __syncthreads();   // xs/hs/us are rewritten by the next tile
  }
}

template <bool MOM, bool DROP>
static cudaError_t launch_fwd_variant(const CellParams& P, int grid, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(cell_fwd_f32_kernel<MOM, DROP>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  cell_fwd_f32_kernel<MOM, DROP><<<grid, NT, smem, st>>>(P);
  return cudaGetLastError();
}

cudaError_t launch_cell_fwd_f32(const CellParams& P, int grid, cudaStream_t st) {
  const size_t smem = smem_fwd_bytes(P.F);
  const bool drop = P.drop_thr != 0;
  if (P.q) return drop ? launch_fwd_variant<true, true>(P, grid, smem, st)
                       : launch_fwd_variant<true, false>(P, grid, smem, st);
  return drop ? launch_fwd_variant<false, true>(P, grid, smem, st)
              : launch_fwd_variant<false, false>(P, grid, smem, st);
}

}  // namespace dlap
