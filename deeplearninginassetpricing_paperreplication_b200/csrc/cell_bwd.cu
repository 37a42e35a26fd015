// Backward per-cell kernel (exact fp32) with activation recompute.  Replaces the autograd
// pass behind src/train.py:84 for everything that is per stock-month:
//   du   = m * dP_t * (R - Rbar_t)                     (demeaning + portfolio return, closed form)
//   dW3 += du.h2 ; dz2 = du W3 [z2>0] ; dW2 += dz2 (x) h1 ; db2 += dz2
//   dz1  = (W2^T dz2) [z1>0] ; dW1x += dz1 (x) x ; da_t += dz1
//   dpre = G_k R m S_t (1 - q^2) ; dWqx += dpre (x) x ; dc_t += dpre
// Nothing saved by the forward is read except per-month / per-stock vectors: the panel is
// streamed again and h1, z2, q are recomputed in shared memory.
#include "cell_f32.cuh"

namespace dlap {

struct SmemBwd {
  float* dzs; float* W2n; float* dps; float* dus; float* cqs;
  float* accw2; float* accw1; float* accwq; float* wred;
};

__host__ __device__ inline size_t smem_bwd_extra_floats(int F) {
  const int Fr = round4(F);
  return (size_t)64 * TILE + 64 * 64 + 8 * TILE + TILE + TILE + 64 * 64 + (size_t)64 * Fr +
         (size_t)8 * Fr + 8 * 72;
}
size_t smem_bwd_bytes(int F) { return smem_fwd_bytes(F) + smem_bwd_extra_floats(F) * 4 + 16; }

__device__ __forceinline__ void carve_bwd(unsigned char* base, int F, SmemBwd& b) {
  const int Fr = round4(F);
  float* p = reinterpret_cast<float*>(
      (reinterpret_cast<uintptr_t>(base) + smem_fwd_bytes(F) + 15) & ~uintptr_t(15));
  b.dzs = p; p += 64 * TILE;
  b.W2n = p; p += 64 * 64;
  b.dps = p; p += 8 * TILE;
  b.dus = p; p += TILE;
  b.cqs = p; p += TILE;
  b.accw2 = p; p += 64 * 64;
  b.accw1 = p; p += 64 * Fr;
  b.accwq = p; p += 8 * Fr;
  b.wred = p;
}

__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
  return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w)));
}

// out[r0+i][c0+j] (4x4 per work item) = sum over the 128 cells of A[row][cell] * B[col][cell];
// A and B are swizzled unit-major tiles (or plain when `a_plain`).  Adds into acc[ld].
template <bool A_PLAIN>
__device__ __forceinline__ void wgrad_4x4(const float* A, const float* B, int ra, int rb,
                                          float* acc, int ld) {
  const float4* A4 = reinterpret_cast<const float4*>(A);
  const float4* B4 = reinterpret_cast<const float4*>(B);
  float r[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) r[i][j] = 0.f;
#pragma unroll 2
  for (int grp = 0; grp < TILE / 4; ++grp) {
    float4 a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      a[i] = A_PLAIN ? A4[(ra + i) * (TILE / 4) + grp] : A4[act_idx4(ra + i, grp)];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = B4[act_idx4(rb + j, grp)];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r[i][j] += dot4(a[i], b[j]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[(ra + i) * ld + rb + j] += r[i][j];
}

template <bool SDF, bool MOM, bool DROP>
__global__ void __launch_bounds__(NT, 1) cell_bwd_f32_kernel(const CellParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmemFwd s;
  SmemBwd b;
  carve_fwd(smem_raw, P.F, s);
  carve_bwd(smem_raw, P.F, b);
  const int F = P.F, Fr = round4(F), Fq = Fr >> 2;
  const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7, warp = tid >> 5, lane = tid & 31;
  const int ntiles = P.T * P.tpm;

  if (tid == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  load_weights_fwd(P, s);
  for (int i = tid; i < 64 * 64; i += NT) b.W2n[i] = __ldg(P.W2 + i);
  for (int i = tid; i < 64 * 64; i += NT) b.accw2[i] = 0.f;
  for (int i = tid; i < 64 * Fr; i += NT) b.accw1[i] = 0.f;
  for (int i = tid; i < 8 * Fr; i += NT) b.accwq[i] = 0.f;
  float dw3_acc[8], db2_acc[8];
#pragma unroll
  for (int o = 0; o < 8; ++o) dw3_acc[o] = db2_acc[o] = 0.f;
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
    if (tid >= 64 && tid < 72) s.cs[tid - 64] = __ldg(P.c_t + (size_t)t * P.months_stride * 8 + tid - 64);
    if (tid == 96) {
      const float* rs = P.row_stats + (size_t)t * RS_W;
      s.sc[0] = month_dP(P, t);
      s.sc[1] = rs[RS_RBAR];
      s.sc[2] = rs[RS_S];
      s.sc[3] = __ldg(P.cot[0]);
      s.sc[4] = P.res_b ? __ldg(P.res_b + t) : 0.f;
      s.sc[5] = rs[RS_MU];
    }
    __syncthreads();
    const int ncell = L.ncell;
    const int nxt = tile + gridDim.x;
    if (nxt < ntiles) {
      L = plan_tile(P, nxt / P.tpm, (nxt % P.tpm) * TILE);
      issue_tile(L, F, s.stage, s.bar);
    }

    // per-cell scalars
    const int cell = tid & (TILE - 1), half = tid >> 7;
    const bool live = cell < ncell;
    const size_t gi = (size_t)t * P.N + n0 + cell;
    if (half == 0) {
      float m = 0.f, r = 0.f;
      if (live) { m = P.mask[gi] ? 1.f : 0.f; r = __ldg(P.R + gi); }
      float du = m * s.sc[0] * (r - s.sc[1]);
      if (P.res_b && live) du = fmaf(m * s.sc[4], __ldg(P.u_fwd + gi) - s.sc[5], du);     // residual loss (src/model.py:435-483)
      b.dus[cell] = du;
      b.cqs[cell] = r * m * s.sc[2];
    }

    float acc[4][8];
    const uint32_t cell0 = (uint32_t)((size_t)t * P.N + n0 + 4 * ty);
    if (SDF) {
      layer1(s, F, ty, tx, acc);
      if (DROP) relu_dropout(P, cell0, 0, tx, acc);   // hs holds h1 AFTER dropout
      store_relu(s.hs, ty, tx, acc);
    }
    float qp[4];
    if (MOM) moment_pre(s, F, qp);
    __syncthreads();                       // hs, dus, cqs visible

    if (MOM) {
      // dpre_k = G_k[n] * R m S_t * (1 - q^2)   (src/model.py:152-153, :428)
      const float cq = b.cqs[cell];
      float dp[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float q = tanhf(qp[i]);
        const float gk = live ? s.sc[3] * __ldg(P.G + (size_t)(n0 + cell) * GS + 4 * half + i) : 0.f;
        dp[i] = gk * cq * (1.f - q * q);
        b.dps[(4 * half + i) * TILE + cell] = dp[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float v = warp_sum(dp[i]);
        if (lane == 0) b.wred[64 * 8 + warp * 4 + i] = v;
      }
    }

    if (SDF) {
      layer2(s, ty, tx, acc);
      if (DROP) relu_dropout(P, cell0, 1, tx, acc);   // acc = h2 after dropout (>0 iff kept & z2>0)
      const float dscale = DROP ? P.drop_scale : 1.f;
      const float4 du4 = reinterpret_cast<const float4*>(b.dus)[ty];
      const float du[4] = {du4.x, du4.y, du4.z, du4.w};
      float4* dz4 = reinterpret_cast<float4*>(b.dzs);
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        const int j = unit_of(tx, o);
        const float w3 = s.w3s[j];
        float dz[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const float z = acc[c][o];
          dw3_acc[o] = fmaf(du[c], fmaxf(z, 0.f), dw3_acc[o]);
          dz[c] = (z > 0.f) ? du[c] * w3 * dscale : 0.f;
          db2_acc[o] += dz[c];
        }
        dz4[act_idx4(j, ty)] = make_float4(dz[0], dz[1], dz[2], dz[3]);
      }
    }
    __syncthreads();                       // dzs, dps, wred(dc) visible

    if (MOM && tid < 8) {
      const int hf = tid >> 2, i = tid & 3;
      float v = 0.f;
      for (int w = 0; w < 4; ++w) v += b.wred[64 * 8 + (hf * 4 + w) * 4 + i];
      P.dadc_part[((size_t)t * P.tpm + jt) * 72 + 64 + tid] = v;
    }
    if (!MOM && tid < 8) P.dadc_part[((size_t)t * P.tpm + jt) * 72 + 64 + tid] = 0.f;

    if (SDF) {
      // dW2[j][k] += sum_c dz2[c][j] h1[c][k]
      wgrad_4x4<false>(b.dzs, s.hs, 4 * (tid >> 4), 4 * (tid & 15), b.accw2, 64);
      // dz1 = (dz2 . W2) [h1 > 0]
#pragma unroll
      for (int o = 0; o < 8; ++o)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[c][o] = 0.f;
      const float4* dz4 = reinterpret_cast<const float4*>(b.dzs);
      const float4* w4 = reinterpret_cast<const float4*>(b.W2n);
#pragma unroll 4
      for (int j = 0; j < 64; ++j) {
        const float4 dv = dz4[act_idx4(j, ty)];
        const float4 w0 = w4[j * 16 + tx], w1 = w4[j * 16 + 8 + tx];
        DLAP_FMA4x8(acc, dv, w0, w1);
      }
      __syncthreads();                     // every read of h1 (wgrad2) done: overwrite in place
      float4* hs4 = reinterpret_cast<float4*>(s.hs);
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        const int j = unit_of(tx, o);
        const float4 h = hs4[act_idx4(j, ty)];
        float4 d;
        const float ds1 = DROP ? P.drop_scale : 1.f;   // h1 > 0 iff kept and z1 > 0
        d.x = h.x > 0.f ? acc[0][o] * ds1 : 0.f; d.y = h.y > 0.f ? acc[1][o] * ds1 : 0.f;
        d.z = h.z > 0.f ? acc[2][o] * ds1 : 0.f; d.w = h.w > 0.f ? acc[3][o] * ds1 : 0.f;
        hs4[act_idx4(j, ty)] = d;
        float sum = (d.x + d.y) + (d.z + d.w);          // da_t partial over this thread's cells
        sum += __shfl_xor_sync(0xffffffffu, sum, 8);
        sum += __shfl_xor_sync(0xffffffffu, sum, 16);
        if (lane < 8) b.wred[warp * 64 + j] = sum;
      }
      __syncthreads();                     // dz1 tile + da partials visible
      if (tid < 64) {
        float v = 0.f;
        for (int w = 0; w < 8; ++w) v += b.wred[w * 64 + tid];
        P.dadc_part[((size_t)t * P.tpm + jt) * 72 + tid] = v;
      }
    } else if (tid < 64) {
      P.dadc_part[((size_t)t * P.tpm + jt) * 72 + tid] = 0.f;
    }

    // dW1x[j][f] += sum_c dz1[c][j] x[c][f]  and  dWqx[k][f] += sum_c dpre[c][k] x[c][f]
    const int work_sdf = SDF ? 16 * Fq : 0, work_all = work_sdf + (MOM ? 2 * Fq : 0);
    for (int w = tid; w < work_all; w += NT) {
      if (w < work_sdf)
        wgrad_4x4<false>(s.hs, s.xs, 4 * (w / Fq), 4 * (w % Fq), b.accw1, Fr);
      else
        wgrad_4x4<true>(b.dps, s.xs, 4 * ((w - work_sdf) / Fq), 4 * ((w - work_sdf) % Fq),
                        b.accwq, Fr);
    }
    __syncthreads();                       // tile buffers are rewritten next iteration
  }

  // ---- per-CTA partials out ----------------------------------------------------------------
  float* wp = P.wpart + (size_t)blockIdx.x * wp_stride(Fr);
  for (int i = tid; i < 64 * 64; i += NT) wp[wp_off_w2() + i] = b.accw2[i];
  for (int i = tid; i < 64 * Fr; i += NT) wp[wp_off_w1() + i] = b.accw1[i];
  for (int i = tid; i < 8 * Fr; i += NT) wp[wp_off_wq(Fr) + i] = b.accwq[i];
  __syncthreads();
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    float a = dw3_acc[o], c = db2_acc[o];
    a += __shfl_xor_sync(0xffffffffu, a, 8);  a += __shfl_xor_sync(0xffffffffu, a, 16);
    c += __shfl_xor_sync(0xffffffffu, c, 8);  c += __shfl_xor_sync(0xffffffffu, c, 16);
    if (lane < 8) {
      b.wred[warp * 64 + unit_of(tx, o)] = a;
      b.dzs[warp * 64 + unit_of(tx, o)] = c;
    }
  }
  __syncthreads();
  if (tid < 64) {
    float a = 0.f, c = 0.f;
    for (int w = 0; w < 8; ++w) { a += b.wred[w * 64 + tid]; c += b.dzs[w * 64 + tid]; }
    wp[wp_off_w3(Fr) + tid] = a;
    wp[wp_off_b2(Fr) + tid] = c;
  }
}

template <bool SDF, bool MOM, bool DROP>
static cudaError_t launch_bwd_variant(const CellParams& P, int grid, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(cell_bwd_f32_kernel<SDF, MOM, DROP>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  cell_bwd_f32_kernel<SDF, MOM, DROP><<<grid, NT, smem, st>>>(P);
  return cudaGetLastError();
}

cudaError_t launch_cell_bwd_f32(const CellParams& P, int grid, bool sdf, bool mom, cudaStream_t st) {
  const size_t smem = smem_bwd_bytes(P.F);
  const bool drop = P.drop_thr != 0 && sdf;
  if (sdf && mom) return drop ? launch_bwd_variant<true, true, true>(P, grid, smem, st)
                              : launch_bwd_variant<true, true, false>(P, grid, smem, st);
  if (sdf) return drop ? launch_bwd_variant<true, false, true>(P, grid, smem, st)
                       : launch_bwd_variant<true, false, false>(P, grid, smem, st);
  return launch_bwd_variant<false, true, false>(P, grid, smem, st);
}

}  // namespace dlap
