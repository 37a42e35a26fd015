// Macro path: everything that is per MONTH rather than per stock-month.
//
//   forward: input projection W_ih.macro_t for all months in parallel (multi-CTA), then one persistent CTA:
//            the serial LSTM recurrence on one warp (one lane per unit, h broadcast by shuffles) and, trailing
//            it on a second warp, the per-month biases that replace the reference's two tile-over-N concats:
//                                  a_t = W1[:,F:].h_t + b1  (src/model.py:253-255, :211)
//                                  c_t = Wq[:,:M].macro_t + bq (src/model.py:514-516, :152, by the projection).
//            The months are published as they complete, so the per-cell forward kernel can run beside it.
//            LSTM semantics: torch.nn.LSTM, gates i,f,g,o (src/model.py:39-45, :68).
//   backward (one persistent CTA): dh_t = W1[:,F:]^T.da_t, BPTT on one warp, then the
//            month-summed weight gradients (LSTM, W1[:,F:], b1, Wq[:,:M], bq).
#include "../../include/dlap.h"
#include "dlap_common.cuh"
#include "macro_role.cuh"
#include "grad_finish.cuh"

namespace dlap {

template <int H>
__global__ void __launch_bounds__(MT, 1) macro_fwd_kernel(const MacroParams P) {
  extern __shared__ __align__(16) float sm[];
  macro_fwd_role<H>(P, sm);
}

// BPTT: one CTA.  Writes dpre[T][4H] (pre-activation gradients) to P.dpre.
// The carried state is (dh_rec, dc) per unit and every step is linear in it:
//   dht = dhe_t + dh_rec ;  dcn = dc + a_c . dht ;  dh_rec' = A_t^T dcn + B_t^T dht ;  dc' = f_t . dcn
// with  A_t[j][k] = Whh_i[j][k] a_i[j] + Whh_f[j][k] a_f[j] + Whh_g[j][k] a_g[j]  and  B_t[j][k] = Whh_o[j][k] a_o[j]
// (a_* = the local derivatives of step t).  Everything that does not depend on the carried state -- dhe_t, the
// a_*, A_t, B_t -- is computed by all threads first; the serial chain on warp 0 is then 2H shuffles + 2H FMAs
// per step and lane, and the gate gradients d{i,f,g} = dcn a_{i,f,g}, do = dht a_o follow in parallel.
// dynamic smem: sv[T][6H] (saved states, later S[T][2H] = dht | dcn) | Y[T][H][REC] | W1m[64][H] | Whh[4H][H]
// record (REC = 8 + 2 HP floats, HP = H rounded up to 4): a_c f dhe . | a_i a_f a_g a_o | A[.][k] | B[.][k]
// With FIN.dadc_part (an unsharded step) the kernel also finishes the per-cell backward kernel's partials, which
// used to be a launch of its own in front of this one: the first gridDim.x - 1 CTAs sum the per-CTA weight-gradient
// partials and the da_t | dc_t partials of every month (grad_finish.cuh) and count themselves off in *FIN.counter;
// the LAST CTA is the recurrence: it stages its tables, waits until the da_t | dc_t rows are complete (the grid is
// co-resident: a cooperative launch, ~50 CTAs), and puts the counter back to zero.
//
// Stock shards over peer memory (FIN.world > 1): what the summing CTAs compute is only this rank's share.  They write
// it into this rank's exchange buffer and count off; the recurrence's CTA waits for all of them, raises this launch's
// flag at every peer and waits for everybody's (the backward pass' one exchange), then releases the summing CTAs,
// which add all ranks' shares in rank order and store the result; the da_t | dc_t ones count off a second time.
__device__ __forceinline__ float ld_peer_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void spin_until(const unsigned int* word, unsigned int value, bool sys) {
  unsigned int v;
  unsigned long long t0 = 0;
  for (;;) {
    if (sys) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
    else asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
    if ((int)(v - value) >= 0) return;
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > 2000000000ull) __trap();      // a peer (or a CTA) that never shows up: no hang
  }
}

template <int H>
__global__ void __launch_bounds__(MT, 1) macro_bptt_kernel(const MacroParams P, const FinishParams FIN) {
  extern __shared__ float sm[];
  const bool peers = FIN.world > 1;
  const unsigned int launch_no = peers ? __ldcg(FIN.gepoch) + 1u : 0u;
  if (blockIdx.x + 1 < gridDim.x) {
    const int stride = wp_stride(round4(FIN.F));
    const int idx = blockIdx.x * MT + threadIdx.x;
    if (peers) {
      float* mine = FIN.gpeer[FIN.rank] + (size_t)(launch_no & 1u) * FIN.gcap;
      const bool w = idx < stride, d = !w && idx - stride < FIN.T * 72;
      if (w) mine[idx] = wpart_sum(FIN, idx);
      else if (d) mine[idx] = dadc_sum(FIN, idx - stride);
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(FIN.counter, 1u);
        spin_until(FIN.gepoch + 1, launch_no, false);          // every rank's shares are in place
      }
      __syncthreads();
      if (w || d) {
        float v = 0.f;
        for (int r = 0; r < FIN.world; ++r) v += ld_peer_f32(FIN.gpeer[r] + (size_t)(launch_no & 1u) * FIN.gcap + idx);
        if (w) wpart_store(FIN, idx, v); else dadc_store(FIN, idx - stride, v);
      }
    } else {
      if (idx < stride) finish_wpart(FIN, idx);
      else if (idx - stride < FIN.T * 72) finish_dadc(FIN, idx - stride);
    }
    if ((int)(blockIdx.x + 1) * MT > stride) {            // this CTA wrote da_t | dc_t elements
      __syncthreads();
      if (threadIdx.x == 0) { __threadfence(); atomicAdd(FIN.counter, 1u); }
    }
    return;
  }
  const int T = P.T;
  constexpr int G4 = 4 * H;
  float* sv_s = sm;                                  // [T][6H]: h, c, i, f, g, o
  constexpr int HP = (H + 3) & ~3, REC = 8 + 2 * HP;
  float* Y = sv_s + (size_t)T * 6 * H;               // [T][H][REC]
  float* w1m = Y + (size_t)T * H * REC;
  float* whh = w1m + D1 * H;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < T * 6 * H; i += MT) sv_s[i] = P.save[i];
  for (int i = tid; i < D1 * H; i += MT) w1m[i] = __ldg(P.W1 + (size_t)(i / H) * P.ldw1 + P.F + (i % H));
  for (int i = tid; i < G4 * H; i += MT) whh[i] = __ldg(P.Whh + i);
  if (peers) {
    if (tid == 0) {                                         // every summing CTA has written this rank's shares
      spin_until(FIN.counter, (unsigned int)FIN.nfin, false);
      atomicExch(FIN.counter, 0u);
    }
    __syncthreads();
    if (tid < FIN.world) {                                  // the exchange: raise our flag at every peer, wait for theirs
      unsigned int* flags_theirs = reinterpret_cast<unsigned int*>(FIN.gpeer[tid] + 2 * (size_t)FIN.gcap) + FIN.rank;
      const unsigned int* flags_mine = reinterpret_cast<const unsigned int*>(FIN.gpeer[FIN.rank] + 2 * (size_t)FIN.gcap) + tid;
      __threadfence_system();
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flags_theirs), "r"(launch_no) : "memory");
      spin_until(flags_mine, launch_no, true);
    }
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(FIN.gepoch + 1), "r"(launch_no) : "memory");
    }
  }
  if (FIN.dadc_part && tid == 0) {                          // the da_t | dc_t rows are being summed by the other CTAs
    spin_until(FIN.counter, (unsigned int)FIN.dadc_ctas, false);
    atomicExch(FIN.counter, 0u);
    if (peers) *FIN.gepoch = launch_no;
  }
  __syncthreads();
  // pass 1, thread = (month, unit): dhe_t = W1[:,F:]^T . da_t and the local derivatives
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, j = i % H;
    const float4* da4 = reinterpret_cast<const float4*>(P.dadc + (size_t)t * 72);
    float v0 = 0.f, v1 = 0.f;
#pragma unroll 4
    for (int j4 = 0; j4 < D1 / 4; ++j4) {
      const float4 d = __ldcg(da4 + j4);                 // may have been written by this grid a moment ago
      v0 = fmaf(w1m[(4 * j4) * H + j], d.x, v0);     v1 = fmaf(w1m[(4 * j4 + 1) * H + j], d.y, v1);
      v0 = fmaf(w1m[(4 * j4 + 2) * H + j], d.z, v0); v1 = fmaf(w1m[(4 * j4 + 3) * H + j], d.w, v1);
    }
    const float* sv = sv_s + (size_t)t * 6 * H;
    const float cc = sv[H + j], gi = sv[2 * H + j], gf = sv[3 * H + j], gg = sv[4 * H + j], go = sv[5 * H + j];
    const float cprev = (t > 0) ? sv_s[(size_t)(t - 1) * 6 * H + H + j] : (P.c0 ? __ldg(P.c0 + j) : 0.f);
    const float tc = tanh_fast(cc);
    float* y = Y + (size_t)i * REC;
    y[0] = go * (1.f - tc * tc);                     // a_c
    y[1] = gf;
    y[2] = v0 + v1;                                  // dhe
    y[4] = gg * gi * (1.f - gi);                     // a_i
    y[5] = cprev * gf * (1.f - gf);                  // a_f
    y[6] = gi * (1.f - gg * gg);                     // a_g
    y[7] = tc * go * (1.f - go);                     // a_o
  }
  __syncthreads();
  // pass 2, thread = (month, unit k): column k of A_t and B_t
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, k = i % H;
    float* y = Y + (size_t)i * REC;
#pragma unroll
    for (int j = 0; j < HP; ++j) {
      float a = 0.f, b = 0.f;
      if (j < H) {
        const float* yj = Y + ((size_t)t * H + j) * REC;
        a = whh[(0 * H + j) * H + k] * yj[4] + whh[(1 * H + j) * H + k] * yj[5] + whh[(2 * H + j) * H + k] * yj[6];
        b = whh[(3 * H + j) * H + k] * yj[7];
      }
      y[8 + j] = a;
      y[8 + HP + j] = b;
    }
  }
  __syncthreads();
  // The chain over the months, as a chunked scan.  One step maps the carried state s = (dh_rec, dc) in R^2H to
  // M_t s + b_t (see above), so a run of steps is again an affine map.  (1) every warp composes the map of its own
  // chunk of months -- lane j < 2H pushes the unit vector e_j through the chunk's steps (column j of the composite
  // matrix), lane 2H pushes the zero vector WITH the inhomogeneous terms (the composite offset); (2) one warp walks
  // the NCH chunk maps from the last chunk to the first and notes the state that enters each chunk; (3) every warp
  // re-runs its chunk from its entry state and writes S[t] = dht | dcn over the dead saved states.  T / NCH + NCH
  // dependent steps instead of T (240 months: 15 + 16 + 15; measured: 27 -> 22 us for the whole kernel, whose
  // rest is staging, the wait for the da_t | dc_t rows and the two passes that build the step records).
  float* S = sv_s;
  constexpr int S2 = 2 * H, NCH = MT / 32;
  float* maps = whh + G4 * H;                        // [NCH][S2 + 1][S2]
  float* entry = maps + NCH * (S2 + 1) * S2;         // [NCH][S2]
  const int L = (T + NCH - 1) / NCH;
  {
    const int t_hi = min(T, (warp + 1) * L) - 1, t_lo = warp * L;
    if (lane <= S2) {
      float dh[H], dc[H];
#pragma unroll
      for (int k = 0; k < H; ++k) { dh[k] = (lane == k) ? 1.f : 0.f; dc[k] = (lane == H + k) ? 1.f : 0.f; }
      const float inhom = (lane == S2) ? 1.f : 0.f;
      for (int t = t_hi; t >= t_lo; --t) {
        float dht[H], dcn[H];
#pragma unroll
        for (int k = 0; k < H; ++k) {
          const float* y = Y + ((size_t)t * H + k) * REC;
          dht[k] = fmaf(inhom, y[2], dh[k]);
          dcn[k] = fmaf(dht[k], y[0], dc[k]);
        }
#pragma unroll
        for (int k = 0; k < H; ++k) {
          const float* y = Y + ((size_t)t * H + k) * REC;
          float r0 = 0.f, r1 = 0.f;
#pragma unroll
          for (int j = 0; j < H; ++j) { r0 = fmaf(y[8 + j], dcn[j], r0); r1 = fmaf(y[8 + HP + j], dht[j], r1); }
          dh[k] = r0 + r1;
          dc[k] = dcn[k] * y[1];
        }
      }
      float* out = maps + ((size_t)warp * (S2 + 1) + lane) * S2;
#pragma unroll
      for (int k = 0; k < H; ++k) { out[k] = dh[k]; out[H + k] = dc[k]; }
    }
  }
  __syncthreads();
  if (warp == 0) {                                   // the state that enters every chunk (the last chunk: zero)
    float s = 0.f;                                   // lane r < 2H holds component r
    for (int c = NCH - 1; c >= 0; --c) {
      if (lane < S2) entry[c * S2 + lane] = s;
      const float* mc = maps + (size_t)c * (S2 + 1) * S2;
      float acc = lane < S2 ? mc[S2 * S2 + lane] : 0.f;          // offset
#pragma unroll
      for (int j = 0; j < S2; ++j) acc = fmaf(lane < S2 ? mc[j * S2 + lane] : 0.f, __shfl_sync(0xffffffffu, s, j), acc);
      s = acc;
    }
  }
  __syncthreads();
  {
    const int t_hi = min(T, (warp + 1) * L) - 1, t_lo = warp * L;
    const int k = lane < H ? lane : 0;
    float dh_rec = entry[warp * S2 + k], dc = entry[warp * S2 + H + k];
    for (int t = t_hi; t >= t_lo; --t) {
      const float* y = Y + ((size_t)t * H + k) * REC;
      const float dht = y[2] + dh_rec;
      const float dcn = fmaf(dht, y[0], dc);
      if (lane < H) { S[t * 2 * H + lane] = dht; S[t * 2 * H + H + lane] = dcn; }
      float r0 = 0.f, r1 = 0.f;
#pragma unroll
      for (int j = 0; j < H; ++j) {
        r0 = fmaf(y[8 + j], __shfl_sync(0xffffffffu, dcn, j), r0);
        r1 = fmaf(y[8 + HP + j], __shfl_sync(0xffffffffu, dht, j), r1);
      }
      dh_rec = r0 + r1;
      dc = dcn * y[1];
    }
  }
  __syncthreads();
  // gate gradients, thread = (month, unit)
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, j = i % H;
    const float dht = S[t * 2 * H + j], dcn = S[t * 2 * H + H + j];
    const float* y = Y + (size_t)i * REC;
    float* dp = P.dpre + (size_t)t * G4;
    dp[j] = dcn * y[4]; dp[H + j] = dcn * y[5]; dp[2 * H + j] = dcn * y[6]; dp[3 * H + j] = dht * y[7];
  }
}

// Month-summed gradients, one warp per output row ("task"), lanes striding over the months, fixed-tree warp
// sums; eight tasks per CTA, the coefficients (dpre | dc_t) staged in shared memory once per CTA:
//   task m < M           dW_ih[r][m] = sum_t dpre[t][r] macro[t][m]  (r < 4H) ; dWq[k][m] = sum_t dc_t[k] macro[t][m]
//   task M + j, j < 64   dW1[j][F+k] = sum_t da_t[j] h_t[k] ; db1[j] = sum_t da_t[j]
//   task M + 64 + r      dW_hh[r][k] = sum_t dpre[t][r] h_{t-1}[k] ; db_ih[r] = db_hh[r] = sum_t dpre[t][r]
//   task M + 64 + 4H + q dbq[q] = sum_t dc_t[q]
// dynamic smem: coef[T][NR + 1]   (NR = 4H + 8; the odd pitch keeps the month-strided reads conflict-free)
__global__ void __launch_bounds__(256) macro_wgrad_kernel(const MacroParams P) {
  extern __shared__ float sm[];
  const int T = P.T, M = P.M, H = P.H, G4 = 4 * H, NR = G4 + KM, CP = NR + 1;
  float* coef = sm;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i0 = tid; i0 < T * NR; i0 += 4 * 256) {         // four independent loads per thread and round
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = min(i0 + u * 256, T * NR - 1), t = i / NR, r = i % NR;
      v[u] = (r < G4) ? (P.have_sdf ? __ldcg(P.dpre + (size_t)t * G4 + r) : 0.f)
                      : (P.have_mom ? __ldg(P.dadc + (size_t)t * 72 + 64 + (r - G4)) : 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      if (i < T * NR) coef[(i / NR) * CP + i % NR] = v[u];
    }
  }
  __syncthreads();
  const int task = blockIdx.x * 8 + warp;
  if (task < M) {
    const int m = task;
    float acc[4 * HMAX + KM];
#pragma unroll
    for (int r = 0; r < 4 * HMAX + KM; ++r) acc[r] = 0.f;
    for (int t0 = lane; t0 < T; t0 += 4 * 32) {              // four months' loads in flight per lane
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(P.macro + (size_t)min(t0 + 32 * u, T - 1) * M + m);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int t = t0 + 32 * u;
        if (t < T) {
          const float* c = coef + t * CP;
#pragma unroll
          for (int r = 0; r < 4 * HMAX + KM; ++r)
            if (r < NR) acc[r] = fmaf(c[r], v[u], acc[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4 * HMAX + KM; ++r) {
      if (r < NR) {
        const float v = warp_sum(acc[r]);
        if (lane == 0) {
          if (r < G4) { if (P.have_sdf) P.gWih[(size_t)r * M + m] = v; }
          else if (P.have_mom) P.gWq[(size_t)(r - G4) * P.ldwq + m] = v;
        }
      }
    }
  } else if (task < M + D1) {
    if (!P.have_sdf) return;
    const int j = task - M;
    float a[HMAX], b = 0.f;
#pragma unroll
    for (int k = 0; k < HMAX; ++k) a[k] = 0.f;
    for (int t = lane; t < T; t += 32) {
      const float d = __ldg(P.dadc + (size_t)t * 72 + j);
      b += d;
#pragma unroll
      for (int k = 0; k < HMAX; ++k)
        if (k < H) a[k] = fmaf(d, __ldg(P.save + (size_t)t * 6 * H + k), a[k]);
    }
    b = warp_sum(b);
#pragma unroll
    for (int k = 0; k < HMAX; ++k) a[k] = warp_sum(a[k]);
    if (lane == 0) {
      P.gb1[j] = b;
#pragma unroll
      for (int k = 0; k < HMAX; ++k)
        if (k < H) P.gW1[(size_t)j * P.ldw1 + P.F + k] = a[k];
    }
  } else if (task < M + D1 + G4) {
    if (!P.have_sdf) return;
    const int r = task - M - D1;
    float a[HMAX], b = 0.f;
#pragma unroll
    for (int k = 0; k < HMAX; ++k) a[k] = 0.f;
    for (int t = lane; t < T; t += 32) {
      const float d = coef[t * CP + r];
      b += d;
#pragma unroll
      for (int k = 0; k < HMAX; ++k)
        if (k < H) {
          const float hp = (t > 0) ? __ldg(P.save + (size_t)(t - 1) * 6 * H + k) : (P.h0 ? __ldg(P.h0 + k) : 0.f);
          a[k] = fmaf(d, hp, a[k]);
        }
    }
    b = warp_sum(b);
#pragma unroll
    for (int k = 0; k < HMAX; ++k) a[k] = warp_sum(a[k]);
    if (lane == 0) {
      P.gbih[r] = b; P.gbhh[r] = b;
#pragma unroll
      for (int k = 0; k < HMAX; ++k)
        if (k < H) P.gWhh[r * H + k] = a[k];
    }
  } else if (task < M + D1 + G4 + KM) {
    if (!P.have_mom) return;
    const int q = task - M - D1 - G4;
    float b = 0.f;
    for (int t = lane; t < T; t += 32) b += coef[t * CP + G4 + q];
    b = warp_sum(b);
    if (lane == 0) P.gbq[q] = b;
  }
}

size_t macro_fwd_smem(int T, int M, int H) { return macro_fwd_smem_bytes(T, M, H); }
size_t macro_bwd_smem(int T, int H) {
  const int rec = 8 + 2 * ((H + 3) & ~3);
  const size_t scan = (size_t)(MT / 32) * (2 * H + 2) * 2 * H;      // chunk maps + entry states of the scan
  return ((size_t)T * 6 * H + (size_t)T * H * rec + (size_t)D1 * H + (size_t)4 * H * H + scan) * 4;
}

template <int H>
static cudaError_t launch_fwd_h(const MacroParams& P, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(macro_fwd_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  macro_fwd_kernel<H><<<1, MT, smem, st>>>(P);
  return cudaGetLastError();
}
template <int H>
static cudaError_t launch_bptt_h(const MacroParams& P, const FinishParams& FIN, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(macro_bptt_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  FinishParams F2 = FIN;
  int extra = 0;
  if (FIN.dadc_part) {
    const int stride = wp_stride(round4(FIN.F)), total = stride + FIN.T * 72;
    extra = (total + MT - 1) / MT;
    F2.dadc_ctas = extra - stride / MT;                     // CTAs whose range reaches past the weight partials
    F2.nfin = extra;
    if (F2.world > 1 && total > F2.gcap) return cudaErrorInvalidValue;
  }
  if (extra > 0 && FIN.cooperative) {
    void* args[] = {(void*)&P, (void*)&F2};
    return cudaLaunchCooperativeKernel((const void*)macro_bptt_kernel<H>, dim3(1 + extra), dim3(MT), args, smem, st);
  }
  macro_bptt_kernel<H><<<1 + extra, MT, smem, st>>>(P, F2);
  return cudaGetLastError();
}
#define DLAP_H_SWITCH(FN, ...)                       \
  switch (P.H) {                                     \
    case 1: return FN<1>(__VA_ARGS__);               \
    case 2: return FN<2>(__VA_ARGS__);               \
    case 3: return FN<3>(__VA_ARGS__);               \
    case 4: return FN<4>(__VA_ARGS__);               \
    case 5: return FN<5>(__VA_ARGS__);               \
    case 6: return FN<6>(__VA_ARGS__);               \
    case 7: return FN<7>(__VA_ARGS__);               \
    case 8: return FN<8>(__VA_ARGS__);               \
    default: return cudaErrorInvalidValue;           \
  }

cudaError_t launch_macro_fwd(const MacroParams& P, cudaStream_t st) {
  const size_t smem = macro_fwd_smem(P.T, P.M, P.H);
  DLAP_H_SWITCH(launch_fwd_h, P, smem, st)
}
static cudaError_t launch_bptt(const MacroParams& P, const FinishParams& FIN, cudaStream_t st) {
  const size_t smem = macro_bwd_smem(P.T, P.H);
  DLAP_H_SWITCH(launch_bptt_h, P, FIN, smem, st)
}
cudaError_t launch_grad_finish(const FinishParams& FIN, cudaStream_t st);      // reduce.cu
// FIN.dadc_part / FIN.wpart non-null: the per-cell kernel's partials have not been summed yet
cudaError_t launch_macro_bwd(const MacroParams& P, const FinishParams& FIN, cudaStream_t st) {
  if (P.have_sdf && P.H > 0) {
    cudaError_t e = launch_bptt(P, FIN, st);
    if (e != cudaSuccess) return e;
  } else if (FIN.dadc_part) {           // no recurrence to hide the sums behind: the plain finishing kernel
    cudaError_t e = launch_grad_finish(FIN, st);
    if (e != cudaSuccess) return e;
  }
  const int NR = 4 * P.H + KM;
  const size_t smem = (size_t)P.T * (NR + 1) * 4;
  cudaError_t e = cudaFuncSetAttribute(macro_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int tasks = P.M + D1 + 4 * P.H + KM;
  macro_wgrad_kernel<<<(tasks + 7) / 8, 256, smem, st>>>(P);
  return cudaGetLastError();
}

// M = 0 (macro_features=None, src/model.py:247-257, :517-518): the per-month biases are just b1 / bq.
__global__ void const_bias_kernel(const float* __restrict__ b1, const float* __restrict__ bq, int T,
                                  float* __restrict__ a_t, float* __restrict__ c_t) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < T * D1) a_t[i] = __ldg(b1 + (i % D1));
  if (i < T * KM) c_t[i] = __ldg(bq + (i % KM));
}
// MacroParams of a series of T months (T may exceed the plan's: a month shard runs the recurrence over the whole
// series and uses its own months' rows of a_t / c_t)
void fill_macro(MacroParams& P, const dlap_plan* plan, int T, const float* macro, const dlap_tensors* p,
                const float* h0, const float* c0) {
  P = MacroParams{};
  P.macro = macro; P.T = T; P.M = plan->dims.M; P.H = plan->dims.M > 0 ? plan->dims.H : 0;
  P.F = plan->dims.F;
  P.Wih = p->w_ih; P.Whh = p->w_hh; P.bih = p->b_ih; P.bhh = p->b_hh;
  P.W1 = p->w1; P.ldw1 = plan->ldw1; P.b1 = p->b1; P.Wq = p->wq; P.ldwq = plan->ldwq; P.bq = p->bq;
  P.h0 = h0; P.c0 = c0;
}

// a_t = b1, c_t = bq for every month (no macro input)
cudaError_t launch_const_bias(const dlap_tensors* p, int T, float* a_t, float* c_t, cudaStream_t st) {
  const int n = T * D1;
  const_bias_kernel<<<(n + 255) / 256, 256, 0, st>>>(p->b1, p->bq, T, a_t, c_t);
  return cudaGetLastError();
}

}  // namespace dlap
