// Macro path: everything that is per MONTH rather than per stock-month.
//
//   forward  (one persistent CTA): input projection W_ih.macro_t for all months in parallel,
//            then the serial LSTM recurrence on one warp (one lane per gate row, h broadcast by
//            shuffles), then the per-month biases that replace the reference's two
//            tile-over-N concats:  a_t = W1[:,F:].h_t + b1  (src/model.py:253-255, :211)
//                                  c_t = Wq[:,:M].macro_t + bq (src/model.py:514-516, :152).
//            LSTM semantics: torch.nn.LSTM, gates i,f,g,o (src/model.py:39-45, :68).
//   backward (one persistent CTA): dh_t = W1[:,F:]^T.da_t, BPTT on one warp, then the
//            month-summed weight gradients (LSTM, W1[:,F:], b1, Wq[:,:M], bq).
#include "../../include/dlap.h"
#include "dlap_common.cuh"

namespace dlap {

struct MacroParams {
  const float* macro;   // [T,M]
  int T, M, H, F;
  const float* Wih;     // [4H,M]
  const float* Whh;     // [4H,H]
  const float* bih;     // [4H]
  const float* bhh;     // [4H]
  const float* W1; int ldw1;   // [64, ldw1]; macro-state columns start at F
  const float* b1;      // [64]
  const float* Wq; int ldwq;   // [8, ldwq]; macro columns 0..M-1
  const float* bq;      // [8]
  const float* h0;      // [H] or null
  const float* c0;      // [H] or null
  // forward outputs
  float* save;          // [T, 6H]: h(H) c(H) i f g o (4H), post-activation
  float* xp;            // [T, 4H] scratch: input projection (shares the dpre buffer)
  float* a_t;           // [T,64]
  float* c_t;           // [T,8]
  float* hn;            // [H]
  float* cn;            // [H]
  // backward
  const float* dadc;    // [T,72]  da_t (64) | dc_t (8), summed over stocks
  float* dpre;          // [T,4H] scratch: LSTM pre-activation gradients
  int have_sdf, have_mom;
  int wgrad_staged;     // macro_wgrad: h_t / da_t staged in shared memory (fits for T up to ~600)
  float* gWih; float* gWhh; float* gbih; float* gbhh;   // LSTM grads
  float* gW1;           // [64, ldw1]  (columns F.. written here)
  float* gb1;           // [64]
  float* gWq;           // [8, ldwq]   (columns 0..M-1 written here)
  float* gbq;           // [8]
};

// exp2/rcp-based activations for the serial recurrence (absolute error ~2e-7, far inside the parity
// tolerance; the recurrence is contractive so it does not build up over the months)
__device__ __forceinline__ float sigmoid_fast(float v) { return __fdividef(1.f, 1.f + __expf(-v)); }
__device__ __forceinline__ float tanh_fast(float v) { return fmaf(2.f, sigmoid_fast(2.f * v), -1.f); }

constexpr int MT = 512;    // threads of the single-CTA macro kernels

// Input projection for every month (multi-CTA, one warp per month):
//   xp[t][r] = W_ih[r].macro_t + b_ih[r] + b_hh[r]   and   c_t[t][q] = Wq[q,:M].macro_t + bq[q]
// dynamic smem: Wih[4H][M] | Wqm[8][M]
template <int H>
__global__ void __launch_bounds__(256) macro_proj_kernel(const MacroParams P) {
  extern __shared__ float sm[];
  const int T = P.T, M = P.M;
  constexpr int G4 = 4 * H;
  float* wih = sm;
  float* wqm = wih + (size_t)G4 * M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < G4 * M; i += 256) wih[i] = __ldg(P.Wih + i);
  for (int i = tid; i < KM * M; i += 256) wqm[i] = __ldg(P.Wq + (size_t)(i / M) * P.ldwq + (i % M));
  __syncthreads();
  const int t = blockIdx.x * 8 + warp;
  if (t >= T) return;
  float acc[G4 + KM];
#pragma unroll
  for (int r = 0; r < G4 + KM; ++r) acc[r] = 0.f;
  const float* row = P.macro + (size_t)t * M;
  for (int k0 = 0; k0 < M; k0 += 128) {              // four independent loads in flight per lane
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = (k0 + u * 32 + lane < M) ? __ldg(row + k0 + u * 32 + lane) : 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + u * 32 + lane;
      if (k < M) {
#pragma unroll
        for (int r = 0; r < G4; ++r) acc[r] = fmaf(v[u], wih[r * M + k], acc[r]);
#pragma unroll
        for (int q = 0; q < KM; ++q) acc[G4 + q] = fmaf(v[u], wqm[q * M + k], acc[G4 + q]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < G4 + KM; ++r) acc[r] = warp_sum(acc[r]);
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < G4; ++r) P.xp[(size_t)t * G4 + r] = acc[r] + __ldg(P.bih + r) + __ldg(P.bhh + r);
#pragma unroll
    for (int q = 0; q < KM; ++q) P.c_t[(size_t)t * KM + q] = acc[G4 + q] + __ldg(P.bq + q);
  }
}

// The serial part: one CTA.  dynamic smem: xp[T][4H]
template <int H>
__global__ void __launch_bounds__(MT, 1) macro_fwd_kernel(const MacroParams P) {
  extern __shared__ float sm[];
  const int T = P.T;
  constexpr int G4 = 4 * H;
  float* xp = sm;
  float* hs = xp + (size_t)T * G4;           // h_t of every month, for the a_t pass
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < T * G4; i += MT) xp[i] = P.xp[i];
  __syncthreads();

  // phase 2: the recurrence on warp 0.  Lane j < H owns unit j and evaluates its four gates itself
  // (four independent FMA chains + four independent activations), so one step costs a single round of
  // shuffles (broadcast of h) instead of two.  The next step's input projection is fetched early.
  if (warp == 0) {
    const int j = lane < H ? lane : 0;
    float whh[4][H];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int k = 0; k < H; ++k) whh[g][k] = __ldg(P.Whh + (g * H + j) * H + k);
    float h = (lane < H && P.h0) ? __ldg(P.h0 + lane) : 0.f;
    float c = (lane < H && P.c0) ? __ldg(P.c0 + lane) : 0.f;
    float xi = xp[j], xf = xp[H + j], xg = xp[2 * H + j], xo = xp[3 * H + j];
    for (int t = 0; t < T; ++t) {
      float pi = xi, pf = xf, pg = xg, po = xo;
      if (t + 1 < T) {
        const float* nx = xp + (t + 1) * G4;
        xi = nx[j]; xf = nx[H + j]; xg = nx[2 * H + j]; xo = nx[3 * H + j];
      }
#pragma unroll
      for (int k = 0; k < H; ++k) {
        const float hk = __shfl_sync(0xffffffffu, h, k);
        pi = fmaf(whh[0][k], hk, pi); pf = fmaf(whh[1][k], hk, pf);
        pg = fmaf(whh[2][k], hk, pg); po = fmaf(whh[3][k], hk, po);
      }
      const float gi = sigmoid_fast(pi), gf = sigmoid_fast(pf), gg = tanh_fast(pg), go = sigmoid_fast(po);
      c = fmaf(gf, c, gi * gg);
      h = go * tanh_fast(c);
      if (lane < H) {
        float* sv = P.save + (size_t)t * 6 * H;
        hs[t * H + lane] = h;
        sv[lane] = h; sv[H + lane] = c;
        sv[2 * H + lane] = gi; sv[3 * H + lane] = gf; sv[4 * H + lane] = gg; sv[5 * H + lane] = go;
      }
    }
    if (lane < H) { P.hn[lane] = h; P.cn[lane] = c; }
  }
  __syncthreads();

  // phase 3: a_t = W1[:,F:F+H].h_t + b1 ; thread = (unit j, month slice), weights in registers
  {
    const int j = tid % D1;
    float w[H];
#pragma unroll
    for (int k = 0; k < H; ++k) w[k] = __ldg(P.W1 + (size_t)j * P.ldw1 + P.F + k);
    const float b = __ldg(P.b1 + j);
    for (int t = tid / D1; t < T; t += MT / D1) {
      float v = b;
#pragma unroll
      for (int k = 0; k < H; ++k) v = fmaf(w[k], hs[t * H + k], v);
      P.a_t[(size_t)t * D1 + j] = v;
    }
  }
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
template <int H>
__global__ void __launch_bounds__(MT, 1) macro_bptt_kernel(const MacroParams P) {
  extern __shared__ float sm[];
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
  __syncthreads();
  // pass 1, thread = (month, unit): dhe_t = W1[:,F:]^T . da_t and the local derivatives
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, j = i % H;
    const float4* da4 = reinterpret_cast<const float4*>(P.dadc + (size_t)t * 72);
    float v0 = 0.f, v1 = 0.f;
#pragma unroll 4
    for (int j4 = 0; j4 < D1 / 4; ++j4) {
      const float4 d = __ldg(da4 + j4);
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
  // serial chain: lane k < H carries (dh_rec_k, dc_k); S[t] = dht | dcn overwrites the dead saved states
  float* S = sv_s;
  if (warp == 0) {
    const int k = lane < H ? lane : 0;
    float dh_rec = 0.f, dc = 0.f;
    constexpr int R4 = REC / 4, NA = HP / 4;
    const float4* y4 = reinterpret_cast<const float4*>(Y);
    float4 c0, ca[NA], cb[NA];
    auto fetch = [&](int t) {
      const size_t o = ((size_t)t * H + k) * R4;
      c0 = y4[o];
#pragma unroll
      for (int u = 0; u < NA; ++u) { ca[u] = y4[o + 2 + u]; cb[u] = y4[o + 2 + NA + u]; }
    };
    fetch(T - 1);
    for (int t = T - 1; t >= 0; --t) {
      const float4 q0 = c0;
      float A[HP], B[HP];
#pragma unroll
      for (int u = 0; u < NA; ++u) {
        A[4 * u] = ca[u].x; A[4 * u + 1] = ca[u].y; A[4 * u + 2] = ca[u].z; A[4 * u + 3] = ca[u].w;
        B[4 * u] = cb[u].x; B[4 * u + 1] = cb[u].y; B[4 * u + 2] = cb[u].z; B[4 * u + 3] = cb[u].w;
      }
      if (t > 0) fetch(t - 1);
      const float dht = q0.z + dh_rec;
      const float dcn = fmaf(dht, q0.x, dc);
      if (lane < H) { S[t * 2 * H + lane] = dht; S[t * 2 * H + H + lane] = dcn; }
      float r0 = 0.f, r1 = 0.f;
#pragma unroll
      for (int j = 0; j < H; ++j) {
        r0 = fmaf(A[j], __shfl_sync(0xffffffffu, dcn, j), r0);
        r1 = fmaf(B[j], __shfl_sync(0xffffffffu, dht, j), r1);
      }
      dh_rec = r0 + r1;
      dc = dcn * q0.y;
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
  for (int i = tid; i < T * NR; i += 256) {
    const int t = i / NR, r = i % NR;
    coef[t * CP + r] = (r < G4) ? (P.have_sdf ? P.dpre[(size_t)t * G4 + r] : 0.f)
                                : (P.have_mom ? __ldg(P.dadc + (size_t)t * 72 + 64 + (r - G4)) : 0.f);
  }
  __syncthreads();
  const int task = blockIdx.x * 8 + warp;
  if (task < M) {
    const int m = task;
    float acc[4 * HMAX + KM];
#pragma unroll
    for (int r = 0; r < 4 * HMAX + KM; ++r) acc[r] = 0.f;
    for (int t = lane; t < T; t += 32) {
      const float v = __ldg(P.macro + (size_t)t * M + m);
      const float* c = coef + t * CP;
#pragma unroll
      for (int r = 0; r < 4 * HMAX + KM; ++r)
        if (r < NR) acc[r] = fmaf(c[r], v, acc[r]);
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

size_t macro_fwd_smem(int T, int M, int H) { return (size_t)T * 5 * H * 4 + 0 * (size_t)M; }
size_t macro_proj_smem(int M, int H) { return ((size_t)4 * H * M + (size_t)KM * M) * 4; }
size_t macro_bwd_smem(int T, int H) {
  const int rec = 8 + 2 * ((H + 3) & ~3);
  return ((size_t)T * 6 * H + (size_t)T * H * rec + (size_t)D1 * H + (size_t)4 * H * H) * 4;
}

template <int H>
static cudaError_t launch_fwd_h(const MacroParams& P, size_t smem, cudaStream_t st) {
  const size_t psm = macro_proj_smem(P.M, H);
  cudaError_t e = cudaFuncSetAttribute(macro_proj_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psm);
  if (e != cudaSuccess) return e;
  macro_proj_kernel<H><<<(P.T + 7) / 8, 256, psm, st>>>(P);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(macro_fwd_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  macro_fwd_kernel<H><<<1, MT, smem, st>>>(P);
  return cudaGetLastError();
}
template <int H>
static cudaError_t launch_bptt_h(const MacroParams& P, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(macro_bptt_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  macro_bptt_kernel<H><<<1, MT, smem, st>>>(P);
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
static cudaError_t launch_bptt(const MacroParams& P, cudaStream_t st) {
  const size_t smem = macro_bwd_smem(P.T, P.H);
  DLAP_H_SWITCH(launch_bptt_h, P, smem, st)
}
cudaError_t launch_macro_bwd(const MacroParams& P, cudaStream_t st) {
  if (P.have_sdf && P.H > 0) {
    cudaError_t e = launch_bptt(P, st);
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

static void fill_macro(MacroParams& P, const dlap_plan* plan, const float* macro, const dlap_tensors* p,
                       const float* h0, const float* c0) {
  P = MacroParams{};
  P.macro = macro; P.T = plan->dims.T; P.M = plan->dims.M; P.H = plan->dims.M > 0 ? plan->dims.H : 0;
  P.F = plan->dims.F;
  P.Wih = p->w_ih; P.Whh = p->w_hh; P.bih = p->b_ih; P.bhh = p->b_hh;
  P.W1 = p->w1; P.ldw1 = plan->ldw1; P.b1 = p->b1; P.Wq = p->wq; P.ldwq = plan->ldwq; P.bq = p->bq;
  P.h0 = h0; P.c0 = c0;
}

cudaError_t launch_macro_fwd_c(const dlap_plan* plan, const float* macro, const dlap_tensors* p, const float* h0,
                               const float* c0, float* save, float* a_t, float* c_t, float* hn, float* cn,
                               float* xp, cudaStream_t st) {
  if (plan->dims.M == 0) {
    const int n = plan->dims.T * D1;
    const_bias_kernel<<<(n + 255) / 256, 256, 0, st>>>(p->b1, p->bq, plan->dims.T, a_t, c_t);
    return cudaGetLastError();
  }
  MacroParams P;
  fill_macro(P, plan, macro, p, h0, c0);
  P.save = save; P.a_t = a_t; P.c_t = c_t; P.hn = hn; P.cn = cn; P.xp = xp;
  return launch_macro_fwd(P, st);
}

cudaError_t launch_macro_bwd_c(const dlap_plan* plan, const float* macro, const dlap_tensors* p, const float* h0,
                               const float* c0, const float* save, const float* dadc, float* dpre, int want_sdf,
                               int want_mom, const dlap_tensors* g, cudaStream_t st) {
  MacroParams P;
  fill_macro(P, plan, macro, p, h0, c0);
  P.save = const_cast<float*>(save); P.dadc = dadc; P.dpre = dpre;
  P.have_sdf = want_sdf; P.have_mom = want_mom;
  P.gWih = g->w_ih; P.gWhh = g->w_hh; P.gbih = g->b_ih; P.gbhh = g->b_hh;
  P.gW1 = g->w1; P.gb1 = g->b1; P.gWq = g->wq; P.gbq = g->bq;
  return launch_macro_bwd(P, st);
}

}  // namespace dlap
