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
  float* a_t;           // [T,64]
  float* c_t;           // [T,8]
  float* hn;            // [H]
  float* cn;            // [H]
  int* months_ready;    // null, or a device counter: months [0, *months_ready) of a_t (and all of c_t) are written
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

// Forward: ONE CTA of 16 warps.
// dynamic smem: Wih[4H][M] | Wqm[8][M] | xp[T][4H] | hs[T][6H] | cts[T][8] | xready[T] | group mbarriers
//   warps 2,3,5,6,7,9,..,15 (eleven; none shares warp 0's scheduler): input projection, two months per warp and
//             round:  xp[t][r] = W_ih[r].macro_t + b_ih[r] + b_hh[r]  and  c_t[t][q] = Wq[q,:M].macro_t + bq[q],
//             both into shared memory, the next month's row of the series already in flight.  The first months
//             are ready a microsecond after the weights are staged, which is what a per-cell kernel waiting
//             beside this one needs.  (Tried and slower to the first month: a thread per month with transposed
//             weights; a separate multi-CTA projection kernel costs a launch and ~10 us before the recurrence
//             starts, but leaves the recurrence alone on its SM -- 56 us for both against 62 us here.)
//   warp 0    the recurrence, as soon as its month's projection is there; every month's state (h c i f g o) goes
//             to shared memory and is announced in groups of PUB months, one mbarrier per group (an arrive costs
//             the serial chain a few cycles; a fence + flag store costs it a memory barrier every time);
//   warp 1    follows warp 0: copies the states to `save`, the c_t rows to c_t, and turns h_t into a_t rows
//             (the only warp that writes global memory month by month).  With months_ready the
//             progress is published device-wide, so that a per-cell forward kernel launched beside this one
//             (another stream) can start on the first months while the later ones are still being computed.
constexpr int PUB = 4;
constexpr int NPROJ = 11;
__device__ __forceinline__ void st_release_cta(int* p, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}

template <int H>
__global__ void __launch_bounds__(MT, 1) macro_fwd_kernel(const MacroParams P) {
  extern __shared__ __align__(16) float sm[];
  const int T = P.T, M = P.M;
  constexpr int G4 = 4 * H;
  float* wih = sm;                           // [4H][M]
  float* wqm = wih + (size_t)G4 * M;         // [8][M]
  float* xp = wqm + (size_t)KM * M;          // [T][4H]
  float* hs = xp + (size_t)T * G4;           // [T][6H] states of every month (the layout of `save`)
  float* cts = hs + (size_t)T * 6 * H;       // [T][8] c_t rows
  int* xready = reinterpret_cast<int*>(cts + (size_t)T * KM);     // [T] month's projection is in xp / cts
  const int ngroup = (T + PUB - 1) / PUB;    // group c = months [PUB c, PUB c + PUB): its states are in hs
  uint64_t* gbar = reinterpret_cast<uint64_t*>(xready + ((T + 1) & ~1));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < ngroup; i += MT) mbar_init(gbar + i, 1);
  if (tid == 0) fence_mbar_init();
  for (int i = tid; i < T; i += MT) xready[i] = 0;
  for (int i = tid; i < G4 * M; i += MT) wih[i] = __ldg(P.Wih + i);
  for (int i = tid; i < KM * M; i += MT) wqm[i] = __ldg(P.Wq + (size_t)(i / M) * P.ldwq + (i % M));
  __syncthreads();

  if (warp == 0) {
    // Lane j < H owns unit j and evaluates its four gates itself (four independent FMA chains + four independent
    // activations), so one step costs a single round of shuffles (broadcast of h) instead of two.  The next
    // step's input projection is fetched early.
    const int j = lane < H ? lane : 0;
    float whh[4][H];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int k = 0; k < H; ++k) whh[g][k] = __ldg(P.Whh + (g * H + j) * H + k);
    float h = (lane < H && P.h0) ? __ldg(P.h0 + lane) : 0.f;
    float c = (lane < H && P.c0) ? __ldg(P.c0 + lane) : 0.f;
    // Months [0, xavail) are known to be projected; the flags are looked at only when the recurrence has caught
    // up with that, 32 months per look, so a step carries no check it would have to wait for as long as the
    // projection warps stay ahead.
    int xavail = 0;
    auto need = [&](int t) {
      while (xavail <= t) {
        const int m = xavail + lane;
        const int f = (m < T) ? *reinterpret_cast<volatile int*>(xready + m) : 0;
        const unsigned b = __ballot_sync(0xffffffffu, f != 0);
        xavail += (b == 0xffffffffu) ? 32 : __ffs(~b) - 1;
      }
      __threadfence_block();                     // the flags before the xp reads
    };
    need(0);
    float xi = xp[j], xf = xp[H + j], xg = xp[2 * H + j], xo = xp[3 * H + j];
    for (int t = 0; t < T; ++t) {
      float pi = xi, pf = xf, pg = xg, po = xo;
      if (t + 1 < T) {
        if (t + 1 >= xavail) need(t + 1);
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
        float* sv = hs + t * 6 * H;
        sv[lane] = h; sv[H + lane] = c;
        sv[2 * H + lane] = gi; sv[3 * H + lane] = gf; sv[4 * H + lane] = gg; sv[5 * H + lane] = go;
      }
      if ((t % PUB) == PUB - 1 || t == T - 1) {
        __syncwarp();                            // the lanes' hs stores before lane 0's arrive (release)
        if (lane == 0) mbar_arrive(gbar + t / PUB);
      }
    }
    if (lane < H) { P.hn[lane] = h; P.cn[lane] = c; }
  } else if (warp == 1) {
    // a_t = W1[:,F:F+H].h_t + b1 ; lane = units lane and lane + 32, weights in registers
    float w[2][H], b[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int unit = lane + 32 * e;
#pragma unroll
      for (int k = 0; k < H; ++k) w[e][k] = __ldg(P.W1 + (size_t)unit * P.ldw1 + P.F + k);
      b[e] = __ldg(P.b1 + unit);
    }
    for (int c = 0; c < ngroup; ++c) {
      mbar_wait(gbar + c, 0);
      const int done = c * PUB, avail = min(T, done + PUB);
      for (int i = done * 6 * H + lane; i < avail * 6 * H; i += 32) P.save[i] = hs[i];
      for (int i = done * KM + lane; i < avail * KM; i += 32) P.c_t[i] = cts[i];
      for (int t = done; t < avail; ++t) {
        float v0 = b[0], v1 = b[1];
#pragma unroll
        for (int k = 0; k < H; ++k) {
          const float hk = hs[t * 6 * H + k];
          v0 = fmaf(w[0][k], hk, v0); v1 = fmaf(w[1][k], hk, v1);
        }
        P.a_t[(size_t)t * D1 + lane] = v0;
        P.a_t[(size_t)t * D1 + 32 + lane] = v1;
      }
      if (P.months_ready) {
        __syncwarp();                            // the lanes' a_t stores before lane 0's fence + release
        if (lane == 0) {
          __threadfence();
          asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(P.months_ready), "r"(avail) : "memory");
        }
      }
    }
  } else if ((warp & 3) != 0) {
    // input projection, months in pairs (2 pp, 2 pp + 1), pp = pw, pw + NPROJ, ... ; lanes stride over the macro
    // series' columns.  Two months per pass: every weight read from shared memory feeds two FMAs (one month per
    // pass kept the SM's shared-memory pipe busy for the first half of the kernel, and the recurrence warp's own
    // loads and stores queued behind it).  The first 256 columns of the NEXT pair's rows are fetched before this
    // pair's products start.
    const int pw = warp - 2 - (warp >> 2);
    float pre[2][8];
    auto fetch = [&](int t0) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float* row = P.macro + (size_t)min(t0 + e, T - 1) * M;
#pragma unroll
        for (int u = 0; u < 8; ++u) pre[e][u] = (u * 32 + lane < M) ? __ldg(row + u * 32 + lane) : 0.f;
      }
    };
    fetch(2 * pw);
    for (int t0 = 2 * pw; t0 < T; t0 += 2 * NPROJ) {
      float acc[2][G4 + KM];
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int r = 0; r < G4 + KM; ++r) acc[e][r] = 0.f;
      float cur[2][8];
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int u = 0; u < 8; ++u) cur[e][u] = pre[e][u];
      fetch(t0 + 2 * NPROJ);
      const float* row0 = P.macro + (size_t)t0 * M;
      const float* row1 = P.macro + (size_t)min(t0 + 1, T - 1) * M;
      for (int k0 = 0; k0 < M; k0 += 128) {              // products and sums in the order k = k0 + 32 u + lane
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int k = k0 + u * 32 + lane;
          if (k < M) {
            const float v0 = (k0 == 0) ? cur[0][u] : (k0 == 128) ? cur[0][4 + u] : __ldg(row0 + k);
            const float v1 = (k0 == 0) ? cur[1][u] : (k0 == 128) ? cur[1][4 + u] : __ldg(row1 + k);
#pragma unroll
            for (int r = 0; r < G4; ++r) {
              const float w = wih[r * M + k];
              acc[0][r] = fmaf(v0, w, acc[0][r]); acc[1][r] = fmaf(v1, w, acc[1][r]);
            }
#pragma unroll
            for (int q = 0; q < KM; ++q) {
              const float w = wqm[q * M + k];
              acc[0][G4 + q] = fmaf(v0, w, acc[0][G4 + q]); acc[1][G4 + q] = fmaf(v1, w, acc[1][G4 + q]);
            }
          }
        }
      }
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int r = 0; r < G4 + KM; ++r) acc[e][r] = warp_sum(acc[e][r]);
      if (lane < 2 && t0 + lane < T) {                   // lane e finishes month t0 + e
        const int t = t0 + lane;
#pragma unroll
        for (int r = 0; r < G4; ++r) xp[t * G4 + r] = (lane ? acc[1][r] : acc[0][r]) + __ldg(P.bih + r) + __ldg(P.bhh + r);
#pragma unroll
        for (int q = 0; q < KM; ++q) cts[t * KM + q] = (lane ? acc[1][G4 + q] : acc[0][G4 + q]) + __ldg(P.bq + q);
        st_release_cta(xready + t, 1);
      }
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

size_t macro_fwd_smem(int T, int M, int H) {
  return ((size_t)(4 * H + KM) * M + (size_t)T * 10 * H + (size_t)T * KM + ((T + 1) & ~1)) * 4 +
         (size_t)((T + PUB - 1) / PUB) * 8;
}
size_t macro_bwd_smem(int T, int H) {
  const int rec = 8 + 2 * ((H + 3) & ~3);
  return ((size_t)T * 6 * H + (size_t)T * H * rec + (size_t)D1 * H + (size_t)4 * H * H) * 4;
}

template <int H>
static cudaError_t launch_fwd_h(const MacroParams& P, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(macro_fwd_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
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
__global__ void set_months_kernel(int* months_ready, int T) { *months_ready = T; }

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
                               float* xp, int* months_ready, cudaStream_t st) {
  if (plan->dims.M == 0) {
    const int n = plan->dims.T * D1;
    const_bias_kernel<<<(n + 255) / 256, 256, 0, st>>>(p->b1, p->bq, plan->dims.T, a_t, c_t);
    if (months_ready) set_months_kernel<<<1, 1, 0, st>>>(months_ready, plan->dims.T);
    return cudaGetLastError();
  }
  MacroParams P;
  fill_macro(P, plan, macro, p, h0, c0);
  P.save = save; P.a_t = a_t; P.c_t = c_t; P.hn = hn; P.cn = cn; (void)xp; P.months_ready = months_ready;
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
