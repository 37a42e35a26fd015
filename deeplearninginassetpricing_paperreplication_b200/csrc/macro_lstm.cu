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

// ---- month-summed gradients ------------------------------------------------------------------------------------
// One warp per output row ("task"), lanes striding over the months, sums in a fixed order:
//   task m < M           dW_ih[r][m] = sum_t dpre[t][r] macro[t][m]  (r < 4H) ; dWq[k][m] = sum_t dc_t[k] macro[t][m]
//   task M + j, j < 64   dW1[j][F+k] = sum_t da_t[j] h_t[k] ; db1[j] = sum_t da_t[j]
//   task M + 64 + r      dW_hh[r][k] = sum_t dpre[t][r] h_{t-1}[k] ; db_ih[r] = db_hh[r] = sum_t dpre[t][r]
//   task M + 64 + 4H + q dbq[q] = sum_t dc_t[q]
// Two acts.  What needs only da_t | dc_t (dWq, dW1's macro-state columns, db1, dbq) is finished in the first, and the
// operands of the second (the task's macro column, h_{t-1}) are fetched into registers; the second act needs the
// LSTM's pre-activation gradients dpre: it stages them in shared memory and is left with FMAs and one reduction.
// As a kernel of its own (macro_wgrad_kernel) the acts simply follow each other; as the second job of the BPTT
// kernel's summing CTAs each act waits for the recurrence CTA's word, so that only the short second act is left
// when the serial chain is through.  da_t | dc_t and dpre are read past L1, never through the read-only path: the
// same grid may have written them a moment ago.
// dynamic smem: dc[T][9] | dpre[T][4H + 1]   (odd pitches: the month-strided reads are conflict-free)
__host__ __device__ constexpr int pow2ceil(int v) { return v <= 1 ? 1 : v <= 2 ? 2 : v <= 4 ? 4 : v <= 8 ? 8 : v <= 16 ? 16 : 32; }
__host__ __device__ constexpr int log2i(int v) { return v <= 1 ? 0 : 1 + log2i(v >> 1); }

// K running sums per lane (K a power of two <= 32) -> their totals over the warp, with K + log2(32/K) shuffles
// instead of 5 K: at every step a lane hands half of its values to its partner and adds the partner's other half.
// Afterwards v[0] of lane l is the total of value number l >> (5 - log2 K).  Fixed order.
template <int K>
__device__ __forceinline__ void lanes_reduce(float (&v)[K], int lane) {
  int off = 16;
#pragma unroll
  for (int k = K; k > 1; k >>= 1, off >>= 1) {
    const int half = k >> 1;
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = upper ? v[i] : v[i + half], keep = upper ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
#pragma unroll
  for (; off >= 1; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
}

template <int H, int NTHR, class WaitDadc, class WaitDpre>
__device__ __forceinline__ void wgrad_tasks(const MacroParams& P, float* sm, int task0, WaitDadc wait_dadc,
                                            WaitDpre wait_dpre) {
  constexpr int G4 = 4 * H, KP = pow2ceil(G4), DP = G4 + 1, VB = 8;
  const int T = P.T, M = P.M, Hr = P.H, G4r = 4 * P.H;      // Hr = 0 (no macro input): only db1 and dbq exist
  float* dcs = sm;                                          // [T][9]
  float* dps = sm + (size_t)T * 9;                          // [T][DP]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int task = task0 + warp;
  const int type = task < M ? 0 : task < M + D1 ? 1 : task < M + D1 + G4r ? 2 : task < M + D1 + G4r + KM ? 3 : 4;
  // ---- first act
  wait_dadc();
  if (P.have_mom)
    for (int i = tid; i < T * KM; i += NTHR) dcs[(i >> 3) * 9 + (i & 7)] = __ldcg(P.dadc + (size_t)(i >> 3) * 72 + 64 + (i & 7));
  __syncthreads();
  float v[VB];                 // type 0: the task's macro column, months lane + 32 i
  float hp[VB][H];             // type 2: h_{t-1}
#pragma unroll
  for (int i = 0; i < VB; ++i) {
    v[i] = 0.f;
#pragma unroll
    for (int k = 0; k < H; ++k) hp[i][k] = 0.f;
  }
  if (type == 0) {
    const int m = task;
#pragma unroll
    for (int i = 0; i < VB; ++i) {
      const int t = lane + 32 * i;
      if (t < T) v[i] = __ldg(P.macro + (size_t)t * M + m);
    }
    if (P.have_mom) {
      float a[KM];
#pragma unroll
      for (int k = 0; k < KM; ++k) a[k] = 0.f;
      for (int base = 0; base < T; base += 32 * VB) {
#pragma unroll
        for (int i = 0; i < VB; ++i) {
          const int t = base + lane + 32 * i;
          if (t < T) {
            const float xv = base == 0 ? v[i] : __ldg(P.macro + (size_t)t * M + m);
#pragma unroll
            for (int k = 0; k < KM; ++k) a[k] = fmaf(dcs[t * 9 + k], xv, a[k]);
          }
        }
      }
      lanes_reduce<KM>(a, lane);
      if ((lane & 3) == 0) P.gWq[(size_t)(lane >> 2) * P.ldwq + m] = a[0];
    }
  } else if (type == 1) {
    if (P.have_sdf) {
      const int j = task - M;
      float a[H], b = 0.f;
#pragma unroll
      for (int k = 0; k < H; ++k) a[k] = 0.f;
#pragma unroll 8
      for (int t = lane; t < T; t += 32) {
        const float d = __ldcg(P.dadc + (size_t)t * 72 + j);
        b += d;
#pragma unroll
        for (int k = 0; k < H; ++k)
          if (k < Hr) a[k] = fmaf(d, __ldg(P.save + (size_t)t * 6 * Hr + k), a[k]);
      }
      b = warp_sum(b);
#pragma unroll
      for (int k = 0; k < H; ++k) a[k] = warp_sum(a[k]);
      if (lane == 0) {
        P.gb1[j] = b;
#pragma unroll
        for (int k = 0; k < H; ++k)
          if (k < Hr) P.gW1[(size_t)j * P.ldw1 + P.F + k] = a[k];
      }
    }
  } else if (type == 2) {
    if (P.have_sdf) {
#pragma unroll
      for (int i = 0; i < VB; ++i) {
        const int t = lane + 32 * i;
        if (t < T) {
#pragma unroll
          for (int k = 0; k < H; ++k)
            hp[i][k] = (t > 0) ? __ldg(P.save + (size_t)(t - 1) * 6 * H + k) : (P.h0 ? __ldg(P.h0 + k) : 0.f);
        }
      }
    }
  } else if (type == 3) {
    if (P.have_mom) {
      const int q = task - M - D1 - G4r;
      float b = 0.f;
      for (int t = lane; t < T; t += 32) b += dcs[t * 9 + q];
      b = warp_sum(b);
      if (lane == 0) P.gbq[q] = b;
    }
  }
  // ---- second act: everything that needs dpre
  if (!P.have_sdf || Hr == 0) return;
  wait_dpre();
  for (int i0 = tid; i0 < T * G4; i0 += 8 * NTHR) {          // eight loads in flight per thread: one round trip
    float ld[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) ld[b] = __ldcg(P.dpre + min(i0 + b * NTHR, T * G4 - 1));
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int i = i0 + b * NTHR;
      if (i < T * G4) dps[(i / G4) * DP + i % G4] = ld[b];
    }
  }
  __syncthreads();
  if (type == 0) {
    const int m = task;
    float a[KP];
#pragma unroll
    for (int r = 0; r < KP; ++r) a[r] = 0.f;
    for (int base = 0; base < T; base += 32 * VB) {
#pragma unroll
      for (int i = 0; i < VB; ++i) {
        const int t = base + lane + 32 * i;
        if (t < T) {
          const float xv = base == 0 ? v[i] : __ldg(P.macro + (size_t)t * M + m);
#pragma unroll
          for (int r = 0; r < G4; ++r) a[r] = fmaf(dps[t * DP + r], xv, a[r]);
        }
      }
    }
    lanes_reduce<KP>(a, lane);
    constexpr int SH = 5 - log2i(KP);
    const int r = lane >> SH;
    if ((lane & ((1 << SH) - 1)) == 0 && r < G4) P.gWih[(size_t)r * M + m] = a[0];
  } else if (type == 2) {
    const int r = task - M - D1;
    float a[H], b = 0.f;
#pragma unroll
    for (int k = 0; k < H; ++k) a[k] = 0.f;
    for (int base = 0; base < T; base += 32 * VB) {
#pragma unroll
      for (int i = 0; i < VB; ++i) {
        const int t = base + lane + 32 * i;
        if (t < T) {
          const float d = dps[t * DP + r];
          b += d;
#pragma unroll
          for (int k = 0; k < H; ++k) {
            const float hv = base == 0 ? hp[i][k]
                                       : ((t > 0) ? __ldg(P.save + (size_t)(t - 1) * 6 * H + k) : 0.f);
            a[k] = fmaf(d, hv, a[k]);
          }
        }
      }
    }
    b = warp_sum(b);
#pragma unroll
    for (int k = 0; k < H; ++k) a[k] = warp_sum(a[k]);
    if (lane == 0) {
      P.gbih[r] = b; P.gbhh[r] = b;
#pragma unroll
      for (int k = 0; k < H; ++k) P.gWhh[r * H + k] = a[k];
    }
  }
}

// H = max(P.H, 1): without macro input (P.H = 0) only db1 and dbq are left
template <int H>
__global__ void __launch_bounds__(256) macro_wgrad_kernel(const MacroParams P) {
  extern __shared__ float sm[];
  wgrad_tasks<H, 256>(P, sm, blockIdx.x * 8, [] {}, [] {});
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
    __nanosleep(100);
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > 2000000000ull) __trap();      // a peer (or a CTA) that never shows up: no hang
  }
}

#define BPTT_STAMP(k)                                                                   \
  do {                                                                                   \
    if (FIN.tl && threadIdx.x == 0) {                                                    \
      unsigned long long now_;                                                           \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));                          \
      FIN.tl[(blockIdx.x + 1 == gridDim.x ? 0 : 16 * (1 + blockIdx.x)) + (k)] = (long long)now_; \
    }                                                                                    \
  } while (0)

template <int H>
__global__ void __launch_bounds__(MT, 1) macro_bptt_kernel(const MacroParams P, const FinishParams FIN) {
  extern __shared__ float sm[];
  BPTT_STAMP(0);
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
    BPTT_STAMP(1);
    if (FIN.fold_wgrad) {
      // Second job: the month-summed weight gradients (what used to be the macro_wgrad launch behind this kernel),
      // sixteen tasks per CTA, paced by two words of the recurrence CTA (rows final; dpre written).  The last CTA
      // through puts the words back to zero.
      wgrad_tasks<H, MT>(
          P, sm, blockIdx.x * (MT / 32),
          [&] {                                     // every da_t | dc_t row is final
            if (threadIdx.x == 0) spin_until(FIN.counter + 3, 1u, false);
            __syncthreads();
            BPTT_STAMP(2);
          },
          [&] {                                     // the chain is through: dpre is written
            if (threadIdx.x == 0) spin_until(FIN.counter + 1, 1u, false);
            __syncthreads();
            BPTT_STAMP(3);
          });
      __syncthreads();
      BPTT_STAMP(4);
      if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(FIN.counter + 2, 1u) == gridDim.x - 2) {
          atomicExch(FIN.counter + 1, 0u); atomicExch(FIN.counter + 2, 0u); atomicExch(FIN.counter + 3, 0u);
        }
      }
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
    if (FIN.fold_wgrad)                                     // the summing CTAs' first act may start
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(FIN.counter + 3), "r"(1u) : "memory");
  }
  __syncthreads();
  BPTT_STAMP(1);
  // pass 1, thread = (month, unit): dhe_t = W1[:,F:]^T . da_t and the local derivatives
  // (measured against a warp per month with the row spread over the lanes and a lane reduction: one round trip
  // instead of two, but the per-month tail on four lanes made the pass 3 us slower)
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, j = i % H;
    const float4* da4 = reinterpret_cast<const float4*>(P.dadc + (size_t)t * 72);
    float v0 = 0.f, v1 = 0.f;
    float4 d4[D1 / 4];                                   // the whole row in flight: one L2 round trip, not four
#pragma unroll
    for (int j4 = 0; j4 < D1 / 4; ++j4) d4[j4] = __ldcg(da4 + j4);      // may have been written by this grid a moment ago
#pragma unroll
    for (int j4 = 0; j4 < D1 / 4; ++j4) {
      const float4 d = d4[j4];
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
  BPTT_STAMP(2);
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
  BPTT_STAMP(3);
  // gate gradients, thread = (month, unit)
  for (int i = tid; i < T * H; i += MT) {
    const int t = i / H, j = i % H;
    const float dht = S[t * 2 * H + j], dcn = S[t * 2 * H + H + j];
    const float* y = Y + (size_t)i * REC;
    float* dp = P.dpre + (size_t)t * G4;
    dp[j] = dcn * y[4]; dp[H + j] = dcn * y[5]; dp[2 * H + j] = dcn * y[6]; dp[3 * H + j] = dht * y[7];
  }
  if (FIN.fold_wgrad) {                              // dpre is complete: release the summing CTAs into their second act
    __syncthreads();
    BPTT_STAMP(4);
    if (tid == 0) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(FIN.counter + 1), "r"(1u) : "memory");
    }
    BPTT_STAMP(5);
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
    if (FIN.fold_wgrad) {                                   // enough CTAs for the month-summed gradients' tasks
      const int per = MT / 32;
      const int tasks = P.M + D1 + 4 * P.H + KM, need = (tasks + per - 1) / per;
      if (extra < need) extra = need;
    }
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
template <int H>
static cudaError_t launch_wgrad_h(const MacroParams& P, cudaStream_t st) {
  const size_t smem = (size_t)P.T * (9 + 4 * H + 1) * 4;
  cudaError_t e = cudaFuncSetAttribute(macro_wgrad_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int tasks = P.M + D1 + 4 * P.H + KM;
  macro_wgrad_kernel<H><<<(tasks + 7) / 8, 256, smem, st>>>(P);
  return cudaGetLastError();
}
static cudaError_t launch_wgrad(const MacroParams& P, cudaStream_t st) {
  if (P.H == 0) return launch_wgrad_h<1>(P, st);
  DLAP_H_SWITCH(launch_wgrad_h, P, st)
}

static long long* g_timeline_bptt = nullptr;
void set_bptt_timeline(long long* p) { g_timeline_bptt = p; }

cudaError_t launch_macro_bwd(const MacroParams& P, const FinishParams& FIN0, cudaStream_t st) {
  FinishParams FIN = FIN0;
  FIN.tl = g_timeline_bptt;
  // the month-summed gradients ride in the BPTT kernel whenever it has summing CTAs beside the recurrence -- and the
  // grid is launched co-resident: those CTAs then wait for the recurrence's CTA, which is the last one of the grid
  FIN.fold_wgrad = (P.have_sdf && P.H > 0 && FIN.dadc_part && FIN.counter && FIN.cooperative) ? 1 : 0;
  if (P.have_sdf && P.H > 0) {
    cudaError_t e = launch_bptt(P, FIN, st);
    if (e != cudaSuccess) return e;
    if (FIN.fold_wgrad) return cudaSuccess;
  } else if (FIN.dadc_part) {           // no recurrence to hide the sums behind: the plain finishing kernel
    cudaError_t e = launch_grad_finish(FIN, st);
    if (e != cudaSuccess) return e;
  }
  return launch_wgrad(P, st);
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
