// Macro path, forward: the LSTM over the macro series and the per-month biases that replace the reference's two
// tile-over-N concats, written as a CTA ROLE: macro_fwd_kernel (macro_lstm.cu) runs it as a kernel of its own, the
// tensor-core per-cell forward kernel (cell_fwd_tc.cu) runs it on one of its CTAs, beside the per-cell work that
// consumes the months as they are published.
#pragma once
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
  float* gWih; float* gWhh; float* gbih; float* gbhh;   // LSTM grads
  float* gW1;           // [64, ldw1]  (columns F.. written here)
  float* gb1;           // [64]
  float* gWq;           // [8, ldwq]   (columns 0..M-1 written here)
  float* gbq;           // [8]
};

// exp2/rcp-based activations for the serial recurrence (absolute error ~2e-7, far inside the parity
// tolerance; the recurrence is contractive so it does not build up over the months)
// Written with the two MUFU instructions directly: __expf / __fdividef wrap them in range fix-ups (an FSETP and two
// predicated FMULs per ex2 on the serial chain of the recurrence) that only matter below 2^-126, where 1 + e is 1
// either way; the scale of the argument is one constant (tanh: -2 log2 e) instead of two dependent multiplies.
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoid_fast(float v) { return rcp_ftz(1.f + ex2_ftz(v * -1.4426950408889634f)); }
__device__ __forceinline__ float tanh_fast(float v) {
  return fmaf(2.f, rcp_ftz(1.f + ex2_ftz(v * -2.8853900817779268f)), -1.f);
}

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
constexpr int PUB = 16;
constexpr int NPROJ = 11;
__device__ __forceinline__ void st_release_cta(int* p, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}

template <int H>
__device__ __forceinline__ void macro_fwd_role(const MacroParams& P, float* sm) {
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

// dynamic shared memory of the role, bytes
__host__ __device__ inline size_t macro_fwd_smem_bytes(int T, int M, int H) {
  return ((size_t)(4 * H + KM) * M + (size_t)T * 10 * H + (size_t)T * KM + ((T + 1) & ~1)) * 4 +
         (size_t)((T + PUB - 1) / PUB) * 8;
}

}  // namespace dlap
