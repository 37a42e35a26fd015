// Everything of the step that is a sum over stocks (per month) or over months (per stock), in ONE kernel whose
// grid is co-resident and steps through its phases with grid-wide barriers:
//   phase 0  per month:  N_t, mu_t, P_t, Rbar_t from u and R          src/model.py:276-278, :363-367
//   phase 1  N_bar, pr_t, S_t, Sharpe (CTA 0 writes them; every CTA recomputes the S_t it needs);
//            per stock over the months of one chunk: sum_t q_k R S_t, sum_t R m S_t, sum_t m     :376, :382, :428
//            and, fused, the demeaned weights w = m (u - mu_t)                                     :271-279
//   phase 2  the month chunks summed into chunk 0 (all CTAs; with month shards this is what gets all-reduced)
//   phase 3  per stock: e_ki = colsum / T_i, per-block partial sums of e^2, and the UNIT loss coefficients
//            Gc_ki = 2 e_ki / (K N T_i), Gu_i = 2 e0_i / (N T_i)        :382-385, :428-431, autograd of them
//   phase 4  L_c, L_u; per month the two sums the backward pass needs:
//            dSc_t = sum_i R sum_k Gc_ki q_kti,  dSu_t = sum_i Gu_i R m   (dL/dS_t = g_c dSc_t + g_u dSu_t)
// A launch runs a contiguous set of phases (a month-sharded step exchanges the per-stock sums between 2 and 3).
// Replaces six launches of round 1 (row_sum, row_finish, col_partial, loss_finish, coeff, ds_rows).
// Every cross-CTA sum has a fixed order (no float atomics): results are reproducible run to run.
// Data written in one phase and read in a later one is read with ld.cg / ld.acquire, never through the
// read-only (nc) path.
#include "../../include/dlap.h"
#include "dlap_common.cuh"
#include "cell_f32.cuh"

namespace dlap {

struct ReduceParams {
  const float* u;            // packed (index != null) or dense [T,N]
  const float* q;            // [8] planes like u: tanh outputs of the moment net, or null
  const float* R;            // [T,N]
  const uint8_t* mask;       // [T,N]
  const int* index;          // valid-cell index or null (dense u, q)
  int T, N;
  int weighted;
  int no_demean;             // normalize_w = False (src/model.py:274): w = m u; mu_t and Rbar_t reported as 0
  int q_dense;               // q is dense [8,T,N] although u is left-packed
  const float* n_bar_in;     // month shards: the series' N_bar; else null
  int n_total;               // stocks in the loss means
  float* row_stats;          // [T,RS_W]
  float* pr;                 // [T]
  float* scalars;            // [SC_W]
  float* w_out;              // [T,N] or null
  double* col_part;          // [nsplit,10,N]
  int nsplit, tchunk;        // month chunks of phase 1; phase 2 sums them into chunk 0, which phase 3 reads
  float* e; float* t_i; float* G;      // [9,N], [N], [N,GS] (stock-major rows: G_0..G_7, G_u, padding)
  double* loss_part;         // [ceil(N/64), 2]
  double* dS2;               // [T,2] or null (no backward pass will follow)
  double* row_part;          // [T,nseg,4] partial sums of the per-month phases
  int nseg;                  // segments a month's cells are cut into in phases 0 and 4
  unsigned int* bar;         // [1 + T], zero between launches: the grid barrier word, then one counter per month
  // month shards: what crosses the wire between phases 2 and 3, in fp32: [10,N] per-stock sums, then the series'
  // portfolio returns (this rank fills its own months, the rest is zero, so that the all-reduce assembles them)
  float* xbuf;               // [10 N + T_series] or null (single device)
  int T_series, month0, month_stride;
  // stock shards (this rank owns a slice of the stocks, all months): the per-month sums are partial sums.  Phase 0
  // leaves the raw sums in row_sums [T,4] (all-reduced by the caller); a launch that starts at phase 1 first turns
  // them into row_stats; phase 4 leaves the sums of e^2 behind the dS sums, dS2 [T,2] + [2] (all-reduced by the
  // caller); phase 5 turns those two into L_c and L_u.
  int stock_shard;
  double* row_sums;
  // stock shards over peer memory (NVLink): every rank's partial sums sit in a buffer all ranks can address,
  // peer[r] = rank r's, laid out as doubles [4 T | 2 T + 2] and then 2 x 8 uint32 flags.  The two all-reduces become
  // "write mine, barrier across the GPUs, add everybody's in rank order" inside this one launch (phases 0..5).
  int world, rank, Tcap;     // Tcap: months the buffers were laid out for (>= T)
  double* peer[8];
  unsigned int* epoch;       // launches so far (device, this rank's own): the value the flags are raised to
  int phase_lo, phase_hi;    // inclusive
  long long* tl;             // diagnostics (dlap_debug_timeline): globaltimer stamps of CTA 0 and of the last CTA, or null
};

namespace {
constexpr int RT = 256;

#define RED_STAMP(k)                                                                                    \
  do {                                                                                                    \
    if (P.tl && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {                   \
      unsigned long long now_;                                                                            \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));                                           \
      P.tl[(blockIdx.x == 0 ? 0 : 32) + (k)] = (long long)now_;                                           \
    }                                                                                                     \
  } while (0)

__device__ __forceinline__ float S_of(const float* rs, int weighted, float n_bar, float* pr_out) {
  const float P = __ldcg(rs + RS_P), nt = __ldcg(rs + RS_NT);
  const float pr = weighted ? (float)((double)P / (double)nt * (double)n_bar) : P;
  if (pr_out) *pr_out = pr;
  return 1.f + pr;
}

// four doubles summed over the CTA in a fixed tree; result in every thread (two barriers)
__device__ __forceinline__ void block_sum4(double (&v)[4], double* red /* [4][8] */) {
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = warp_sum(v[k]);
  __syncthreads();
  if (l == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) red[k * 8 + w] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < RT / 32; ++i) s += red[k * 8 + i];
    v[k] = s;
  }
}
// Barrier across the GPUs of a stock-sharded step, entered right after a grid barrier (this rank's partial sums are
// written): CTA 0 raises slot `b` of every peer's flag block to this launch's number, and EVERY CTA waits until all
// peers have raised this rank's flags (they sit in this rank's own memory, so the polling is local) -- no second
// grid barrier.  Flags only ever grow, so a rank that is one launch ahead cannot be mistaken for one that has
// arrived.  A peer that never shows up (its process died) turns into a trap after 2 s instead of a hang.
__device__ __forceinline__ void peer_barrier(const ReduceParams& P, int b, unsigned int value) {
  const int r = threadIdx.x;
  if (r < P.world) {
    const size_t flag_off = (size_t)(6 * P.Tcap + 2) * sizeof(double) + (size_t)b * 8 * sizeof(unsigned int);
    if (blockIdx.x == 0) {
      __threadfence_system();
      unsigned int* theirs = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(P.peer[r]) + flag_off) + P.rank;
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(value) : "memory");
    }
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(reinterpret_cast<const char*>(P.peer[P.rank]) + flag_off) + r;
    unsigned int v;
    unsigned long long t0 = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
      if ((int)(v - value) >= 0) break;
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 2000000000ull) __trap();
    }
  }
  __syncthreads();
}
__device__ __forceinline__ double ld_peer(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
}  // namespace

__global__ void __launch_bounds__(RT, 4) step_reduce_kernel(const ReduceParams P) {
  extern __shared__ float sm[];            // phase 1: S_t then mu_t of this CTA's month chunk
  __shared__ double red[32];
  const int tid = threadIdx.x, lane = tid & 31;
  const int T = P.T, N = P.N;
  const size_t plane = (size_t)T * N;
  unsigned int nsync = 0;
  const bool peers = P.stock_shard && P.world > 1;
  const unsigned int launch_no = peers ? __ldcg(P.epoch) + 1u : 0u;

  RED_STAMP(0);
  // ---------------------------------------------------------------- phase 0: row sums
  // A month is cut into nseg segments of its cells so that every CTA has work; the CTA that finishes a month's
  // last segment (a counter per month) adds the segments' partial sums in segment order.
  if (P.phase_lo <= 0 && P.phase_hi >= 0) {
    const int nseg = P.nseg;
    if (P.xbuf && blockIdx.x == gridDim.x - 1)
      for (int t = tid; t < P.T_series; t += RT) P.xbuf[(size_t)(KM + 2) * N + t] = 0.f;
    for (int unit = blockIdx.x; unit < T * nseg; unit += gridDim.x) {
      const int t = unit / nseg, seg = unit % nseg;
      double v[4] = {0.0, 0.0, 0.0, 0.0};          // count, sum u, sum u R, sum R
      const size_t base = (size_t)t * N;
      if (P.index) {
        const int cnt = __ldg(P.index + idx_off_cnt(T) + t);
        const int per = (cnt + nseg - 1) / nseg, j0 = seg * per, j1 = min(cnt, j0 + per);
        const int* vrow = P.index + idx_off_vidx(T) + base;
        for (int jb = j0 + tid; jb < j1; jb += 8 * RT) {        // stock numbers and u of eight cells, then their R
          int ii[8];
          float uu[8], rr[8];
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            const int j = min(jb + b * RT, j1 - 1);
            ii[b] = __ldg(vrow + j);
            uu[b] = __ldcg(P.u + base + j);
          }
#pragma unroll
          for (int b = 0; b < 8; ++b) rr[b] = __ldg(P.R + base + ii[b]);
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            if (jb + b * RT < j1) {
              v[1] += (double)uu[b]; v[2] += (double)uu[b] * (double)rr[b]; v[3] += (double)rr[b];
            }
          }
        }
        v[0] = tid == 0 ? (double)max(j1 - j0, 0) : 0.0;
      } else {
        const int per = (N + nseg - 1) / nseg, i0 = seg * per, i1 = min(N, i0 + per);
#pragma unroll 4
        for (int i = i0 + tid; i < i1; i += RT) {
          const bool live = P.mask[base + i] != 0;
          const float uv = __ldcg(P.u + base + i), r = __ldg(P.R + base + i);
          v[0] += live ? 1.0 : 0.0; v[1] += live ? (double)uv : 0.0;
          v[2] += live ? (double)uv * (double)r : 0.0; v[3] += live ? (double)r : 0.0;
        }
      }
      block_sum4(v, red);
      if (tid == 0) {
        bool last = true;
        if (nseg > 1) {
          double* mine = P.row_part + ((size_t)t * nseg + seg) * 4;
          mine[0] = v[0]; mine[1] = v[1]; mine[2] = v[2]; mine[3] = v[3];
          __threadfence();
          last = atomicAdd(P.bar + 1 + t, 1u) == (unsigned int)(nseg - 1);
          if (last) {
            __threadfence();
            v[0] = v[1] = v[2] = v[3] = 0.0;
            for (int sg = 0; sg < nseg; ++sg) {
              const double* pp = P.row_part + ((size_t)t * nseg + sg) * 4;
              v[0] += __ldcg(pp); v[1] += __ldcg(pp + 1); v[2] += __ldcg(pp + 2); v[3] += __ldcg(pp + 3);
            }
            atomicExch(P.bar + 1 + t, 0u);
          }
        }
        if (last && P.stock_shard) {             // partial sums of this rank's stocks: all-reduced by the caller, or
          double* o = (peers ? P.peer[P.rank] : P.row_sums) + (size_t)t * 4;       // added up over peer memory below
          o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
        } else if (last) {
          const double nt = fmax(v[0], 1.0);
          const float mu = P.no_demean ? 0.f : (float)(v[1] / nt);
          float* rs = P.row_stats + (size_t)t * RS_W;
          rs[RS_NT] = (float)nt;
          rs[RS_MU] = mu;
          rs[RS_P] = (float)(v[2] - (double)mu * v[3]);
          rs[RS_RBAR] = P.no_demean ? 0.f : (float)(v[3] / nt);      // what the backward subtracts from R
          rs[RS_SR] = (float)v[3];
        }
      }
    }
    RED_STAMP(1);
    if (P.phase_hi >= 1) grid_sync(P.bar, ++nsync);
    RED_STAMP(2);
  }

  // ---------------------------------------------------------------- phase 1: N_bar, S_t; column partial sums
  if (P.phase_lo <= 1 && P.phase_hi >= 1) {
    if (peers) peer_barrier(P, 0, launch_no);     // this rank's partial sums are written: wait for everybody's
    if (P.stock_shard && (P.phase_lo == 1 || peers)) {     // the months' sums over ALL stocks: row_stats from them
      for (int t = blockIdx.x * RT + tid; t < T; t += gridDim.x * RT) {
        double v[4];
        if (peers) {
          v[0] = v[1] = v[2] = v[3] = 0.0;
          for (int r = 0; r < P.world; ++r) {            // rank order: every rank gets the same bits
            const double* q = P.peer[r] + (size_t)t * 4;
            v[0] += ld_peer(q); v[1] += ld_peer(q + 1); v[2] += ld_peer(q + 2); v[3] += ld_peer(q + 3);
          }
        } else {
          const double* q = P.row_sums + (size_t)t * 4;
          v[0] = q[0]; v[1] = q[1]; v[2] = q[2]; v[3] = q[3];
        }
        const double nt = fmax(v[0], 1.0);
        const float mu = P.no_demean ? 0.f : (float)(v[1] / nt);
        float* rs = P.row_stats + (size_t)t * RS_W;
        rs[RS_NT] = (float)nt;
        rs[RS_MU] = mu;
        rs[RS_P] = (float)(v[2] - (double)mu * v[3]);
        rs[RS_RBAR] = P.no_demean ? 0.f : (float)(v[3] / nt);
        rs[RS_SR] = (float)v[3];
      }
      grid_sync(P.bar, ++nsync);
    }
    double nsum = 0.0;
    for (int t = tid; t < T; t += RT) nsum += (double)__ldcg(P.row_stats + (size_t)t * RS_W + RS_NT);
    nsum = block_sum(nsum, red);
    const float n_bar = P.n_bar_in ? __ldg(P.n_bar_in) : (float)(nsum / (double)T);
    if (blockIdx.x == 0) {               // pr_t, S_t, Sharpe (src/model.py:360, :367, :380, :551)
      double s1 = 0.0;
      for (int t = tid; t < T; t += RT) {
        float* rs = P.row_stats + (size_t)t * RS_W;
        float pr;
        const float S = S_of(rs, P.weighted, n_bar, &pr);
        rs[RS_PR] = pr; rs[RS_S] = S;
        if (P.xbuf) P.xbuf[(size_t)(KM + 2) * N + P.month0 + (size_t)t * P.month_stride] = pr;
        else P.pr[t] = pr;
        s1 += (double)pr;
      }
      s1 = block_sum(s1, red);
      const double mean = s1 / (double)T;
      double s2 = 0.0;
      for (int t = tid; t < T; t += RT) {
        float pr;
        S_of(P.row_stats + (size_t)t * RS_W, P.weighted, n_bar, &pr);
        const double d = (double)pr - mean;
        s2 += d * d;
      }
      s2 = block_sum(s2, red);
      if (tid == 0) {
        P.scalars[SC_NBAR] = n_bar;
        P.scalars[SC_ZERO] = 0.f;          // what a NULL loss cotangent points at in dlap_cell_bwd
        // (a month shard only knows its own months here: the series' Sharpe ratio follows the exchange, phase 3)
        if (!P.xbuf) P.scalars[SC_SHARPE] = (T > 1) ? (float)(mean / (sqrt(s2 / (double)(T - 1)) + 1e-8)) : nanf("");
      }
    }
    RED_STAMP(3);
    const int nb = (N + RT - 1) / RT, nw = (N + 31) >> 5;
    const int tchunk = P.tchunk;
    for (int unit = blockIdx.x; unit < nb * P.nsplit; unit += gridDim.x) {
      const int bx = unit % nb, by = unit / nb;
      const int t0 = by * tchunk, t1 = min(T, t0 + tchunk);
      __syncthreads();
      for (int t = t0 + tid; t < t1; t += RT) {
        const float* rs = P.row_stats + (size_t)t * RS_W;
        sm[t - t0] = S_of(rs, P.weighted, n_bar, nullptr);
        sm[tchunk + t - t0] = __ldcg(rs + RS_MU);
      }
      __syncthreads();
      const int i = bx * RT + tid;
      const bool in = i < N;                   // whole warps stay: the packed position comes from a ballot
      const int* rank = P.index ? P.index + idx_off_rank(T, N) + (i >> 5) : nullptr;
      // Sums over the chunk's months (at most a few dozen terms) in fp32, handed on as doubles: the chunks are added
      // in fp64 (phase 2).  Doubles here cost ten conversions and ten double adds per stock-month on the quarter-rate
      // pipe of an issue-bound loop (measured: 26 -> 22 us for this phase); the reference sums all T terms in fp32.
      float acc[KM + 2];
#pragma unroll
      for (int k = 0; k < KM + 2; ++k) acc[k] = 0.f;
      // Latency-bound: per month a mask byte and a rank word, then (at a position that depends on them) u, R and
      // eight q values.  Months go in batches of MB: the batch's keys are already in registers (fetched while the
      // previous batch's values were in flight), all of its value loads are issued before anything is stored -- a
      // store to w_out in between would pin every later load behind it, the pointers may alias as far as the
      // compiler knows -- so a batch costs ONE round trip.  Masked lanes load from a clamped in-bounds position and
      // drop the values.
      constexpr int MB = 2;
      const size_t ic = in ? (size_t)i : (size_t)N - 1;
      const unsigned below = (1u << lane) - 1u;
      uint8_t mk[MB];
      int rk[MB];
      auto load_keys = [&](int tb) {
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          const int tt = min(tb + b, t1 - 1);
          mk[b] = P.mask[(size_t)tt * N + ic];
          rk[b] = P.index ? __ldg(rank + (size_t)tt * nw) : 0;
        }
      };
      load_keys(t0);
      for (int t = t0; t < t1; t += MB) {
        bool live[MB];
        uint32_t gi[MB], vi[MB];                 // T * N < 2^32 (dlap_make_plan)
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          const int tt = min(t + b, t1 - 1);
          live[b] = in && (t + b < t1) && mk[b] != 0;
          const unsigned bal = __ballot_sync(0xffffffffu, live[b]);
          gi[b] = (uint32_t)((size_t)tt * N + ic);
          uint32_t v = P.index ? (uint32_t)((size_t)tt * N) + (uint32_t)rk[b] + (uint32_t)__popc(bal & below) : gi[b];
          if ((size_t)v >= plane) v = (uint32_t)(plane - 1);      // a masked lane behind the last listed stock of the last month
          vi[b] = v;
        }
        float uv[MB], r[MB], qv[MB][KM];
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          uv[b] = P.w_out ? __ldcg(P.u + vi[b]) : 0.f;
          r[b] = __ldg(P.R + gi[b]);
#pragma unroll
          for (int k = 0; k < KM; ++k) qv[b][k] = P.q ? __ldcg(P.q + k * plane + (P.q_dense ? gi[b] : vi[b])) : 0.f;
        }
        if (t + MB < t1) load_keys(t + MB);      // the next batch's keys ride with this batch's values
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          if (t + b < t1) {
            const int tl = t + b - t0;
            const float S = sm[tl];
            if (P.w_out && in) P.w_out[gi[b]] = live[b] ? (uv[b] - sm[tchunk + tl]) : 0.f;
            acc[KM] += live[b] ? (r[b] * S) : 0.f;
            acc[KM + 1] += live[b] ? 1.f : 0.f;                     // a count: exact
#pragma unroll
            for (int k = 0; k < KM; ++k) acc[k] += live[b] ? ((qv[b][k] * r[b]) * S) : 0.f;
          }
        }
      }
      if (in) {
        double* out = P.col_part + (size_t)by * (KM + 2) * N + i;
#pragma unroll
        for (int k = 0; k < KM + 2; ++k) out[(size_t)k * N] = (double)acc[k];
      }
    }
    RED_STAMP(4);
    if (P.tl && threadIdx.x == 0) {                      // every CTA: when its phase-1 work was done, and on which SM
      unsigned long long now_; unsigned int smid_;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));
      P.tl[64 + blockIdx.x] = (long long)now_;
      P.tl[64 + 640 + blockIdx.x] = (long long)smid_;
    }
    if (P.phase_hi >= 2 && !(P.phase_hi == 2 && P.nsplit == 1 && !P.xbuf)) grid_sync(P.bar, ++nsync);
    RED_STAMP(5);
  }

  // ---------------------------------------------------------------- phase 2: month chunks -> chunk 0
  // (every CTA takes part: 10 N sums of nsplit terms.  Measured against letting phase 3 add the chunks itself, one
  // thread per stock and one barrier less: 66 us against 71 for the whole kernel at the paper shape)
  if (P.phase_lo <= 2 && P.phase_hi >= 2 && (P.nsplit > 1 || P.xbuf)) {
    const size_t n10 = (size_t)(KM + 2) * N;
    for (size_t idx = (size_t)blockIdx.x * RT + tid; idx < n10; idx += (size_t)gridDim.x * RT) {
      double s = __ldcg(P.col_part + idx);
      for (int sp = 1; sp < P.nsplit; ++sp) s += __ldcg(P.col_part + (size_t)sp * n10 + idx);
      if (P.xbuf) P.xbuf[idx] = (float)s; else P.col_part[idx] = s;
    }
  }

  RED_STAMP(6);
  if (P.phase_lo <= 2 && P.phase_hi >= 3 && (P.nsplit > 1 || P.xbuf)) grid_sync(P.bar, ++nsync);
  RED_STAMP(7);

  // ---------------------------------------------------------------- phase 3: e, T_i, unit coefficients, sum e^2
  if (P.phase_lo <= 3 && P.phase_hi >= 3) {
    const int nblk = (N + RT - 1) / RT;
    for (int blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
      const int i = blk * RT + tid;
      double sc = 0.0, su = 0.0;
      if (i < N) {
        double v[KM + 2];
#pragma unroll
        for (int k = 0; k < KM + 2; ++k)
          v[k] = P.xbuf ? (double)__ldcg(P.xbuf + (size_t)k * N + i) : __ldcg(P.col_part + (size_t)k * N + i);
        const double ti = fmax(v[KM + 1], 1.0);
        const float tif = (float)ti;
        P.t_i[i] = tif;
        const float inv = 1.f / ((float)P.n_total * tif);
        float g[GS];
#pragma unroll
        for (int k = 0; k <= KM; ++k) {
          const float e = (float)(v[k] / ti);
          P.e[(size_t)k * N + i] = e;
          // autograd of mean_k mean_i e^2 and of mean_i e0^2 for a unit cotangent
          g[k] = k < KM ? ((2.f / KM) * e) * inv : (2.f * e) * inv;
          if (k < KM) sc += (double)e * (double)e; else su += (double)e * (double)e;
        }
        g[9] = g[10] = g[11] = 0.f;
        float4* go = reinterpret_cast<float4*>(P.G + (size_t)i * GS);
        go[0] = make_float4(g[0], g[1], g[2], g[3]);
        go[1] = make_float4(g[4], g[5], g[6], g[7]);
        go[2] = make_float4(g[8], g[9], g[10], g[11]);
      }
      __syncthreads();
      sc = block_sum(sc, red);
      su = block_sum(su, red);
      if (tid == 0) { P.loss_part[2 * blk] = sc; P.loss_part[2 * blk + 1] = su; }
    }
    if (P.xbuf && blockIdx.x == gridDim.x - 1) {        // the series' portfolio returns are complete now: Sharpe ratio
      const float* prs = P.xbuf + (size_t)(KM + 2) * N;
      const int Ts = P.T_series;
      double s1 = 0.0;
      for (int t = tid; t < Ts; t += RT) { const float pr = __ldcg(prs + t); P.pr[t] = pr; s1 += (double)pr; }
      __syncthreads();
      s1 = block_sum(s1, red);
      const double mean = s1 / (double)Ts;
      double s2 = 0.0;
      for (int t = tid; t < Ts; t += RT) { const double d = (double)__ldcg(prs + t) - mean; s2 += d * d; }
      s2 = block_sum(s2, red);
      if (tid == 0) P.scalars[SC_SHARPE] = (Ts > 1) ? (float)(mean / (sqrt(s2 / (double)(Ts - 1)) + 1e-8)) : nanf("");
    }
    RED_STAMP(8);
    if (P.phase_hi >= 4) grid_sync(P.bar, ++nsync);
    RED_STAMP(9);
  }

  // ---------------------------------------------------------------- phase 4: losses; per-month dS sums
  if (P.phase_lo <= 4 && P.phase_hi >= 4) {
    if (blockIdx.x == gridDim.x - 1) {
      const int nblk = (N + RT - 1) / RT;
      double a = 0.0, b = 0.0;
      for (int j = tid; j < nblk; j += RT) { a += __ldcg(P.loss_part + 2 * j); b += __ldcg(P.loss_part + 2 * j + 1); }
      __syncthreads();
      a = block_sum(a, red);
      b = block_sum(b, red);
      if (tid == 0 && P.stock_shard) {       // this rank's stocks only: the sums travel behind the dS sums
        double* ds = peers ? P.peer[P.rank] + 4 * (size_t)P.Tcap : P.dS2;
        ds[2 * (size_t)T] = a; ds[2 * (size_t)T + 1] = b;
      } else if (tid == 0) {
        // means run over ALL columns, dead ones included (src/model.py:385)
        P.scalars[SC_LC] = (float)(a / ((double)KM * (double)P.n_total));
        P.scalars[SC_LU] = (float)(b / (double)P.n_total);
      }
    }
    if (P.dS2) {
      const int nseg = P.nseg;
      for (int unit = blockIdx.x; unit < T * nseg; unit += gridDim.x) {
        const int t = unit / nseg, seg = unit % nseg;
        const size_t base = (size_t)t * N;
        double v[4] = {0.0, 0.0, 0.0, 0.0};
        auto cell = [&](int i, size_t vi, bool live) {
          const float r = __ldg(P.R + base + i);
          const float4* gi = reinterpret_cast<const float4*>(P.G + (size_t)i * GS);
          float c = 0.f;
          if (P.q) {
            const float4 ga = __ldcg(gi), gb = __ldcg(gi + 1);
            const float gk[KM] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
#pragma unroll
            for (int k = 0; k < KM; ++k) c = fmaf(gk[k], __ldcg(P.q + k * plane + vi) * r, c);
          }
          const float uu = __ldcg(P.G + (size_t)i * GS + KM) * r;
          v[0] += live ? (double)c : 0.0;
          v[1] += live ? (double)uu : 0.0;
        };
        if (P.index) {
          const int cnt = __ldg(P.index + idx_off_cnt(T) + t);
          const int per = (cnt + nseg - 1) / nseg, j0 = seg * per, j1 = min(cnt, j0 + per);
          const int* vrow = P.index + idx_off_vidx(T) + base;
          // the stock numbers of eight of this thread's cells first (one round trip), then the cells in pairs: the
          // gathers of G and R depend on them, and a loop that fetched number and values cell by cell paid two
          // dependent round trips per pair of cells
          for (int jb = j0 + tid; jb < j1; jb += 8 * RT) {
            int ii[8];
#pragma unroll
            for (int b = 0; b < 8; ++b) ii[b] = __ldg(vrow + min(jb + b * RT, j1 - 1));
#pragma unroll
            for (int b = 0; b < 8; ++b) {
              const int j = jb + b * RT;
              if (j < j1) cell(ii[b], P.q_dense ? base + ii[b] : base + j, true);
            }
          }
        } else {
          const int per = (N + nseg - 1) / nseg, i0 = seg * per, i1 = min(N, i0 + per);
#pragma unroll 2
          for (int i = i0 + tid; i < i1; i += RT) cell(i, base + i, P.mask[base + i] != 0);
        }
        block_sum4(v, red);
        if (tid == 0) {
          bool last = true;
          if (nseg > 1) {
            double* mine = P.row_part + ((size_t)t * nseg + seg) * 4;
            mine[0] = v[0]; mine[1] = v[1];
            __threadfence();
            last = atomicAdd(P.bar + 1 + t, 1u) == (unsigned int)(nseg - 1);
            if (last) {
              __threadfence();
              v[0] = v[1] = 0.0;
              for (int sg = 0; sg < nseg; ++sg) {
                const double* pp = P.row_part + ((size_t)t * nseg + sg) * 4;
                v[0] += __ldcg(pp); v[1] += __ldcg(pp + 1);
              }
              atomicExch(P.bar + 1 + t, 0u);
            }
          }
          if (last) {
            double* ds = peers ? P.peer[P.rank] + 4 * (size_t)P.Tcap : P.dS2;
            ds[2 * t] = v[0]; ds[2 * t + 1] = v[1];
          }
        }
      }
    }
  }
  // ---------------------------------------------------------------- phase 5 (stock shards): L_c, L_u
  if (peers && P.phase_hi >= 5) {                 // the second exchange: dS sums and sums of e^2, over peer memory
    grid_sync(P.bar, ++nsync);
    peer_barrier(P, 1, launch_no);
    for (int i = blockIdx.x * RT + tid; i < 2 * T + 2; i += gridDim.x * RT) {
      double v = 0.0;
      for (int r = 0; r < P.world; ++r) v += ld_peer(P.peer[r] + 4 * (size_t)P.Tcap + i);
      P.dS2[i] = v;
      if (i == 2 * T) P.scalars[SC_LC] = (float)(v / ((double)KM * (double)P.n_total));
      if (i == 2 * T + 1) P.scalars[SC_LU] = (float)(v / (double)P.n_total);
    }
  } else if (P.phase_lo <= 5 && P.phase_hi >= 5 && P.stock_shard && blockIdx.x == 0 && tid == 0) {
    P.scalars[SC_LC] = (float)(P.dS2[2 * (size_t)T] / ((double)KM * (double)P.n_total));
    P.scalars[SC_LU] = (float)(P.dS2[2 * (size_t)T + 1] / (double)P.n_total);
  }
  RED_STAMP(10);
  if (peers && blockIdx.x == 0 && tid == 0) *P.epoch = launch_no;      // read by every CTA at its start, long ago
  grid_exit(P.bar, nsync);
}

static long long* g_timeline_reduce = nullptr;
void set_reduce_timeline(long long* p) { g_timeline_reduce = p; }

static bool peers_in_use(const ReduceParams& P) { return P.stock_shard && P.world > 1; }

// co-resident CTAs of the kernel on `device` with `sms` SMs in use and this much dynamic shared memory
static int reduce_ctas_per_sm(size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, step_reduce_kernel, RT, smem) != cudaSuccess) per_sm = 0;
  return per_sm;
}

cudaError_t launch_step_reduce(ReduceParams P, int sms, bool cooperative, cudaStream_t st) {
  const size_t smem = 2 * (size_t)P.tchunk * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(step_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int per_sm = reduce_ctas_per_sm(smem);
  if (per_sm < 1) return cudaErrorLaunchOutOfResources;
  if (per_sm > 4) per_sm = 4;
  // no more CTAs than any phase has work for
  if (P.phase_lo == 5 && !peers_in_use(P)) {          // two divisions
    step_reduce_kernel<<<1, RT, smem, st>>>(P);
    return cudaGetLastError();
  }
  const int nb = (P.N + RT - 1) / RT;
  int want = nb * P.nsplit;
  if (want < P.T) want = P.T;
  int grid = per_sm * sms;
  // months cut into segments so that the per-month phases fill the grid (at most 8: the scratch is sized for that)
  int nseg = P.row_part ? grid / (P.T > 0 ? P.T : 1) : 1;
  if (nseg > 8) nseg = 8;
  if (nseg < 1) nseg = 1;
  P.nseg = nseg;
  if (want < P.T * nseg) want = P.T * nseg;
  if (grid > want) grid = want;
  if (grid < 1) grid = 1;
  void* args[] = {(void*)&P};
  if (cooperative)
    return cudaLaunchCooperativeKernel((const void*)step_reduce_kernel, dim3(grid), dim3(RT), args, smem, st);
  step_reduce_kernel<<<grid, RT, smem, st>>>(P);
  return cudaGetLastError();
}

cudaError_t launch_step_reduce_c(const dlap_plan* plan, const float* u, const float* q, const float* R, const uint8_t* mask,
                                 const int* index, int weighted, int flags, const float* n_bar_in, int phase_lo, int phase_hi,
                                 float* row_stats, float* pr, float* scalars, float* w_out, double* col_part, float* e,
                                 float* t_i, float* G, double* loss_part, double* dS2, double* row_part,
                                 unsigned int* bar, float* xbuf, int T_series, int month0, int month_stride,
                                 double* row_sums, int n_total, const dlap_peers* peers, cudaStream_t st) {
  ReduceParams P{};
  if (peers) {
    P.world = peers->world; P.rank = peers->rank; P.epoch = peers->epoch; P.Tcap = peers->capacity;
    for (int r = 0; r < peers->world && r < 8; ++r) P.peer[r] = static_cast<double*>(peers->buf[r]);
  }
  P.stock_shard = (flags & DLAP_REDUCE_STOCK_SHARD) != 0; P.row_sums = row_sums;
  P.row_part = row_part;
  P.xbuf = xbuf; P.T_series = T_series; P.month0 = month0; P.month_stride = month_stride;
  P.u = u; P.q = q; P.R = R; P.mask = mask; P.index = index;
  P.T = plan->dims.T; P.N = plan->dims.N;
  P.weighted = weighted; P.n_bar_in = n_bar_in; P.n_total = n_total > 0 ? n_total : plan->dims.N;
  P.no_demean = (flags & DLAP_REDUCE_NO_DEMEAN) != 0; P.q_dense = (flags & DLAP_REDUCE_Q_DENSE) != 0;
  P.row_stats = row_stats; P.pr = pr; P.scalars = scalars; P.w_out = w_out;
  P.col_part = col_part; P.nsplit = plan->col_splits; P.tchunk = (plan->dims.T + plan->col_splits - 1) / plan->col_splits;
  P.e = e; P.t_i = t_i; P.G = G; P.loss_part = loss_part; P.dS2 = dS2; P.bar = bar;
  P.phase_lo = phase_lo; P.phase_hi = phase_hi;
  P.tl = g_timeline_reduce;
  return launch_step_reduce(P, plan->sms, plan->cooperative != 0, st);
}

}  // namespace dlap
