// Row (per month over N) and column (per stock over T) reductions, loss and gradient finishing
// kernels.  All HBM-bound or tiny; every cross-CTA sum is done in a fixed order (no float
// atomics) so that results are run-to-run reproducible like the reference's CPU path.
#include <mutex>

#include "../../include/dlap.h"
#include "dlap_common.cuh"
#include "cell_f32.cuh"

namespace dlap {

// ---- row finish ------------------------------------------------------------------------------
// N_t = clamp(sum m, 1), mu_t = sum(m u)/N_t                      src/model.py:276-278
// P_t = sum m (u - mu_t) R = sum(m u R) - mu_t sum(m R)           src/model.py:363,367
// pr_t = P_t / N_t * N_bar (weighted_loss) ; S_t = 1 + pr_t       src/model.py:360,367,380
// sharpe = mean(pr) / (std_unbiased(pr) + 1e-8)                   src/model.py:551
// stage 1 (multi-CTA): one warp per month sums the per-tile partial slots in a fixed tree
// `index` != null (tensor-core kernels): the partials of month t are the 4 slots of each of its valid-cell
// tiles, tile_start[t] .. tile_start[t+1]; otherwise `tpm` slots per month.
__global__ void __launch_bounds__(256) row_sum_kernel(const double* __restrict__ row_part, int T, int tpm,
                                                      const int* __restrict__ index, float* __restrict__ row_stats) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * 8 + warp;
  if (t >= T) return;
  size_t first = (size_t)t * tpm;
  if (index) {
    const int* ts = index + idx_off_ts();
    first = (size_t)ts[t] * 4;
    tpm = (ts[t + 1] - ts[t]) * 4;
  }
  double c = 0, su = 0, sur = 0, sr = 0;
  const double4* p = reinterpret_cast<const double4*>(row_part) + first;
  for (int j = lane; j < tpm; j += 32) {
    const double4 v = p[j];
    c += v.x; su += v.y; sur += v.z; sr += v.w;
  }
  c = warp_sum(c); su = warp_sum(su); sur = warp_sum(sur); sr = warp_sum(sr);
  if (lane == 0) {
    const double nt = fmax(c, 1.0);
    const float mu = (float)(su / nt);
    float* rs = row_stats + (size_t)t * RS_W;
    rs[RS_NT] = (float)nt;
    rs[RS_MU] = mu;
    rs[RS_P] = (float)(sur - (double)mu * sr);
    rs[RS_RBAR] = (float)(sr / nt);
  }
}

// stage 2 (one CTA): N_bar, pr_t, S_t, Sharpe from the T per-month rows
__global__ void __launch_bounds__(1024) row_finish_kernel(
    int T, int weighted, const float* __restrict__ n_bar_in, float* __restrict__ row_stats,
    float* __restrict__ pr_out, float* __restrict__ scalars) {
  __shared__ double red[32];
  const int tid = threadIdx.x;
  double nsum = 0.0;
  for (int t = tid; t < T; t += blockDim.x) nsum += (double)row_stats[(size_t)t * RS_W + RS_NT];
  nsum = block_sum(nsum, red);
  const float n_bar = n_bar_in ? *n_bar_in : (float)(nsum / (double)T);
  double s1 = 0.0;
  for (int t = tid; t < T; t += blockDim.x) {
    float* rs = row_stats + (size_t)t * RS_W;
    const float pr = weighted ? (float)((double)rs[RS_P] / (double)rs[RS_NT] * (double)n_bar) : rs[RS_P];
    rs[RS_PR] = pr;
    rs[RS_S] = 1.f + pr;
    pr_out[t] = pr;
    s1 += pr;
  }
  s1 = block_sum(s1, red);
  const double mean = s1 / (double)T;
  double s2 = 0.0;
  for (int t = tid; t < T; t += blockDim.x) {
    const double d = (double)pr_out[t] - mean;
    s2 += d * d;
  }
  s2 = block_sum(s2, red);
  if (tid == 0) {
    scalars[SC_NBAR] = n_bar;
    scalars[SC_SHARPE] = (T > 1) ? (float)(mean / (sqrt(s2 / (double)(T - 1)) + 1e-8)) : nanf("");
  }
}

// ---- column partial sums + weights ---------------------------------------------------------------
// per stock i over the months of one T-chunk:  sum_t g_k S_t (k<8), sum_t R m S_t, sum_t m
// (src/model.py:376,382,428) and, fused, w = m (u - mu_t) (src/model.py:271-279).
// `index` != null: u and g are left-packed (tensor-core kernels, cell_f32.cuh), else dense [T,N].
__global__ void __launch_bounds__(256, 4) col_partial_kernel(
    const float* __restrict__ g, const float* __restrict__ R, const uint8_t* __restrict__ mask,
    const float* __restrict__ u, const float* __restrict__ row_stats, const int* __restrict__ index, int T, int N,
    int tchunk, double* __restrict__ col_part, float* __restrict__ w_out) {
  extern __shared__ float sm[];            // S_t then mu_t for this chunk
  const int t0 = blockIdx.y * tchunk, t1 = min(T, t0 + tchunk);
  for (int t = t0 + threadIdx.x; t < t1; t += blockDim.x) {
    sm[t - t0] = row_stats[(size_t)t * RS_W + RS_S];
    sm[tchunk + t - t0] = row_stats[(size_t)t * RS_W + RS_MU];
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
  const bool in = i < N;                   // whole warps stay: the packed position comes from a ballot
  const int nw = (N + 31) >> 5;
  const int* rank = index ? index + idx_off_rank(T, N) + (i >> 5) : nullptr;
  const size_t plane = (size_t)T * N;
  double acc[KM + 2];
#pragma unroll
  for (int k = 0; k < KM + 2; ++k) acc[k] = 0.0;
  // The kernel is latency-bound (two dependent round trips per month: mask, then values), so the loop body is
  // branch-free -- masked lanes load from a clamped in-bounds position and their values are dropped by a
  // select -- and the unrolled months overlap their loads.
  const size_t ic = in ? (size_t)i : (size_t)N - 1;
#pragma unroll 4
  for (int t = t0; t < t1; ++t) {
    const size_t gi = (size_t)t * N + ic;
    const float S = sm[t - t0];
    // u and g are defined at unmasked cells only (the tensor-core kernels never visit the others)
    const bool live = in && mask[gi] != 0;
    const unsigned b = __ballot_sync(0xffffffffu, live);
    size_t vi = gi;
    if (index) vi = (size_t)t * N + __ldg(rank + (size_t)t * nw) + __popc(b & ((1u << lane) - 1u));
    if (vi >= plane) vi = plane - 1;       // a masked lane behind the last listed stock of the last month
    const float uv = (w_out || !g) ? __ldg(u + vi) : 0.f;
    const float r = R ? __ldg(R + gi) : 0.f;
    float gv[KM];
#pragma unroll
    for (int k = 0; k < KM; ++k) gv[k] = g ? __ldg(g + k * plane + vi) : 0.f;
    if (w_out && in) w_out[gi] = live ? (uv - sm[tchunk + t - t0]) : 0.f;
    acc[KM] += live ? (double)(r * S) : 0.0;
    acc[KM + 1] += live ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < KM; ++k) acc[k] += live ? (double)(gv[k] * S) : 0.0;
  }
  if (!in) return;
  double* out = col_part + (size_t)blockIdx.y * (KM + 2) * N + i;
#pragma unroll
  for (int k = 0; k < KM + 2; ++k) out[(size_t)k * N] = acc[k];
}

// ---- loss finish -----------------------------------------------------------------------------
// e_ki = colsum / T_i ; L_c = mean_k mean_i e_ki^2 ; L_u = mean_i e0_i^2   src/model.py:382-385,428-431
// 256 threads = 64 stocks x 4 groups of chunks: group q sums chunks q, q+4, ... in order, the four partial sums
// are then added in order through shared memory (fixed summation order, four times the loads in flight).
__global__ void __launch_bounds__(256) loss_finish_kernel(
    const double* __restrict__ col_part, int nsplit, int N, int n_total, float* __restrict__ e_out,
    float* __restrict__ ti_out, double* __restrict__ loss_part, unsigned int* __restrict__ counter,
    float* __restrict__ scalars) {
  __shared__ double red[32];
  __shared__ double part[3][KM + 2][64];
  __shared__ bool last;
  const int il = threadIdx.x & 63, q = threadIdx.x >> 6;
  const int i = blockIdx.x * 64 + il;
  double sc = 0.0, su = 0.0;
  double v[KM + 2];
#pragma unroll
  for (int k = 0; k < KM + 2; ++k) v[k] = 0.0;
  if (i < N) {
    for (int s = q; s < nsplit; s += 4)
#pragma unroll
      for (int k = 0; k < KM + 2; ++k) v[k] += col_part[((size_t)s * (KM + 2) + k) * N + i];
  }
  if (q > 0) {
#pragma unroll
    for (int k = 0; k < KM + 2; ++k) part[q - 1][k][il] = v[k];
  }
  __syncthreads();
  if (q == 0 && i < N) {
#pragma unroll
    for (int k = 0; k < KM + 2; ++k) v[k] = ((v[k] + part[0][k][il]) + part[1][k][il]) + part[2][k][il];
    const double ti = fmax(v[KM + 1], 1.0);
    ti_out[i] = (float)ti;
#pragma unroll
    for (int k = 0; k <= KM; ++k) {
      const float e = (float)(v[k] / ti);
      e_out[(size_t)k * N + i] = e;
      if (k < KM) sc += (double)e * (double)e; else su += (double)e * (double)e;
    }
  }
  sc = block_sum(sc, red);
  su = block_sum(su, red);
  if (threadIdx.x == 0) {
    loss_part[2 * blockIdx.x] = sc;
    loss_part[2 * blockIdx.x + 1] = su;
    __threadfence();
    last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last && threadIdx.x < 32) {
    __threadfence();
    double a = 0.0, b = 0.0;
    for (int j = threadIdx.x; j < (int)gridDim.x; j += 32) {
      a += ((volatile double*)loss_part)[2 * j];
      b += ((volatile double*)loss_part)[2 * j + 1];
    }
    a = warp_sum(a); b = warp_sum(b);
    if (threadIdx.x == 0) {
      // means run over ALL columns, dead ones included (src/model.py:385); n_total = global N
      scalars[SC_LC] = (float)(a / ((double)KM * (double)n_total));
      scalars[SC_LU] = (float)(b / (double)n_total);
      *counter = 0u;
    }
  }
}

// ---- backward: loss cotangents -> per-stock coefficients -------------------------------------------
// G_ki = gc * 2 e_ki / (K N T_i) ; G0_i = gu * 2 e0_i / (N T_i)          
__global__ void coeff_kernel(const float* __restrict__ e, const float* __restrict__ ti, int N,
                             const float* __restrict__ gc, const float* __restrict__ gu,
                             float* __restrict__ G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float c = gc ? *gc : 0.f, uu = gu ? *gu : 0.f;
  const float inv = 1.f / ((float)N * ti[i]);
#pragma unroll
  for (int k = 0; k < KM; ++k) G[(size_t)k * N + i] = c * (2.f / KM) * e[(size_t)k * N + i] * inv;
  G[(size_t)KM * N + i] = uu * 2.f * e[(size_t)KM * N + i] * inv;
}

// dS_t partial = sum_i ( sum_k G_ki g_kti + G0_i R m ).  `index` != null: g is left-packed and the month's
// list of unmasked stocks is walked (contiguous g, no mask read); else dense.
__global__ void __launch_bounds__(256) ds_rows_kernel(
    const float* __restrict__ g, const float* __restrict__ R, const uint8_t* __restrict__ mask,
    const float* __restrict__ G, const int* __restrict__ index, int T, int N, int nchunk,
    double* __restrict__ dS_part) {
  __shared__ double red[32];
  const int t = blockIdx.x, sp = blockIdx.y;
  const size_t plane = (size_t)T * N;
  double acc = 0.0;
  if (index) {
    const int cnt = index[idx_off_cnt(T) + t];
    const int chunk = (cnt + (int)gridDim.y - 1) / (int)gridDim.y;
    const int j0 = sp * chunk, j1 = min(cnt, j0 + chunk);
    const int* vrow = index + idx_off_vidx(T) + (size_t)t * N;
    for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
      const int i = __ldg(vrow + j);
      const size_t vi = (size_t)t * N + j;
      float v = __ldg(G + (size_t)KM * N + i) * __ldg(R + (size_t)t * N + i);
      if (g) {
#pragma unroll
        for (int k = 0; k < KM; ++k) v = fmaf(__ldg(G + (size_t)k * N + i), __ldg(g + k * plane + vi), v);
      }
      acc += (double)v;
    }
  } else {
    const int i0 = sp * nchunk, i1 = min(N, i0 + nchunk);
    for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
      const size_t gi = (size_t)t * N + i;
      if (!mask[gi]) continue;                         // g is defined at unmasked cells only
      float v = __ldg(G + (size_t)KM * N + i) * __ldg(R + gi);
      if (g) {
#pragma unroll
        for (int k = 0; k < KM; ++k) v = fmaf(__ldg(G + (size_t)k * N + i), __ldg(g + k * plane + gi), v);
      }
      acc += (double)v;
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) dS_part[(size_t)t * gridDim.y + sp] = acc;
}

// ---- backward: sum per-CTA / per-tile partials into the reference's gradient layouts ----------------
__global__ void __launch_bounds__(256) grad_finish_kernel(
    const float* __restrict__ wpart, int ncta, int F, int have_sdf, int have_mom,
    const float* __restrict__ dadc_part, int T, int tpm /* partial slots per month */,
    const int* __restrict__ index /* != null: (month, CTA) segment layout of the tensor-core kernel */,
    float* __restrict__ gW1, int ldw1, float* __restrict__ gW2, float* __restrict__ gb2,
    float* __restrict__ gW3, float* __restrict__ gb3, float* __restrict__ gWq, int ldwq, int wq_off,
    float* __restrict__ dadc) {
  const int Fr = round4(F), stride = wp_stride(Fr);
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < stride) {
    // four running sums (CTAs c = u mod 4, ascending), sixteen loads in flight: the kernel is one L2 round trip
    // per batch of loads and nothing else
    float v4[4] = {0.f, 0.f, 0.f, 0.f};
    int c = 0;
    for (; c + 16 <= ncta; c += 16) {
      float ld[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) ld[u] = __ldg(wpart + (size_t)(c + u) * stride + idx);
#pragma unroll
      for (int u = 0; u < 16; ++u) v4[u & 3] += ld[u];
    }
    for (; c + 4 <= ncta; c += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) v4[u] += wpart[(size_t)(c + u) * stride + idx];
    }
    for (; c < ncta; ++c) v4[0] += wpart[(size_t)c * stride + idx];
    const float v = (v4[0] + v4[1]) + (v4[2] + v4[3]);
    if (idx < wp_off_w1()) {
      if (have_sdf) gW2[idx] = v;
    } else if (idx < wp_off_wq(Fr)) {
      const int j = (idx - wp_off_w1()) / Fr, f = (idx - wp_off_w1()) % Fr;
      if (have_sdf && f < F) gW1[(size_t)j * ldw1 + f] = v;
    } else if (idx < wp_off_w3(Fr)) {
      const int k = (idx - wp_off_wq(Fr)) / Fr, f = (idx - wp_off_wq(Fr)) % Fr;
      if (have_mom && f < F) gWq[(size_t)k * ldwq + wq_off + f] = v;
    } else if (idx < wp_off_b2(Fr)) {
      if (have_sdf) gW3[idx - wp_off_w3(Fr)] = v;
    } else if (idx < wp_off_b2(Fr) + 64) {
      if (have_sdf) gb2[idx - wp_off_b2(Fr)] = v;
    }
    // d b3 = sum du = dP_t * sum m (R - Rbar_t) = 0 identically (demeaning, src/model.py:279)
    if (idx == 0 && have_sdf) gb3[0] = 0.f;
  } else {
    const int r = idx - stride;
    if (r < T * 72) {
      const int t = r / 72, c = r % 72;
      float v8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const float* p = dadc_part + (size_t)t * tpm * 72 + c;
      if (index) {             // one partial per CTA that owns tiles of this month (consecutive CTAs)
        const int* ts = index + idx_off_ts();
        const int chunk = tc_chunk(index[0], ncta), a = ts[t], b = ts[t + 1];
        tpm = b > a ? (b - 1) / chunk - a / chunk + 1 : 0;
      }
      int j = 0;
      for (; j + 8 <= tpm; j += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u) v8[u] += p[(size_t)(j + u) * 72];
      }
      for (; j < tpm; ++j) v8[0] += p[(size_t)j * 72];
      dadc[r] = ((v8[0] + v8[1]) + (v8[2] + v8[3])) + ((v8[4] + v8[5]) + (v8[6] + v8[7]));
    }
  }
}

// ---- inference: weights, optionally per-month |w|-normalised -----------------------------------------
// w = m (u - mu_t) ; normalised: w / clamp(sum_i |w| m, 1e-8)          src/model.py:581-592
__global__ void __launch_bounds__(256) weights_kernel(
    const float* __restrict__ u, const uint8_t* __restrict__ mask, const float* __restrict__ row_stats,
    const int* __restrict__ index /* != null: u is left-packed */, int T, int N, int normalized,
    float* __restrict__ w_out) {
  __shared__ double red[32];
  const int t = blockIdx.x, lane = threadIdx.x & 31;
  const float mu = row_stats[(size_t)t * RS_W + RS_MU];
  const size_t base = (size_t)t * N;
  const int* rank = index ? index + idx_off_rank(T, N) + (size_t)t * ((N + 31) >> 5) : nullptr;
  // value of u for stock i (whole warps call it: the packed position comes from a ballot)
  auto u_at = [&](int i, bool& live) {
    live = i < N && mask[base + i] != 0;
    const unsigned b = __ballot_sync(0xffffffffu, live);
    if (!live) return 0.f;
    const size_t vi = index ? base + __ldg(rank + (i >> 5)) + __popc(b & ((1u << lane) - 1u)) : base + i;
    return __ldg(u + vi);
  };
  float denom = 1.f;
  if (normalized) {
    double s = 0.0;
    for (int i0 = 0; i0 < N; i0 += blockDim.x) {
      bool live;
      const float v = u_at(i0 + threadIdx.x, live);
      if (live) s += (double)fabsf(v - mu);
    }
    s = block_sum(s, red);
    denom = fmaxf((float)s, 1e-8f);
  }
  for (int i0 = 0; i0 < N; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    bool live;
    const float v = u_at(i, live);
    const float w = live ? (v - mu) : 0.f;
    if (i < N) w_out[base + i] = normalized ? w / denom : w;
  }
}

// ---- launchers ---------------------------------------------------------------------------------------
cudaError_t launch_row_finish(const double* row_part, int T, int tpm, const int* index, int weighted,
                              const float* n_bar_in, float* row_stats, float* pr_out, float* scalars,
                              cudaStream_t st) {
  row_sum_kernel<<<(T + 7) / 8, 256, 0, st>>>(row_part, T, tpm, index, row_stats);
  row_finish_kernel<<<1, T < 1024 ? ((T + 31) / 32) * 32 : 1024, 0, st>>>(T, weighted, n_bar_in, row_stats, pr_out, scalars);
  return cudaGetLastError();
}
cudaError_t launch_col_partial(const float* g, const float* R, const uint8_t* mask, const float* u,
                               const float* row_stats, const int* index, int T, int N, int nsplit,
                               double* col_part, float* w_out, cudaStream_t st) {
  const int tchunk = (T + nsplit - 1) / nsplit;
  dim3 grid((N + 255) / 256, nsplit);
  col_partial_kernel<<<grid, 256, 2 * tchunk * sizeof(float), st>>>(g, R, mask, u, row_stats, index, T, N, tchunk,
                                                                    col_part, w_out);
  return cudaGetLastError();
}
cudaError_t launch_loss_finish(const double* col_part, int nsplit, int N, int n_total, float* e_out,
                               float* ti_out, double* loss_part, unsigned int* counter, float* scalars,
                               cudaStream_t st) {
  loss_finish_kernel<<<(N + 63) / 64, 256, 0, st>>>(col_part, nsplit, N, n_total, e_out, ti_out,
                                                      loss_part, counter, scalars);
  return cudaGetLastError();
}
cudaError_t launch_coeff(const float* e, const float* ti, int N, const float* gc, const float* gu, float* G,
                         cudaStream_t st) {
  coeff_kernel<<<(N + 255) / 256, 256, 0, st>>>(e, ti, N, gc, gu, G);
  return cudaGetLastError();
}
cudaError_t launch_ds_rows(const float* g, const float* R, const uint8_t* mask, const float* G, const int* index,
                           int T, int N, int nsplit, double* dS_part, cudaStream_t st) {
  const int nchunk = (N + nsplit - 1) / nsplit;
  dim3 grid(T, nsplit);
  ds_rows_kernel<<<grid, 256, 0, st>>>(g, R, mask, G, index, T, N, nchunk, dS_part);
  return cudaGetLastError();
}
cudaError_t launch_grad_finish(const float* wpart, int ncta, int F, int have_sdf, int have_mom,
                               const float* dadc_part, int T, int tpm, const int* index, float* gW1,
                               int ldw1, float* gW2,
                               float* gb2, float* gW3, float* gb3, float* gWq, int ldwq, int wq_off,
                               float* dadc, cudaStream_t st) {
  const int total = wp_stride(round4(F)) + T * 72;
  grad_finish_kernel<<<(total + 255) / 256, 256, 0, st>>>(wpart, ncta, F, have_sdf, have_mom, dadc_part, T,
                                                          tpm, index, gW1, ldw1, gW2, gb2, gW3, gb3, gWq, ldwq,
                                                          wq_off, dadc);
  return cudaGetLastError();
}
cudaError_t launch_weights(const float* u, const uint8_t* mask, const float* row_stats, const int* index, int T,
                           int N, int normalized, float* w_out, cudaStream_t st) {
  weights_kernel<<<T, 256, 0, st>>>(u, mask, row_stats, index, T, N, normalized, w_out);
  return cudaGetLastError();
}

// ---- loss finish without caller-provided scratch -------------------------------------------------------
// Some hosts of this library are not allowed to enter the CUDA allocator once they are running: a
// service that was given a fixed, fully pre-partitioned slice of device memory at start-up; an
// evaluation loop with a latency budget, for which cudaMalloc/cudaFree are off limits because they
// take driver-wide locks and synchronise the device; or glue code that simply has no allocator handle
// (it receives raw pointers from another runtime and must not create memory of its own).  Such callers
// cannot produce the two scratch buffers dlap_loss_finish wants, so this variant keeps them in a
// fixed-size array that lives inside the loaded library image, once per GPU, until the process exits.
//
// The array name below is only a HOST-SIDE HANDLE in host code.  Passing it to a kernel would hand the
// kernel a host address; the device address has to be asked for with cudaGetSymbolAddress, per device.
namespace {
constexpr int kNoAllocMaxStocks = DLAP_NOALLOC_MAX_STOCKS;
constexpr int kNoAllocBlocks = (kNoAllocMaxStocks + 63) / 64;
constexpr size_t kLossPartBytes = (size_t)kNoAllocBlocks * 2 * sizeof(double);
constexpr size_t kCounterOffset = kLossPartBytes;                 // 8-byte aligned, holds one uint32
constexpr size_t kNoAllocBytes = kLossPartBytes + 256;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[kMaxDevices] = {nullptr};

// Device address of the array on the CURRENT device; resolved on the first call per device.
int noalloc_scratch(unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return DLAP_E_NODEVICE;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return DLAP_E_NODEVICE;
    g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return DLAP_OK;
}
}  // namespace
}  // namespace dlap

extern "C" int dlap_loss_finish_noalloc(const dlap_plan* plan, const double* col_part, int nsplit, float* e,
                                        float* t_i, float* scalars, void* stream) {
  using namespace dlap;
  if (!plan || !col_part || !e || !t_i || !scalars) return DLAP_E_NULL;
  const int N = plan->dims.N;
  if (N < 1 || nsplit < 1) return DLAP_E_SHAPE;
  if (N > kNoAllocMaxStocks) return DLAP_E_LIMIT;     // the built-in scratch is sized for this many stocks
  unsigned char* base = nullptr;
  int rc = noalloc_scratch(&base);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  double* loss_part = reinterpret_cast<double*>(base);
  unsigned int* counter = reinterpret_cast<unsigned int*>(base + kCounterOffset);
  // The array keeps whatever the previous call left behind.  loss_part is fully overwritten before it
  // is read; the block counter is the one thing that must start at zero, so clear it on the stream.
  cudaError_t ce = cudaMemsetAsync(counter, 0, sizeof(unsigned int), st);
  if (ce != cudaSuccess) return (int)ce;
  ce = launch_loss_finish(col_part, nsplit, N, N, e, t_i, loss_part, counter, scalars, st);
  return ce == cudaSuccess ? DLAP_OK : (int)ce;
}
