// Valid-cell index of a panel (dlap_panel_index): integer work only, bit-exact by construction.
// Layout in cell_f32.cuh.  The mask contract is the reference's (src/data_loader.py:50-65): a bool
// byte per stock-month, nonzero = the cell takes part.
#include "../../include/dlap.h"
#include "cell_f32.cuh"

namespace dlap {

// One CTA per month: stable compaction of the stocks with mask != 0 (ascending stock number).
__global__ void __launch_bounds__(1024) index_rows_kernel(const uint8_t* __restrict__ mask, int N,
                                                          int* __restrict__ cnt, int* __restrict__ vidx,
                                                          int* __restrict__ rank) {
  __shared__ int wsum[32];
  const int t = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint8_t* row = mask + (size_t)t * N;
  int* out = vidx + (size_t)t * N;
  int base = 0;
  for (int n0 = 0; n0 < N; n0 += 1024) {
    const int n = n0 + tid;
    const bool v = n < N && row[n] != 0;
    const unsigned b = __ballot_sync(0xffffffffu, v);
    if (lane == 0) wsum[warp] = __popc(b);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
      const int c = wsum[w];
      before += (w < warp) ? c : 0;
      total += c;
    }
    if (v) out[base + before + __popc(b & ((1u << lane) - 1u))] = n;
    if (lane == 0 && n < N) rank[(size_t)t * ((N + 31) >> 5) + (n >> 5)] = base + before;
    base += total;
    __syncthreads();
  }
  if (tid == 0) cnt[t] = base;
}

// One CTA: exclusive prefix of the per-month tile counts, and the month of every tile.
__global__ void __launch_bounds__(1024) index_tiles_kernel(int T, int* __restrict__ index, int* __restrict__ tmonth) {
  __shared__ int red[32];
  int* ts = index + idx_off_ts();
  const int* cnt = index + idx_off_cnt(T);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int carry = 0;
  for (int t0 = 0; t0 < T; t0 += 1024) {
    const int t = t0 + tid;
    const int c = t < T ? cnt[t] : 0;
    const int nt = (c + TILE - 1) / TILE;
    int v = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += n;
    }
    if (lane == 31) red[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int w = red[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += n;
      }
      red[lane] = w;
    }
    __syncthreads();
    const int first = carry + (warp ? red[warp - 1] : 0) + v - nt;
    if (t < T) {
      ts[t] = first;
      for (int k = 0; k < nt; ++k) tmonth[first + k] = t;
    }
    carry += red[31];
    __syncthreads();
  }
  if (tid == 0) { ts[T] = carry; index[0] = carry; }
}

// ---- panel ingest straight from pinned host memory ------------------------------------------------------
// Only the rows of unmasked stock-months cross the bus: every warp walks valid-cell tiles and copies their
// characteristic rows from the (device-mapped) host panel to the same place in the device panel, 16 rows in
// flight per warp.  One warp per SM already saturates PCIe (measured on the B200 box: 4.15 ms for the 40 %
// dense paper-shape panel against 7.96 ms for a DMA copy of all of it).
template <bool WIDE>
__global__ void __launch_bounds__(256) panel_pull_kernel(const float* __restrict__ xh, const int* __restrict__ index,
                                                         int T, int N, int F, float* __restrict__ xd) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
  const int ntiles = index[0];
  const int* ts = index + idx_off_ts();
  const int* cnt = index + idx_off_cnt(T);
  const int* vidx = index + idx_off_vidx(T);
  for (int tile = warp; tile < ntiles; tile += nwarp) {
    const int t = index[idx_off_tmonth(T, N) + tile], j0 = (tile - ts[t]) * TILE, ncell = min(TILE, cnt[t] - j0);
    const size_t month = (size_t)t * N * F;
    for (int b = 0; b < ncell; b += 16) {
      const int rows = min(16, ncell - b);
      const int n_lane = vidx[(size_t)t * N + j0 + b + min(lane, rows - 1)];     // past the end: repeat the last row
      if (WIDE) {
        const int k = min(lane, (F >> 1) - 1);
        unsigned long long r[16];
        uint32_t off[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          off[j] = (uint32_t)__shfl_sync(0xffffffffu, n_lane, j) * (uint32_t)F;
          r[j] = __ldg(reinterpret_cast<const unsigned long long*>(xh + month + off[j]) + k);
        }
        if (2 * lane < F) {
#pragma unroll
          for (int j = 0; j < 16; ++j) reinterpret_cast<unsigned long long*>(xd + month + off[j])[k] = r[j];
        }
      } else {
        for (int j = 0; j < rows; ++j) {
          const size_t o = month + (size_t)__shfl_sync(0xffffffffu, n_lane, j) * F;
          for (int c = lane; c < F; c += 32) xd[o + c] = __ldg(xh + o + c);
        }
      }
    }
  }
}

cudaError_t launch_panel_pull(const float* x_host, const int* index, int T, int N, int F, float* x_dev, int sms,
                              bool clear, cudaStream_t st) {
  if (clear) {                                                             // masked rows are delivered as zeros
    cudaError_t e = cudaMemsetAsync(x_dev, 0, (size_t)T * N * F * 4, st);
    if (e != cudaSuccess) return e;
  }
  const bool wide = (F % 2 == 0) && ((reinterpret_cast<uintptr_t>(x_host) | reinterpret_cast<uintptr_t>(x_dev)) & 7) == 0;
  if (wide) panel_pull_kernel<true><<<sms, 256, 0, st>>>(x_host, index, T, N, F, x_dev);
  else panel_pull_kernel<false><<<sms, 256, 0, st>>>(x_host, index, T, N, F, x_dev);
  return cudaGetLastError();
}

cudaError_t launch_panel_index(const uint8_t* mask, int T, int N, int* index, cudaStream_t st) {
  index_rows_kernel<<<T, 1024, 0, st>>>(mask, N, index + idx_off_cnt(T), index + idx_off_vidx(T),
                                        index + idx_off_rank(T, N));
  index_tiles_kernel<<<1, 1024, 0, st>>>(T, index, index + idx_off_tmonth(T, N));
  return cudaGetLastError();
}

}  // namespace dlap
