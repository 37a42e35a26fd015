// Valid-cell index of a panel (dlap_panel_index): integer work only, bit-exact by construction.
// Layout in cell_f32.cuh.  The mask contract is the reference's (src/data_loader.py:50-65): a bool
// byte per stock-month, nonzero = the cell takes part.
#include "../../include/dlap.h"
#include "cell_f32.cuh"

namespace dlap {

// One CTA per month: stable compaction of the stocks with mask != 0 (ascending stock number).
__global__ void __launch_bounds__(1024) index_rows_kernel(const uint8_t* __restrict__ mask, int N,
                                                          int* __restrict__ cnt, int* __restrict__ vidx) {
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
    base += total;
    __syncthreads();
  }
  if (tid == 0) cnt[t] = base;
}

// One CTA: exclusive prefix of the per-month tile counts.
__global__ void __launch_bounds__(1024) index_tiles_kernel(int T, int* __restrict__ index) {
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
    if (t < T) ts[t] = first;
    carry += red[31];
    __syncthreads();
  }
  if (tid == 0) { ts[T] = carry; index[0] = carry; }
}

cudaError_t launch_panel_index(const uint8_t* mask, int T, int N, int* index, cudaStream_t st) {
  index_rows_kernel<<<T, 1024, 0, st>>>(mask, N, index + idx_off_cnt(T), index + idx_off_vidx(T));
  index_tiles_kernel<<<1, 1024, 0, st>>>(T, index);
  return cudaGetLastError();
}

}  // namespace dlap
