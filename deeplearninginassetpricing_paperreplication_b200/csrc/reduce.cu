// Stage-level kernels next to the fused reduction of step_reduce.cu: the loss stage on its own (per-stock sums ->
// e, T_i, L_c, L_u; also behind the no-allocation entry point), the stand-alone sum of the backward kernels'
// partials (month shards exchange right after it), and the inference weights.  Every cross-CTA sum is done in a
// fixed order (no float atomics) so that results are run-to-run reproducible like the reference's CPU path.
#include <mutex>

#include "../../include/dlap.h"
#include "dlap_common.cuh"
#include "cell_f32.cuh"
#include "grad_finish.cuh"

namespace dlap {

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

// ---- backward: sum per-CTA / per-tile partials into the reference's gradient layouts (grad_finish.cuh) --------
__global__ void __launch_bounds__(256) grad_finish_kernel(const FinishParams P) {
  const int stride = wp_stride(round4(P.F));
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < stride) {
    if (P.wpart) finish_wpart(P, idx);
  } else if (idx - stride < P.T * 72) {
    finish_dadc(P, idx - stride);
  }
}

// ---- inference: weights, optionally per-month |w|-normalised -----------------------------------------
// w = m (u - mu_t), mu_t = sum_i m u / clamp(sum_i m, 1); normalised: w / clamp(sum_i |w| m, 1e-8)
// (src/model.py:271-279, :581-592).  One CTA per month; `index` != null: u is left-packed.
__global__ void __launch_bounds__(256) weights_kernel(
    const float* __restrict__ u, const uint8_t* __restrict__ mask, const int* __restrict__ index, int T, int N,
    int normalized, int demean, float* __restrict__ w_out) {
  __shared__ double red[32];
  const int t = blockIdx.x, lane = threadIdx.x & 31;
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
  if (index) {
    // u is left-packed: the month's listed cells are the contiguous run u[base .. base + c).  Everything is a pass over
    // that run with several loads in flight -- the mean, the sum of |u - mu|, and the weights themselves, scattered to
    // their stocks (vidx) over a row that was zero-filled first (the barrier orders the two stores to a listed cell).
    // No mask byte, rank word or ballot per stock: the stock-by-stock loop below needed three dependent round trips for
    // each of its 3 x N / 256 iterations (measured at the paper shape: 60 us; this form: see DESIGN.md 4.2).
    const int c = __ldg(index + idx_off_cnt(T) + t);
    const float* ur = u + base;
    const int* vrow = index + idx_off_vidx(T) + base;
    float4* z4 = reinterpret_cast<float4*>(w_out + base);
    const bool al = (reinterpret_cast<uintptr_t>(w_out + base) & 15) == 0;
    if (al) {
      for (int i = threadIdx.x; i < (N >> 2); i += blockDim.x) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int i = (N & ~3) + threadIdx.x; i < N; i += blockDim.x) w_out[base + i] = 0.f;
    } else {
      for (int i = threadIdx.x; i < N; i += blockDim.x) w_out[base + i] = 0.f;
    }
    double su = 0.0;
#pragma unroll 8
    for (int j = threadIdx.x; j < c; j += blockDim.x) su += (double)__ldg(ur + j);
    su = block_sum(su, red);
    const float mu = demean ? (float)(su / fmax((double)c, 1.0)) : 0.f;
    float denom = 1.f;
    if (normalized) {
      double s = 0.0;
#pragma unroll 8
      for (int j = threadIdx.x; j < c; j += blockDim.x) s += (double)fabsf(__ldg(ur + j) - mu);
      s = block_sum(s, red);
      denom = fmaxf((float)s, 1e-8f);
    }
    __syncthreads();                             // the zero-fill before the listed cells' values
    for (int j0 = threadIdx.x; j0 < c; j0 += 8 * blockDim.x) {
      int ii[8];
      float uu[8];
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        const int j = min(j0 + b * (int)blockDim.x, c - 1);
        ii[b] = __ldg(vrow + j);
        uu[b] = __ldg(ur + j);
      }
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        if (j0 + b * (int)blockDim.x < c) {
          const float w = uu[b] - mu;
          w_out[base + ii[b]] = normalized ? w / denom : w;
        }
      }
    }
    return;
  }
  double su = 0.0, cnt = 0.0;
  for (int i0 = 0; i0 < N; i0 += blockDim.x) {
    bool live;
    const float v = u_at(i0 + threadIdx.x, live);
    if (live) { su += (double)v; cnt += 1.0; }
  }
  su = block_sum(su, red);
  cnt = block_sum(cnt, red);
  const float mu = demean ? (float)(su / fmax(cnt, 1.0)) : 0.f;
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
cudaError_t launch_loss_finish(const double* col_part, int nsplit, int N, int n_total, float* e_out,
                               float* ti_out, double* loss_part, unsigned int* counter, float* scalars,
                               cudaStream_t st) {
  loss_finish_kernel<<<(N + 63) / 64, 256, 0, st>>>(col_part, nsplit, N, n_total, e_out, ti_out,
                                                      loss_part, counter, scalars);
  return cudaGetLastError();
}
cudaError_t launch_grad_finish(const FinishParams& P, cudaStream_t st) {
  const int total = wp_stride(round4(P.F)) + P.T * 72;
  grad_finish_kernel<<<(total + 255) / 256, 256, 0, st>>>(P);
  return cudaGetLastError();
}
cudaError_t launch_weights(const float* u, const uint8_t* mask, const int* index, int T, int N, int normalized,
                           int demean, float* w_out, cudaStream_t st) {
  weights_kernel<<<T, 256, 0, st>>>(u, mask, index, T, N, normalized, demean, w_out);
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
