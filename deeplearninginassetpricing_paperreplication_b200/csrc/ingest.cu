// Loader stage on the device: what AssetPricingDataset.__init__ does with numpy once per file
// (reference src/data_loader.py:42-99), from the raw array of the .npz to the step's inputs.
//   panel_ingest : raw[T*N, F+1] (column 0 the return, 1.. the characteristics; fp32 or fp64; device memory or
//                  page-locked host memory the device can address) -> mask, zero-filled R and x.  HBM-bound
//                  byte work: 4(F+1) B read + 4(F+1)+1 B written per stock-month, every access coalesced through a
//                  shared-memory transpose of ROWS rows (the raw row pitch, 188 B at F = 46, is only 4-B aligned).
//   macro_zscore : the macro series' per-column mean / std / (x - mean) / std in numpy's own fp32 summation order
//                  (row after row, or pairwise per column, whichever numpy uses for the loader's array), bit-exact.
#include <cuda_runtime.h>
#include <stdint.h>

namespace dlap {

constexpr int ROWS = 128;          // stock-months per chunk
constexpr int IT = 256;            // threads per CTA

template <typename RawT> __device__ __forceinline__ float raw_to_f32(RawT v);
template <> __device__ __forceinline__ float raw_to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float raw_to_f32<double>(double v) { return __double2float_rn(v); }   // .astype(np.float32)

// Streaming loads: the raw array is read once (and may live across PCIe).
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// e / w for e < 2^14 (one chunk) with inv = ceil(2^32 / w): one multiply instead of a division per element.
__device__ __forceinline__ void row_col(int e, int w, unsigned inv, int& r, int& c) {
  r = w > 1 ? (int)__umulhi((unsigned)e, inv) : e;
  c = e - r * w;
}

template <typename RawT>
__global__ void __launch_bounds__(IT) panel_ingest_kernel(const RawT* __restrict__ raw, long long cells, int F, float thr,
                                                          unsigned inv_w, unsigned inv_f, float* __restrict__ x,
                                                          float* __restrict__ R, uint8_t* __restrict__ mask) {
  extern __shared__ float tile[];                    // [ROWS][pitch], pitch odd: a thread scanning its row hits 32 banks
  __shared__ float keep[ROWS];
  const int W = F + 1, pitch = W | 1;
  const long long chunks = (cells + ROWS - 1) / ROWS;
  for (long long ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
    const long long row0 = ch * ROWS;
    const int rows = (int)((cells - row0 < ROWS) ? (cells - row0) : ROWS);
    const int n_in = rows * W;
    const RawT* src = raw + row0 * W;                // 16-B aligned for fp32: row0 is a multiple of 128
    if (sizeof(RawT) == 4) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      const int n4 = n_in >> 2;
      for (int i = threadIdx.x; i < n4; i += IT) {
        const float4 v = ld_stream4(s4 + i);
        const float vv[4] = {v.x, v.y, v.z, v.w};
        int r, c;
        row_col(i << 2, W, inv_w, r, c);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          tile[r * pitch + c] = vv[k];
          if (++c == W) { c = 0; ++r; }
        }
      }
      for (int e = (n4 << 2) + threadIdx.x; e < n_in; e += IT) {
        int r, c;
        row_col(e, W, inv_w, r, c);
        tile[r * pitch + c] = raw_to_f32<RawT>(src[e]);
      }
    } else {
      for (int e = threadIdx.x; e < n_in; e += IT) {
        int r, c;
        row_col(e, W, inv_w, r, c);
        tile[r * pitch + c] = raw_to_f32<RawT>(src[e]);
      }
    }
    __syncthreads();
    // src/data_loader.py:53-57: valid = R > thr and not NaN and every characteristic > thr (a NaN fails '>').
    if (threadIdx.x < rows) {
      const float* t = tile + threadIdx.x * pitch;
      bool ok = t[0] > thr;
      for (int c = 1; c < W; ++c) ok = ok && (t[c] > thr);
      keep[threadIdx.x] = ok ? 1.f : 0.f;
      R[row0 + threadIdx.x] = ok ? t[0] : 0.f;       // np.where(mask, returns, 0.0), src/data_loader.py:60
      mask[row0 + threadIdx.x] = ok ? 1 : 0;
    }
    __syncthreads();
    // np.where(mask[:, :, None], features, 0.0), src/data_loader.py:61-65; rows*F floats, contiguous in x
    float* dst = x + row0 * F;                       // 16-B aligned: row0 multiple of 128
    const int n_out = rows * F, n4o = n_out >> 2;
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int i = threadIdx.x; i < n4o; i += IT) {
      float vv[4];
      int r, c;
      row_col(i << 2, F, inv_f, r, c);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        vv[k] = keep[r] != 0.f ? tile[r * pitch + 1 + c] : 0.f;
        if (++c == F) { c = 0; ++r; }
      }
      d4[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
    }
    for (int e = (n4o << 2) + threadIdx.x; e < n_out; e += IT) {
      int r, c;
      row_col(e, F, inv_f, r, c);
      dst[e] = keep[r] != 0.f ? tile[r * pitch + 1 + c] : 0.f;
    }
    __syncthreads();
  }
}

cudaError_t launch_panel_ingest(const void* raw, int raw_f64, long long cells, int F, float thr, float* x, float* R,
                                uint8_t* mask, int ctas, cudaStream_t st) {
  if (cells == 0) return cudaSuccess;
  const int pitch = (F + 1) | 1;
  const size_t smem = (size_t)ROWS * pitch * sizeof(float);
  const long long chunks = (cells + ROWS - 1) / ROWS;
  const int grid = (int)(chunks < ctas ? chunks : ctas);
  const unsigned inv_w = (unsigned)(((1ull << 32) + F) / (unsigned)(F + 1));        // ceil(2^32 / (F+1))
  const unsigned inv_f = F > 1 ? (unsigned)(((1ull << 32) + F - 1) / (unsigned)F) : 0xffffffffu;
  if (raw_f64)
    panel_ingest_kernel<double><<<grid, IT, smem, st>>>((const double*)raw, cells, F, thr, inv_w, inv_f, x, R, mask);
  else
    panel_ingest_kernel<float><<<grid, IT, smem, st>>>((const float*)raw, cells, F, thr, inv_w, inv_f, x, R, mask);
  return cudaGetLastError();
}

// Column m of the series, term t: x_t, or (x_t - mu)^2 for the variance; each operation rounded on its own.
struct ColTerm {
  const float* col; size_t stride; bool sq; float mu;
  __device__ __forceinline__ float operator()(int t) const {
    const float v = col[(size_t)t * stride];
    if (!sq) return v;
    const float d = __fsub_rn(v, mu);
    return __fmul_rn(d, d);
  }
};

// numpy's pairwise summation of terms [t0, t0+n) (the scalar loop of add.reduce along a contiguous axis).
__device__ float pairwise_sum(const ColTerm& a, int t0, int n) {
  if (n < 8) {
    float r = 0.f;
    for (int i = 0; i < n; ++i) r = __fadd_rn(r, a(t0 + i));
    return r;
  }
  if (n <= 128) {
    float r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = a(t0 + j);
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], a(t0 + i + j));
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, a(t0 + i));
    return res;
  }
  int n2 = n / 2;
  n2 -= n2 % 8;
  const float lo = pairwise_sum(a, t0, n2);
  return __fadd_rn(lo, pairwise_sum(a, t0 + n2, n - n2));
}

__device__ __forceinline__ float column_sum(const ColTerm& a, int T, bool pairwise) {
  if (pairwise) return pairwise_sum(a, 0, T);
  float s = 0.f;
  for (int t = 0; t < T; ++t) s = __fadd_rn(s, a(t));
  return s;
}

// One thread per macro column.  fp32 throughout, no contraction into FMAs:
//   mean = sum_t x_t / T;  std = sqrt(sum_t (x_t - mean)^2 / T) + 1e-8       src/data_loader.py:78-79
//   out  = (x - mean) / std                                                    src/data_loader.py:84
// The sums run in numpy's order for the array the loader holds at that point: row after row for the file's own
// C-contiguous [T,M] array; pairwise per column after `macro_features[:, macro_idx]` (src/data_loader.py:72-73), whose
// result is column-major.
__global__ void __launch_bounds__(128) macro_zscore_kernel(const float* __restrict__ raw, int T, int M, float* mean,
                                                           float* stdv, int stats_given, int pairwise,
                                                           float* __restrict__ out) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  float mu, sd;
  if (stats_given) { mu = mean[m]; sd = stdv[m]; }
  else {
    ColTerm term{raw + m, (size_t)M, false, 0.f};
    mu = __fdiv_rn(column_sum(term, T, pairwise != 0), (float)T);
    term.sq = true; term.mu = mu;
    sd = __fadd_rn(__fsqrt_rn(__fdiv_rn(column_sum(term, T, pairwise != 0), (float)T)), 1e-8f);
    mean[m] = mu; stdv[m] = sd;
  }
  for (int t = 0; t < T; ++t) out[(size_t)t * M + m] = __fdiv_rn(__fsub_rn(raw[(size_t)t * M + m], mu), sd);
}

cudaError_t launch_macro_zscore(const float* raw, int T, int M, float* mean, float* stdv, int stats_given, int pairwise,
                                float* out, cudaStream_t st) {
  if (T == 0 || M == 0) return cudaSuccess;
  macro_zscore_kernel<<<(M + 127) / 128, 128, 0, st>>>(raw, T, M, mean, stdv, stats_given, pairwise, out);
  return cudaGetLastError();
}

}  // namespace dlap
