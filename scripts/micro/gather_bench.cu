// Micro-benchmark: how fast can one persistent CTA per SM gather scattered 184-B rows of a [T,N,46] fp32 panel?
// Variants: warps per CTA, rows in flight per warp, load width.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <int ROWS>
__global__ void __launch_bounds__(1024, 1) gather_kernel(const float* __restrict__ x, const int* __restrict__ rows, int nrows,
                                                        int F, float* __restrict__ out) {
  // rows: global row numbers (t*N+n), ascending; CTA b takes a contiguous range; warp w takes chunks of ROWS rows
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const long long per = (nrows + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per, hi = min((long long)nrows, lo + per);
  float acc = 0.f;
  for (long long r0 = lo + warp * ROWS; r0 < hi; r0 += (long long)nw * ROWS) {
    float2 v[ROWS];
    const int mine = (r0 + lane < hi && lane < ROWS) ? rows[r0 + lane] : -1;
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      const int nr = max(__shfl_sync(0xffffffffu, mine, j), 0);      // unconditional loads on clamped addresses:
      v[j] = __ldg(reinterpret_cast<const float2*>(x + (size_t)nr * F) + min(lane, F / 2 - 1));   // no predicated-load + select chains
    }
#pragma unroll
    for (int j = 0; j < ROWS; ++j) acc += v[j].x + v[j].y;
  }
  if (acc == 123.456f) out[0] = acc;
}

// two rows per instruction: lanes 0-15 fetch 16-B chunks of row A, lanes 16-31 of row B (aligned spans)
template <int INSTR>
__global__ void __launch_bounds__(512, 1) gather16_kernel(const float* __restrict__ x, const int* __restrict__ rows, int nrows,
                                                          int F, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const long long per = (nrows + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per, hi = min((long long)nrows, lo + per);
  const int chunks = (8 + 4 * F + 15) / 16, c = min(lane & 15, chunks - 1), half = lane >> 4;
  float acc = 0.f;
  for (long long r0 = lo + warp * 2 * INSTR; r0 < hi; r0 += (long long)nw * 2 * INSTR) {
    float4 v[INSTR];
    const int mine = (r0 + lane < hi && lane < 2 * INSTR) ? rows[r0 + lane] : -1;
#pragma unroll
    for (int j = 0; j < INSTR; ++j) {
      const int nr = max(__shfl_sync(0xffffffffu, mine, 2 * j + half), 0);
      const uintptr_t a = reinterpret_cast<uintptr_t>(x + (size_t)nr * F) & ~uintptr_t(15);
      v[j] = __ldg(reinterpret_cast<const float4*>(a) + c);
    }
#pragma unroll
    for (int j = 0; j < INSTR; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
  }
  if (acc == 123.456f) out[0] = acc;
}

// per-row TMA bulk copies (192 B from the 16-B aligned span) issued by W warps; every warp owns a private
// staging area of 32 rows x 2 stages and its own mbarriers
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(1024, 1) tma_rows_kernel(const float* __restrict__ x, const int* __restrict__ rows, int nrows,
                                                           int F, float* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(sm);          // [nw][2]
  float* stage = reinterpret_cast<float*>(sm + 1024) + (size_t)warp * 2 * 32 * 52;
  if (lane == 0) for (int s = 0; s < 2; ++s)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + warp * 2 + s)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const long long per = (nrows + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per, hi = min((long long)nrows, lo + per);
  float acc = 0.f;
  int it = 0;
  for (long long r0 = lo + warp * 32; r0 < hi + (long long)nw * 32; r0 += (long long)nw * 32, ++it) {
    if (r0 < hi) {                                  // issue batch `it` into stage it&1
      const int st = it & 1;
      const int nr = (r0 + lane < hi) ? rows[r0 + lane] : -1;
      const unsigned cnt = __popc(__ballot_sync(0xffffffffu, nr >= 0));
      if (lane == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + warp * 2 + st)), "r"(cnt * 192u) : "memory");
      __syncwarp();
      if (nr >= 0) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(x + (size_t)nr * F) & ~uintptr_t(15);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                     "r"(smem_u32(stage + (st * 32 + lane) * 52)), "l"(a), "r"(192u), "r"(smem_u32(bars + warp * 2 + st)) : "memory");
      }
    }
    if (it > 0) {                                   // consume batch it-1
      const int st = (it - 1) & 1, par = ((it - 1) >> 1) & 1;
      asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::
                   "r"(smem_u32(bars + warp * 2 + st)), "r"(par) : "memory");
      acc += stage[(st * 32 + lane) * 52 + 3];
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
  }
  if (acc == 123.456f) out[0] = acc;
}

// same, but ONE lane per warp issues the 32 copies of a batch from indices it loads itself (no per-lane
// issue loop around the uniform-datapath copy instruction)
__global__ void __launch_bounds__(1024, 1) tma_rows1_kernel(const float* __restrict__ x, const int* __restrict__ rows, int nrows,
                                                            int F, float* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(sm);
  float* stage = reinterpret_cast<float*>(sm + 1024) + (size_t)warp * 2 * 32 * 52;
  if (lane == 0) for (int s = 0; s < 2; ++s)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + warp * 2 + s)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const long long per = (nrows + gridDim.x - 1) / gridDim.x;
  const long long lo = (long long)blockIdx.x * per, hi = min((long long)nrows, lo + per);
  float acc = 0.f;
  int it = 0;
  for (long long r0 = lo + warp * 32; r0 < hi + (long long)nw * 32; r0 += (long long)nw * 32, ++it) {
    if (r0 < hi) {
      const int st = it & 1;
      if (lane == 0) {
        const int cnt = (int)min((long long)32, hi - r0);
        const unsigned bar = smem_u32(bars + warp * 2 + st);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(cnt * 192u) : "memory");
        int n[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) n[j] = (j < cnt) ? rows[r0 + j] : -1;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (n[j] >= 0) {
            const uintptr_t a = reinterpret_cast<uintptr_t>(x + (size_t)n[j] * F) & ~uintptr_t(15);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                         "r"(smem_u32(stage + (st * 32 + j) * 52)), "l"(a), "r"(192u), "r"(bar) : "memory");
          }
        }
      }
      __syncwarp();
    }
    if (it > 0) {
      const int st = (it - 1) & 1, par = ((it - 1) >> 1) & 1;
      asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::
                   "r"(smem_u32(bars + warp * 2 + st)), "r"(par) : "memory");
      acc += stage[(st * 32 + lane) * 52 + 3];
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
  }
  if (acc == 123.456f) out[0] = acc;
}

int main(int argc, char** argv) {
  const int T = 240, N = 10000, F = 46;
  const size_t cells = (size_t)T * N;
  const bool host = argc > 1;                 // any argument: the panel stays in pinned host memory (zero-copy pull)
  float* x;
  if (host) {
    CK(cudaHostAlloc(&x, cells * F * 4, cudaHostAllocDefault)); memset(x, 0, cells * F * 4);
    float* d; CK(cudaMalloc(&d, cells * F * 4));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int it = 0; it < 2; ++it) {
      cudaEventRecord(a); CK(cudaMemcpyAsync(d, x, cells * F * 4, cudaMemcpyHostToDevice)); cudaEventRecord(b);
      CK(cudaEventSynchronize(b)); float ms; cudaEventElapsedTime(&ms, a, b);
      printf("dense H2D copy of the panel: %.3f ms  %.1f GB/s\n", ms, cells * F * 4 / ms / 1e6);
    }
    cudaFree(d);
  } else { CK(cudaMalloc(&x, cells * F * 4)); CK(cudaMemset(x, 0, cells * F * 4)); }
  float* out; CK(cudaMalloc(&out, 4));
  for (double dens : {0.4}) {
    std::vector<int> rows;
    srand(1);
    for (size_t c = 0; c < cells; ++c) if (rand() < dens * RAND_MAX) rows.push_back((int)c);
    int* d; CK(cudaMalloc(&d, rows.size() * 4)); CK(cudaMemcpy(d, rows.data(), rows.size() * 4, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int warps : {1, 2, 4, 8, 16}) {
      auto run = [&](auto kern, int R) {
        for (int it = 0; it < 3; ++it) kern<<<148, warps * 32>>>(x, d, (int)rows.size(), F, out);
        cudaEventRecord(e0);
        for (int it = 0; it < 5; ++it) kern<<<148, warps * 32>>>(x, d, (int)rows.size(), F, out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
        printf("density %.1f warps %2d rows/warp %2d: %.3f ms  %.0f GB/s useful\n", dens, warps, R, ms,
               rows.size() * 184.0 / ms / 1e6);
      };
      {
        const size_t smb = 1024 + (size_t)warps * 2 * 32 * 52 * 4;
        if (smb <= 227 * 1024) {
          CK(cudaFuncSetAttribute(tma_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smb));
          for (int it = 0; it < 3; ++it) tma_rows_kernel<<<148, warps * 32, smb>>>(x, d, (int)rows.size(), F, out);
          cudaEventRecord(e0);
          for (int it = 0; it < 5; ++it) tma_rows_kernel<<<148, warps * 32, smb>>>(x, d, (int)rows.size(), F, out);
          cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
          float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
          printf("density %.1f warps %2d TMA rows (2 stages x 32 rows per warp): %.3f ms  %.0f GB/s useful\n", dens, warps, ms,
                 rows.size() * 184.0 / ms / 1e6);
          CK(cudaFuncSetAttribute(tma_rows1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smb));
          for (int it = 0; it < 3; ++it) tma_rows1_kernel<<<148, warps * 32, smb>>>(x, d, (int)rows.size(), F, out);
          cudaEventRecord(e0);
          for (int it = 0; it < 5; ++it) tma_rows1_kernel<<<148, warps * 32, smb>>>(x, d, (int)rows.size(), F, out);
          cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
          cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
          printf("density %.1f warps %2d TMA rows, one issuing lane per warp:     %.3f ms  %.0f GB/s useful\n", dens, warps, ms,
                 rows.size() * 184.0 / ms / 1e6);
        }
      }
      run(gather_kernel<8>, 8); run(gather_kernel<16>, 16);
      if (warps <= 16) { run(gather16_kernel<4>, 1008); run(gather16_kernel<8>, 1016); run(gather16_kernel<16>, 1032); }
    }
    cudaFree(d);
  }
  return 0;
}
