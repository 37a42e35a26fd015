// Self-test of the tcgen05 plumbing: one CTA computes D[128,N] = A.B^T from operand images that
// the caller lays out, with caller-chosen descriptor strides and majors.  tests/ uses it to pin
// the descriptor conventions of tc_sm100.cuh against numpy before the fused kernels rely on them.
#include "../../include/dlap.h"
#include "tc_sm100.cuh"

namespace dlap {

struct ProbeParams {
  const uint8_t* a_img; const uint8_t* b_img; int a_bytes, b_bytes;
  int M, N, K;
  int a_mn, b_mn;                        // 1 = MN-major operand
  uint32_t a_lbo, a_sbo, a_kstep;        // bytes; kstep = start-address advance per 16 K
  uint32_t b_lbo, b_sbo, b_kstep;
  int m_valid;                           // rows of D to write out
  float* D;                              // [128, N]
  int rr2;                               // timing mode: alternate two accumulators
  int repeat;                            // > 1: timing mode, the K loop is issued `repeat` times
  long long* cycles;                     // timing mode: SM cycles from first issue to completion
};

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const ProbeParams P) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* a_s = sm;
  uint8_t* b_s = sm + ((P.a_bytes + 127) & ~127);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < P.a_bytes / 16; i += 128)
    reinterpret_cast<uint4*>(a_s)[i] = reinterpret_cast<const uint4*>(P.a_img)[i];
  for (int i = tid; i < P.b_bytes / 16; i += 128)
    reinterpret_cast<uint4*>(b_s)[i] = reinterpret_cast<const uint4*>(P.b_img)[i];
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  fence_proxy_async();                   // generic-proxy smem writes -> visible to the tensor core
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = tmem_slot;
  if (warp == 0 && tc::elect_one()) {
    const uint32_t idesc = tc::idesc_bf16(P.M, P.N, P.a_mn, P.b_mn);
    uint64_t ad = tc::smem_desc(smem_u32(a_s), P.a_lbo, P.a_sbo);
    uint64_t bd = tc::smem_desc(smem_u32(b_s), P.b_lbo, P.b_sbo);
    const long long t0 = clock64();
    for (int rep = 0; rep < (P.repeat > 1 ? P.repeat : 1); ++rep) {
      uint64_t a2 = ad, b2 = bd;
      for (int k = 0; k < P.K / 16; ++k) {
        // timing mode with repeat < 0: round-robin over two accumulators (columns 0 and 128)
        tc::umma_bf16(tbase + ((P.rr2 && (k & 1)) ? 128u : 0u), a2, b2, idesc, (k > 1) || (rep > 0) || (!P.rr2 && k > 0));
        a2 = tc::desc_advance(a2, P.a_kstep);
        b2 = tc::desc_advance(b2, P.b_kstep);
      }
    }
    tc::umma_commit(&bar);
    mbar_wait(&bar, 0);
    if (P.cycles) *P.cycles = clock64() - t0;
  }
  mbar_wait(&bar, 0);
  tc::fence_after_sync();
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < P.N; c0 += 8) {
    float v[8];
    tc::tmem_ld8(tc::tmem_addr(tbase, warp * 32, c0), v);
    tc::tmem_wait_ld();
    if (row < P.m_valid)
      for (int i = 0; i < 8; ++i) P.D[(size_t)row * P.N + c0 + i] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tbase, 256);
}

// TS variant: A[128][K] (row-major bf16 in global) is written to TMEM by its owning lanes, B from smem.
__global__ void __launch_bounds__(128, 1) umma_ts_probe_kernel(const ProbeParams P) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* b_s = sm;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < P.b_bytes / 16; i += 128)
    reinterpret_cast<uint4*>(b_s)[i] = reinterpret_cast<const uint4*>(P.b_img)[i];
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tbase = tmem_slot;
  const uint32_t a_col = 256;                        // A lives at columns 256.., D at 0..
  {   // row = tid: K bf16 values -> K/2 packed columns
    const uint32_t* arow = reinterpret_cast<const uint32_t*>(P.a_img) + (size_t)tid * (P.K / 2);
    for (int c0 = 0; c0 < P.K / 2; c0 += 8) {
      uint32_t r[8];
      for (int i = 0; i < 8; ++i) r[i] = arow[c0 + i];
      tc::tmem_st8(tc::tmem_addr(tbase, warp * 32, a_col + c0), r);
    }
    tc::tmem_wait_st();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  if (tid == 0) {
    const uint32_t idesc = tc::idesc_bf16(P.M, P.N, 0, P.b_mn);
    uint64_t bd = tc::smem_desc(smem_u32(b_s), P.b_lbo, P.b_sbo);
    const long long t0 = clock64();
    for (int rep = 0; rep < (P.repeat > 1 ? P.repeat : 1); ++rep) {
      uint64_t b2 = bd;
      for (int k = 0; k < P.K / 16; ++k) {
        tc::umma_bf16_ts(tbase, tbase + a_col + 8 * k, b2, idesc, (k > 0) || (rep > 0));
        b2 = tc::desc_advance(b2, P.b_kstep);
      }
    }
    tc::umma_commit(&bar);
    mbar_wait(&bar, 0);
    if (P.cycles) *P.cycles = clock64() - t0;
  }
  mbar_wait(&bar, 0);
  tc::fence_after_sync();
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < P.N; c0 += 8) {
    float v[8];
    tc::tmem_ld8(tc::tmem_addr(tbase, warp * 32, c0), v);
    tc::tmem_wait_ld();
    for (int i = 0; i < 8; ++i) P.D[(size_t)row * P.N + c0 + i] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tbase, 512);
}

}  // namespace dlap

extern "C" int dlap_umma_ts_selftest(const void* a_rowmajor, const void* b_img, int b_bytes, int N, int K,
                                     int b_mn_major, int b_lbo, int b_sbo, int b_kstep, int repeat, float* D,
                                     long long* cycles, void* stream) {
  using namespace dlap;
  if (!a_rowmajor || !b_img || !D) return DLAP_E_NULL;
  if (N < 16 || N > 256 || (N & 15) || K < 16 || (K & 15) || K > 256 || repeat < 1 || (b_bytes & 15)) return DLAP_E_SHAPE;
  ProbeParams P{};
  P.a_img = (const uint8_t*)a_rowmajor; P.b_img = (const uint8_t*)b_img; P.b_bytes = b_bytes;
  P.M = 128; P.N = N; P.K = K; P.b_mn = b_mn_major; P.b_lbo = b_lbo; P.b_sbo = b_sbo; P.b_kstep = b_kstep;
  P.D = D; P.repeat = repeat; P.cycles = cycles;
  const size_t smem = b_bytes + 256;
  cudaError_t e = cudaFuncSetAttribute(umma_ts_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  umma_ts_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(P);
  e = cudaGetLastError();
  return e == cudaSuccess ? DLAP_OK : (int)e;
}

extern "C" int dlap_umma_selftest(const void* a_img, int a_bytes, const void* b_img, int b_bytes, int M, int N,
                                  int K, int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep,
                                  int b_lbo, int b_sbo, int b_kstep, float* D, void* stream) {
  using namespace dlap;
  if (!a_img || !b_img || !D) return DLAP_E_NULL;
  if ((M != 64 && M != 128) || N < 8 || N > 256 || (N & 7) || K < 16 || (K & 15)) return DLAP_E_SHAPE;
  if ((a_bytes & 15) || (b_bytes & 15) || a_bytes + b_bytes > 200 * 1024) return DLAP_E_SHAPE;
  ProbeParams P;
  P.a_img = (const uint8_t*)a_img; P.b_img = (const uint8_t*)b_img; P.a_bytes = a_bytes; P.b_bytes = b_bytes;
  P.M = M; P.N = N; P.K = K; P.a_mn = a_mn_major; P.b_mn = b_mn_major;
  P.a_lbo = a_lbo; P.a_sbo = a_sbo; P.a_kstep = a_kstep; P.b_lbo = b_lbo; P.b_sbo = b_sbo; P.b_kstep = b_kstep;
  P.m_valid = 128; P.D = D; P.repeat = 1; P.cycles = nullptr; P.rr2 = 0;
  const size_t smem = ((a_bytes + 127) & ~127) + b_bytes + 128;
  cudaError_t e = cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(P);
  e = cudaGetLastError();
  return e == cudaSuccess ? DLAP_OK : (int)e;
}

// Timing twin of dlap_umma_selftest: issues the same K loop `repeat` times back to back and reports the SM
// cycles until the last MMA has completed (host pointer-free: *cycles is a device int64).
extern "C" int dlap_umma_cycles(const void* a_img, int a_bytes, const void* b_img, int b_bytes, int M, int N,
                                int K, int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep,
                                int b_lbo, int b_sbo, int b_kstep, int repeat, float* D, long long* cycles,
                                void* stream) {
  using namespace dlap;
  if (!a_img || !b_img || !D || !cycles) return DLAP_E_NULL;
  if ((M != 64 && M != 128) || N < 8 || N > 256 || (N & 7) || K < 16 || (K & 15) || repeat == 0) return DLAP_E_SHAPE;
  if ((a_bytes & 15) || (b_bytes & 15) || a_bytes + b_bytes > 200 * 1024) return DLAP_E_SHAPE;
  ProbeParams P;
  P.a_img = (const uint8_t*)a_img; P.b_img = (const uint8_t*)b_img; P.a_bytes = a_bytes; P.b_bytes = b_bytes;
  P.M = M; P.N = N; P.K = K; P.a_mn = a_mn_major; P.b_mn = b_mn_major;
  P.a_lbo = a_lbo; P.a_sbo = a_sbo; P.a_kstep = a_kstep; P.b_lbo = b_lbo; P.b_sbo = b_sbo; P.b_kstep = b_kstep;
  P.m_valid = 128; P.D = D; P.repeat = repeat < 0 ? -repeat : repeat; P.cycles = cycles; P.rr2 = repeat < 0;
  const size_t smem = ((a_bytes + 127) & ~127) + b_bytes + 128;
  cudaError_t e = cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(P);
  e = cudaGetLastError();
  return e == cudaSuccess ? DLAP_OK : (int)e;
}
