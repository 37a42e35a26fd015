// Shared device helpers for the adversarial-SDF kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef __CUDA_ARCH__
#define DLAP_HOST 1
#endif

namespace dlap {

// Fixed architecture of the fast path (reference defaults, src/model.py:304-344).
constexpr int D1 = 64;        // first hidden width   (hidden_dim[0])
constexpr int D2 = 64;        // second hidden width  (hidden_dim[1])
constexpr int KM = 8;         // moments              (num_condition_moment)
constexpr int GS = 12;        // row pitch of the per-stock loss coefficients G [N,GS]: G_0..G_7, G_u, 3 floats of padding
constexpr int TILE = 128;     // stock-months per tile; a tile never straddles a month
constexpr int NT = 256;       // threads per CTA of the per-cell kernels
constexpr int FMAX = 64;      // characteristics supported by the smem layout
constexpr int HMAX = 8;       // LSTM units: 4*H gate rows must fit one warp

// row_stats[t][*] written by row_finish
enum { RS_NT = 0, RS_MU = 1, RS_P = 2, RS_PR = 3, RS_S = 4, RS_RBAR = 5, RS_SR = 6 /* sum_i m R */, RS_W = 8 };
// scalars[*] written by row_finish / loss_finish
enum { SC_LC = 0, SC_LU = 1, SC_SHARPE = 2, SC_NBAR = 3, SC_ZERO = 7 /* always 0.0 */, SC_W = 8 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier + 1-D bulk copy (TMA engine; SASS: UBLKCP) -------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "W_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra D_%=;\n"
      "bra W_%=;\n"
      "D_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared, size and both addresses multiples of 16 B
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// the same for data that is read ONCE per kernel (the panel's X images): the L2 lines are marked evict-first, so the
// stream does not push out what the next kernel wants to find there (u, q: 35 MB written beside 190 MB streamed)
__device__ __forceinline__ void bulk_g2s_stream(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  uint64_t policy;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// ---- smem tile layouts -----------------------------------------------------------------
// Activation tiles are stored "unit-major": row = unit (k), 128 cells per row, the 4-cell
// groups XOR-swizzled by the row so that (a) a thread storing its 4 cells for units
// 4*tx+i (tx = 0..7) and (b) threads reading rows 4*tk+i or 8*tk+i all hit distinct banks.
__device__ __forceinline__ int swz(int k) { return ((k >> 2) & 7) ^ ((k >> 5) & 1); }
__device__ __forceinline__ int act_idx(int k, int cell) {
  return k * TILE + ((((cell >> 2) ^ swz(k))) << 2) + (cell & 3);
}
// index of the float4 holding cells 4*grp .. 4*grp+3 of unit k
__device__ __forceinline__ int act_idx4(int k, int grp) { return k * (TILE / 4) + (grp ^ swz(k)); }

// ---- reductions ------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block sum (fixed tree).  `red` needs blockDim.x/32 slots; result valid in
// every thread.  Contains two __syncthreads.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  T s = 0;
  for (int i = 0; i < nw; ++i) s += red[i];
  return s;
}

__device__ __forceinline__ float ld_nc(const float* p) { return __ldg(p); }

// ---- grid-wide barrier ---------------------------------------------------------------------
// For kernels whose whole grid is co-resident (cooperative launch, or a grid no larger than what the device
// holds at once).  *bar is zero when the kernel starts; the k-th barrier of the launch (k = 1, 2, ...) opens
// when the word reaches k * gridDim.x.  Data written before the barrier is visible after it to loads that do
// not go through the read-only path (plain, ld.cg, ld.acquire).
__device__ __forceinline__ void grid_sync(unsigned int* bar, unsigned int k) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    const unsigned int target = k * gridDim.x;
    unsigned int v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
    } while (v < target);
  }
  __syncthreads();
}
// The last thing such a kernel does (nsync = barriers it has passed): the last CTA to get here puts the word back
// to zero for the next launch on the stream.
__device__ __forceinline__ void grid_exit(unsigned int* bar, unsigned int nsync) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int old = atomicAdd(bar, 1u);
    if (old == (nsync + 1u) * gridDim.x - 1u) atomicExch(bar, 0u);
  }
}

}  // namespace dlap
