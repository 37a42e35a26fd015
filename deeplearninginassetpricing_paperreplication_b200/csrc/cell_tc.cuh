// Tensor-core variant of the per-cell kernels (precision "bf16x3").
//
// Every fp32 contraction D = A.B^T of the SDF MLP / moment net is issued as three tcgen05 BF16 MMAs
// on split operands  a = a_hi + a_lo  (|a - a_hi - a_lo| <= 2^-18 |a|):
//     D = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi            (fp32 accumulation in TMEM)
// which keeps ~fp32 accuracy (measured against the oracle in tests/) at 3x the MMA count.
//
// Shared-memory operand images use the no-swizzle canonical layout [group of 8][row][8] (tc_sm100.cuh).
//   X  [6][128][8]   characteristics of the tile; k = 46 holds 1.0 (the "ones column"), k = 47 = 0
//   W1 [6][80][8]    rows 0..63 = W1[:, :F], rows 64..71 = Wq[:, M:], rows 72..79 = 0;
//                    column k = 46 carries the per-month bias a_t | c_t, so both first-layer biases
//                    are added by the MMA itself
//   H1 [8][128][8]   relu(z1) written by the epilogue warps, A operand of layer 2
//   W2 [8][64][8]
#pragma once
#include "cell_f32.cuh"
#include "tc_sm100.cuh"

namespace dlap {
namespace tcl {

constexpr int KG1 = 6;                  // k-groups of layer 1 (K = 48)
constexpr int N1 = 80;                  // layer-1 MMA N: 64 hidden + 8 moments + 8 pad
constexpr int KG2 = 8;                  // k-groups of layer 2 (K = 64)
constexpr uint32_t X_IMG = KG1 * TILE * 16;      // 12288 B per hi or lo image
constexpr uint32_t W1_IMG = KG1 * N1 * 16;       //  7680 B
constexpr uint32_t H_IMG = KG2 * TILE * 16;      // 16384 B
constexpr uint32_t W2_IMG = KG2 * 64 * 16;       //  8192 B
constexpr int ONES_K = 46;              // index of the ones column inside X (needs F <= 46)

// fill the static part of a W1 image pair (hi, lo) from global parameters
__device__ __forceinline__ void fill_w1_images(const CellParams& P, uint8_t* hi, uint8_t* lo, int tid, int nthr) {
  for (int i = tid; i < KG1 * N1 * 4; i += nthr) {      // one bf16x2 pair per step
    const int g = i / (N1 * 4), n = (i / 4) % N1, pr = i % 4;
    const int k0 = g * 8 + pr * 2;
    float v[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int k = k0 + e;
      float w = 0.f;
      if (k < P.F) {
        if (n < 64) w = __ldg(P.W1 + (size_t)n * P.ldw1 + k);
        else if (n < 72 && P.Wq) w = __ldg(P.Wq + (size_t)(n - 64) * P.ldwq + P.wq_off + k);
      }
      v[e] = w;
    }
    uint32_t h, l;
    tc::split2(v[0], v[1], h, l);
    reinterpret_cast<uint32_t*>(hi)[i] = h;
    reinterpret_cast<uint32_t*>(lo)[i] = l;
  }
}
// W2 image: element (row n = out unit j, k = in unit) = W2[j][k]
__device__ __forceinline__ void fill_w2_images(const CellParams& P, uint8_t* hi, uint8_t* lo, int tid, int nthr) {
  for (int i = tid; i < KG2 * 64 * 4; i += nthr) {
    const int g = i / (64 * 4), n = (i / 4) % 64, pr = i % 4;
    const int k0 = g * 8 + pr * 2;
    uint32_t h, l;
    tc::split2(__ldg(P.W2 + n * 64 + k0), __ldg(P.W2 + n * 64 + k0 + 1), h, l);
    reinterpret_cast<uint32_t*>(hi)[i] = h;
    reinterpret_cast<uint32_t*>(lo)[i] = l;
  }
}

// Dropout on N consecutive hidden units U0.. of one stock-month (N even): v[k] <- v[k] * (0 | scale).
// Same counter-based generator and keying as relu_dropout() of the fp32 kernels (cell_f32.cuh), so the
// two precisions drop exactly the same units for a given seed, forward and backward.
template <int N>
__device__ __forceinline__ void tc_dropout(const CellParams& P, uint32_t cell_global, int layer, int U0, float* v) {
  const uint32_t key = mix32(cell_global ^ P.seed_lo) ^ P.seed_hi;
#pragma unroll
  for (int k = 0; k < N; k += 2) {
    const uint32_t r = mix32(key + (uint32_t)(layer * 32 + ((U0 + k) >> 1)) * 0x9E3779B9u);
    v[k] = ((r & 0xffffu) >= P.drop_thr) ? v[k] * P.drop_scale : 0.f;
    v[k + 1] = ((r >> 16) >= P.drop_thr) ? v[k + 1] * P.drop_scale : 0.f;
  }
}

// three-term split product for one K = 16 step
__device__ __forceinline__ void umma3(uint32_t d, uint64_t ah, uint64_t al, uint64_t bh, uint64_t bl,
                                      uint32_t idesc, bool first) {
  tc::umma_bf16(d, ah, bh, idesc, first ? 0u : 1u);
  tc::umma_bf16(d, ah, bl, idesc, 1u);
  tc::umma_bf16(d, al, bh, idesc, 1u);
}

}  // namespace tcl
}  // namespace dlap
