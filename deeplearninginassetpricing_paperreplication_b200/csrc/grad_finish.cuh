// Fixed-order sums of the per-cell backward kernels' partials into the reference's gradient layouts: shared by
// grad_finish_kernel (reduce.cu; a launch of its own, month shards exchange right after it) and by the BPTT kernel
// (macro_lstm.cu), whose extra CTAs do the same work beside the serial recurrence.
#pragma once
#include "cell_f32.cuh"

namespace dlap {

struct FinishParams {
  const float* wpart; int ncta, F, have_sdf, have_mom;
  const float* dadc_part; int T, tpm;      // tpm: partial slots per month (dense-tile layout)
  const int* index;                        // != null: (month, CTA) segment layout of the tensor-core kernel
  float* gW1; int ldw1; float* gW2; float* gb2; float* gW3; float* gb3; float* gWq; int ldwq, wq_off;
  float* dadc; int dadc_stride;            // row t of this panel goes to row t * dadc_stride of dadc
  // BPTT kernel only (macro_lstm.cu): CTAs that sum da_t | dc_t elements count themselves off here
  unsigned int* counter; int dadc_ctas; int cooperative;      // counter: four words, zero between launches
  long long* tl;                           // diagnostics (dlap_debug_timeline): globaltimer stamps, or null
  int fold_wgrad;                          // the summing CTAs go on to the month-summed gradients (macro_lstm.cu)
  // stock shards over peer memory (BPTT kernel only): the sums of this rank's partials are this rank's share; every
  // rank writes its shares into gpeer[rank] (two halves of gcap floats, used alternately, then 8 uint32 flags), a flag
  // barrier across the GPUs, and every rank adds everybody's shares in rank order.  gepoch[0]: launches so far,
  // gepoch[1]: the "go" word the summing CTAs wait on.
  int world, rank, gcap, nfin;
  float* gpeer[8];
  unsigned int* gepoch;
};

// element idx < wp_stride(round4(F)) of the per-CTA weight-gradient partials, summed over the CTAs
__device__ __forceinline__ float wpart_sum(const FinishParams& P, int idx) {
  const int Fr = round4(P.F), stride = wp_stride(Fr), ncta = P.ncta;
  // four running sums (CTAs c = u mod 4, ascending), sixteen loads in flight: one L2 round trip per batch of loads
  float v4[4] = {0.f, 0.f, 0.f, 0.f};
  int c = 0;
  for (; c + 16 <= ncta; c += 16) {
    float ld[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) ld[u] = __ldg(P.wpart + (size_t)(c + u) * stride + idx);
#pragma unroll
    for (int u = 0; u < 16; ++u) v4[u & 3] += ld[u];
  }
  for (; c + 4 <= ncta; c += 4) {
#pragma unroll
    for (int u = 0; u < 4; ++u) v4[u] += P.wpart[(size_t)(c + u) * stride + idx];
  }
  for (; c < ncta; ++c) v4[0] += P.wpart[(size_t)c * stride + idx];
  return (v4[0] + v4[1]) + (v4[2] + v4[3]);
}
// ... and stored where the reference keeps that gradient
__device__ __forceinline__ void wpart_store(const FinishParams& P, int idx, float v) {
  const int Fr = round4(P.F), F = P.F;
  if (idx < wp_off_w1()) {
    if (P.have_sdf) P.gW2[idx] = v;
  } else if (idx < wp_off_wq(Fr)) {
    const int j = (idx - wp_off_w1()) / Fr, f = (idx - wp_off_w1()) % Fr;
    if (P.have_sdf && f < F) P.gW1[(size_t)j * P.ldw1 + f] = v;
  } else if (idx < wp_off_w3(Fr)) {
    const int k = (idx - wp_off_wq(Fr)) / Fr, f = (idx - wp_off_wq(Fr)) % Fr;
    if (P.have_mom && f < F) P.gWq[(size_t)k * P.ldwq + P.wq_off + f] = v;
  } else if (idx < wp_off_b2(Fr)) {
    if (P.have_sdf) P.gW3[idx - wp_off_w3(Fr)] = v;
  } else if (idx < wp_off_b2(Fr) + 64) {
    if (P.have_sdf) P.gb2[idx - wp_off_b2(Fr)] = v;
  }
  // d b3 = sum du = dP_t * sum m (R - Rbar_t) = 0 identically (demeaning, src/model.py:279)
  if (idx == 0 && P.have_sdf) P.gb3[0] = 0.f;
}
__device__ __forceinline__ void finish_wpart(const FinishParams& P, int idx) { wpart_store(P, idx, wpart_sum(P, idx)); }

// element r < T * 72 of da_t | dc_t, summed over the month's partial slots
__device__ __forceinline__ float dadc_sum(const FinishParams& P, int r) {
  const int t = r / 72, c = r % 72;
  int tpm = P.tpm;
  float v8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const float* p = P.dadc_part + (size_t)t * tpm * 72 + c;
  if (P.index) {             // one partial per CTA that owns tiles of this month (consecutive CTAs)
    const int* ts = P.index + idx_off_ts();
    const int chunk = tc_chunk(P.index[0], P.ncta), a = ts[t], b = ts[t + 1];
    tpm = b > a ? (b - 1) / chunk - a / chunk + 1 : 0;
  }
  int j = 0;
  for (; j + 8 <= tpm; j += 8) {
#pragma unroll
    for (int u = 0; u < 8; ++u) v8[u] += p[(size_t)(j + u) * 72];
  }
  for (; j < tpm; ++j) v8[0] += p[(size_t)j * 72];
  return ((v8[0] + v8[1]) + (v8[2] + v8[3])) + ((v8[4] + v8[5]) + (v8[6] + v8[7]));
}
__device__ __forceinline__ void dadc_store(const FinishParams& P, int r, float v) {
  P.dadc[(size_t)(r / 72) * P.dadc_stride * 72 + r % 72] = v;
}
__device__ __forceinline__ void finish_dadc(const FinishParams& P, int r) { dadc_store(P, r, dadc_sum(P, r)); }

}  // namespace dlap
