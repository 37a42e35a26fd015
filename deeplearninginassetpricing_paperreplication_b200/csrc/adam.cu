// clip_grad_norm_ + Adam.step() of the training loop (src/train.py:87-94) over the model's few tensors (12,233
// parameters in all) in ONE launch of one CTA: sum of squares of every gradient in a fixed order, the clipping
// coefficient, then torch.optim.Adam's update (the arithmetic of its default single-/multi-tensor paths, op by op:
// lerp for exp_avg, mul + addcmul for exp_avg_sq, sqrt / sqrt(bias_correction2) + eps, addcdiv).  The step counts
// -- one per tensor, as in torch: a parameter that had no gradient in some step lags behind -- live in device memory
// so that the launch can be replayed from a CUDA graph.
#include "../../include/dlap.h"
#include "dlap_common.cuh"

namespace dlap {

constexpr int AT = 1024;

__global__ void __launch_bounds__(AT, 1) clip_adam_kernel(const dlap_adam_tensors A, float max_norm, float lr,
                                                          float beta1, float beta2, float eps,
                                                          float* __restrict__ norm_out) {
  __shared__ double red[32];
  const int tid = threadIdx.x;
  // ---- total gradient norm (torch.nn.utils.clip_grad_norm_: 2-norm over all gradients)
  double ss = 0.0;
  for (int k = 0; k < A.count; ++k) {
    const float* g = A.grad[k];
    for (long long i = tid; i < A.numel[k]; i += AT) { const double v = (double)g[i]; ss += v * v; }
  }
  ss = block_sum(ss, red);
  const float total = (float)sqrt(ss);
  float coef = 1.f;
  if (max_norm > 0.f) coef = fminf(max_norm / (total + 1e-6f), 1.f);
  // ---- Adam
  const float w1 = 1.f - beta1, w2 = 1.f - beta2;
  for (int k = 0; k < A.count; ++k) {
    float* p = A.param[k]; float* g = A.grad[k]; float* m = A.exp_avg[k]; float* v = A.exp_avg_sq[k];
    const long long t = *reinterpret_cast<const long long*>(A.step[k]) + 1;
    const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
    const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
    for (long long i = tid; i < A.numel[k]; i += AT) {
      float gi = g[i];
      if (max_norm > 0.f) { gi *= coef; g[i] = gi; }                 // clip_grad_norm_ scales the gradients in place
      // roundings placed where torch's kernels place them (each of its ops is a kernel of its own, whose
      // a + alpha * x contracts to one fma)
      const float mi = __fmaf_rn(w1, __fsub_rn(gi, m[i]), m[i]);     // exp_avg.lerp_(grad, 1 - beta1)
      const float vi = __fmaf_rn(w2, __fmul_rn(gi, gi), __fmul_rn(v[i], beta2));   // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
      m[i] = mi; v[i] = vi;
      const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), bc2_sqrt), eps);
      p[i] = __fmaf_rn(-step_size, __fdiv_rn(mi, denom), p[i]);      // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
  }
  __syncthreads();
  if (tid < A.count) *reinterpret_cast<long long*>(A.step[tid]) += 1;
  if (tid == 0 && norm_out) *norm_out = total;
}

}  // namespace dlap

extern "C" int dlap_clip_adam(const dlap_adam_tensors* tensors, float max_norm, float lr, float beta1, float beta2,
                              float eps, float* grad_norm, void* stream) {
  using namespace dlap;
  if (!tensors) return DLAP_E_NULL;
  if (tensors->count < 0 || tensors->count > DLAP_ADAM_MAX_TENSORS) return DLAP_E_SHAPE;
  for (int k = 0; k < tensors->count; ++k) {
    if (!tensors->param[k] || !tensors->grad[k] || !tensors->exp_avg[k] || !tensors->exp_avg_sq[k] || !tensors->step[k])
      return DLAP_E_NULL;
    if (reinterpret_cast<uintptr_t>(tensors->step[k]) & 7) return DLAP_E_ALIGN;
    if (tensors->numel[k] < 0) return DLAP_E_SHAPE;
  }
  if (!(lr >= 0.f) || !(beta1 >= 0.f && beta1 < 1.f) || !(beta2 >= 0.f && beta2 < 1.f) || !(eps >= 0.f)) return DLAP_E_SHAPE;
  clip_adam_kernel<<<1, AT, 0, (cudaStream_t)stream>>>(*tensors, max_norm, lr, beta1, beta2, eps, grad_norm);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DLAP_OK : (int)e;
}
