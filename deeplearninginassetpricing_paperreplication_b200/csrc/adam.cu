// clip_grad_norm_ + Adam.step() of the training loop (src/train.py:87-94) over the model's few tensors (12,233
// parameters in all) in ONE launch of one CTA: sum of squares of every gradient in a fixed order, the clipping
// coefficient, then torch.optim.Adam's update (the arithmetic of its default single-/multi-tensor paths, op by op:
// lerp for exp_avg, mul + addcmul for exp_avg_sq, sqrt / sqrt(bias_correction2) + eps, addcdiv).  The step count
// lives in device memory so that the launch can be replayed from a CUDA graph.
#include "../../include/dlap.h"
#include "dlap_common.cuh"

namespace dlap {

constexpr int AT = 1024;

__global__ void __launch_bounds__(AT, 1) clip_adam_kernel(const dlap_adam_tensors A, float max_norm, float lr,
                                                          float beta1, float beta2, float eps,
                                                          long long* __restrict__ step, float* __restrict__ norm_out) {
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
  const long long t = *step + 1;
  const double bc1 = 1.0 - pow((double)beta1, (double)t), bc2 = 1.0 - pow((double)beta2, (double)t);
  const float step_size = (float)((double)lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  const float w1 = 1.f - beta1, w2 = 1.f - beta2;
  for (int k = 0; k < A.count; ++k) {
    float* p = A.param[k]; float* g = A.grad[k]; float* m = A.exp_avg[k]; float* v = A.exp_avg_sq[k];
    for (long long i = tid; i < A.numel[k]; i += AT) {
      float gi = g[i];
      if (max_norm > 0.f) { gi *= coef; g[i] = gi; }                 // clip_grad_norm_ scales the gradients in place
      const float mi = m[i] + w1 * (gi - m[i]);                      // exp_avg.lerp_(grad, 1 - beta1)
      const float vi = v[i] * beta2 + w2 * gi * gi;                  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
      m[i] = mi; v[i] = vi;
      const float denom = sqrtf(vi) / bc2_sqrt + eps;
      p[i] = p[i] - step_size * (mi / denom);                        // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
  }
  __syncthreads();
  if (tid == 0) {
    *step = t;
    if (norm_out) *norm_out = total;
  }
}

}  // namespace dlap

extern "C" int dlap_clip_adam(const dlap_adam_tensors* tensors, float max_norm, float lr, float beta1, float beta2,
                              float eps, int64_t* step, float* grad_norm, void* stream) {
  using namespace dlap;
  if (!tensors || !step) return DLAP_E_NULL;
  if (tensors->count < 0 || tensors->count > DLAP_ADAM_MAX_TENSORS) return DLAP_E_SHAPE;
  for (int k = 0; k < tensors->count; ++k) {
    if (!tensors->param[k] || !tensors->grad[k] || !tensors->exp_avg[k] || !tensors->exp_avg_sq[k]) return DLAP_E_NULL;
    if (tensors->numel[k] < 0) return DLAP_E_SHAPE;
  }
  if (!(lr >= 0.f) || !(beta1 >= 0.f && beta1 < 1.f) || !(beta2 >= 0.f && beta2 < 1.f) || !(eps >= 0.f)) return DLAP_E_SHAPE;
  clip_adam_kernel<<<1, AT, 0, (cudaStream_t)stream>>>(*tensors, max_norm, lr, beta1, beta2, eps,
                                                        reinterpret_cast<long long*>(step), grad_norm);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DLAP_OK : (int)e;
}
