// The step right after the backward: the Adam update of the parameter groups and the densification
// statistics of the frame.
//
// Stands in for `gaussians.optimizer.step()` (train.py:192-194; the optimizer is torch.optim.Adam(l, lr=0.0, eps=1e-15)
// over ~20 tensors in 8-9 groups, scene/gaussian_model.py:439-452) and for the per-frame bookkeeping of train.py:166-167,
// 183-184 / scene/gaussian_model.py:756-760 (max_radii2D, add_densification_stats).  torch runs Adam as ~12 foreach
// launches over all tensors (lerp, mul, addcmul, sqrt, div, add, addcdiv ...), each a full pass over parameters, gradients
// and both moments: ~11 passes of 4 B per element where one pass of 28 B/element (read p, g, m, v; write p, m, v)
// suffices.  Here every tensor of the step is updated by ONE launch; HBM bound.
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_ADAM_MAX_TENSORS 48
#define GP_ADAM_THREADS 256
#define GP_ADAM_ELEMS_PER_BLOCK (GP_ADAM_THREADS * 16)

struct AdamTensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  float step_size;    // lr / (1 - beta1^step)
  float inv_bc2_sqrt; // 1 / sqrt(1 - beta2^step)
  int first_block;    // blocks [first_block, next.first_block) work on this tensor
};
struct AdamBatch {
  AdamTensor t[GP_ADAM_MAX_TENSORS];
  int count;
};

// Element update in the operation order of torch.optim.Adam's kernels (torch/optim/adam.py, _single_tensor_adam /
// _multi_tensor_adam; no weight decay, no amsgrad, as the reference configures it):
//   exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
//   denom = exp_avg_sq.sqrt() / sqrt(bias_correction2) + eps; param.addcdiv_(exp_avg, denom, value = -step_size)
__device__ __forceinline__ void adam_element(float& p, const float g, float& m, float& v, const float w1, const float beta2,
                                             const float w2, const float eps, const float step_size,
                                             const float inv_bc2_sqrt) {
  m = m + w1 * (g - m);
  v = v * beta2;
  v = v + (w2 * g) * g;
  const float denom = sqrtf(v) * inv_bc2_sqrt + eps;
  p = p + (-step_size) * (m / denom);
}

static __global__ void __launch_bounds__(GP_ADAM_THREADS)
gp_adam_kernel(const __grid_constant__ AdamBatch b, float w1, float beta2, float w2, float eps) {
  // which tensor does this block belong to? (count <= 48: a short scan of the parameter struct)
  int k = 0;
  while (k + 1 < b.count && (int)blockIdx.x >= b.t[k + 1].first_block) k++;
  const AdamTensor& t = b.t[k];
  const long long base = (long long)((int)blockIdx.x - t.first_block) * GP_ADAM_ELEMS_PER_BLOCK;
  const long long end = min(t.n, base + GP_ADAM_ELEMS_PER_BLOCK);
  const bool vec = ((((size_t)t.p | (size_t)t.g | (size_t)t.m | (size_t)t.v) & 15) == 0);
  if (vec) {
    for (long long i = base + 4 * (long long)threadIdx.x; i + 3 < end; i += 4 * GP_ADAM_THREADS) {
      float4 p = *reinterpret_cast<const float4*>(t.p + i);
      const float4 g = __ldg(reinterpret_cast<const float4*>(t.g + i));
      float4 m = *reinterpret_cast<const float4*>(t.m + i);
      float4 v = *reinterpret_cast<const float4*>(t.v + i);
      adam_element(p.x, g.x, m.x, v.x, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      adam_element(p.y, g.y, m.y, v.y, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      adam_element(p.z, g.z, m.z, v.z, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      adam_element(p.w, g.w, m.w, v.w, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      *reinterpret_cast<float4*>(t.p + i) = p;
      *reinterpret_cast<float4*>(t.m + i) = m;
      *reinterpret_cast<float4*>(t.v + i) = v;
    }
    // tail of the tensor (n not a multiple of 4): the last block's first threads
    const long long tail = end & ~3LL;
    if (end == t.n && tail + threadIdx.x < end && tail >= base) {
      const long long i = tail + threadIdx.x;
      float p = t.p[i], m = t.m[i], v = t.v[i];
      adam_element(p, t.g[i], m, v, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      t.p[i] = p; t.m[i] = m; t.v[i] = v;
    }
  } else {
    for (long long i = base + threadIdx.x; i < end; i += GP_ADAM_THREADS) {
      float p = t.p[i], m = t.m[i], v = t.v[i];
      adam_element(p, t.g[i], m, v, w1, beta2, w2, eps, t.step_size, t.inv_bc2_sqrt);
      t.p[i] = p; t.m[i] = m; t.v[i] = v;
    }
  }
}

extern "C" int gp_adam_step(int count, const gp_adam_tensor* tensors, double beta1, double beta2, double eps,
                            gp_stream_t stream_) {
  if (count < 0 || (count > 0 && !tensors)) return GP_ERR_BAD_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  // the hyper-parameters arrive as doubles (Python floats in the reference): 1 - 0.999f would already be off by 1.3e-5
  const float w1 = (float)(1.0 - beta1), w2 = (float)(1.0 - beta2);
  int done = 0;
  while (done < count) {
    AdamBatch b = {};
    int blocks = 0;
    while (done < count && b.count < GP_ADAM_MAX_TENSORS) {
      const gp_adam_tensor& s = tensors[done++];
      if (s.n < 0 || s.step < 1) return GP_ERR_BAD_ARG;
      if (s.n == 0) continue;
      if (!s.param || !s.grad || !s.exp_avg || !s.exp_avg_sq) return GP_ERR_BAD_ARG;
      AdamTensor& t = b.t[b.count++];
      t.p = s.param; t.g = s.grad; t.m = s.exp_avg; t.v = s.exp_avg_sq; t.n = s.n;
      // bias corrections in double on the host, like the Python floats of torch/optim/adam.py
      const double bc1 = 1.0 - pow(beta1, (double)s.step), bc2 = 1.0 - pow(beta2, (double)s.step);
      t.step_size = (float)((double)s.lr / bc1);
      t.inv_bc2_sqrt = 1.0f / (float)sqrt(bc2);
      t.first_block = blocks;
      blocks += (int)((s.n + GP_ADAM_ELEMS_PER_BLOCK - 1) / GP_ADAM_ELEMS_PER_BLOCK);
    }
    if (blocks == 0) continue;
    GP_KERNEL("gp_adam_kernel", stream, gp_adam_kernel<<<blocks, GP_ADAM_THREADS, 0, stream>>>(b, w1, (float)beta2, w2, (float)eps));
    GP_LAUNCH_CHECK();
  }
  return GP_OK;
}

// max_radii2D[vis] = max(max_radii2D[vis], radii[vis]); add_densification_stats(viewspace_points.grad, vis) with
// vis = radii > 0 (train.py:166-167, scene/gaussian_model.py:756-760): one pass, 36 B per Gaussian.
static __global__ void __launch_bounds__(256)
gp_densification_stats_kernel(int P, const float* __restrict__ vgrad, const int* __restrict__ radii,
                              float* __restrict__ max_radii2D, float* __restrict__ accum, float* __restrict__ denom,
                              float* __restrict__ accum_max) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  if (r <= 0) return;
  if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
  const float gx = vgrad[3 * (size_t)i], gy = vgrad[3 * (size_t)i + 1];
  const float g = sqrtf(gx * gx + gy * gy);
  accum[i] += g;
  denom[i] += 1.0f;
  if (accum_max) {
    const float m = accum_max[i];
    accum_max[i] = g > m ? g : m;
  }
}

extern "C" int gp_densification_stats(int P, const float* viewspace_grad, const int* radii, float* max_radii2D,
                                      float* xyz_gradient_accum, float* denom, float* xyz_gradient_accum_max,
                                      gp_stream_t stream_) {
  if (P < 0 || (P > 0 && (!viewspace_grad || !radii || !xyz_gradient_accum || !denom))) return GP_ERR_BAD_ARG;
  if (P == 0) return GP_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  GP_KERNEL("gp_densification_stats_kernel", stream,
            gp_densification_stats_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, viewspace_grad, radii, max_radii2D,
                                                                              xyz_gradient_accum, denom,
                                                                              xyz_gradient_accum_max));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
