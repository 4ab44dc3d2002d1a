// Fused image loss: L1 + SSIM (11x11 Gaussian window, sigma 1.5), forward and backward.
//
// Reference path replaced (the step right after the rasterizer in a train iteration):
// train.py:105-108  loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))
// utils/loss_utils.py:54-100  l1_loss, gaussian, create_window, ssim, _ssim  (five grouped 11x11 conv2d with zero
// padding + ~15 elementwise kernels forward, and their autograd backward).
//
// Here: one forward kernel (separable window evaluated in shared memory on a 26x26 halo tile per 16x16 output tile,
// the five local moments, the SSIM map, |x - y|, a deterministic two-stage sum, and the three derivative maps the
// backward needs) and one backward kernel (the same separable window over the derivative maps).  Per pixel-channel
// the forward reads 8 B and writes 12 B, the backward reads 20 B and writes 4 B.
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_IL_T 16              // output tile edge
#define GP_IL_R 5               // window radius
#define GP_IL_H (GP_IL_T + 2 * GP_IL_R)   // 26: halo tile edge

__constant__ float gp_il_w[11];
static bool gp_il_w_ready[64];

static int gp_il_init_weights() {
  int dev = 0;
  GP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return GP_ERR_UNSUPPORTED;
  if (gp_il_w_ready[dev]) return GP_OK;
  // utils/loss_utils.py:60-62: exp(-(x - 5)^2 / (2 * 1.5^2)) as float32, normalised by their float32 sum
  float g[11];
  float s = 0.f;
  for (int i = 0; i < 11; i++) {
    g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5));
    s += g[i];
  }
  for (int i = 0; i < 11; i++) g[i] = g[i] / s;
  GP_CUDA_TRY(cudaMemcpyToSymbol(gp_il_w, g, sizeof(g)));
  gp_il_w_ready[dev] = true;
  return GP_OK;
}

// forward: per block partial sums of |x-y| and of the SSIM map; derivative maps for the backward
__global__ void __launch_bounds__(GP_IL_T * GP_IL_T)
gp_image_loss_fwd_kernel(int C, int H, int W, const float* __restrict__ img, const float* __restrict__ gt,
                         float* __restrict__ dmaps /* [3][C][H][W] or NULL */, float2* __restrict__ partial) {
  __shared__ float s_x[GP_IL_H][GP_IL_H + 1];
  __shared__ float s_y[GP_IL_H][GP_IL_H + 1];
  __shared__ float s_h[5][GP_IL_H][GP_IL_T + 1];   // horizontally filtered rows of x, y, xx, yy, xy
  __shared__ float2 s_red[GP_IL_T * GP_IL_T / 32];
  const int tx = threadIdx.x % GP_IL_T, ty = threadIdx.x / GP_IL_T;
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * GP_IL_T, y0 = blockIdx.y * GP_IL_T;
  const size_t plane = (size_t)H * W;
  const float* ip = img + (size_t)c * plane;
  const float* gp = gt + (size_t)c * plane;
  for (int e = threadIdx.x; e < GP_IL_H * GP_IL_H; e += GP_IL_T * GP_IL_T) {
    const int ly = e / GP_IL_H, lx = e - ly * GP_IL_H;
    const int gx = x0 + lx - GP_IL_R, gy = y0 + ly - GP_IL_R;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;     // zero padding (conv2d padding=5)
    s_x[ly][lx] = in ? __ldg(ip + (size_t)gy * W + gx) : 0.f;
    s_y[ly][lx] = in ? __ldg(gp + (size_t)gy * W + gx) : 0.f;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < GP_IL_H * GP_IL_T; e += GP_IL_T * GP_IL_T) {
    const int ly = e / GP_IL_T, lx = e - ly * GP_IL_T;
    float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float wv = gp_il_w[k];
      const float xv = s_x[ly][lx + k], yv = s_y[ly][lx + k];
      a += wv * xv;
      b += wv * yv;
      aa += wv * xv * xv;
      bb += wv * yv * yv;
      ab += wv * xv * yv;
    }
    s_h[0][ly][lx] = a; s_h[1][ly][lx] = b; s_h[2][ly][lx] = aa; s_h[3][ly][lx] = bb; s_h[4][ly][lx] = ab;
  }
  __syncthreads();
  float mu1 = 0.f, mu2 = 0.f, ex2 = 0.f, ey2 = 0.f, exy = 0.f;
#pragma unroll
  for (int k = 0; k < 11; k++) {
    const float wv = gp_il_w[k];
    mu1 += wv * s_h[0][ty + k][tx];
    mu2 += wv * s_h[1][ty + k][tx];
    ex2 += wv * s_h[2][ty + k][tx];
    ey2 += wv * s_h[3][ty + k][tx];
    exy += wv * s_h[4][ty + k][tx];
  }
  const int px = x0 + tx, py = y0 + ty;
  const bool inside = px < W && py < H;
  float l1 = 0.f, ss = 0.f;
  if (inside) {
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
    const float sigma1_sq = ex2 - mu1_sq, sigma2_sq = ey2 - mu2_sq, sigma12 = exy - mu1_mu2;
    const float A = 2.f * mu1_mu2 + C1, B = 2.f * sigma12 + C2;
    const float Cc = mu1_sq + mu2_sq + C1, Dd = sigma1_sq + sigma2_sq + C2;
    ss = (A * B) / (Cc * Dd);
    l1 = fabsf(s_x[ty + GP_IL_R][tx + GP_IL_R] - s_y[ty + GP_IL_R][tx + GP_IL_R]);
    if (dmaps) {
      const float inv = 1.f / (Cc * Dd);
      const size_t o = (size_t)c * plane + (size_t)py * W + px;
      const size_t cm = (size_t)C * plane;
      // d ssim / d mu1 (through A, B, Cc, Dd), d / d E[x^2], d / d E[xy]
      dmaps[o] = 2.f * mu2 * (B - A) * inv - 2.f * mu1 * A * B * (Dd - Cc) * inv * inv;
      dmaps[cm + o] = -A * B * inv / Dd;
      dmaps[2 * cm + o] = 2.f * A * inv;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = make_float2(l1, ss);
  __syncthreads();
  if (threadIdx.x == 0) {
    float2 t = make_float2(0.f, 0.f);
    for (int j = 0; j < GP_IL_T * GP_IL_T / 32; j++) {
      t.x += s_red[j].x;
      t.y += s_red[j].y;
    }
    partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
  }
}

// deterministic final sum: sums[0] = mean |x - y|, sums[1] = mean SSIM
__global__ void __launch_bounds__(1024) gp_image_loss_sum_kernel(int n, const float2* __restrict__ partial, double inv_count,
                                                                 float* __restrict__ sums) {
  __shared__ double s_a[32], s_b[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) {
    a += (double)partial[i].x;
    b += (double)partial[i].y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s_a[threadIdx.x >> 5] = a;
    s_b[threadIdx.x >> 5] = b;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0.0, tb = 0.0;
    for (int j = 0; j < 32; j++) {
      ta += s_a[j];
      tb += s_b[j];
    }
    sums[0] = (float)(ta * inv_count);
    sums[1] = (float)(tb * inv_count);
  }
}

// backward: grad_img = gscale * ( w_l1 * sign(x - y) / n + w_ssim / n * (conv(dmu1) + 2 x conv(dEx2) + y conv(dExy)) )
__global__ void __launch_bounds__(GP_IL_T * GP_IL_T)
gp_image_loss_bwd_kernel(int C, int H, int W, const float* __restrict__ img, const float* __restrict__ gt,
                         const float* __restrict__ dmaps, float w_l1, float w_ssim, const float* __restrict__ gscale,
                         float inv_count, float* __restrict__ grad_img) {
  __shared__ float s_m[3][GP_IL_H][GP_IL_H + 1];
  __shared__ float s_h[3][GP_IL_H][GP_IL_T + 1];
  const int tx = threadIdx.x % GP_IL_T, ty = threadIdx.x / GP_IL_T;
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * GP_IL_T, y0 = blockIdx.y * GP_IL_T;
  const size_t plane = (size_t)H * W, cm = (size_t)C * plane;
  for (int e = threadIdx.x; e < GP_IL_H * GP_IL_H; e += GP_IL_T * GP_IL_T) {
    const int ly = e / GP_IL_H, lx = e - ly * GP_IL_H;
    const int gx = x0 + lx - GP_IL_R, gy = y0 + ly - GP_IL_R;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
    const size_t o = (size_t)c * plane + (size_t)gy * W + gx;
#pragma unroll
    for (int m = 0; m < 3; m++) s_m[m][ly][lx] = in ? __ldg(dmaps + m * cm + o) : 0.f;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < GP_IL_H * GP_IL_T; e += GP_IL_T * GP_IL_T) {
    const int ly = e / GP_IL_T, lx = e - ly * GP_IL_T;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float wv = gp_il_w[k];
      a0 += wv * s_m[0][ly][lx + k];
      a1 += wv * s_m[1][ly][lx + k];
      a2 += wv * s_m[2][ly][lx + k];
    }
    s_h[0][ly][lx] = a0; s_h[1][ly][lx] = a1; s_h[2][ly][lx] = a2;
  }
  __syncthreads();
  const int px = x0 + tx, py = y0 + ty;
  if (px >= W || py >= H) return;
  float g0 = 0.f, g1 = 0.f, g2 = 0.f;
#pragma unroll
  for (int k = 0; k < 11; k++) {
    const float wv = gp_il_w[k];
    g0 += wv * s_h[0][ty + k][tx];
    g1 += wv * s_h[1][ty + k][tx];
    g2 += wv * s_h[2][ty + k][tx];
  }
  const size_t o = (size_t)c * plane + (size_t)py * W + px;
  const float x = img[o], y = gt[o];
  const float d = x - y;
  const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);   // torch.abs backward: sign(0) = 0
  const float gs = gscale ? *gscale : 1.f;
  grad_img[o] = gs * inv_count * (w_l1 * sgn + w_ssim * (g0 + 2.f * x * g1 + y * g2));
}

extern "C" size_t gp_image_loss_ws_bytes(int C, int H, int W) {
  if (C <= 0 || H <= 0 || W <= 0) return 0;
  const size_t blocks = (size_t)C * ((H + GP_IL_T - 1) / GP_IL_T) * ((W + GP_IL_T - 1) / GP_IL_T);
  return gp_align_up(blocks * sizeof(float2), 256);
}

extern "C" int gp_image_loss_forward(int C, int H, int W, const float* img, const float* gt, float* dmaps, void* ws,
                                     float* sums, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0 || !img || !gt || !ws || !sums) return GP_ERR_BAD_ARG;
  int rc = gp_il_init_weights();
  if (rc) return rc;
  dim3 grid((W + GP_IL_T - 1) / GP_IL_T, (H + GP_IL_T - 1) / GP_IL_T, C);
  if (grid.z > 65535) return GP_ERR_UNSUPPORTED;
  const int nblocks = (int)(grid.x * grid.y * grid.z);
  GP_KERNEL("gp_image_loss_fwd_kernel", stream,
            gp_image_loss_fwd_kernel<<<grid, GP_IL_T * GP_IL_T, 0, stream>>>(C, H, W, img, gt, dmaps, (float2*)ws));
  GP_KERNEL("gp_image_loss_sum_kernel", stream,
            gp_image_loss_sum_kernel<<<1, 1024, 0, stream>>>(nblocks, (const float2*)ws, 1.0 / ((double)C * H * W), sums));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_image_loss_backward(int C, int H, int W, const float* img, const float* gt, const float* dmaps,
                                      float w_l1, float w_ssim, const float* grad_scale, float* grad_img,
                                      gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0 || !img || !gt || !dmaps || !grad_img) return GP_ERR_BAD_ARG;
  int rc = gp_il_init_weights();
  if (rc) return rc;
  dim3 grid((W + GP_IL_T - 1) / GP_IL_T, (H + GP_IL_T - 1) / GP_IL_T, C);
  if (grid.z > 65535) return GP_ERR_UNSUPPORTED;
  GP_KERNEL("gp_image_loss_bwd_kernel", stream,
            gp_image_loss_bwd_kernel<<<grid, GP_IL_T * GP_IL_T, 0, stream>>>(C, H, W, img, gt, dmaps, w_l1, w_ssim, grad_scale,
                                                                           (float)(1.0 / ((double)C * H * W)), grad_img));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
