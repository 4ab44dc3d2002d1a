// Per-Gaussian blending of the K nearest keypoints' translations and rotations (sparse form).
//
// Reference path replaced (scene/gaussian_model.py): fill_nearest (:214-229: two softmaxes over K
// logits scattered into dense zeros[N,Kp]), the two dense products weights_r @ delta_q and
// weights_xyz @ delta_xyz (:272-273), xyz + delta (:285), get_rotation_ = normalize(quat_mul(
// normalize(delta_q), _rotation)) (:286, :314-315; quat_mul: utils/camera_utils.py:158-170),
// exp(_scaling) (:287) and sigmoid(_opacity) (:304).  The dense [N,Kp] matrices (2 x 2 GB at 1M
// Gaussians x 512 keypoints) are never materialised: each thread reads its K indices and 2K logits,
// gathers the <= K keypoint deltas from shared memory and writes 48 B.  The same kernel serves the
// extrapolation loop of eval.py:125-141 (pass kpts[i] - kpts as delta_xyz and kpts_rotation[i] as
// delta_q).
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_BLEND_THREADS 256
#define GP_BLEND_KMAX 16

__device__ __forceinline__ float4 quat_mul_ref(const float4 q1, const float4 q2) {
  // (w,x,y,z); returns q1 (x) q2 with the operand order of utils/camera_utils.py:165-169
  float4 r;
  r.x = q2.x * q1.x - q2.y * q1.y - q2.z * q1.z - q2.w * q1.w;
  r.y = q2.y * q1.x + q2.x * q1.y + q2.w * q1.z - q2.z * q1.w;
  r.z = q2.z * q1.x - q2.w * q1.y + q2.x * q1.z + q2.y * q1.w;
  r.w = q2.w * q1.x + q2.z * q1.y - q2.y * q1.z + q2.x * q1.w;
  return r;
}

__device__ __forceinline__ float qnorm(const float4 q) { return sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w); }

// softmax over K values (torch: exp(x - max) / sum)
template <int KMAX>
__device__ __forceinline__ void softmax_k(const float* __restrict__ logit, int K, float* w) {
  float m = -__int_as_float(0x7f800000);
#pragma unroll
  for (int k = 0; k < KMAX; k++)
    if (k < K) m = fmaxf(m, logit[k]);
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < KMAX; k++)
    if (k < K) {
      w[k] = expf(logit[k] - m);
      s += w[k];
    }
  const float inv = 1.0f / s;
#pragma unroll
  for (int k = 0; k < KMAX; k++)
    if (k < K) w[k] *= inv;
}

// One Gaussian's 2K logits and K neighbour indices into registers.  With the reference's nearest_num = 6 the rows are
// 48 and 24 bytes: three 128-bit and three 64-bit loads instead of 18 scalar ones.
template <int KMAX, bool VEC6>
__device__ __forceinline__ void load_row(int i, int K, const float* __restrict__ logits, const int* __restrict__ idx32,
                                         const long long* __restrict__ idx64, int Kp, float* lg, int* jj) {
  if (VEC6) {
    const float4* l4 = reinterpret_cast<const float4*>(logits + (size_t)i * 12);
    const float4 a = __ldg(l4), b = __ldg(l4 + 1), c = __ldg(l4 + 2);
    lg[0] = a.x; lg[1] = a.y; lg[2] = a.z; lg[3] = a.w; lg[4] = b.x; lg[5] = b.y;
    lg[KMAX + 0] = b.z; lg[KMAX + 1] = b.w; lg[KMAX + 2] = c.x; lg[KMAX + 3] = c.y; lg[KMAX + 4] = c.z; lg[KMAX + 5] = c.w;
    const int2* j2 = reinterpret_cast<const int2*>(idx32 + (size_t)i * 6);
    const int2 p = __ldg(j2), q = __ldg(j2 + 1), r = __ldg(j2 + 2);
    jj[0] = p.x; jj[1] = p.y; jj[2] = q.x; jj[3] = q.y; jj[4] = r.x; jj[5] = r.y;
  } else {
#pragma unroll
    for (int k = 0; k < KMAX; k++) {
      if (k < K) {
        lg[k] = logits[(size_t)i * 2 * K + k];
        lg[KMAX + k] = logits[(size_t)i * 2 * K + K + k];
        jj[k] = idx32 ? idx32[(size_t)i * K + k] : (int)idx64[(size_t)i * K + k];
      }
    }
  }
#pragma unroll
  for (int k = 0; k < KMAX; k++)
    if (k < K && (jj[k] < 0 || jj[k] >= Kp)) jj[k] = -1;
}

// s_kp layout: [Kp][8] = (dx, dy, dz, 0, qw, qx, qy, qz)
__device__ __forceinline__ void stage_keypoints(float* s_kp, int Kp, const float* __restrict__ kp_dxyz,
                                                const float* __restrict__ kp_dq) {
  for (int e = threadIdx.x; e < Kp * 8; e += blockDim.x) {
    const int j = e >> 3, c = e & 7;
    float v = 0.f;
    if (c < 3) v = kp_dxyz[(size_t)j * 3 + c];
    else if (c >= 4) v = kp_dq[(size_t)j * 4 + (c - 4)];
    s_kp[e] = v;
  }
}

// VEC6: K == 6 with aligned rows (the host checks): compile-time K, vector row loads.  Otherwise K at run time.
template <bool VEC6>
__global__ void __launch_bounds__(GP_BLEND_THREADS, VEC6 ? 4 : 2)
gp_blend_fwd_kernel(int N, int Kp, int Kr, const float* __restrict__ logits, const int* __restrict__ idx32,
                    const long long* __restrict__ idx64, const float* __restrict__ kp_dxyz,
                    const float* __restrict__ kp_dq, const float* __restrict__ xyz, const float* __restrict__ rot,
                    const float* __restrict__ scaling, const float* __restrict__ opacity, float* __restrict__ xyz_t,
                    float* __restrict__ q_t, float* __restrict__ s_t, float* __restrict__ o_t) {
  extern __shared__ __align__(16) float s_kp[];
  stage_keypoints(s_kp, Kp, kp_dxyz, kp_dq);
  __syncthreads();
  const int K = VEC6 ? 6 : Kr;
  for (int i = blockIdx.x * GP_BLEND_THREADS + threadIdx.x; i < N; i += gridDim.x * GP_BLEND_THREADS) {
  float wx[GP_BLEND_KMAX], wr[GP_BLEND_KMAX];
  float lgr[2 * GP_BLEND_KMAX];
  int jr[GP_BLEND_KMAX];
  load_row<GP_BLEND_KMAX, VEC6>(i, K, logits, idx32, idx64, Kp, lgr, jr);
  softmax_k<GP_BLEND_KMAX>(lgr, K, wx);
  softmax_k<GP_BLEND_KMAX>(lgr + GP_BLEND_KMAX, K, wr);
  float dx = 0.f, dy = 0.f, dz = 0.f;
  float4 dq = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < GP_BLEND_KMAX; k++) {
    if (k < K) {
      const int j = jr[k];
      if (j >= 0) {
        const float4 a = *(const float4*)(s_kp + 8 * j);
        const float4 b = *(const float4*)(s_kp + 8 * j + 4);
        dx += wx[k] * a.x; dy += wx[k] * a.y; dz += wx[k] * a.z;
        dq.x += wr[k] * b.x; dq.y += wr[k] * b.y; dq.z += wr[k] * b.z; dq.w += wr[k] * b.w;
      }
    }
  }
  xyz_t[(size_t)i * 3] = xyz[(size_t)i * 3] + dx;
  xyz_t[(size_t)i * 3 + 1] = xyz[(size_t)i * 3 + 1] + dy;
  xyz_t[(size_t)i * 3 + 2] = xyz[(size_t)i * 3 + 2] + dz;
  // F.normalize: v / max(|v|, 1e-12)
  const float n1 = fmaxf(qnorm(dq), 1e-12f);
  const float4 dqn = make_float4(dq.x / n1, dq.y / n1, dq.z / n1, dq.w / n1);
  const float4 r = *(const float4*)(rot + (size_t)i * 4);
  const float4 m = quat_mul_ref(dqn, r);
  const float n2 = fmaxf(qnorm(m), 1e-12f);
  *(float4*)(q_t + (size_t)i * 4) = make_float4(m.x / n2, m.y / n2, m.z / n2, m.w / n2);
  if (s_t) {
    s_t[(size_t)i * 3] = expf(scaling[(size_t)i * 3]);
    s_t[(size_t)i * 3 + 1] = expf(scaling[(size_t)i * 3 + 1]);
    s_t[(size_t)i * 3 + 2] = expf(scaling[(size_t)i * 3 + 2]);
  }
  if (o_t) o_t[i] = 1.0f / (1.0f + expf(-opacity[i]));
  }
}

// d/dv of normalize(v) = v / max(|v|, eps), applied to an upstream gradient g
__device__ __forceinline__ float4 normalize_bwd(const float4 v, const float4 g) {
  const float n = qnorm(v);
  if (n < 1e-12f) return make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
  const float inv = 1.0f / n;
  const float4 u = make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
  const float d = u.x * g.x + u.y * g.y + u.z * g.z + u.w * g.w;
  return make_float4((g.x - u.x * d) * inv, (g.y - u.y * d) * inv, (g.z - u.z * d) * inv, (g.w - u.w * d) * inv);
}

// Backward of gp_blend_fwd_kernel.  Keypoint gradients (Kp x 7) are accumulated in shared memory per
// block and flushed with one atomicAdd per touched entry; g_kp_* must be zeroed by the caller.
template <bool VEC6>
__global__ void __launch_bounds__(GP_BLEND_THREADS, VEC6 ? 4 : 2)
gp_blend_bwd_kernel(int N, int Kp, int Kr, const float* __restrict__ logits, const int* __restrict__ idx32,
                    const long long* __restrict__ idx64, const float* __restrict__ kp_dxyz,
                    const float* __restrict__ kp_dq, const float* __restrict__ rot, const float* __restrict__ s_t,
                    const float* __restrict__ o_t, const float* __restrict__ g_xyz_t, const float* __restrict__ g_q_t,
                    const float* __restrict__ g_s_t, const float* __restrict__ g_o_t, float* __restrict__ g_logits,
                    float* __restrict__ g_xyz, float* __restrict__ g_rot, float* __restrict__ g_scaling,
                    float* __restrict__ g_opacity, float* __restrict__ g_kp_dxyz, float* __restrict__ g_kp_dq,
                    const bool acc) {
  // acc: the per-Gaussian outputs are ADDED to what the arrays hold (a running sum over the samples of a gradient-
  // accumulation step: no scratch copy and no separate add pass); the keypoint outputs are always plain sums
  extern __shared__ __align__(16) float smem[];
  float* s_kp = smem;            // [Kp][8]
  float* s_g = smem + Kp * 8;    // [Kp][9] gradient accumulators (7 used; odd stride spreads the banks)
  stage_keypoints(s_kp, Kp, kp_dxyz, kp_dq);
  for (int e = threadIdx.x; e < Kp * 9; e += blockDim.x) s_g[e] = 0.f;
  __syncthreads();
  const int K = VEC6 ? 6 : Kr;
  for (int i = blockIdx.x * GP_BLEND_THREADS + threadIdx.x; i < N; i += gridDim.x * GP_BLEND_THREADS) {
    float wx[GP_BLEND_KMAX], wr[GP_BLEND_KMAX];
    float lgr[2 * GP_BLEND_KMAX];
    int jj[GP_BLEND_KMAX];
    load_row<GP_BLEND_KMAX, VEC6>(i, K, logits, idx32, idx64, Kp, lgr, jj);
    softmax_k<GP_BLEND_KMAX>(lgr, K, wx);
    softmax_k<GP_BLEND_KMAX>(lgr + GP_BLEND_KMAX, K, wr);
    float4 dq = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < GP_BLEND_KMAX; k++) {
      if (k < K) {
        const int j = jj[k];
        if (j >= 0) {
          const float4 b = *(const float4*)(s_kp + 8 * j + 4);
          dq.x += wr[k] * b.x; dq.y += wr[k] * b.y; dq.z += wr[k] * b.z; dq.w += wr[k] * b.w;
        }
      }
    }
    // recompute the forward chain
    const float n1 = fmaxf(qnorm(dq), 1e-12f);
    const float4 dqn = make_float4(dq.x / n1, dq.y / n1, dq.z / n1, dq.w / n1);
    const float4 r = *(const float4*)(rot + (size_t)i * 4);
    const float4 m = quat_mul_ref(dqn, r);
    // backward: q_t = normalize(m)
    const float4 gq = g_q_t ? *(const float4*)(g_q_t + (size_t)i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 gm = normalize_bwd(m, gq);
    // m = quat_mul(q1 = dqn, q2 = r): gradients w.r.t. both operands
    const float4 q1 = dqn, q2 = r;
    float4 g1, g2;
    g1.x = gm.x * q2.x + gm.y * q2.y + gm.z * q2.z + gm.w * q2.w;
    g1.y = -gm.x * q2.y + gm.y * q2.x - gm.z * q2.w + gm.w * q2.z;
    g1.z = -gm.x * q2.z + gm.y * q2.w + gm.z * q2.x - gm.w * q2.y;
    g1.w = -gm.x * q2.w - gm.y * q2.z + gm.z * q2.y + gm.w * q2.x;
    g2.x = gm.x * q1.x + gm.y * q1.y + gm.z * q1.z + gm.w * q1.w;
    g2.y = -gm.x * q1.y + gm.y * q1.x + gm.z * q1.w - gm.w * q1.z;
    g2.z = -gm.x * q1.z - gm.y * q1.w + gm.z * q1.x + gm.w * q1.y;
    g2.w = -gm.x * q1.w + gm.y * q1.z - gm.z * q1.y + gm.w * q1.x;
    if (g_rot) {
      float4* o = (float4*)(g_rot + (size_t)i * 4);
      if (acc) {
        const float4 old = *o;
        g2.x += old.x; g2.y += old.y; g2.z += old.z; g2.w += old.w;
      }
      *o = g2;
    }
    const float4 gdq = normalize_bwd(dq, g1);   // w.r.t. the blended (un-normalised) delta quaternion
    const float gx = g_xyz_t ? g_xyz_t[(size_t)i * 3] : 0.f, gy = g_xyz_t ? g_xyz_t[(size_t)i * 3 + 1] : 0.f,
                gz = g_xyz_t ? g_xyz_t[(size_t)i * 3 + 2] : 0.f;
    if (g_xyz) {
      float* o = g_xyz + (size_t)i * 3;
      o[0] = acc ? o[0] + gx : gx; o[1] = acc ? o[1] + gy : gy; o[2] = acc ? o[2] + gz : gz;
    }
    // weights and keypoints
    float gwx[GP_BLEND_KMAX], gwr[GP_BLEND_KMAX];
    float sx = 0.f, sr = 0.f;
#pragma unroll
    for (int k = 0; k < GP_BLEND_KMAX; k++) {
      if (k < K) {
        gwx[k] = 0.f; gwr[k] = 0.f;
        const int j = jj[k];
        if (j >= 0) {
          const float4 a = *(const float4*)(s_kp + 8 * j);
          const float4 b = *(const float4*)(s_kp + 8 * j + 4);
          gwx[k] = gx * a.x + gy * a.y + gz * a.z;
          gwr[k] = gdq.x * b.x + gdq.y * b.y + gdq.z * b.z + gdq.w * b.w;
          float* ga = s_g + 9 * j;
          atomicAdd(ga + 0, wx[k] * gx); atomicAdd(ga + 1, wx[k] * gy); atomicAdd(ga + 2, wx[k] * gz);
          atomicAdd(ga + 3, wr[k] * gdq.x); atomicAdd(ga + 4, wr[k] * gdq.y);
          atomicAdd(ga + 5, wr[k] * gdq.z); atomicAdd(ga + 6, wr[k] * gdq.w);
        }
        sx += wx[k] * gwx[k];
        sr += wr[k] * gwr[k];
      }
    }
    if (g_logits && VEC6) {
      float4* o4 = reinterpret_cast<float4*>(g_logits + (size_t)i * 12);
      float4 a = make_float4(wx[0] * (gwx[0] - sx), wx[1] * (gwx[1] - sx), wx[2] * (gwx[2] - sx), wx[3] * (gwx[3] - sx));
      float4 b = make_float4(wx[4] * (gwx[4] - sx), wx[5] * (gwx[5] - sx), wr[0] * (gwr[0] - sr), wr[1] * (gwr[1] - sr));
      float4 c = make_float4(wr[2] * (gwr[2] - sr), wr[3] * (gwr[3] - sr), wr[4] * (gwr[4] - sr), wr[5] * (gwr[5] - sr));
      if (acc) {
        const float4 oa = o4[0], ob = o4[1], oc = o4[2];
        a.x += oa.x; a.y += oa.y; a.z += oa.z; a.w += oa.w;
        b.x += ob.x; b.y += ob.y; b.z += ob.z; b.w += ob.w;
        c.x += oc.x; c.y += oc.y; c.z += oc.z; c.w += oc.w;
      }
      o4[0] = a; o4[1] = b; o4[2] = c;
    } else if (g_logits) {
#pragma unroll
      for (int k = 0; k < GP_BLEND_KMAX; k++) {
        if (k < K) {
          float* o = g_logits + (size_t)i * 2 * K;
          o[k] = (acc ? o[k] : 0.f) + wx[k] * (gwx[k] - sx);
          o[K + k] = (acc ? o[K + k] : 0.f) + wr[k] * (gwr[k] - sr);
        }
      }
    }
    if (g_scaling) {
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const float v = g_s_t ? g_s_t[(size_t)i * 3 + c] * s_t[(size_t)i * 3 + c] : 0.f;
        g_scaling[(size_t)i * 3 + c] = acc ? g_scaling[(size_t)i * 3 + c] + v : v;
      }
    }
    if (g_opacity) {
      const float o = o_t[i];
      const float v = g_o_t ? g_o_t[i] * o * (1.0f - o) : 0.f;
      g_opacity[i] = acc ? g_opacity[i] + v : v;
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < Kp * 7; e += blockDim.x) {
    const int j = e / 7, c = e - j * 7;
    const float v = s_g[9 * j + c];
    if (v != 0.f) {
      if (c < 3) atomicAdd(g_kp_dxyz + (size_t)j * 3 + c, v);
      else atomicAdd(g_kp_dq + (size_t)j * 4 + (c - 3), v);
    }
  }
}

// Dense [N,Kp] weight matrices for API parity (return_weights=True, eval.py:125-126; weights_sum, :263).
__global__ void __launch_bounds__(GP_BLEND_THREADS)
gp_dense_weights_kernel(int N, int Kp, int K, const float* __restrict__ logits, const int* __restrict__ idx32,
                        const long long* __restrict__ idx64, float* __restrict__ w_xyz, float* __restrict__ w_r,
                        float* __restrict__ w_sum) {
  const int i = blockIdx.x * GP_BLEND_THREADS + threadIdx.x;
  if (i >= N) return;
  float wx[GP_BLEND_KMAX], wr[GP_BLEND_KMAX];
  const float* lg = logits + (size_t)i * 2 * K;
  softmax_k<GP_BLEND_KMAX>(lg, K, wx);
  softmax_k<GP_BLEND_KMAX>(lg + K, K, wr);
#pragma unroll
  for (int k = 0; k < GP_BLEND_KMAX; k++) {
    if (k < K) {
      const int j = idx32 ? idx32[(size_t)i * K + k] : (int)idx64[(size_t)i * K + k];
      if (j >= 0 && j < Kp) {
        if (w_xyz) w_xyz[(size_t)i * Kp + j] = wx[k];
        if (w_r) w_r[(size_t)i * Kp + j] = wr[k];
        if (w_sum) w_sum[(size_t)i * Kp + j] = fabsf(wx[k]) + fabsf(wr[k]);
      }
    }
  }
}

// persistent grid: `per_sm` blocks per SM (what registers and shared memory allow), each staging the keypoint table once
// and flushing its keypoint gradients once; the kernels are latency bound, so resident warps are what hides the loads
static int gp_blend_grid(int N, int per_sm) {
  const int need = (N + GP_BLEND_THREADS - 1) / GP_BLEND_THREADS;
  return need < 148 * per_sm ? need : 148 * per_sm;
}

static int blend_check(int N, int Kp, int K) {
  if (N < 0 || Kp < 0 || K < 1) return GP_ERR_BAD_ARG;
  if (K > GP_BLEND_KMAX) return GP_ERR_UNSUPPORTED;
  if ((size_t)Kp * 17 * 4 > 200 * 1024) return GP_ERR_UNSUPPORTED;  // keypoint table must fit in shared memory
  return GP_OK;
}

extern "C" int gp_blend_forward(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                                const float* kp_delta_xyz, const float* kp_delta_q, const float* xyz, const float* rot,
                                const float* scaling, const float* opacity, float* xyz_t, float* q_t, float* s_t,
                                float* o_t, gp_stream_t stream_) {
  int rc = blend_check(N, Kp, K);
  if (rc) return rc;
  if (N == 0) return GP_OK;
  if (!logits || (!idx32 && !idx64) || !xyz || !rot || !xyz_t || !q_t) return GP_ERR_BAD_ARG;
  if ((((size_t)rot) | ((size_t)q_t)) & 15) return GP_ERR_BAD_ARG;   // quaternion rows move as 128-bit words
  if (Kp > 0 && (!kp_delta_xyz || !kp_delta_q)) return GP_ERR_BAD_ARG;
  if ((s_t && !scaling) || (o_t && !opacity)) return GP_ERR_BAD_ARG;
  const size_t smem = (size_t)Kp * 8 * 4;
  const bool vec6 = K == 6 && idx32 && ((((size_t)logits) & 15) == 0) && ((((size_t)idx32) & 7) == 0);
  auto kernel = vec6 ? gp_blend_fwd_kernel<true> : gp_blend_fwd_kernel<false>;
  if (smem > 48 * 1024)
    GP_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  GP_KERNEL("gp_blend_fwd_kernel", (cudaStream_t)stream_, kernel<<<gp_blend_grid(N, vec6 ? 4 : 2), GP_BLEND_THREADS, smem, (cudaStream_t)stream_>>>(
      N, Kp, K, logits, idx32, idx64, kp_delta_xyz, kp_delta_q, xyz, rot, scaling, opacity, xyz_t, q_t, s_t, o_t));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

static int blend_backward_impl(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                               const float* kp_delta_xyz, const float* kp_delta_q, const float* rot, const float* s_t,
                               const float* o_t, const float* g_xyz_t, const float* g_q_t, const float* g_s_t,
                               const float* g_o_t, float* g_logits, float* g_xyz, float* g_rot, float* g_scaling,
                               float* g_opacity, float* g_kp_delta_xyz, float* g_kp_delta_q, bool acc,
                               gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = blend_check(N, Kp, K);
  if (rc) return rc;
  if (!g_kp_delta_xyz || !g_kp_delta_q) return GP_ERR_BAD_ARG;
  GP_CUDA_TRY(cudaMemsetAsync(g_kp_delta_xyz, 0, (size_t)Kp * 3 * 4, stream));
  GP_CUDA_TRY(cudaMemsetAsync(g_kp_delta_q, 0, (size_t)Kp * 4 * 4, stream));
  if (N == 0) return GP_OK;
  if (!logits || (!idx32 && !idx64) || !rot || !kp_delta_xyz || !kp_delta_q) return GP_ERR_BAD_ARG;
  if ((((size_t)rot) | ((size_t)g_q_t) | ((size_t)g_rot)) & 15) return GP_ERR_BAD_ARG;   // 128-bit quaternion rows (NULL is fine)
  if ((g_scaling && g_s_t && !s_t) || (g_opacity && !o_t)) return GP_ERR_BAD_ARG;
  const size_t smem = (size_t)Kp * 17 * 4;
  const bool vec6 = K == 6 && idx32 && ((((size_t)logits) & 15) == 0) && ((((size_t)idx32) & 7) == 0) &&
                    ((((size_t)g_logits) & 15) == 0);
  auto kernel = vec6 ? gp_blend_bwd_kernel<true> : gp_blend_bwd_kernel<false>;
  if (smem > 48 * 1024)
    GP_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  GP_KERNEL("gp_blend_bwd_kernel", stream, kernel<<<gp_blend_grid(N, vec6 ? 4 : 2), GP_BLEND_THREADS, smem, stream>>>(
      N, Kp, K, logits, idx32, idx64, kp_delta_xyz, kp_delta_q, rot, s_t, o_t, g_xyz_t, g_q_t, g_s_t, g_o_t, g_logits,
      g_xyz, g_rot, g_scaling, g_opacity, g_kp_delta_xyz, g_kp_delta_q, acc));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_blend_backward(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                                 const float* kp_delta_xyz, const float* kp_delta_q, const float* rot, const float* s_t,
                                 const float* o_t, const float* g_xyz_t, const float* g_q_t, const float* g_s_t,
                                 const float* g_o_t, float* g_logits, float* g_xyz, float* g_rot, float* g_scaling,
                                 float* g_opacity, float* g_kp_delta_xyz, float* g_kp_delta_q, gp_stream_t stream) {
  return blend_backward_impl(N, Kp, K, logits, idx32, idx64, kp_delta_xyz, kp_delta_q, rot, s_t, o_t, g_xyz_t, g_q_t,
                             g_s_t, g_o_t, g_logits, g_xyz, g_rot, g_scaling, g_opacity, g_kp_delta_xyz, g_kp_delta_q,
                             false, stream);
}

extern "C" int gp_blend_backward_accumulate(int N, int Kp, int K, const float* logits, const int* idx32,
                                            const long long* idx64, const float* kp_delta_xyz, const float* kp_delta_q,
                                            const float* rot, const float* s_t, const float* o_t, const float* g_xyz_t,
                                            const float* g_q_t, const float* g_s_t, const float* g_o_t, float* g_logits,
                                            float* g_xyz, float* g_rot, float* g_scaling, float* g_opacity,
                                            float* g_kp_delta_xyz, float* g_kp_delta_q, gp_stream_t stream) {
  return blend_backward_impl(N, Kp, K, logits, idx32, idx64, kp_delta_xyz, kp_delta_q, rot, s_t, o_t, g_xyz_t, g_q_t,
                             g_s_t, g_o_t, g_logits, g_xyz, g_rot, g_scaling, g_opacity, g_kp_delta_xyz, g_kp_delta_q,
                             true, stream);
}

extern "C" int gp_blend_dense_weights(int N, int Kp, int K, const float* logits, const int* idx32,
                                      const long long* idx64, float* w_xyz, float* w_r, float* w_sum,
                                      gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = blend_check(N, Kp, K);
  if (rc) return rc;
  if (N == 0 || Kp == 0) return GP_OK;
  if (!logits || (!idx32 && !idx64)) return GP_ERR_BAD_ARG;
  const size_t bytes = (size_t)N * Kp * 4;
  if (w_xyz) GP_CUDA_TRY(cudaMemsetAsync(w_xyz, 0, bytes, stream));
  if (w_r) GP_CUDA_TRY(cudaMemsetAsync(w_r, 0, bytes, stream));
  if (w_sum) GP_CUDA_TRY(cudaMemsetAsync(w_sum, 0, bytes, stream));
  GP_KERNEL("gp_dense_weights_kernel", stream, gp_dense_weights_kernel<<<(N + GP_BLEND_THREADS - 1) / GP_BLEND_THREADS, GP_BLEND_THREADS, 0, stream>>>(
      N, Kp, K, logits, idx32, idx64, w_xyz, w_r, w_sum));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
