// K nearest keypoints per Gaussian (fixed radius), brute force with the keypoints resident in
// shared memory.
//
// Reference path replaced: get_nearest_mask (scene/gaussian_model.py:110-125) ->
// frnn.frnn_grid_points (submodules/FRNN/frnn/frnn.py:210-347) -> insert_points / prefix_sum /
// counting_sort (twice) -> FindNbrsNDKernelV0 (D = 35, "hybird") or FindNbrsNDKernelV2<3,6> ("3D")
// (frnn/csrc/grid/find_nbrs.cu:244-329, 418-504).  With the radii this repo uses (1e3, 1e8) the
// reference grid is a single cell, i.e. an exhaustive scan with the top-K in global memory; here the
// <= 512 keypoints (<= 72 KB at D = 35) are staged once per block in shared memory, the query point
// and the top-K live in registers, and the five set-up kernels and their host syncs disappear.
//
// Parity: squared distances are accumulated exactly as the reference kernel does
// (diff = ref - query; acc = fma(diff, diff, acc), d = 0..D-1 in order), so the distances are
// bit-identical and on tie-free inputs the indices are too.  Exact ties are broken by the smaller
// keypoint index (the reference's tie order depends on an atomicAdd race).
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_KNN_THREADS 128

template <int KMAX>
struct TopK {
  float d[KMAX];
  int i[KMAX];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int k = 0; k < KMAX; k++) {
      d[k] = __int_as_float(0x7f800000);  // +inf
      i[k] = -1;
    }
  }
  // keep the K smallest (d, index) pairs in ascending order; candidates arrive in ascending index
  // order, so a strict '<' keeps the smaller index first among equal distances.
  __device__ __forceinline__ void insert(float dist, int idx, int K) {
#pragma unroll
    for (int k = 0; k < KMAX; k++)
      if (k == K - 1) {
        d[k] = dist;
        i[k] = idx;
      }
#pragma unroll
    for (int k = KMAX - 1; k > 0; k--) {
      if (k < K && d[k] < d[k - 1]) {
        const float td = d[k];
        d[k] = d[k - 1];
        d[k - 1] = td;
        const int ti = i[k];
        i[k] = i[k - 1];
        i[k - 1] = ti;
      }
    }
  }
  __device__ __forceinline__ float worst(int K) const {
    float wv = d[0];
#pragma unroll
    for (int k = 0; k < KMAX; k++)
      if (k == K - 1) wv = d[k];
    return wv;
  }
};

// Query = [xyz (3) | feat (F) * amplify] assembled on the fly when `q_feat` is given, otherwise the
// D floats of q_points.  Same for the keypoints.  D is a template parameter for the two shapes the
// repo uses (3 and 35); D == 0 selects the generic runtime-D path (query re-read from L1).
template <int D, int KMAX>
__global__ void __launch_bounds__(GP_KNN_THREADS)
gp_knn_kernel(int N, int Kp, int Drt, int K, const float* __restrict__ q_points, const float* __restrict__ q_feat,
              const float* __restrict__ r_points, const float* __restrict__ r_feat, int F, float amplify, float r2,
              int kp_chunk, long long* __restrict__ idx64, int* __restrict__ idx32, float* __restrict__ dists) {
  extern __shared__ __align__(16) float s_ref[];  // [kp_chunk][DP]
  constexpr int DD = D > 0 ? D : 1;
  const int Dn = D > 0 ? D : Drt;
  const int DP = (Dn + 3) & ~3;
  const int q = blockIdx.x * GP_KNN_THREADS + threadIdx.x;
  const bool valid = q < N;
  const int qc = valid ? q : (N - 1);

  float qv[DD];
  if (D > 0) {
#pragma unroll
    for (int d = 0; d < DD; d++) {
      if (q_feat != nullptr)
        qv[d] = d < 3 ? q_points[(size_t)qc * 3 + d] : __fmul_rn(q_feat[(size_t)qc * F + (d - 3)], amplify);
      else
        qv[d] = q_points[(size_t)qc * Dn + d];
    }
  }
  TopK<KMAX> top;
  top.init();
  float worst = __int_as_float(0x7f800000);
  int filled = 0;

  for (int base = 0; base < Kp; base += kp_chunk) {
    const int cnt = min(kp_chunk, Kp - base);
    __syncthreads();
    for (int e = threadIdx.x; e < cnt * DP; e += GP_KNN_THREADS) {
      const int j = e / DP, d = e - j * DP;
      float v = 0.f;
      if (d < Dn) {
        if (r_feat != nullptr)
          v = d < 3 ? r_points[(size_t)(base + j) * 3 + d] : __fmul_rn(r_feat[(size_t)(base + j) * F + (d - 3)], amplify);
        else
          v = r_points[(size_t)(base + j) * Dn + d];
      }
      s_ref[e] = v;
    }
    __syncthreads();
    for (int j = 0; j < cnt; j++) {
      const float* rp = s_ref + j * DP;
      float acc = 0.f;
      if (D > 0) {
#pragma unroll
        for (int d4 = 0; d4 < (DD + 3) / 4; d4++) {
          const float4 r4 = *(const float4*)(rp + 4 * d4);
          const float rr[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int d = 4 * d4 + u;
            if (d < DD) {
              const float diff = rr[u] - qv[d];
              acc = __fmaf_rn(diff, diff, acc);
            }
          }
        }
      } else {
        for (int d = 0; d < Dn; d++) {
          float qd;
          if (q_feat != nullptr)
            qd = d < 3 ? q_points[(size_t)qc * 3 + d] : __fmul_rn(q_feat[(size_t)qc * F + (d - 3)], amplify);
          else
            qd = q_points[(size_t)qc * Dn + d];
          const float diff = rp[d] - qd;
          acc = __fmaf_rn(diff, diff, acc);
        }
      }
      if (acc <= r2 && (filled < K || acc < worst)) {
        top.insert(acc, base + j, K);
        if (filled < K) filled++;
        worst = filled < K ? __int_as_float(0x7f800000) : top.worst(K);
      }
    }
  }
  if (!valid) return;
#pragma unroll
  for (int k = 0; k < KMAX; k++) {
    if (k < K) {
      const bool has = k < filled;
      if (idx64) idx64[(size_t)q * K + k] = has ? (long long)top.i[k] : -1ll;
      if (idx32) idx32[(size_t)q * K + k] = has ? top.i[k] : -1;
      if (dists) dists[(size_t)q * K + k] = has ? top.d[k] : -1.0f;
    }
  }
}

template <int D, int KMAX>
static int launch_knn(int N, int Kp, int Drt, int K, const float* q_points, const float* q_feat, const float* r_points,
                      const float* r_feat, int F, float amplify, float r2, long long* idx64, int* idx32, float* dists,
                      cudaStream_t stream) {
  const int Dn = D > 0 ? D : Drt;
  const int DP = (Dn + 3) & ~3;
  const size_t max_smem = 200 * 1024;
  int kp_chunk = Kp;
  if ((size_t)kp_chunk * DP * 4 > max_smem) kp_chunk = (int)(max_smem / ((size_t)DP * 4));
  if (kp_chunk < 1) kp_chunk = 1;
  const size_t smem = (size_t)kp_chunk * DP * 4;
  auto kern = gp_knn_kernel<D, KMAX>;
  if (smem > 48 * 1024) GP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  GP_KERNEL("gp_knn_kernel", stream, kern<<<(N + GP_KNN_THREADS - 1) / GP_KNN_THREADS, GP_KNN_THREADS, smem, stream>>>(
      N, Kp, Drt, K, q_points, q_feat, r_points, r_feat, F, amplify, r2, kp_chunk, idx64, idx32, dists));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

static int knn_dispatch(int N, int Kp, int D, int K, const float* q_points, const float* q_feat, const float* r_points,
                        const float* r_feat, int F, float amplify, float r, long long* idx64, int* idx32, float* dists,
                        cudaStream_t stream) {
  if (N < 0 || Kp < 0 || D < 1 || K < 1) return GP_ERR_BAD_ARG;
  if (K > 32) return GP_ERR_UNSUPPORTED;
  if (N == 0) return GP_OK;
  const float r2 = r * r;  // the reference passes r*r computed in float (frnn.py:176)
#define GP_KNN_GO(DV, KM) \
  return launch_knn<DV, KM>(N, Kp, D, K, q_points, q_feat, r_points, r_feat, F, amplify, r2, idx64, idx32, dists, stream)
  if (K <= 8) {
    if (D == 3) GP_KNN_GO(3, 8);
    if (D == 35) GP_KNN_GO(35, 8);
    GP_KNN_GO(0, 8);
  } else {
    if (D == 3) GP_KNN_GO(3, 32);
    if (D == 35) GP_KNN_GO(35, 32);
    GP_KNN_GO(0, 32);
  }
#undef GP_KNN_GO
}

extern "C" int gp_knn_points(int N, int Kp, int D, int K, float r, const float* points1, const float* points2,
                             long long* idxs, float* dists, gp_stream_t stream) {
  if (N > 0 && (!points1 || !idxs)) return GP_ERR_BAD_ARG;
  if (Kp > 0 && !points2) return GP_ERR_BAD_ARG;
  return knn_dispatch(N, Kp, D, K, points1, nullptr, points2, nullptr, 0, 1.0f, r, idxs, nullptr, dists,
                      (cudaStream_t)stream);
}

extern "C" int gp_knn_keypoints(int N, int Kp, int F, int K, float r, float feature_amplify, const float* xyz,
                                const float* motion_feature, const float* kp_xyz, const float* kp_feature,
                                long long* idx64, int* idx32, float* dists, gp_stream_t stream) {
  if (N > 0 && !xyz) return GP_ERR_BAD_ARG;
  if (Kp > 0 && !kp_xyz) return GP_ERR_BAD_ARG;
  if (F < 0 || (F > 0 && (!motion_feature || !kp_feature))) return GP_ERR_BAD_ARG;
  if (!idx64 && !idx32) return GP_ERR_BAD_ARG;
  if (F == 0)
    return knn_dispatch(N, Kp, 3, K, xyz, nullptr, kp_xyz, nullptr, 0, 1.0f, r, idx64, idx32, dists, (cudaStream_t)stream);
  return knn_dispatch(N, Kp, 3 + F, K, xyz, motion_feature, kp_xyz, kp_feature, F, feature_amplify, r, idx64, idx32,
                      dists, (cudaStream_t)stream);
}
