// Pieces shared by the forward and backward blend kernels.
#pragma once
#include "common.cuh"
#include "gmath.cuh"

// Which of the eight 8x4-pixel warp patches of a 16x16 tile can a Gaussian contribute to?
//
// A pair (pixel, Gaussian) is skipped by the blend loops unless power <= 0 and opacity * exp(power) >= 1/255
// (reference forward.cu:345-352, backward.cu:497-504), i.e. unless the pixel lies in the ellipse
// d^T Q d <= 2 ln(255 opacity) around the projected centre (Q = conic).  Its axis-aligned bounding box is
// |dx| <= sqrt(2 tau cov_xx), |dy| <= sqrt(2 tau cov_yy) with cov = Q^-1.  Bit w of the result is set when that
// box (inflated by 0.1 % + 0.01 px against rounding) meets patch w = 2 * row + col.  Patches whose bit is clear
// would be skipped pixel by pixel anyway, so dropping them changes no result; it only removes work.  The
// reference's tile list is built from a circle of radius 3 sigma_max and therefore holds many such pairs.
__device__ __forceinline__ unsigned gp_patch_mask(const float4 co, const float cx, const float cy, const float tile_x0,
                                                  const float tile_y0) {
  const float o = co.w;
  if (o < 1.0f / 255.0f) return 0u;  // alpha <= opacity < 1/255 everywhere (NaN opacity is not culled, like fminf)
  const float a = co.x, b = co.y, c = co.z;
  const float det = a * c - b * b;
  if (!(det > 0.f) || !(a > 0.f) || !(c > 0.f)) return 0xFFu;  // not a proper ellipse: no culling
  const float k = 2.0f * logf(255.0f * o) / det;
  const float hx = sqrtf(k * c) * 1.001f + 0.01f;
  const float hy = sqrtf(k * a) * 1.001f + 0.01f;
  if (!(hx < 1e30f) || !(hy < 1e30f)) return 0xFFu;
  const float xlo = cx - hx - tile_x0, xhi = cx + hx - tile_x0;  // relative to the tile origin
  const float ylo = cy - hy - tile_y0, yhi = cy + hy - tile_y0;
  unsigned colm = 0u, rowm = 0u;
  if (xhi >= 0.f && xlo <= 7.f) colm |= 1u;
  if (xhi >= 8.f && xlo <= 15.f) colm |= 2u;
#pragma unroll
  for (int r = 0; r < 4; r++)
    if (yhi >= 4.f * r && ylo <= 4.f * r + 3.f) rowm |= 1u << r;
  unsigned m = 0u;
#pragma unroll
  for (int r = 0; r < 4; r++)
    if (rowm & (1u << r)) m |= colm << (2 * r);
  return m;
}

// ------------------------------------------------------------------------------------------------
// SH rows staged through shared memory (preprocess forward and backward)
// ------------------------------------------------------------------------------------------------
#define GP_SH_ROW 52  // floats per staged SH row in shared memory: 48 + 4 pad (16-byte aligned rows, conflict-free LDS.128)
#define GP_SH_TILE (32 * GP_SH_ROW)  // floats of shared memory per warp tile

// Split SH storage: the reference keeps coefficient 0 (_features_dc [P,1,3]) and the other 15 (_features_rest
// [P,15,3]) as two parameters and concatenates them every frame (scene/gaussian_model.py:155-158).  In split mode a
// warp's tile holds flat copies of its 32 rows of either array ([32][3] at float 0, [32][45] at float 96): the copies
// are plain 128-bit moves on both sides, and lane r reads its row at stride 45 (odd: conflict-free 32-bit accesses).
// The concatenated tensor and the split of its gradient then never exist in HBM.
#define GP_SH_SPLIT_REST 96

// One Gaussian's coefficients, readable as sh[k] (k a compile-time constant at every use).
template <bool SPLIT>
struct GpShRow {
  const float* a;  // interleaved: the whole row; split: coefficient 0
  const float* b;  // split: coefficients 1..
  __device__ __forceinline__ V3 operator[](int k) const {
    const float* q = (SPLIT && k > 0) ? b + 3 * (k - 1) : a + 3 * k;
    return V3{q[0], q[1], q[2]};
  }
};
template <bool SPLIT>
struct GpShRowOut {
  float* a;
  float* b;
  __device__ __forceinline__ void set(int k, const V3 v) const {
    float* q = (SPLIT && k > 0) ? b + 3 * (k - 1) : a + 3 * k;
    q[0] = v.x; q[1] = v.y; q[2] = v.z;
  }
};

// flat copies of n floats by one warp; both sides 16-byte aligned
__device__ __forceinline__ void gp_warp_copy_in(float* dst_s, const float* __restrict__ src, int n, int lane) {
  const int n4 = n >> 2;
#pragma unroll 4
  for (int i = lane; i < n4; i += 32) reinterpret_cast<float4*>(dst_s)[i] = __ldg(reinterpret_cast<const float4*>(src) + i);
  for (int e = (n4 << 2) + lane; e < n; e += 32) dst_s[e] = __ldg(src + e);
}
__device__ __forceinline__ void gp_warp_copy_out(float* __restrict__ dst, const float* src_s, int n, int lane) {
  const int n4 = n >> 2;
#pragma unroll 4
  for (int i = lane; i < n4; i += 32) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src_s)[i];
  for (int e = (n4 << 2) + lane; e < n; e += 32) dst[e] = src_s[e];
}
