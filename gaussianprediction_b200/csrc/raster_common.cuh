// Pieces shared by the forward and backward blend kernels.
#pragma once
#include "common.cuh"

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
