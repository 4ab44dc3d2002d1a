// Small fixed-size vector/matrix helpers for the per-Gaussian kernels.
//
// The reference evaluates its 3x3 products through a column-major matrix library; to get the same
// FP32 roundings (radii, tile rectangles and sort keys must be bit-exact) the
// products below keep the same summation order: out[col][row] = a[0][row]*b[col][0] +
// a[1][row]*b[col][1] + a[2][row]*b[col][2], left to right, and nvcc's default FMA contraction is
// left on.  Storage is column-major: c[col][row].
#pragma once
#include "common.cuh"

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(const V3& a, const V3& b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(const V3& a, const V3& b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(float s, const V3& a) { return V3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator*(const V3& a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator*(const V3& a, const V3& b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ V3 operator/(const V3& a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ V3& operator+=(V3& a, const V3& b) {
  a.x += b.x;
  a.y += b.y;
  a.z += b.z;
  return a;
}
__device__ __forceinline__ float dot3(const V3& a, const V3& b) {
  V3 t = a * b;
  return t.x + t.y + t.z;
}
__device__ __forceinline__ float len3(const V3& a) { return sqrtf(dot3(a, a)); }

struct M3 {
  float c[3][3];  // c[col][row]
};
// Fill in column order (first three arguments are column 0).
__device__ __forceinline__ M3 m3(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1,
                                 float c2) {
  M3 m;
  m.c[0][0] = a0; m.c[0][1] = a1; m.c[0][2] = a2;
  m.c[1][0] = b0; m.c[1][1] = b1; m.c[1][2] = b2;
  m.c[2][0] = c0; m.c[2][1] = c1; m.c[2][2] = c2;
  return m;
}
__device__ __forceinline__ M3 m3mul(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
  return r;
}
__device__ __forceinline__ M3 m3t(const M3& a) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) r.c[j][i] = a.c[i][j];
  return r;
}
__device__ __forceinline__ M3 m3scale(float s, const M3& a) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) r.c[j][i] = a.c[j][i] * s;
  return r;
}

// Row-vector * 4x4 (matrices arrive transposed, reference auxiliary.h:58-77).
__device__ __forceinline__ V3 xform4x3(const V3& p, const float* m) {
  return V3{m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform4x4(const V3& p, const float* m) {
  return float4{m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
__device__ __forceinline__ V3 xformvec4x3T(const V3& p, const float* m) {
  return V3{m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
            m[8] * p.x + m[9] * p.y + m[10] * p.z};
}

// ndc -> pixel in double, as the reference does (auxiliary.h:41-44 has unsuffixed literals).
__device__ __forceinline__ float ndc2pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle covered by a disc (reference auxiliary.h:46-56).
__device__ __forceinline__ void tile_rect(float px, float py, int max_radius, unsigned gx, unsigned gy, uint2& rmin,
                                          uint2& rmax) {
  rmin.x = min(gx, (unsigned)max(0, (int)((px - max_radius) / GP_TILE_X)));
  rmin.y = min(gy, (unsigned)max(0, (int)((py - max_radius) / GP_TILE_Y)));
  rmax.x = min(gx, (unsigned)max(0, (int)((px + max_radius + GP_TILE_X - 1) / GP_TILE_X)));
  rmax.y = min(gy, (unsigned)max(0, (int)((py + max_radius + GP_TILE_Y - 1) / GP_TILE_Y)));
}

__device__ const float GP_SH_C0 = 0.28209479177387814f;
__device__ const float GP_SH_C1 = 0.4886025119029199f;
__device__ const float GP_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                        -1.0925484305920792f, 0.5462742152960396f};
__device__ const float GP_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                        0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                        -0.5900435899266435f};
