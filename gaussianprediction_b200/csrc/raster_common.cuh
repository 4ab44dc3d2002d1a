// Pieces shared by the forward and backward blend kernels.
#pragma once
#include "common.cuh"
#include "gmath.cuh"

// Which of the eight 8x4-pixel warp patches of a 16x16 tile can a Gaussian contribute to?
//
// A pair (pixel, Gaussian) is skipped by the blend loops unless power <= 0 and opacity * exp(power) >= 1/255
// (reference forward.cu:345-352, backward.cu:497-504), i.e. unless the pixel lies in the ellipse
// d^T Q d <= 2 tau, tau = ln(255 opacity), around the projected centre (Q = conic).  Its axis-aligned bounding box is
// |dx| <= sqrt(2 tau c / det), |dy| <= sqrt(2 tau a / det).  Bit w of the result is set when that box meets patch
// w = 2 * row + col.  Patches whose bit is clear would be skipped pixel by pixel anyway, so dropping them changes no
// result; it only removes work (the reference's tile list comes from a circle of radius 3 sigma_max and holds many
// such pairs).
//
// The box must contain every pixel the blend loops ACCEPT IN FLOATING POINT, so it is inflated by bounds on their
// rounding (u = 2^-24), not by a guess:
//   * acceptance means fl(o * fl(exp(power_c))) >= 1/255 with expf good to 2 ulp: power_c >= -(tau + 5u');
//     logf(255 o) is off by at most 2u(1 + tau): tau is replaced by tau + 1e-6 (covers both, tau <= 5.6).
//   * power_c = fma(fma(a dx, dx, (c dy) dy), -0.5, -((b dx) dy)) carries <= 4 roundings on terms whose absolute sum is
//     <= a dx^2 + c dy^2 <= 2 rho * |power| with rho = (a c + b^2) / det (largest generalised eigenvalue of diag(a, c)
//     against Q): |power_c - power| <= 16 u rho |power|.
//   * det_c = fl(a c - b^2) is off by <= 3 u rho relative (cancellation of two rounded products).
//   Both relative errors enter the squared half-extents once: factor (1 + 32 u rho) <= 1 + 2e-6 rho, plus 1e-5 for
//   the divisions, square roots and the products below.  rho is itself computed from det_c, so a Gaussian whose
//   rho_c exceeds 1e5 (relative bound 0.2: an extremely elongated, strongly rotated splat) is not culled at all.
//   * centre +- half-extent minus the tile origin: three roundings at magnitude <= 2^13 px -> 0.002 px added.
// The box of one Gaussian (xlo, xhi, ylo, yhi in pixels) -- the part of the mask that does not depend on the tile.
// "No culling" is the box of everything, "never visible" an empty box.
__device__ __forceinline__ float4 gp_patch_extent(const float4 co, const float cx, const float cy) {
  const float4 all = make_float4(-1e30f, 1e30f, -1e30f, 1e30f);
  const float o = co.w;
  if (o < 1.0f / 255.0f) return make_float4(1e30f, -1e30f, 1e30f, -1e30f);  // alpha <= opacity < 1/255 everywhere (NaN opacity is not culled, like fminf)
  const float a = co.x, b = co.y, c = co.z;
  const float det = a * c - b * b;
  if (!(det > 0.f) || !(a > 0.f) || !(c > 0.f)) return all;  // not a proper ellipse: no culling
  const float rho = (a * c + b * b) / det;
  if (!(rho < 1e5f)) return all;
  const float k = 2.0f * (logf(255.0f * o) + 1e-6f) / det * (1.0f + 1e-5f + 2e-6f * rho);
  const float hx = sqrtf(k * c) + 0.002f;
  const float hy = sqrtf(k * a) + 0.002f;
  if (!(hx < 1e30f) || !(hy < 1e30f) || !(fabsf(cx) < 8192.f) || !(fabsf(cy) < 8192.f)) return all;
  return make_float4(cx - hx, cx + hx, cy - hy, cy + hy);
}

// Bit w of the result: does the box meet patch w = 2 * row + col of the tile with origin (tile_x0, tile_y0)?
__device__ __forceinline__ unsigned gp_patch_mask_of(const float4 ext, const float tile_x0, const float tile_y0) {
  const float xlo = ext.x - tile_x0, xhi = ext.y - tile_x0;  // relative to the tile origin
  const float ylo = ext.z - tile_y0, yhi = ext.w - tile_y0;
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

__device__ __forceinline__ unsigned gp_patch_mask(const float4 co, const float cx, const float cy, const float tile_x0,
                                                  const float tile_y0) {
  return gp_patch_mask_of(gp_patch_extent(co, cx, cy), tile_x0, tile_y0);
}

// ------------------------------------------------------------------------------------------------
// Explicit shared-space accesses (32-bit shared addresses).  A generic pointer into shared memory makes the compiler
// rebuild the window base (S2R CgaCtaId + LEA) at the uses; the blend kernels keep one opaque base register instead.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 gp_lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ float gp_lds32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t gp_lds32u(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void gp_sts32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void gp_sts32u(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void gp_sts128(uint32_t a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ uint4 gp_lds128u(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}

// shared -> shared copy of one word by the lanes whose predicate holds (predicated, not branched)
__device__ __forceinline__ void gp_copy_word_if(uint32_t src, uint32_t dst, uint32_t p) {
  asm volatile("{\n\t.reg .pred q;\n\t.reg .b32 v;\n\tsetp.ne.u32 q, %2, 0;\n\t@q ld.shared.b32 v, [%0];\n\t@q st.shared.b32 [%1], v;\n\t}"
               ::"r"(src), "r"(dst), "r"(p) : "memory");
}

// Asynchronous global -> shared copies (LDGSTS): the blend kernels stage the records of the NEXT batch of a tile's list
// while the current one is being blended, so the gather latency (DRAM at 2 M Gaussians) is off the critical path.
__device__ __forceinline__ void gp_cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void gp_cp_async8(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void gp_cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void gp_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void gp_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Entries of the sorted instance list.  The patch mask of a (tile, Gaussian) pair depends only on the pair, so it is
// computed ONCE, where the instance is emitted (gp_duplicate_kernel), and travels through the tile sort in the top
// 8 bits of the 32-bit value next to the 24-bit Gaussian index.  Frames with more than 2^24 Gaussians keep plain
// indices and every mask reads as 0xFF (no culling, same results).
#define GP_LIST_ID_BITS 24
__host__ __device__ __forceinline__ bool gp_list_packed(int P) { return P <= (1 << GP_LIST_ID_BITS); }
__device__ __forceinline__ uint32_t gp_list_id(uint32_t v, bool packed) { return packed ? (v & ((1u << GP_LIST_ID_BITS) - 1u)) : v; }
__device__ __forceinline__ uint32_t gp_list_mask(uint32_t v, bool packed) { return packed ? (v >> GP_LIST_ID_BITS) : 0xFFu; }

__device__ __forceinline__ uint32_t gp_smem_base(const void* p) {
  uint32_t sb = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("" : "+r"(sb));   // opaque to the optimiser: stays in a register
  return sb;
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
// dst[0..n) += src_s[0..n) by one warp (same alignment rules as gp_warp_copy_out)
__device__ __forceinline__ void gp_warp_add_out(float* __restrict__ dst, const float* src_s, int n, int lane) {
  const int n4 = n >> 2;
#pragma unroll 4
  for (int i = lane; i < n4; i += 32) {
    float4 v = reinterpret_cast<const float4*>(src_s)[i];
    const float4 o = reinterpret_cast<float4*>(dst)[i];
    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
    reinterpret_cast<float4*>(dst)[i] = v;
  }
  for (int e = (n4 << 2) + lane; e < n; e += 32) dst[e] += src_s[e];
}

__device__ __forceinline__ void gp_warp_copy_out(float* __restrict__ dst, const float* src_s, int n, int lane) {
  const int n4 = n >> 2;
#pragma unroll 4
  for (int i = lane; i < n4; i += 32) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src_s)[i];
  for (int e = (n4 << 2) + lane; e < n; e += 32) dst[e] = src_s[e];
}
