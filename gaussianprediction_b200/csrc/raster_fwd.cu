// Rasterizer forward for sm_100a: preprocess -> depth sort -> tile-count scan -> duplicate ->
// tile sort -> tile ranges -> alpha / median-depth blending.
//
// Reference path replaced: CudaRasterizer::Rasterizer::forward
// (submodules/diff-gaussian-rasterization-w-depth/cuda_rasterizer/rasterizer_impl.cu:198-342),
// kernels forward.cu:155-256 (preprocessCUDA), rasterizer_impl.cu:70-138 (duplicateWithKeys,
// identifyTileRanges), forward.cu:261-398 (renderCUDA).
//
// Differences in structure (results are identical, see tests/test_raster_parity.py):
//  * The reference sorts R 64-bit (tile|depth) keys (6 radix passes over R).  Here the P Gaussians
//    are sorted once by depth bits (ties by index, as a stable sort of the reference's emission
//    order gives), instances are emitted in that order, and only the tile id (<= 2 passes) is sorted
//    over R.  A stable sort by tile of a depth-ordered list is the (tile, depth, index) order.
//  * No host round trip: R lives in the workspace header, every R-sized kernel has a fixed grid.
//  * The blend kernel stages colour and depth in shared memory with the conic, so the inner loop
//    touches no global memory.
#include "gmath.cuh"
#include "raster_common.cuh"
#include "raster_state.cuh"

// ------------------------------------------------------------------------------------------------
// preprocess
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ void cov3d_from_scale_rot(const V3 scale, float mod, const float4 rot, float* cov3D) {
  // rot is used as given: this fork does NOT normalise the quaternion (reference forward.cu:127).
  M3 S = m3(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
  S.c[0][0] = mod * scale.x;
  S.c[1][1] = mod * scale.y;
  S.c[2][2] = mod * scale.z;
  const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
  M3 R = m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
            2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
            2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  M3 M = m3mul(S, R);
  M3 Sigma = m3mul(m3t(M), M);
  cov3D[0] = Sigma.c[0][0];
  cov3D[1] = Sigma.c[0][1];
  cov3D[2] = Sigma.c[0][2];
  cov3D[3] = Sigma.c[1][1];
  cov3D[4] = Sigma.c[1][2];
  cov3D[5] = Sigma.c[2][2];
}

__device__ __forceinline__ float3 cov2d_from_cov3d(const V3& mean, float focal_x, float focal_y, float tan_fovx,
                                                   float tan_fovy, const float* cov3D, const float* view) {
  V3 t = xform4x3(mean, view);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  M3 J = m3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
            -(focal_y * t.y) / (t.z * t.z), 0, 0, 0);
  M3 Wm = m3(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  M3 T = m3mul(Wm, J);
  M3 Vrk = m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
  M3 cov = m3mul(m3mul(m3t(T), m3t(Vrk)), T);
  cov.c[0][0] += 0.3f;  // low-pass: at least one pixel wide
  cov.c[1][1] += 0.3f;
  return float3{cov.c[0][0], cov.c[0][1], cov.c[1][1]};
}

template <bool SPLIT>
__device__ __forceinline__ V3 sh_to_rgb(int deg, const V3& pos, const V3& campos, const GpShRow<SPLIT> sh,
                                        unsigned& clamp_bits) {
  V3 dir = pos - campos;
  dir = dir / len3(dir);
  V3 result = GP_SH_C0 * sh[0];
  if (deg > 0) {
    const float x = dir.x, y = dir.y, z = dir.z;
    result = result - GP_SH_C1 * y * sh[1] + GP_SH_C1 * z * sh[2] - GP_SH_C1 * x * sh[3];
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
      result = result + GP_SH_C2[0] * xy * sh[4] + GP_SH_C2[1] * yz * sh[5] +
               GP_SH_C2[2] * (2.0f * zz - xx - yy) * sh[6] + GP_SH_C2[3] * xz * sh[7] +
               GP_SH_C2[4] * (xx - yy) * sh[8];
      if (deg > 2) {
        result = result + GP_SH_C3[0] * y * (3.0f * xx - yy) * sh[9] + GP_SH_C3[1] * xy * z * sh[10] +
                 GP_SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
                 GP_SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                 GP_SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13] + GP_SH_C3[5] * z * (xx - yy) * sh[14] +
                 GP_SH_C3[6] * x * (xx - 3.0f * yy) * sh[15];
      }
    }
  }
  result.x += 0.5f;
  result.y += 0.5f;
  result.z += 0.5f;
  clamp_bits = (result.x < 0 ? 1u : 0u) | (result.y < 0 ? 2u : 0u) | (result.z < 0 ? 4u : 0u);
  return V3{fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
}

// Per-Gaussian part of the preprocess; sh_row holds this Gaussian's 3*M SH coefficients (global memory or the
// warp's staging tile in shared memory).
template <bool SPLIT>
__device__ __forceinline__ void
gp_preprocess_body(int idx, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
                   float scale_modifier, const float* __restrict__ rotations, const float* __restrict__ opacities,
                   const GpShRow<SPLIT> sh_row, const float* __restrict__ cov3D_precomp,
                   const float* __restrict__ colors_precomp, const float* __restrict__ viewmatrix,
                   const float* __restrict__ projmatrix, const float* __restrict__ cam_pos, int W, int H,
                   float tan_fovx, float tan_fovy, float focal_x, float focal_y, int prefiltered, unsigned grid_x,
                   unsigned grid_y, GpGeom g, int* __restrict__ radii) {
  int radius_out = 0;
  unsigned tiles_out = 0;
  unsigned key_out = 0xFFFFFFFFu;

  const V3 p_orig = v3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
  // near culling exactly as the reference's in_frustum (auxiliary.h:139-163)
  const V3 p_view = xform4x3(p_orig, viewmatrix);
  bool alive = true;
  if (p_view.z <= 0.2f) {
    if (prefiltered) __trap();
    alive = false;
  }
  if (alive) {
    const float4 p_hom = xform4x4(p_orig, projmatrix);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const V3 p_proj = v3(p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w);

    float cov3D_local[6];
    const float* cov3D;
    if (cov3D_precomp != nullptr) {
      cov3D = cov3D_precomp + (size_t)idx * 6;
    } else {
      const V3 sc = v3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
      const float4 rot = *(const float4*)(rotations + 4 * (size_t)idx);
      cov3d_from_scale_rot(sc, scale_modifier, rot, cov3D_local);
#pragma unroll
      for (int k = 0; k < 6; k++) g.cov3D[(size_t)idx * 6 + k] = cov3D_local[k];
      cov3D = cov3D_local;
    }
    const float3 cov = cov2d_from_cov3d(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, viewmatrix);
    const float det = (cov.x * cov.z - cov.y * cov.y);
    if (det != 0.0f) {
      const float det_inv = 1.f / det;
      const float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
      const float mid = 0.5f * (cov.x + cov.z);
      const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
      const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
      const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
      const float2 point_image = {ndc2pix(p_proj.x, W), ndc2pix(p_proj.y, H)};
      uint2 rect_min, rect_max;
      tile_rect(point_image.x, point_image.y, (int)my_radius, grid_x, grid_y, rect_min, rect_max);
      const unsigned ntiles = (rect_max.x - rect_min.x) * (rect_max.y - rect_min.y);
      if (ntiles != 0) {
        if (colors_precomp == nullptr) {
          unsigned cb;
          const V3 c = sh_to_rgb<SPLIT>(D, p_orig, v3(cam_pos[0], cam_pos[1], cam_pos[2]), sh_row, cb);
          g.rgb[idx] = make_float4(c.x, c.y, c.z, 0.f);
          g.clamped[idx] = (uint8_t)cb;
        }
        g.xy_depth[idx] = make_float4(point_image.x, point_image.y, p_view.z, 0.f);
        g.conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z, opacities[idx]);
        radius_out = (int)my_radius;
        tiles_out = ntiles;
        key_out = __float_as_uint(p_view.z);
      }
    }
  }
  radii[idx] = radius_out;
  g.tiles_touched[idx] = tiles_out;
  g.dkeys[0][idx] = key_out;
  g.dvals[0][idx] = (uint32_t)idx;
}

// Thread per Gaussian.  The 192-byte SH rows of a warp's 32 consecutive Gaussians (6 KB, the bulk of the 311 B per
// Gaussian this kernel moves) are fetched with fully coalesced 128-bit loads into a shared-memory tile (rows padded
// to 52 floats) and read from there, instead of one 32-byte sector per lane per load.
template <bool SPLIT>
__global__ void __launch_bounds__(256)
gp_preprocess_kernel(int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
                     float scale_modifier, const float* __restrict__ rotations, const float* __restrict__ opacities,
                     const float* __restrict__ shs, const float* __restrict__ shs_rest,
                     const float* __restrict__ cov3D_precomp, const float* __restrict__ colors_precomp,
                     const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                     const float* __restrict__ cam_pos, int W, int H, float tan_fovx, float tan_fovy, float focal_x,
                     float focal_y, int prefiltered, unsigned grid_x, unsigned grid_y, GpGeom g,
                     int* __restrict__ radii, int staged) {
  extern __shared__ __align__(16) float s_tiles[];  // staged: [8 warps][GP_SH_TILE]
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (blockIdx.x == 0) gp_rs_reset_by_block(g.radix);  // the depth sort that follows needs zeroed histograms
  if (idx == 0) {
    g.hdr[GP_HDR_NUM_POINTS] = P;
    g.hdr[GP_HDR_NUM_RENDERED] = 0;
    g.hdr[GP_HDR_OVERFLOW] = 0;
    g.hdr[GP_HDR_NUM_SORTED] = 0;
    g.hdr[GP_HDR_SCAN_TICKET] = 0;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int warp_base = idx - lane;
  GpShRow<SPLIT> sh_row{shs ? shs + (size_t)idx * M * 3 : nullptr, nullptr};
  if (staged && warp_base < P) {
    float* tin = s_tiles + (size_t)w * GP_SH_TILE;
    const int rows = min(32, P - warp_base);
    if (SPLIT) {  // coefficient 0 and coefficients 1..15 come from two arrays (raster_common.cuh)
      gp_warp_copy_in(tin, shs + (size_t)warp_base * 3, rows * 3, lane);
      gp_warp_copy_in(tin + GP_SH_SPLIT_REST, shs_rest + (size_t)warp_base * 45, rows * 45, lane);
      sh_row = GpShRow<SPLIT>{tin + lane * 3, tin + GP_SH_SPLIT_REST + lane * 45};
    } else {
      const float4* src = (const float4*)(shs + (size_t)warp_base * 48);
#pragma unroll
      for (int it = 0; it < 12; it++) {
        const int i4 = it * 32 + lane;
        const int row = i4 / 12, c4 = i4 - row * 12;
        if (row < rows) *(float4*)(tin + row * GP_SH_ROW + 4 * c4) = __ldg(src + i4);
      }
      sh_row = GpShRow<SPLIT>{tin + lane * GP_SH_ROW, nullptr};
    }
    __syncwarp();
  }
  if (idx < P)
    gp_preprocess_body<SPLIT>(idx, D, M, means3D, scales, scale_modifier, rotations, opacities, sh_row, cov3D_precomp,
                              colors_precomp, viewmatrix, projmatrix, cam_pos, W, H, tan_fovx, tan_fovy, focal_x,
                              focal_y, prefiltered, grid_x, grid_y, g, radii);
}

// ------------------------------------------------------------------------------------------------
// scan of tiles touched in depth order, and instance emission
// ------------------------------------------------------------------------------------------------

// exclusive scan of block_sums[nb] in place by ONE block of NT threads, total -> header.
// 8 values per thread and round: the loads of a round are independent (one L2 round trip for 2048 values instead of
// one per 256 -- at 2 M Gaussians the serial version cost ~35 us, most of gp_block_sum_kernel); the values go through
// shared memory so that both the global accesses and the per-thread runs of 8 are conflict-free (stride 9).
template <int NT>
__device__ __forceinline__ void gp_spine_scan(int nb, uint32_t* __restrict__ block_sums, int* __restrict__ hdr,
                                              long long capacity) {
  constexpr int IT = 8;
  __shared__ uint32_t s_v[NT * IT + NT];
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry_s;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  if (t == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += NT * IT) {
    uint32_t r[IT];
#pragma unroll
    for (int k = 0; k < IT; k++) {
      const int i = base + k * NT + t;
      r[k] = i < nb ? __ldcg(block_sums + i) : 0u;
    }
#pragma unroll
    for (int k = 0; k < IT; k++) {
      const int e = k * NT + t;
      s_v[e + e / IT] = r[k];
    }
    __syncthreads();
    unsigned long long v = 0ull;
#pragma unroll
    for (int k = 0; k < IT; k++) v += s_v[t * (IT + 1) + k];
    unsigned long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    unsigned long long wbase = 0;
    for (int j = 0; j < w; j++) wbase += warp_tot[j];
    const unsigned long long carry = carry_s;
    unsigned long long run = carry + wbase + incl - v;
#pragma unroll
    for (int k = 0; k < IT; k++) {
      const uint32_t x = s_v[t * (IT + 1) + k];
      s_v[t * (IT + 1) + k] = (uint32_t)(run > 0xFFFFFFFFull ? 0xFFFFFFFFull : run);
      run += x;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IT; k++) {
      const int i = base + k * NT + t;
      const int e = k * NT + t;
      if (i < nb) block_sums[i] = s_v[e + e / IT];
    }
    if (t == NT - 1) carry_s = carry + wbase + incl;
    __syncthreads();
  }
  if (t == 0) {
    const unsigned long long total = carry_s;
    const bool over = total > (unsigned long long)capacity || total > 0x7FFFFFFFull;
    hdr[GP_HDR_NUM_RENDERED] = (int)(total > 0x7FFFFFFFull ? 0x7FFFFFFFull : total);
    hdr[GP_HDR_OVERFLOW] = over ? 1 : 0;
    hdr[GP_HDR_NUM_SORTED] = over ? 0 : (int)total;
  }
}

// block b: sum of tiles_touched over sorted positions [b*256, b*256+256).  The block that finishes last (a counter in
// the header, zeroed by the preprocess kernel) then scans the block sums: one launch instead of two.
__global__ void __launch_bounds__(GP_SCAN_BLOCK, 8)
gp_block_sum_kernel(int P, const uint32_t* __restrict__ sorted_idx, const uint32_t* __restrict__ tiles_touched,
                    uint32_t* __restrict__ block_sums, int* __restrict__ hdr, GpRadixTemp tile_radix) {
  __shared__ uint32_t ws[GP_SCAN_BLOCK / 32];
  __shared__ bool s_last;
  const int s = blockIdx.x * GP_SCAN_BLOCK + threadIdx.x;
  uint32_t v = s < P ? tiles_touched[sorted_idx[s]] : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int j = 0; j < GP_SCAN_BLOCK / 32; j++) t += ws[j];
    block_sums[blockIdx.x] = t;
    __threadfence();                                           // the sum is visible before the ticket is
    s_last = atomicAdd(hdr + GP_HDR_SCAN_TICKET, 1) == (int)gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  gp_rs_reset_by_block(tile_radix);                            // the tile sort of gp_raster_render starts from zeroed counters
  __threadfence();                                             // and every other block's sum before it is read here
  gp_spine_scan<GP_SCAN_BLOCK>((int)gridDim.x, block_sums, hdr, 0x7FFFFFFFll);
}

// Emits (tile id, Gaussian index) for every tile of every visible Gaussian, in depth order.
// One block per 256 sorted Gaussians; the block's instances are produced cooperatively (one
// instance per thread per step) so a huge splat does not serialise on one thread as in the
// reference's per-Gaussian loop (rasterizer_impl.cu:98-109).
__global__ void __launch_bounds__(GP_SCAN_BLOCK)
gp_duplicate_kernel(int P, const uint32_t* __restrict__ sorted_idx, const uint32_t* __restrict__ tiles_touched,
                    const float4* __restrict__ xy_depth, const float4* __restrict__ conic_opacity,
                    const int* __restrict__ radii, const uint32_t* __restrict__ block_excl, int* hdr, unsigned grid_x,
                    unsigned grid_y, uint32_t* __restrict__ tkeys, uint32_t* __restrict__ tvals, GpRadixTemp tile_radix,
                    int passes, long long capacity, uint2* __restrict__ ranges, int tiles) {
  __shared__ uint32_t s_hist[GP_RS_MAX_PASSES][GP_RS_RADIX];  // digit counts of the tile ids this block emits
  __shared__ uint32_t s_off[GP_SCAN_BLOCK + 1];  // exclusive offsets inside the block
  __shared__ uint32_t s_g[GP_SCAN_BLOCK];
  __shared__ uint32_t s_rx[GP_SCAN_BLOCK], s_ry[GP_SCAN_BLOCK], s_rw[GP_SCAN_BLOCK];
  __shared__ float4 s_ext[GP_SCAN_BLOCK];  // 1/255-alpha boxes of the block's Gaussians: the patch mask of every emitted
                                           // instance is computed here, once (raster_common.cuh)
  __shared__ uint32_t ws[GP_SCAN_BLOCK / 32];
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const bool packed = gp_list_packed(P);
  const int s = blockIdx.x * GP_SCAN_BLOCK + t;
  const uint32_t gidx_early = s < P ? sorted_idx[s] : 0u;   // in flight together with the header word below
  // does the frame fit the binning buffers?  Every block decides for itself from R (no block reads the flag another
  // one writes); block 0 publishes the decision for the kernels that follow.  Tiles without instances keep range (0,0).
  const long long R = hdr[GP_HDR_NUM_RENDERED];
  // (one plain pointer for the header: block 0 writes two of its words while every block reads a third)
  // R >= 2^30 is refused: the radix sort's look-back words keep a 30-bit count next to 2 flag bits (radix_sort.cuh)
  const bool over = R > capacity || R >= GP_MAX_INSTANCES;
  if (blockIdx.x == 0 && t == 0) {
    hdr[GP_HDR_OVERFLOW] = over ? 1 : 0;
    hdr[GP_HDR_NUM_SORTED] = over ? 0 : (int)R;
  }
  for (int i = blockIdx.x * GP_SCAN_BLOCK + t; i < tiles; i += gridDim.x * GP_SCAN_BLOCK) ranges[i] = make_uint2(0u, 0u);
  if (over) return;
  // the tile sort that follows: its status words are cleared and its digit histograms built here, where the keys
  // are produced, instead of by a separate pass over the R keys (the counters were zeroed by gp_block_sum_kernel)
  gp_rs_clear_status(tile_radix, (int)R, passes);
  for (int e = t; e < GP_RS_MAX_PASSES * GP_RS_RADIX; e += GP_SCAN_BLOCK) (&s_hist[0][0])[e] = 0u;
  // One dependent round trip after the sorted index, not three: everything that is known per Gaussian is loaded at once
  // (the ~2 % culled Gaussians load a record nobody uses).
  uint32_t gidx = 0, tt = 0;
  int rad = 0;
  float4 xyd = make_float4(0.f, 0.f, 0.f, 0.f), co = make_float4(0.f, 0.f, 0.f, 0.f);
  if (s < P) {
    gidx = gidx_early;
    tt = tiles_touched[gidx];
    xyd = xy_depth[gidx];
    rad = radii[gidx];
    if (packed) co = conic_opacity[gidx];
  }
  uint2 rmin = {0, 0}, rmax = {0, 0};
  if (tt) {
    tile_rect(xyd.x, xyd.y, rad, grid_x, grid_y, rmin, rmax);
    if (packed) s_ext[t] = gp_patch_extent(co, xyd.x, xyd.y);
  }
  s_g[t] = gidx;
  s_rx[t] = rmin.x;
  s_ry[t] = rmin.y;
  s_rw[t] = rmax.x - rmin.x;
  uint32_t incl = tt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  if (lane == 31) ws[w] = incl;
  __syncthreads();
  uint32_t wbase = 0;
  for (int j = 0; j < w; j++) wbase += ws[j];
  s_off[t] = wbase + incl - tt;
  if (t == GP_SCAN_BLOCK - 1) s_off[GP_SCAN_BLOCK] = wbase + incl;
  __syncthreads();
  const uint32_t total = s_off[GP_SCAN_BLOCK];
  const uint32_t base = block_excl[blockIdx.x];
  for (uint32_t e = t; e < total; e += GP_SCAN_BLOCK) {
    // largest j with s_off[j] <= e
    int lo = 0, hi = GP_SCAN_BLOCK;
#pragma unroll
    for (int it = 0; it < 8; it++) {
      const int mid = (lo + hi) >> 1;
      if (s_off[mid] <= e) lo = mid; else hi = mid;
    }
    const uint32_t local = e - s_off[lo];
    const uint32_t rw = s_rw[lo];
    const uint32_t ty = s_ry[lo] + local / rw;
    const uint32_t tx = s_rx[lo] + local % rw;
    const uint32_t key = ty * grid_x + tx;
    tkeys[base + e] = key;
    uint32_t val = s_g[lo];
    if (packed) val |= gp_patch_mask_of(s_ext[lo], (float)(tx * GP_TILE_X), (float)(ty * GP_TILE_Y)) << GP_LIST_ID_BITS;
    tvals[base + e] = val;
    for (int p = 0; p < passes; p++) atomicAdd(&s_hist[p][(key >> (8 * p)) & 0xFFu], 1u);
  }
  __syncthreads();
  gp_rs_flush_histogram(tile_radix, s_hist, passes);
}

__global__ void __launch_bounds__(256)
gp_tile_ranges_kernel(const int* __restrict__ hdr, const uint32_t* __restrict__ tkeys, uint2* __restrict__ ranges,
                      GpRadixTemp tile_radix) {
  // the tile sort is over: leave its counters zeroed, so that rendering the same preprocessed frame again (e.g. with a
  // larger capacity after an overflow) starts clean without another preprocess
  if (blockIdx.x == 0) gp_rs_reset_by_block(tile_radix);
  const int L = hdr[GP_HDR_NUM_SORTED];
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < L; idx += gridDim.x * blockDim.x) {
    const uint32_t cur = tkeys[idx];
    if (idx == 0) {
      ranges[cur].x = 0;
    } else {
      const uint32_t prev = tkeys[idx - 1];
      if (cur != prev) {
        ranges[prev].y = idx;
        ranges[cur].x = idx;
      }
    }
    if (idx == L - 1) ranges[cur].y = L;
  }
}

// ------------------------------------------------------------------------------------------------
// blending
// ------------------------------------------------------------------------------------------------

// One 16x16 tile per block; warp w owns an 8x4 pixel patch (lane = 8*row + col) so that the 32
// pixels of a warp are spatially compact.  Batches of 256 list entries are staged in shared memory:
// conic+opacity, centre+depth, colour and index (52 B each) plus one word of patch bits per (warp, 32-entry chunk), in
// TWO buffers: while batch i is
// blended, the records of batch i + 1 are already on their way (cp.async; the list entries themselves are read
// one batch earlier still), so a round costs one barrier and no exposed gather latency -- at 2 M Gaussians a
// warp finds only a few Gaussians of a batch on its patch, and the gather was ~40 % of the kernel's stall samples.
// The per-pair arithmetic is written with explicit round-to-nearest intrinsics in the exact sequence the
// reference build executes (forward.cu:341-379 as contracted by nvcc), because median depth / tidx / n_contrib
// are discrete picks that flip on a 1-ulp change.
#define GP_FW_CO 0
#define GP_FW_XYD (GP_TILE_PIX * 16)
#define GP_FW_RGB (2 * GP_TILE_PIX * 16)
#define GP_FW_ID (3 * GP_TILE_PIX * 16)
#define GP_FW_BITS (GP_FW_ID + GP_TILE_PIX * 4)   // [warp][chunk of 32 entries]: which entries of the chunk meet the warp's patch
#define GP_FW_BUF (GP_FW_BITS + 8 * 8 * 4)

__global__ void __launch_bounds__(GP_TILE_PIX)
gp_render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                     unsigned grid_x, const float4* __restrict__ xy_depth, const float4* __restrict__ conic_opacity,
                     const float4* __restrict__ rgb4, const float* __restrict__ colors_precomp,
                     const float* __restrict__ bg_color, const int* __restrict__ hdr, bool packed,
                     float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                     float* __restrict__ out_depth, int* __restrict__ out_tidx) {
  // one shared block addressed through an opaque 32-bit base (raster_common.cuh): 2 x [co | xyd | rgb | id | patch bits]
  __shared__ __align__(16) unsigned char s_raw[2 * GP_FW_BUF];
  const uint32_t sb = gp_smem_base(s_raw);

  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  const unsigned tile = blockIdx.x;
  const unsigned tile_x = tile % grid_x, tile_y = tile / grid_x;
  const unsigned px = tile_x * GP_TILE_X + (w & 1) * 8 + (lane & 7);
  const unsigned py = tile_y * GP_TILE_Y + (w >> 1) * 4 + (lane >> 3);
  const bool inside = px < (unsigned)W && py < (unsigned)H;
  const unsigned pix_id = (unsigned)W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  // A frame that did not fit the caller's capacity is not rendered, but its outputs are DEFINED: the empty image
  // (background, depth sentinel, no Gaussian), so a caller that misses the overflow flag never reads stale memory.
  // Tile ranges are all (0,0) in that case (gp_duplicate_kernel), so the loop below simply does not run.
  const uint2 range = hdr[GP_HDR_OVERFLOW] ? make_uint2(0u, 0u) : ranges[tile];
  const unsigned n = range.y - range.x;
  const int rounds = (int)((n + GP_TILE_PIX - 1) / GP_TILE_PIX);
  bool done = !inside;

  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f;
  float Dm = 15.0f;  // median-depth sentinel of this fork (forward.cu:310)
  int tid = -1;

  // this thread's entry of a batch: records -> shared memory (asynchronously), index and patch mask beside them
  auto stage = [&](const uint32_t buf, const uint32_t v, const bool valid) {
    uint32_t m = 0u;
    if (valid) {
      const uint32_t id = gp_list_id(v, packed);
      m = gp_list_mask(v, packed);
      gp_cp_async16(buf + GP_FW_CO + t * 16, conic_opacity + id);
      gp_cp_async16(buf + GP_FW_XYD + t * 16, xy_depth + id);
      if (colors_precomp != nullptr) {
        const float* c = colors_precomp + 3 * (size_t)id;
        gp_cp_async4(buf + GP_FW_RGB + t * 16, c);
        gp_cp_async4(buf + GP_FW_RGB + t * 16 + 4, c + 1);
        gp_cp_async4(buf + GP_FW_RGB + t * 16 + 8, c + 2);
      } else {
        gp_cp_async16(buf + GP_FW_RGB + t * 16, rgb4 + id);
      }
      gp_sts32u(buf + GP_FW_ID + t * 4, id);
    }
    // The staging warp w holds chunk w of the batch: it transposes its 32 patch masks into one 32-bit word per
    // consumer warp, so that a consumer reads its eight chunk words with one load and skips empty chunks outright (at
    // 2 M Gaussians a warp finds two or three entries of a 256-entry batch on its patch).
    unsigned mine = 0u;
#pragma unroll
    for (int ww = 0; ww < 8; ww++) {
      const unsigned b = __ballot_sync(0xffffffffu, (m >> ww) & 1u);
      if (lane == ww) mine = b;
    }
    if (lane < 8) gp_sts32u(buf + GP_FW_BITS + (lane * 8 + w) * 4, mine);
    gp_cp_async_commit();
  };

  uint32_t v_next = (unsigned)t < n ? point_list[range.x + t] : 0u;
  stage(sb, v_next, (unsigned)t < n);
  v_next = (unsigned)t + GP_TILE_PIX < n ? point_list[range.x + GP_TILE_PIX + t] : 0u;
  for (int i = 0; i < rounds; i++) {
    gp_cp_async_wait_all();
    // batch i is complete for every thread past this barrier, and nobody reads batch i - 1 any more
    const int num_done = __syncthreads_count(done);
    if (num_done == GP_TILE_PIX) break;
    const uint32_t cur = sb + (uint32_t)(i & 1) * GP_FW_BUF;
    if (i + 1 < rounds) {
      const unsigned k1 = (unsigned)(i + 1) * GP_TILE_PIX + t;
      stage(sb + (uint32_t)((i + 1) & 1) * GP_FW_BUF, v_next, k1 < n);
      v_next = k1 + GP_TILE_PIX < n ? point_list[range.x + k1 + GP_TILE_PIX] : 0u;
    }
    const unsigned chunk_bits = lane < 8 ? gp_lds32u(cur + GP_FW_BITS + (w * 8 + lane) * 4) : 0u;
    unsigned chunks = __ballot_sync(0xffffffffu, chunk_bits != 0u);
    while (chunks) {
      const int c = __ffs(chunks) - 1;
      chunks &= chunks - 1;
      if (__all_sync(0xffffffffu, done)) break;
      const int c0 = c * 32;
      unsigned bits = __shfl_sync(0xffffffffu, chunk_bits, c);
      while (bits) {
        const int j = c0 + __ffs(bits) - 1;
        bits &= bits - 1;
        if (done) continue;
        const uint32_t aj = cur + j * 16;
        const float4 xyd = gp_lds128(aj + GP_FW_XYD);
        const float4 co = gp_lds128(aj + GP_FW_CO);
        const float dx = xyd.x - pixfx;
        const float dy = xyd.y - pixfy;
        const float qa = __fmul_rn(co.x, dx);
        const float qb = __fmul_rn(__fmul_rn(co.z, dy), dy);
        const float qs = __fmaf_rn(qa, dx, qb);
        const float qc = __fmul_rn(__fmul_rn(co.y, dx), dy);
        const float power = __fmaf_rn(qs, -0.5f, -qc);
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, __fmul_rn(co.w, expf(power)));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = __fmul_rn(T, __fadd_rn(1.0f, -alpha));
        if (test_T < 0.0001f) {
          done = true;
          continue;
        }
        const float4 col = gp_lds128(aj + GP_FW_RGB);
        C0 = __fmaf_rn(T, __fmul_rn(alpha, col.x), C0);
        C1 = __fmaf_rn(T, __fmul_rn(alpha, col.y), C1);
        C2 = __fmaf_rn(T, __fmul_rn(alpha, col.z), C2);
        if (T > 0.5f && test_T < 0.5f) {
          Dm = xyd.z;
          tid = (int)gp_lds32u(cur + GP_FW_ID + j * 4);
        }
        T = test_T;
        last_contributor = (uint32_t)(i * GP_TILE_PIX + j + 1);  // position in the tile list, 1-based
      }
    }
  }
  gp_cp_async_wait_all();
  if (inside) {
    final_T[pix_id] = T;
    n_contrib[pix_id] = last_contributor;
    const size_t hw = (size_t)H * W;
    out_color[pix_id] = __fmaf_rn(bg_color[0], T, C0);
    out_color[hw + pix_id] = __fmaf_rn(bg_color[1], T, C1);
    out_color[2 * hw + pix_id] = __fmaf_rn(bg_color[2], T, C2);
    out_depth[pix_id] = Dm;
    out_tidx[pix_id] = tid;
  }
}

__global__ void gp_check_frustum_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                        uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const V3 p = v3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
  const V3 pv = xform4x3(p, view);
  present[idx] = pv.z > 0.2f ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------

extern "C" int gp_raster_sizes(int P, int width, int height, long long capacity, size_t* geom_bytes,
                               size_t* binning_bytes, size_t* image_bytes) {
  if (P < 0 || width < 0 || height < 0 || capacity < 0) return GP_ERR_BAD_ARG;
  if (capacity >= GP_MAX_INSTANCES || P >= GP_MAX_INSTANCES) return GP_ERR_LIMIT;   // 30-bit counts in the sort's look-back words
  if (geom_bytes) *geom_bytes = gp_carve_geom(nullptr, P).bytes;
  if (binning_bytes) *binning_bytes = gp_carve_binning(nullptr, capacity).bytes;
  if (image_bytes) *image_bytes = gp_carve_image(nullptr, width, height).bytes;
  return GP_OK;
}

extern "C" int gp_raster_preprocess(const gp_raster_cfg* cfg, const float* means3D, const float* shs,
                                    const float* colors_precomp, const float* opacities, const float* scales,
                                    const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                    const float* projmatrix, const float* campos, void* geom_ws, int* radii,
                                    gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!cfg || !geom_ws || !radii) return GP_ERR_BAD_ARG;
  const int P = cfg->P;
  if (P <= 0 || cfg->width <= 0 || cfg->height <= 0) return GP_ERR_BAD_ARG;
  if ((shs == nullptr) == (colors_precomp == nullptr)) return GP_ERR_BAD_ARG;
  const bool has_sr = scales != nullptr && rotations != nullptr;
  if (has_sr == (cov3D_precomp != nullptr)) return GP_ERR_BAD_ARG;
  if (shs && (cfg->sh_degree < 0 || cfg->sh_degree > 3 || cfg->sh_coeffs < (cfg->sh_degree + 1) * (cfg->sh_degree + 1)))
    return GP_ERR_BAD_ARG;
  GpGeom g = gp_carve_geom(geom_ws, P);
  const int W = cfg->width, H = cfg->height;
  const float focal_y = H / (2.0f * cfg->tanfovy);
  const float focal_x = W / (2.0f * cfg->tanfovx);
  const unsigned gx = (W + GP_TILE_X - 1) / GP_TILE_X, gy = (H + GP_TILE_Y - 1) / GP_TILE_Y;
  const int staged = (shs != nullptr && cfg->sh_coeffs == 16 && ((((size_t)shs) | ((size_t)cfg->sh_rest)) & 15) == 0) ? 1 : 0;
  if (cfg->sh_rest && !staged) return GP_ERR_UNSUPPORTED;  // split SH storage: 16 coefficients, 16-byte aligned arrays
  const size_t pp_smem = staged ? (size_t)8 * GP_SH_TILE * sizeof(float) : 0;
  auto kernel = cfg->sh_rest ? gp_preprocess_kernel<true> : gp_preprocess_kernel<false>;
  if (pp_smem > 48 * 1024)
    GP_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pp_smem));
  GP_KERNEL("gp_preprocess_kernel", stream, kernel<<<(P + 255) / 256, 256, pp_smem, stream>>>(
      P, cfg->sh_degree, cfg->sh_coeffs, means3D, scales, cfg->scale_modifier, rotations, opacities, shs, cfg->sh_rest,
      cov3D_precomp, colors_precomp, viewmatrix, projmatrix, campos, W, H, cfg->tanfovx, cfg->tanfovy, focal_x, focal_y,
      cfg->prefiltered, gx, gy, g, radii, staged));
  GP_LAUNCH_CHECK();
  // depth sort: P keys, all 32 bits; result lands back in buffer 0 (4 passes)
  gp_radix_sort_pairs(g.dkeys, g.dvals, g.hdr + GP_HDR_NUM_POINTS, 0, 32, g.radix, stream, /*reset=*/false);
  GP_LAUNCH_CHECK();
  const int nb = (P + GP_SCAN_BLOCK - 1) / GP_SCAN_BLOCK;
  // block sums + (in the last block to finish) their scan; publishes R.  The capacity check follows in
  // gp_raster_render (first thing gp_duplicate_kernel does), once the caller has sized the binning workspace.
  GpRadixTemp tile_counters = {g.tile_hist, g.tile_tickets, nullptr, 0};
  GP_KERNEL("gp_block_sum_kernel", stream,
            gp_block_sum_kernel<<<nb, GP_SCAN_BLOCK, 0, stream>>>(P, g.dvals[0], g.tiles_touched, g.block_sums, g.hdr, tile_counters));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_raster_num_rendered(const void* geom_ws, gp_stream_t stream_, int* num_rendered) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!geom_ws || !num_rendered) return GP_ERR_BAD_ARG;
  GP_CUDA_TRY(cudaMemcpyAsync(num_rendered, (const int*)geom_ws + GP_HDR_NUM_RENDERED, sizeof(int),
                              cudaMemcpyDeviceToHost, stream));
  GP_CUDA_TRY(cudaStreamSynchronize(stream));
  return GP_OK;
}

extern "C" int gp_raster_render(const gp_raster_cfg* cfg, const float* bg, const float* colors_precomp, void* geom_ws,
                                void* binning_ws, long long capacity, void* image_ws, const int* radii,
                                float* out_color, float* out_depth, int* out_tidx, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!cfg || !bg || !geom_ws || !binning_ws || !image_ws || !radii || !out_color || !out_depth || !out_tidx)
    return GP_ERR_BAD_ARG;
  const int P = cfg->P, W = cfg->width, H = cfg->height;
  if (P <= 0 || W <= 0 || H <= 0 || capacity < 0) return GP_ERR_BAD_ARG;
  if (capacity >= GP_MAX_INSTANCES || P >= GP_MAX_INSTANCES) return GP_ERR_LIMIT;
  GpGeom g = gp_carve_geom(geom_ws, P);
  GpBinning b = gp_carve_binning(binning_ws, capacity);
  GpImage im = gp_carve_image(image_ws, W, H);
  const unsigned gx = (W + GP_TILE_X - 1) / GP_TILE_X, gy = (H + GP_TILE_Y - 1) / GP_TILE_Y;
  const int nb = (P + GP_SCAN_BLOCK - 1) / GP_SCAN_BLOCK;

  const int bits = gp_tile_bits(W, H);
  b.radix.hist = g.tile_hist;        // counters of the tile sort live in the geometry workspace (zeroed by the
  b.radix.tickets = g.tile_tickets;  // preprocess stage); the status words are the binning workspace's own
  GP_KERNEL("gp_duplicate_kernel", stream, gp_duplicate_kernel<<<nb, GP_SCAN_BLOCK, 0, stream>>>(P, g.dvals[0], g.tiles_touched, g.xy_depth, g.conic_opacity, radii, g.block_sums,
                                                        g.hdr, gx, gy, b.tkeys[0], b.tvals[0], b.radix, (bits + 7) / 8,
                                                        capacity, im.ranges, (int)(gx * gy)));
  GP_LAUNCH_CHECK();
  const int sel = gp_radix_sort_pairs(b.tkeys, b.tvals, g.hdr + GP_HDR_NUM_SORTED, 0, bits, b.radix, stream, /*reset=*/false,
                                      /*prepared=*/true);
  GP_LAUNCH_CHECK();
  GP_KERNEL("gp_tile_ranges_kernel", stream, gp_tile_ranges_kernel<<<148 * 4, 256, 0, stream>>>(g.hdr, b.tkeys[sel], im.ranges, b.radix));
  GP_LAUNCH_CHECK();
  GP_KERNEL("gp_render_fwd_kernel", stream, gp_render_fwd_kernel<<<gx * gy, GP_TILE_PIX, 0, stream>>>(im.ranges, b.tvals[sel], W, H, gx, g.xy_depth,
                                                           g.conic_opacity, g.rgb, colors_precomp, bg, g.hdr,
                                                           gp_list_packed(P), im.final_T, im.n_contrib, out_color, out_depth, out_tidx));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                               uint8_t* present, gp_stream_t stream_) {
  (void)projmatrix;  // the reference computes the projection but only tests view-space z (auxiliary.h:154)
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return GP_ERR_BAD_ARG;
  if (P == 0) return GP_OK;
  GP_KERNEL("gp_check_frustum_kernel", (cudaStream_t)stream_, gp_check_frustum_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, means3D, viewmatrix, present));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------------------------------------
// state export for parity tests
// ------------------------------------------------------------------------------------------------

__global__ void gp_export_geom_kernel(int P, GpGeom g, float* depths,
                                      float* means2D, float* cov3D, float* conic_opacity, float* rgb,
                                      uint32_t* tiles_touched, uint8_t* clamped) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const bool vis = g.tiles_touched[i] != 0;
  const float4 xyd = vis ? g.xy_depth[i] : make_float4(0, 0, 0, 0);
  const float4 co = vis ? g.conic_opacity[i] : make_float4(0, 0, 0, 0);
  const float4 c = vis ? g.rgb[i] : make_float4(0, 0, 0, 0);
  if (depths) depths[i] = xyd.z;
  if (means2D) {
    means2D[2 * i] = xyd.x;
    means2D[2 * i + 1] = xyd.y;
  }
  if (cov3D)
    for (int k = 0; k < 6; k++) cov3D[6 * (size_t)i + k] = g.cov3D[6 * (size_t)i + k];
  if (conic_opacity) {
    conic_opacity[4 * (size_t)i] = co.x;
    conic_opacity[4 * (size_t)i + 1] = co.y;
    conic_opacity[4 * (size_t)i + 2] = co.z;
    conic_opacity[4 * (size_t)i + 3] = co.w;
  }
  if (rgb) {
    rgb[3 * (size_t)i] = c.x;
    rgb[3 * (size_t)i + 1] = c.y;
    rgb[3 * (size_t)i + 2] = c.z;
  }
  if (tiles_touched) tiles_touched[i] = g.tiles_touched[i];
  if (clamped) {
    const unsigned cb = vis ? g.clamped[i] : 0;
    clamped[3 * (size_t)i] = cb & 1;
    clamped[3 * (size_t)i + 1] = (cb >> 1) & 1;
    clamped[3 * (size_t)i + 2] = (cb >> 2) & 1;
  }
}

__global__ void gp_export_binning_kernel(const int* __restrict__ hdr, const uint32_t* __restrict__ tkeys,
                                         const uint32_t* __restrict__ tvals, const float4* __restrict__ xy_depth,
                                         uint32_t* point_list, uint64_t* keys, bool packed) {
  const int L = hdr[GP_HDR_NUM_SORTED];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < L; i += gridDim.x * blockDim.x) {
    const uint32_t gidx = gp_list_id(tvals[i], packed);
    if (point_list) point_list[i] = gidx;
    if (keys) keys[i] = ((uint64_t)tkeys[i] << 32) | (uint64_t)__float_as_uint(xy_depth[gidx].z);
  }
}

extern "C" int gp_raster_export_state(const gp_raster_cfg* cfg, const void* geom_ws, const void* image_ws,
                                      float* depths, float* means2D, float* cov3D, float* conic_opacity, float* rgb,
                                      uint32_t* tiles_touched, uint8_t* clamped, uint32_t* ranges, float* final_T,
                                      uint32_t* n_contrib, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!cfg || !geom_ws) return GP_ERR_BAD_ARG;
  const int P = cfg->P, W = cfg->width, H = cfg->height;
  GpGeom g = gp_carve_geom(const_cast<void*>(geom_ws), P);
  GP_KERNEL("gp_export_geom_kernel", stream, gp_export_geom_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g, depths, means2D, cov3D, conic_opacity, rgb,
                                                            tiles_touched, clamped));
  GP_LAUNCH_CHECK();
  if (image_ws) {
    GpImage im = gp_carve_image(const_cast<void*>(image_ws), W, H);
    const size_t npix = (size_t)W * H;
    const size_t tiles = (size_t)((W + GP_TILE_X - 1) / GP_TILE_X) * ((H + GP_TILE_Y - 1) / GP_TILE_Y);
    if (ranges) GP_CUDA_TRY(cudaMemcpyAsync(ranges, im.ranges, tiles * sizeof(uint2), cudaMemcpyDeviceToDevice, stream));
    if (final_T) GP_CUDA_TRY(cudaMemcpyAsync(final_T, im.final_T, npix * 4, cudaMemcpyDeviceToDevice, stream));
    if (n_contrib) GP_CUDA_TRY(cudaMemcpyAsync(n_contrib, im.n_contrib, npix * 4, cudaMemcpyDeviceToDevice, stream));
  }
  return GP_OK;
}

extern "C" int gp_raster_export_binning(const gp_raster_cfg* cfg, const void* geom_ws, const void* binning_ws,
                                        long long capacity, uint32_t* point_list, uint64_t* point_list_keys,
                                        gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!cfg || !geom_ws || !binning_ws) return GP_ERR_BAD_ARG;
  GpGeom g = gp_carve_geom(const_cast<void*>(geom_ws), cfg->P);
  GpBinning b = gp_carve_binning(const_cast<void*>(binning_ws), capacity);
  const int sel = gp_tile_sort_sel(cfg->width, cfg->height);
  GP_KERNEL("gp_export_binning_kernel", stream, gp_export_binning_kernel<<<148 * 4, 256, 0, stream>>>(g.hdr, b.tkeys[sel], b.tvals[sel], g.xy_depth, point_list,
                                                       point_list_keys, gp_list_packed(cfg->P)));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
