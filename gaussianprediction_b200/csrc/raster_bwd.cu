// Rasterizer backward for sm_100a.
//
// Reference path replaced: CudaRasterizer::Rasterizer::backward
// (submodules/diff-gaussian-rasterization-w-depth/cuda_rasterizer/rasterizer_impl.cu:346-439):
// renderCUDA backward (backward.cu:399-557), computeCov2DCUDA (:144-274), preprocessCUDA (:346-396).
//
// Structure:
//  * Blend backward: same back-to-front traversal per pixel as the reference (T is rebuilt by the
//    same division), but the reference's 9 global atomicAdd per pixel x Gaussian pair are replaced
//    by a warp-level transposing reduction (14 shuffles per Gaussian per warp), shared-memory
//    accumulation per tile, and three 128-bit vector reductions (red.global.add.v4.f32) per
//    tile x Gaussian instance into one packed 48-byte record per Gaussian.
//  * Per-Gaussian backward: the reference's two kernels are fused; the packed record is read once
//    and every output gradient is written exactly once, including zeros for culled Gaussians, so no
//    output tensor needs a memset (the reference zero-fills 300 B per Gaussian per call,
//    rasterize_points.cu:153-161).
#include <stdlib.h>
#include <atomic>
#include <mutex>
#include "gmath.cuh"
#include "raster_common.cuh"
#include "raster_state.cuh"

// packed per-Gaussian accumulator filled by the blend backward: 12 floats
//   [0] dL/dmean2D.x  [1] dL/dmean2D.y  [2] dL/dconic.x  [3] dL/dconic.y
//   [4] dL/dconic.w   [5] dL/dopacity   [6] dL/dr        [7] dL/dg
//   [8] dL/db         [9..11] unused
#define GP_GREC 12
#define GP_STATIC_MAX_DEVICES_ATTR 64
#define GP_SACC 9

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Sum v[0..7] over the 32 lanes with a transposing butterfly: after the call lane L (L % 4 == 0)
// holds the warp total of component L / 4.  9 shuffles instead of 40.
__device__ __forceinline__ float warp_transpose_reduce8(float (&v)[8], int lane) {
  const bool h16 = lane & 16;
  float w4[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float send = h16 ? v[i] : v[i + 4];
    const float keep = h16 ? v[i + 4] : v[i];
    w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  const bool h8 = lane & 8;
  float w2[2];
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const float send = h8 ? w4[i] : w4[i + 2];
    const float keep = h8 ? w4[i + 2] : w4[i];
    w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  const bool h4 = lane & 4;
  const float send = h4 ? w2[0] : w2[1];
  const float keep = h4 ? w2[1] : w2[0];
  float r = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  r += __shfl_xor_sync(0xffffffffu, r, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;  // component index = ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1)
}

__global__ void __launch_bounds__(GP_TILE_PIX, 4)
gp_render_bwd_butterfly_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                     unsigned grid_x, const float* __restrict__ bg_color, const float4* __restrict__ xy_depth,
                     const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb4,
                     const float* __restrict__ colors_precomp, const float* __restrict__ final_Ts,
                     const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                     const int* __restrict__ hdr, bool packed, float* __restrict__ grec) {
  __shared__ float4 s_co[GP_TILE_PIX];
  __shared__ float4 s_xyd[GP_TILE_PIX];
  __shared__ float4 s_rgb[GP_TILE_PIX];
  __shared__ int s_id[GP_TILE_PIX];
  __shared__ float s_acc[GP_TILE_PIX * GP_SACC];  // [Gaussian in batch][9 sums]; odd stride: conflict-free both ways
  __shared__ unsigned s_mask[GP_TILE_PIX];

  if (hdr[GP_HDR_OVERFLOW]) return;
  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  const unsigned tile = blockIdx.x;
  const unsigned tile_x = tile % grid_x, tile_y = tile / grid_x;
  const unsigned px = tile_x * GP_TILE_X + (w & 1) * 8 + (lane & 7);
  const unsigned py = tile_y * GP_TILE_Y + (w >> 1) * 4 + (lane >> 3);
  const bool inside = px < (unsigned)W && py < (unsigned)H;
  const unsigned pix_id = (unsigned)W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  const uint2 range = ranges[tile];
  const int rounds = ((range.y - range.x + GP_TILE_PIX - 1) / GP_TILE_PIX);
  int toDo = range.y - range.x;

  const float T_final = inside ? final_Ts[pix_id] : 0.f;
  float T = T_final;
  const int total = toDo;
  const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;

  float accum0 = 0.f, accum1 = 0.f, accum2 = 0.f;
  float dpx0 = 0.f, dpx1 = 0.f, dpx2 = 0.f;
  if (inside) {
    const size_t hw = (size_t)H * W;
    dpx0 = dL_dpixels[pix_id];
    dpx1 = dL_dpixels[hw + pix_id];
    dpx2 = dL_dpixels[2 * hw + pix_id];
  }
  float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
  const float ddelx_dx = 0.5f * W;
  const float ddely_dy = 0.5f * H;
  const float bg_dot_dpixel = bg_color[0] * dpx0 + bg_color[1] * dpx1 + bg_color[2] * dpx2;

  // the largest list position any pixel of this warp still has to visit
  unsigned warp_last = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));

  const float tile_fx0 = (float)(tile_x * GP_TILE_X), tile_fy0 = (float)(tile_y * GP_TILE_Y);
  for (int i = 0; i < rounds; i++, toDo -= GP_TILE_PIX) {
    __syncthreads();
    const int progress = i * GP_TILE_PIX + t;
    unsigned pmask = 0u;
    if (range.x + progress < range.y) {
      const int id = (int)gp_list_id(point_list[range.y - progress - 1], packed);
      s_id[t] = id;
      const float4 co = __ldg(conic_opacity + id);
      const float4 xyd = __ldg(xy_depth + id);
      s_co[t] = co;
      s_xyd[t] = xyd;
      if (colors_precomp != nullptr)
        s_rgb[t] = make_float4(colors_precomp[3 * (size_t)id], colors_precomp[3 * (size_t)id + 1],
                               colors_precomp[3 * (size_t)id + 2], 0.f);
      else
        s_rgb[t] = __ldg(rgb4 + id);
      pmask = gp_patch_mask(co, xyd.x, xyd.y, tile_fx0, tile_fy0);
    }
    s_mask[t] = pmask;
#pragma unroll
    for (int k = 0; k < GP_SACC; k++) s_acc[k * GP_TILE_PIX + t] = 0.f;
    __syncthreads();

    const int nb = min(GP_TILE_PIX, toDo);
    // elements of this batch in front of index (total - warp_last) lie behind the last contributor of every pixel
    // of this warp: start at the first 32-chunk that can matter
    const int first = total - (int)warp_last - i * GP_TILE_PIX;
    for (int c0 = max(first, 0) & ~31; c0 < nb; c0 += 32) {
      unsigned bits = __ballot_sync(0xffffffffu, (s_mask[c0 + lane] >> w) & 1u);
      while (bits) {
      const int j = c0 + __ffs(bits) - 1;
      bits &= bits - 1;
      // 0-based position of this Gaussian in the tile's front-to-back list
      const uint32_t contributor = (uint32_t)(total - 1 - (i * GP_TILE_PIX + j));
      if (contributor >= warp_last) continue;  // warp-uniform: nobody in this warp reaches it
      bool active = inside && contributor < last_contributor;
      const float4 xyd = s_xyd[j];
      const float4 co = s_co[j];
      const float dx = xyd.x - pixfx;
      const float dy = xyd.y - pixfy;
      const float qa = __fmul_rn(co.x, dx);
      const float qb = __fmul_rn(__fmul_rn(co.z, dy), dy);
      const float qs = __fmaf_rn(qa, dx, qb);
      const float qc = __fmul_rn(__fmul_rn(co.y, dx), dy);
      const float power = __fmaf_rn(qs, -0.5f, -qc);
      active = active && !(power > 0.0f);
      const float G = expf(power);
      const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
      active = active && !(alpha < 1.0f / 255.0f);
      if (!__any_sync(0xffffffffu, active)) continue;

      float v[8];
      float v_op = 0.f;
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = 0.f;
      if (active) {
        // one correctly rounded reciprocal serves both divisions by (1 - alpha) (reference backward.cu:503,533)
        const float inv1ma = __frcp_rn(1.f - alpha);
        T = T * inv1ma;
        const float dchannel_dcolor = alpha * T;
        const float4 col = s_rgb[j];
        float dL_dalpha = 0.0f;
        accum0 = last_alpha * lc0 + (1.f - last_alpha) * accum0;
        lc0 = col.x;
        dL_dalpha += (col.x - accum0) * dpx0;
        accum1 = last_alpha * lc1 + (1.f - last_alpha) * accum1;
        lc1 = col.y;
        dL_dalpha += (col.y - accum1) * dpx1;
        accum2 = last_alpha * lc2 + (1.f - last_alpha) * accum2;
        lc2 = col.z;
        dL_dalpha += (col.z - accum2) * dpx2;
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final * inv1ma) * bg_dot_dpixel;

        const float dL_dG = co.w * dL_dalpha;
        const float gdx = G * dx;
        const float gdy = G * dy;
        const float dG_ddelx = -gdx * co.x - gdy * co.y;
        const float dG_ddely = -gdy * co.z - gdx * co.y;
        v[0] = dL_dG * dG_ddelx * ddelx_dx;
        v[1] = dL_dG * dG_ddely * ddely_dy;
        v[2] = -0.5f * gdx * dx * dL_dG;
        v[3] = -0.5f * gdx * dy * dL_dG;
        v[4] = -0.5f * gdy * dy * dL_dG;
        v[5] = dchannel_dcolor * dpx0;
        v[6] = dchannel_dcolor * dpx1;
        v[7] = dchannel_dcolor * dpx2;
        v_op = G * dL_dalpha;
      }
      float* acc = s_acc + j * GP_SACC;   // slots: 0..4 = v[0..4], 5 = opacity, 6..8 = colour
      const unsigned amask = __ballot_sync(0xffffffffu, active);
      if (__popc(amask) <= 2) {
        // one or two pixels of this patch see the Gaussian: adding them directly is cheaper than a 32-lane reduction
        if (active) {
          atomicAdd(acc + 0, v[0]); atomicAdd(acc + 1, v[1]); atomicAdd(acc + 2, v[2]);
          atomicAdd(acc + 3, v[3]); atomicAdd(acc + 4, v[4]); atomicAdd(acc + 5, v_op);
          atomicAdd(acc + 6, v[5]); atomicAdd(acc + 7, v[6]); atomicAdd(acc + 8, v[7]);
        }
      } else {
        const float r = warp_transpose_reduce8(v, lane);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v_op += __shfl_xor_sync(0xffffffffu, v_op, o);
        if ((lane & 3) == 0) {
          const int comp = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
          atomicAdd(acc + (comp < 5 ? comp : comp + 1), r);
        }
        if (lane == 1) atomicAdd(acc + 5, v_op);
      }
      }
    }
    __syncthreads();
    if (t < nb) {
      float* dst = grec + (size_t)s_id[t] * GP_GREC;
      const float* ac = s_acc + t * GP_SACC;
      const float a0 = ac[0], a1 = ac[1], a2 = ac[2], a3 = ac[3], a4 = ac[4], a5 = ac[5], a6 = ac[6], a7 = ac[7], a8 = ac[8];
      if (a0 != 0.f || a1 != 0.f || a2 != 0.f || a3 != 0.f) red_add_v4(dst, a0, a1, a2, a3);
      if (a4 != 0.f || a5 != 0.f || a6 != 0.f || a7 != 0.f) red_add_v4(dst + 4, a4, a5, a6, a7);
      if (a8 != 0.f) atomicAdd(dst + 8, a8);
    }
  }
}


// ------------------------------------------------------------------------------------------------
// Blend backward, two-phase ("pixel pass, then splat pass") -- the kernel gp_raster_backward launches.
//
// The butterfly kernel above spends more instructions on summing 9 values over the 32 pixels of a warp than on the
// blend math itself (ncu, profiles/r1_ncu_full.md: 14 SHFL + 16 SEL per visit, float atomicAdd on shared memory =
// CAS spin loops, 31 % bank-conflicted wavefronts).  Here the reduction disappears:
//   phase 1 (lane = pixel, as before): walk the tile's list back to front, rebuild T by division, and for every
//           (patch, Gaussian) visit with at least one contributing pixel store TWO scalars per pixel in a warp-private
//           shared-memory slot:  Wg = G * dL/dalpha  and  U = alpha * T.
//           dL/dalpha uses ONE scalar recurrence instead of the reference's three colour accumulators:
//           dL/dalpha_i = T_i (c_i . dpx) - S_i / (1 - alpha_i),  S_i = sum_{k behind i} (c_k . dpx) alpha_k T_k + T_final (bg . dpx)
//           (the same quantity as backward.cu:506-527, regrouped).
//   phase 2 (every 16 visits; lane = (slot, half)): lane (s, h) owns the Gaussian of slot s and walks the 16 pixels of
//           patch rows 2h, 2h+1, accumulating in registers the six moments of Wg about the Gaussian's centre
//           (1, dx, dy, dx^2, dx dy, dy^2) and sum U dpx_c.  All nine reference gradients are linear in those:
//             dL/dmean2D = -o (a Mx + b My) W/2, -o (c My + b Mx) H/2;  dL/dconic = -o/2 (Mxx, Mxy, Myy);  dL/dopacity = M0.
//           One shuffle per value joins the two halves; three red.global.add.v4.f32 per visit add the result to the
//           Gaussian's packed record.  No shared-memory atomics, no cross-warp accumulator, conflict-free rows (stride 33;
//           interleaving Wg and U as 64-bit pairs was measured slower: the splat pass then spills).
// ------------------------------------------------------------------------------------------------
#define GP_BW_SLOTS 16
#define GP_BW_STRIDE 33
#define GP_BW_BATCH 128
// dynamic shared memory, byte offsets (explicit shared-space addressing: a generic pointer into dynamic shared memory
// costs S2R CgaCtaId + LEA at every access, ~8 % of the round-1 kernel's instructions).
// Two batch buffers [co | x, y, id, - | rgb | patch bits] of 128 entries, the pixels' dL/dpixel, the warps' (Wg, U) slots
// and the warps' slot tables (what phase 2 needs of each slot's Gaussian): 54,528 B, four blocks per SM.
#define GP_BW_B_CO 0
#define GP_BW_B_XYD (GP_BW_BATCH * 16)
#define GP_BW_B_RGB (2 * GP_BW_BATCH * 16)
#define GP_BW_B_BITS (3 * GP_BW_BATCH * 16)    // [warp][chunk of 32 entries]: which entries of the chunk meet the warp's patch
#define GP_BW_BUF (GP_BW_B_BITS + 8 * 4 * 4)
#define GP_BW_DPX (2 * GP_BW_BUF)
#define GP_BW_W (GP_BW_DPX + GP_TILE_PIX * 16)
#define GP_BW_WARP_BYTES (2 * GP_BW_SLOTS * GP_BW_STRIDE * 4)
#define GP_BW_TAB (GP_BW_W + 8 * GP_BW_WARP_BYTES)
#define GP_BW_TAB_BYTES (GP_BW_SLOTS * 32)
#define GP_BW_SMEM (GP_BW_TAB + 8 * GP_BW_TAB_BYTES)

__global__ void __launch_bounds__(GP_TILE_PIX, 4)
gp_render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                     unsigned grid_x, const float* __restrict__ bg_color, const float4* __restrict__ xy_depth,
                     const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb4,
                     const float* __restrict__ colors_precomp, const float* __restrict__ final_Ts,
                     const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                     const int* __restrict__ hdr, bool packed, float* __restrict__ grec) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  __shared__ unsigned s_last[8];
  const uint32_t sb = gp_smem_base(s_raw);

  if (hdr[GP_HDR_OVERFLOW]) return;
  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  const unsigned tile = blockIdx.x;
  const unsigned tile_x = tile % grid_x, tile_y = tile / grid_x;
  const unsigned px0 = tile_x * GP_TILE_X + (w & 1) * 8, py0 = tile_y * GP_TILE_Y + (w >> 1) * 4;  // patch origin
  const unsigned px = px0 + (lane & 7);
  const unsigned py = py0 + (lane >> 3);
  const bool inside = px < (unsigned)W && py < (unsigned)H;
  const unsigned pix_id = (unsigned)W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  const uint2 range = ranges[tile];

  const float T_final = inside ? final_Ts[pix_id] : 0.f;
  float T = T_final;
  const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
  float dpx0 = 0.f, dpx1 = 0.f, dpx2 = 0.f;
  if (inside) {
    const size_t hw = (size_t)H * W;
    dpx0 = dL_dpixels[pix_id];
    dpx1 = dL_dpixels[hw + pix_id];
    dpx2 = dL_dpixels[2 * hw + pix_id];
  }
  gp_sts128(sb + GP_BW_DPX + t * 16, make_float4(dpx0, dpx1, dpx2, 0.f));   // read back by this warp only (phase 2)
  float S = T_final * (bg_color[0] * dpx0 + bg_color[1] * dpx1 + bg_color[2] * dpx2);

  unsigned warp_last = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
  // Entries behind the last contributor of EVERY pixel of the tile are never visited: the walk (and the staging)
  // starts at the tile's last contributor instead of the end of its list.
  if (lane == 0) s_last[w] = warp_last;
  __syncthreads();
  unsigned block_last = 0u;
#pragma unroll
  for (int k = 0; k < 8; k++) block_last = max(block_last, s_last[k]);
  const int total = (int)min(block_last, range.y - range.x);   // list positions 0 .. total-1 matter, walked back to front
  const int rounds = (total + GP_BW_BATCH - 1) / GP_BW_BATCH;
  const uint32_t* list_end = point_list + range.x + total;     // entry k (from the back) is list_end[-1 - k]

  // phase-2 identity of this lane
  const int slot = lane & (GP_BW_SLOTS - 1), half = lane >> 4;
  const uint32_t bufW = sb + GP_BW_W + w * GP_BW_WARP_BYTES;                       // [slot][33] Wg, then [slot][33] U
  const uint32_t rowW = bufW + (slot * GP_BW_STRIDE + half * 16) * 4;
  const uint32_t rowD = sb + GP_BW_DPX + (w * 32 + half * 16) * 16;
  const uint32_t tab = sb + GP_BW_TAB + w * GP_BW_TAB_BYTES;                       // [slot] {conic, opacity | x, y, id, -}
  // lanes 0-6 copy one word each of a visited Gaussian's record into the slot table (no divergent branch)
  const uint32_t tsrc = lane < 4 ? GP_BW_B_CO + 4 * lane : GP_BW_B_XYD + 4 * (lane - 4);
  uint32_t tw = tab + lane * 4;
  const uint32_t tcopy = lane < 7 ? 1u : 0u;
  const float fx0 = (float)px0, fy0 = (float)(py0 + 2 * half);
  const float half_w = 0.5f * W, half_h = 0.5f * H;
  int nslots = 0;                    // warp-uniform
  uint32_t wr = bufW + lane * 4;     // where this lane's (Wg, U) of the next visit go

  // Slots outlive the batch their Gaussian came from (its conic, centre and index are copied into the warp's slot
  // table at the visit), so every splat pass but the last one of a tile runs with all 16 slots filled -- at 2 M
  // Gaussians a warp meets only two or three Gaussians per batch.
  auto splat_pass = [&]() {
    __syncwarp();
    const uint4 tb = gp_lds128u(tab + slot * 32 + 16);
    const float bx = __uint_as_float(tb.x) - fx0, by = __uint_as_float(tb.y) - fy0;
    float m0 = 0.f, mx = 0.f, my = 0.f, mxx = 0.f, mxy = 0.f, myy = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) {
      const float Wg = gp_lds32(rowW + i * 4), U = gp_lds32(rowW + GP_BW_SLOTS * GP_BW_STRIDE * 4 + i * 4);
      const float4 d = gp_lds128(rowD + i * 16);
      const float dx = bx - (float)(i & 7), dy = by - (float)(i >> 3);
      const float wx = Wg * dx, wy = Wg * dy;
      m0 += Wg;
      mx += wx;
      my += wy;
      mxx = fmaf(wx, dx, mxx);
      mxy = fmaf(wx, dy, mxy);
      myy = fmaf(wy, dy, myy);
      c0 = fmaf(U, d.x, c0);
      c1 = fmaf(U, d.y, c1);
      c2 = fmaf(U, d.z, c2);
    }
    m0 += __shfl_xor_sync(0xffffffffu, m0, 16);
    mx += __shfl_xor_sync(0xffffffffu, mx, 16);
    my += __shfl_xor_sync(0xffffffffu, my, 16);
    mxx += __shfl_xor_sync(0xffffffffu, mxx, 16);
    mxy += __shfl_xor_sync(0xffffffffu, mxy, 16);
    myy += __shfl_xor_sync(0xffffffffu, myy, 16);
    c0 += __shfl_xor_sync(0xffffffffu, c0, 16);
    c1 += __shfl_xor_sync(0xffffffffu, c1, 16);
    c2 += __shfl_xor_sync(0xffffffffu, c2, 16);
    if (half == 0 && slot < nslots) {
      const float4 co = gp_lds128(tab + slot * 32);
      float* dst = grec + (size_t)tb.z * GP_GREC;
      const float no = -co.w;
      red_add_v4(dst, no * (co.x * mx + co.y * my) * half_w, no * (co.z * my + co.y * mx) * half_h, 0.5f * no * mxx,
                 0.5f * no * mxy);
      red_add_v4(dst + 4, 0.5f * no * myy, m0, c0, c1);
      atomicAdd(dst + 8, c2);
    }
    __syncwarp();
    nslots = 0;
    wr = bufW + lane * 4;
    tw = tab + lane * 4;
  };

  // Staging, double-buffered: the 128 entries of a batch are copied by the 256 threads (threads 0-127: conic+opacity,
  // centre, index, mask of entry t; threads 128-255: the colour of entry t-128) with cp.async, one batch AHEAD of the
  // blend; the list entries themselves are read one batch earlier still.  One barrier per batch.  The depth is not
  // needed here, so the index takes its place next to the centre.
  const int e = t & (GP_BW_BATCH - 1);
  const bool first_half = t < GP_BW_BATCH;
  auto stage = [&](const uint32_t buf, const uint32_t v, const bool valid) {
    if (valid) {
      const uint32_t id = gp_list_id(v, packed);
      if (first_half) {
        gp_cp_async16(buf + GP_BW_B_CO + e * 16, conic_opacity + id);
        gp_cp_async8(buf + GP_BW_B_XYD + e * 16, xy_depth + id);
        gp_sts32u(buf + GP_BW_B_XYD + e * 16 + 8, id);
      } else if (colors_precomp != nullptr) {
        const float* c = colors_precomp + 3 * (size_t)id;
        gp_cp_async4(buf + GP_BW_B_RGB + e * 16, c);
        gp_cp_async4(buf + GP_BW_B_RGB + e * 16 + 4, c + 1);
        gp_cp_async4(buf + GP_BW_B_RGB + e * 16 + 8, c + 2);
      } else {
        gp_cp_async16(buf + GP_BW_B_RGB + e * 16, rgb4 + id);
      }
    }
    if (first_half) {   // warp w (< 4) holds chunk w: one word of patch bits per consumer warp (see the forward kernel)
      const unsigned m = valid ? gp_list_mask(v, packed) : 0u;
      unsigned mine = 0u;
#pragma unroll
      for (int ww = 0; ww < 8; ww++) {
        const unsigned b = __ballot_sync(0xffffffffu, (m >> ww) & 1u);
        if (lane == ww) mine = b;
      }
      if (lane < 8) gp_sts32u(buf + GP_BW_B_BITS + (lane * 4 + w) * 4, mine);
    }
    gp_cp_async_commit();
  };

  uint32_t v_next = e < total ? list_end[-1 - e] : 0u;
  stage(sb, v_next, e < total);
  v_next = e + GP_BW_BATCH < total ? list_end[-1 - e - GP_BW_BATCH] : 0u;
  for (int i = 0; i < rounds; i++) {
    gp_cp_async_wait_all();
    __syncthreads();   // batch i has landed for every thread, and nobody reads batch i - 1 any more
    const uint32_t cur = sb + (uint32_t)(i & 1) * GP_BW_BUF;
    if (i + 1 < rounds) {
      const int k1 = (i + 1) * GP_BW_BATCH + e;
      stage(sb + (uint32_t)((i + 1) & 1) * GP_BW_BUF, v_next, k1 < total);
      v_next = k1 + GP_BW_BATCH < total ? list_end[-1 - k1 - GP_BW_BATCH] : 0u;
    }

    // list position (0-based, front to back) of batch element j is base - j
    const int base = total - 1 - i * GP_BW_BATCH;
    // elements of this batch in front of index (total - warp_last) lie behind the last contributor of every pixel
    // of this warp: start at the first 32-chunk that can matter
    const int first = total - (int)warp_last - i * GP_BW_BATCH;
    const int start = max(first, 0) >> 5;
    const unsigned chunk_bits = lane < 4 ? gp_lds32u(cur + GP_BW_B_BITS + (w * 4 + lane) * 4) : 0u;
    unsigned chunks = __ballot_sync(0xffffffffu, chunk_bits != 0u) & (start < 4 ? (0xFu << start) : 0u);
    while (chunks) {
      const int c0 = (__ffs(chunks) - 1) * 32;
      const unsigned chunk_word = __shfl_sync(0xffffffffu, chunk_bits, c0 >> 5);
      chunks &= chunks - 1;
      unsigned bits = chunk_word;
      while (bits) {
        const int j = c0 + __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t contributor = (uint32_t)(base - j);
        if (contributor >= warp_last) continue;  // warp-uniform: nobody in this warp reaches it
        bool active = inside && contributor < last_contributor;
        const uint32_t aj = cur + j * 16;
        const float4 xyd = gp_lds128(aj + GP_BW_B_XYD);
        const float4 co = gp_lds128(aj + GP_BW_B_CO);
        const float dx = xyd.x - pixfx;
        const float dy = xyd.y - pixfy;
        const float qa = __fmul_rn(co.x, dx);
        const float qb = __fmul_rn(__fmul_rn(co.z, dy), dy);
        const float qs = __fmaf_rn(qa, dx, qb);
        const float qc = __fmul_rn(__fmul_rn(co.y, dx), dy);
        const float power = __fmaf_rn(qs, -0.5f, -qc);
        active = active && !(power > 0.0f);
        // exp() through ex2.approx (2 ulp, 6 instructions fewer than expf): the gradients' bar is 1e-3 relative; a
        // pair within 3e-7 relative of the alpha >= 1/255 threshold may be classified differently from the forward
        // pass, which changes that pixel's chain by <= 0.4 % -- far below the bar (tests/test_raster_parity.py).
        float G;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(power * 1.4426950408889634f));
        const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
        active = active && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, active)) continue;

        // Branch-free: every lane computes, inactive lanes keep their state and store zeros.  One reciprocal serves
        // both divisions by (1 - alpha) (reference backward.cu:503,533); 1 - alpha >= 0.01, so the approximate
        // reciprocal (2 ulp) needs no special cases.
        float inv1ma;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv1ma) : "f"(1.f - alpha));
        const float Tn = T * inv1ma;
        const float4 col = gp_lds128(aj + GP_BW_B_RGB);
        const float cd = col.x * dpx0 + col.y * dpx1 + col.z * dpx2;
        const float dL_dalpha = Tn * cd - S * inv1ma;
        const float Un = alpha * Tn;
        T = active ? Tn : T;
        S = active ? fmaf(cd, Un, S) : S;
        gp_sts32(wr, active ? G * dL_dalpha : 0.f);
        gp_sts32(wr + GP_BW_SLOTS * GP_BW_STRIDE * 4, active ? Un : 0.f);
        wr += GP_BW_STRIDE * 4;
        // what phase 2 needs of this Gaussian (conic, opacity, centre, index) goes into the slot table
        gp_copy_word_if(aj + tsrc, tw, tcopy);
        tw += 32;
        if (++nslots == GP_BW_SLOTS) splat_pass();
      }
    }
  }
  if (nslots) splat_pass();
}

// ------------------------------------------------------------------------------------------------
// per-Gaussian backward (fused cov2D + projection + SH + cov3D)
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ V3 dnormvdv3(const V3& v, const V3& dv) {
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  V3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

// SH backward (reference backward.cu:20-139): writes dL/dsh for the active degree and returns the
// view-direction part of dL/dmean.
template <bool SPLIT>
__device__ __forceinline__ V3 sh_backward(int deg, int M, const V3& pos, const V3& campos, const GpShRow<SPLIT> sh,
                                          unsigned clamp_bits, V3 dL_dRGB, const GpShRowOut<SPLIT> dL_dsh) {
  const V3 dir_orig = pos - campos;
  const V3 dir = dir_orig / len3(dir_orig);
  dL_dRGB.x *= (clamp_bits & 1u) ? 0.f : 1.f;
  dL_dRGB.y *= (clamp_bits & 2u) ? 0.f : 1.f;
  dL_dRGB.z *= (clamp_bits & 4u) ? 0.f : 1.f;
  V3 dRGBdx = v3(0, 0, 0), dRGBdy = v3(0, 0, 0), dRGBdz = v3(0, 0, 0);
  const float x = dir.x, y = dir.y, z = dir.z;
  const float xx = x * x, yy = y * y, zz = z * z;
  const float xy = x * y, yz = y * z, xz = x * z;

  // Every read of the coefficients comes before the first write of a gradient: the staged kernel keeps both in ONE
  // shared-memory row per Gaussian (the gradient row overwrites the coefficient row), which halves its shared memory.
  if (deg > 0) {
    dRGBdx = -GP_SH_C1 * sh[3];
    dRGBdy = -GP_SH_C1 * sh[1];
    dRGBdz = GP_SH_C1 * sh[2];
    if (deg > 1) {
      dRGBdx += GP_SH_C2[0] * y * sh[4] + GP_SH_C2[2] * 2.f * -x * sh[6] + GP_SH_C2[3] * z * sh[7] +
                GP_SH_C2[4] * 2.f * x * sh[8];
      dRGBdy += GP_SH_C2[0] * x * sh[4] + GP_SH_C2[1] * z * sh[5] + GP_SH_C2[2] * 2.f * -y * sh[6] +
                GP_SH_C2[4] * 2.f * -y * sh[8];
      dRGBdz += GP_SH_C2[1] * y * sh[5] + GP_SH_C2[2] * 2.f * 2.f * z * sh[6] + GP_SH_C2[3] * x * sh[7];
      if (deg > 2) {
        dRGBdx += (GP_SH_C3[0] * sh[9] * 3.f * 2.f * xy + GP_SH_C3[1] * sh[10] * yz + GP_SH_C3[2] * sh[11] * -2.f * xy +
                   GP_SH_C3[3] * sh[12] * -3.f * 2.f * xz + GP_SH_C3[4] * sh[13] * (-3.f * xx + 4.f * zz - yy) +
                   GP_SH_C3[5] * sh[14] * 2.f * xz + GP_SH_C3[6] * sh[15] * 3.f * (xx - yy));
        dRGBdy += (GP_SH_C3[0] * sh[9] * 3.f * (xx - yy) + GP_SH_C3[1] * sh[10] * xz +
                   GP_SH_C3[2] * sh[11] * (-3.f * yy + 4.f * zz - xx) + GP_SH_C3[3] * sh[12] * -3.f * 2.f * yz +
                   GP_SH_C3[4] * sh[13] * -2.f * xy + GP_SH_C3[5] * sh[14] * -2.f * yz +
                   GP_SH_C3[6] * sh[15] * -3.f * 2.f * xy);
        dRGBdz += (GP_SH_C3[1] * sh[10] * xy + GP_SH_C3[2] * sh[11] * 4.f * 2.f * yz +
                   GP_SH_C3[3] * sh[12] * 3.f * (2.f * zz - xx - yy) + GP_SH_C3[4] * sh[13] * 4.f * 2.f * xz +
                   GP_SH_C3[5] * sh[14] * (xx - yy));
      }
    }
  }

  dL_dsh.set(0, GP_SH_C0 * dL_dRGB);
  int written = 1;
  if (deg > 0) {
    dL_dsh.set(1, (-GP_SH_C1 * y) * dL_dRGB);
    dL_dsh.set(2, (GP_SH_C1 * z) * dL_dRGB);
    dL_dsh.set(3, (-GP_SH_C1 * x) * dL_dRGB);
    written = 4;
    if (deg > 1) {
      dL_dsh.set(4, (GP_SH_C2[0] * xy) * dL_dRGB);
      dL_dsh.set(5, (GP_SH_C2[1] * yz) * dL_dRGB);
      dL_dsh.set(6, (GP_SH_C2[2] * (2.f * zz - xx - yy)) * dL_dRGB);
      dL_dsh.set(7, (GP_SH_C2[3] * xz) * dL_dRGB);
      dL_dsh.set(8, (GP_SH_C2[4] * (xx - yy)) * dL_dRGB);
      written = 9;
      if (deg > 2) {
        dL_dsh.set(9, (GP_SH_C3[0] * y * (3.f * xx - yy)) * dL_dRGB);
        dL_dsh.set(10, (GP_SH_C3[1] * xy * z) * dL_dRGB);
        dL_dsh.set(11, (GP_SH_C3[2] * y * (4.f * zz - xx - yy)) * dL_dRGB);
        dL_dsh.set(12, (GP_SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dL_dRGB);
        dL_dsh.set(13, (GP_SH_C3[4] * x * (4.f * zz - xx - yy)) * dL_dRGB);
        dL_dsh.set(14, (GP_SH_C3[5] * z * (xx - yy)) * dL_dRGB);
        dL_dsh.set(15, (GP_SH_C3[6] * x * (xx - 3.f * yy)) * dL_dRGB);
        written = 16;
      }
    }
  }
  for (int k = written; k < M; k++) dL_dsh.set(k, v3(0.f, 0.f, 0.f));  // coefficients above the active degree
  const V3 dL_ddir = v3(dot3(dRGBdx, dL_dRGB), dot3(dRGBdy, dL_dRGB), dot3(dRGBdz, dL_dRGB));
  return dnormvdv3(dir_orig, dL_ddir);
}


// Per-Gaussian part of the backward.  sh_row / dsh_row: this Gaussian's 3*M SH coefficients and where their
// gradients go (global memory, or the warp's staging tile in shared memory).
template <bool SPLIT>
__device__ __forceinline__ void
gp_preprocess_bwd_body(int idx, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                       const GpShRow<SPLIT> sh_row, const uint8_t* __restrict__ clamped,
                       const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                       const float* __restrict__ cov3Ds, const float* __restrict__ view,
                       const float* __restrict__ proj, float h_x, float h_y, float tan_fovx, float tan_fovy,
                       const float* __restrict__ campos, const float* __restrict__ grec,
                       float* __restrict__ dL_dmean2D_out, float* __restrict__ dL_dcolors_out,
                       float* __restrict__ dL_dopacity_out, float* __restrict__ dL_dmeans_out,
                       float* __restrict__ dL_dcov3D_out, const GpShRowOut<SPLIT> dsh_row,
                       float* __restrict__ dL_dscale_out, float* __restrict__ dL_drot_out) {
  const bool visible = radii[idx] > 0;

  float rec[GP_GREC];
  if (visible) {
    const float4* r4 = (const float4*)(grec + (size_t)idx * GP_GREC);
    const float4 a = r4[0], b = r4[1], c = r4[2];
    rec[0] = a.x; rec[1] = a.y; rec[2] = a.z; rec[3] = a.w;
    rec[4] = b.x; rec[5] = b.y; rec[6] = b.z; rec[7] = b.w;
    rec[8] = c.x; rec[9] = 0.f; rec[10] = 0.f; rec[11] = 0.f;
  } else {
#pragma unroll
    for (int k = 0; k < GP_GREC; k++) rec[k] = 0.f;
  }
  // gradients that are plain copies of the accumulator
  dL_dmean2D_out[3 * (size_t)idx] = rec[0];
  dL_dmean2D_out[3 * (size_t)idx + 1] = rec[1];
  dL_dmean2D_out[3 * (size_t)idx + 2] = 0.f;
  dL_dopacity_out[idx] = rec[5];
  if (dL_dcolors_out) {
    dL_dcolors_out[3 * (size_t)idx] = rec[6];
    dL_dcolors_out[3 * (size_t)idx + 1] = rec[7];
    dL_dcolors_out[3 * (size_t)idx + 2] = rec[8];
  }

  V3 dL_dmean = v3(0.f, 0.f, 0.f);
  float dL_dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (!visible) {
    dL_dmeans_out[3 * (size_t)idx] = 0.f;
    dL_dmeans_out[3 * (size_t)idx + 1] = 0.f;
    dL_dmeans_out[3 * (size_t)idx + 2] = 0.f;
    if (dL_dcov3D_out)
      for (int k = 0; k < 6; k++) dL_dcov3D_out[6 * (size_t)idx + k] = 0.f;
    if (dsh_row.a)
      for (int k = 0; k < M; k++) dsh_row.set(k, v3(0.f, 0.f, 0.f));
    if (dL_dscale_out) {
      for (int k = 0; k < 3; k++) dL_dscale_out[3 * (size_t)idx + k] = 0.f;
      for (int k = 0; k < 4; k++) dL_drot_out[4 * (size_t)idx + k] = 0.f;
    }
    return;
  }

  const V3 mean = v3(means3D[3 * (size_t)idx], means3D[3 * (size_t)idx + 1], means3D[3 * (size_t)idx + 2]);

  // ---- 2D covariance / conic backward (reference computeCov2DCUDA, backward.cu:144-274) ----
  {
    const float* cov3D = cov3Ds + 6 * (size_t)idx;
    const float dconx = rec[2], dcony = rec[3], dconw = rec[4];
    V3 t = xform4x3(mean, view);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;
    const float x_grad_mul = txtz < -limx || txtz > limx ? 0.f : 1.f;
    const float y_grad_mul = tytz < -limy || tytz > limy ? 0.f : 1.f;

    const M3 J = m3(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z), 0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z), 0, 0, 0);
    const M3 Wm = m3(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    const M3 Vrk = m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    const M3 T = m3mul(Wm, J);
    M3 cov2D = m3mul(m3mul(m3t(T), m3t(Vrk)), T);
    const float a = cov2D.c[0][0] += 0.3f;
    const float b = cov2D.c[0][1];
    const float c = cov2D.c[1][1] += 0.3f;
    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dconx + 2 * b * c * dcony + (denom - a * c) * dconw);
      dL_dc = denom2inv * (-a * a * dconw + 2 * a * b * dcony + (denom - a * c) * dconx);
      dL_db = denom2inv * 2 * (b * c * dconx - (denom + 2 * b * b) * dcony + a * b * dconw);
      dL_dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
      dL_dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
      dL_dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
      dL_dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db +
                   2 * T.c[1][0] * T.c[1][1] * dL_dc;
      dL_dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db +
                   2 * T.c[1][0] * T.c[1][2] * dL_dc;
      dL_dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db +
                   2 * T.c[1][1] * T.c[1][2] * dL_dc;
    }
    const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
    const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
    const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
    const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;
    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                         (2 * h_y * t.y) * tz3 * dL_dJ12;
    dL_dmean = xformvec4x3T(v3(dL_dtx, dL_dty, dL_dtz), view);
  }

  // ---- projection part of dL/dmean (reference backward.cu:367-387) ----
  {
    const float4 m_hom = xform4x4(mean, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    const float gx = rec[0], gy = rec[1];
    V3 d;
    d.x = (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
    d.y = (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
    d.z = (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
    dL_dmean += d;
  }

  // ---- SH backward ----
  if (sh_row.a) {
    const V3 dm = sh_backward<SPLIT>(D, M, mean, v3(campos[0], campos[1], campos[2]), sh_row, clamped[idx],
                              v3(rec[6], rec[7], rec[8]), dsh_row);
    dL_dmean += dm;
  }
  dL_dmeans_out[3 * (size_t)idx] = dL_dmean.x;
  dL_dmeans_out[3 * (size_t)idx + 1] = dL_dmean.y;
  dL_dmeans_out[3 * (size_t)idx + 2] = dL_dmean.z;
  if (dL_dcov3D_out)
    for (int k = 0; k < 6; k++) dL_dcov3D_out[6 * (size_t)idx + k] = dL_dcov[k];

  // ---- 3D covariance backward to scale / raw quaternion (reference backward.cu:278-341) ----
  if (scales) {
    const float4 q = *(const float4*)(rotations + 4 * (size_t)idx);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const M3 R = m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                    2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                    2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    const V3 s = v3(scale_modifier * scales[3 * (size_t)idx], scale_modifier * scales[3 * (size_t)idx + 1],
                    scale_modifier * scales[3 * (size_t)idx + 2]);
    M3 S = m3(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S.c[0][0] = s.x;
    S.c[1][1] = s.y;
    S.c[2][2] = s.z;
    const M3 Mm = m3mul(S, R);
    const M3 dL_dSigma = m3(dL_dcov[0], 0.5f * dL_dcov[1], 0.5f * dL_dcov[2], 0.5f * dL_dcov[1], dL_dcov[3],
                            0.5f * dL_dcov[4], 0.5f * dL_dcov[2], 0.5f * dL_dcov[4], dL_dcov[5]);
    const M3 dL_dM = m3mul(m3scale(2.0f, Mm), dL_dSigma);
    const M3 Rt = m3t(R);
    M3 dL_dMt = m3t(dL_dM);
    dL_dscale_out[3 * (size_t)idx] =
        Rt.c[0][0] * dL_dMt.c[0][0] + Rt.c[0][1] * dL_dMt.c[0][1] + Rt.c[0][2] * dL_dMt.c[0][2];
    dL_dscale_out[3 * (size_t)idx + 1] =
        Rt.c[1][0] * dL_dMt.c[1][0] + Rt.c[1][1] * dL_dMt.c[1][1] + Rt.c[1][2] * dL_dMt.c[1][2];
    dL_dscale_out[3 * (size_t)idx + 2] =
        Rt.c[2][0] * dL_dMt.c[2][0] + Rt.c[2][1] * dL_dMt.c[2][1] + Rt.c[2][2] * dL_dMt.c[2][2];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      dL_dMt.c[0][k] *= s.x;
      dL_dMt.c[1][k] *= s.y;
      dL_dMt.c[2][k] *= s.z;
    }
    float4 dq;
    dq.x = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) +
           2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
    dq.y = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) +
           2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
    dq.z = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) +
           2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
    dq.w = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) +
           2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
    *(float4*)(dL_drot_out + 4 * (size_t)idx) = dq;  // w.r.t. the raw (un-normalised) quaternion
  }
}

// Thread per Gaussian.  The SH coefficients (read) and their gradients (written) are 192-byte rows: accessed
// thread-per-row they cost one 32-byte sector per lane per instruction.  Each warp therefore moves the 6 KB of its
// 32 consecutive Gaussians with fully coalesced 128-bit accesses through a shared-memory tile (rows padded to 52
// floats so that the per-lane row reads are conflict-free).
// Four blocks per SM (64 registers, 53 KB of tiles): the kernel waits on memory, and with separate in / out tiles only
// two blocks fitted (0.33 ms at 2 M Gaussians; 0.24 ms now).
template <bool SPLIT>
__global__ void __launch_bounds__(256, 4)
gp_preprocess_bwd_kernel(int P, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                         const float* __restrict__ shs, const float* __restrict__ shs_rest,
                         const uint8_t* __restrict__ clamped,
                         const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                         const float* __restrict__ cov3Ds, const float* __restrict__ view,
                         const float* __restrict__ proj, float h_x, float h_y, float tan_fovx, float tan_fovy,
                         const float* __restrict__ campos, const float* __restrict__ grec,
                         float* __restrict__ dL_dmean2D_out, float* __restrict__ dL_dcolors_out,
                         float* __restrict__ dL_dopacity_out, float* __restrict__ dL_dmeans_out,
                         float* __restrict__ dL_dcov3D_out, float* __restrict__ dL_dsh_out,
                         float* __restrict__ dL_dsh_rest_out, float* __restrict__ dL_dscale_out,
                         float* __restrict__ dL_drot_out, int staged, int sh_accumulate) {
  extern __shared__ __align__(16) float s_tiles[];  // staged: [8 warps][GP_SH_TILE], coefficients in / gradients out IN PLACE
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int warp_base = idx - lane;
  const int rows = min(32, P - warp_base);
  GpShRow<SPLIT> sh_row{shs ? shs + (size_t)idx * M * 3 : nullptr, nullptr};
  GpShRowOut<SPLIT> dsh_row{dL_dsh_out ? dL_dsh_out + (size_t)idx * M * 3 : nullptr, nullptr};
  float* tin = s_tiles + (size_t)w * GP_SH_TILE;
  float* tout = tin;   // sh_backward reads every coefficient before it writes the first gradient
  if (staged && rows > 0) {
    if (SPLIT) {  // two arrays in, two arrays out (raster_common.cuh)
      gp_warp_copy_in(tin, shs + (size_t)warp_base * 3, rows * 3, lane);
      gp_warp_copy_in(tin + GP_SH_SPLIT_REST, shs_rest + (size_t)warp_base * 45, rows * 45, lane);
      sh_row = GpShRow<SPLIT>{tin + lane * 3, tin + GP_SH_SPLIT_REST + lane * 45};
      dsh_row = GpShRowOut<SPLIT>{tout + lane * 3, tout + GP_SH_SPLIT_REST + lane * 45};
    } else {
      const float4* src = (const float4*)(shs + (size_t)warp_base * 48);
#pragma unroll
      for (int it = 0; it < 12; it++) {
        const int i4 = it * 32 + lane;
        const int row = i4 / 12, c4 = i4 - row * 12;
        if (row < rows) *(float4*)(tin + row * GP_SH_ROW + 4 * c4) = __ldg(src + i4);
      }
      sh_row = GpShRow<SPLIT>{tin + lane * GP_SH_ROW, nullptr};
      dsh_row = GpShRowOut<SPLIT>{tout + lane * GP_SH_ROW, nullptr};
    }
    __syncwarp();
  }
  if (idx < P)
    gp_preprocess_bwd_body<SPLIT>(idx, D, M, means3D, radii, sh_row, clamped, scales, rotations, scale_modifier, cov3Ds,
                                  view, proj, h_x, h_y, tan_fovx, tan_fovy, campos, grec, dL_dmean2D_out,
                                  dL_dcolors_out, dL_dopacity_out, dL_dmeans_out, dL_dcov3D_out, dsh_row, dL_dscale_out,
                                  dL_drot_out);
  if (staged && rows > 0) {
    __syncwarp();
    if (SPLIT) {
      if (sh_accumulate) {   // running sum over the samples of a gradient-accumulation step (gp_b200.h)
        gp_warp_add_out(dL_dsh_out + (size_t)warp_base * 3, tout, rows * 3, lane);
        gp_warp_add_out(dL_dsh_rest_out + (size_t)warp_base * 45, tout + GP_SH_SPLIT_REST, rows * 45, lane);
      } else {
        gp_warp_copy_out(dL_dsh_out + (size_t)warp_base * 3, tout, rows * 3, lane);
        gp_warp_copy_out(dL_dsh_rest_out + (size_t)warp_base * 45, tout + GP_SH_SPLIT_REST, rows * 45, lane);
      }
    } else {
      float4* dst = (float4*)(dL_dsh_out + (size_t)warp_base * 48);
#pragma unroll
      for (int it = 0; it < 12; it++) {
        const int i4 = it * 32 + lane;
        const int row = i4 / 12, c4 = i4 - row * 12;
        if (row < rows) {
          float4 v = *(const float4*)(tout + row * GP_SH_ROW + 4 * c4);
          if (sh_accumulate) {
            const float4 o = dst[i4];
            v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
          }
          dst[i4] = v;
        }
      }
    }
  }
}

extern "C" size_t gp_raster_backward_ws_bytes(int P) { return gp_align_up((size_t)(P > 0 ? P : 1) * GP_GREC * 4, 256); }

extern "C" int gp_raster_backward(const gp_raster_cfg* cfg, const float* bg, const float* means3D, const float* shs,
                                  const float* colors_precomp, const float* scales, const float* rotations,
                                  const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                  const float* campos, const int* radii, const void* geom_ws, const void* binning_ws,
                                  long long capacity, const void* image_ws, const float* dL_dpix, void* grad_ws, float* dL_dmeans2D,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                  float* dL_dsh, float* dL_dscales, float* dL_drots, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!cfg || !bg || !means3D || !radii || !geom_ws || !binning_ws || !image_ws || !dL_dpix || !grad_ws ||
      !dL_dmeans2D || !dL_dopacity || !dL_dmeans3D)
    return GP_ERR_BAD_ARG;
  const int P = cfg->P, W = cfg->width, H = cfg->height;
  if (P <= 0 || W <= 0 || H <= 0 || capacity < 0) return GP_ERR_BAD_ARG;
  if ((shs == nullptr) == (colors_precomp == nullptr)) return GP_ERR_BAD_ARG;
  if (shs && !dL_dsh) return GP_ERR_BAD_ARG;
  if (colors_precomp && !dL_dcolors) return GP_ERR_BAD_ARG;
  const bool has_sr = scales != nullptr && rotations != nullptr;
  if (has_sr == (cov3D_precomp != nullptr)) return GP_ERR_BAD_ARG;
  if (has_sr && (!dL_dscales || !dL_drots)) return GP_ERR_BAD_ARG;
  if (!has_sr && !dL_dcov3D) return GP_ERR_BAD_ARG;

  GpGeom g = gp_carve_geom(const_cast<void*>(geom_ws), P);
  GpBinning bn = gp_carve_binning(const_cast<void*>(binning_ws), capacity);
  GpImage im = gp_carve_image(const_cast<void*>(image_ws), W, H);
  const unsigned gx = (W + GP_TILE_X - 1) / GP_TILE_X, gy = (H + GP_TILE_Y - 1) / GP_TILE_Y;
  const float focal_y = H / (2.0f * cfg->tanfovy);
  const float focal_x = W / (2.0f * cfg->tanfovx);
  float* grec = (float*)grad_ws;
  GP_CUDA_TRY(cudaMemsetAsync(grec, 0, (size_t)P * GP_GREC * 4, stream));
  const uint32_t* point_list = bn.tvals[gp_tile_sort_sel(W, H)];
  static const bool use_butterfly = getenv("GP_RENDER_BWD_BUTTERFLY") != nullptr;   // A/B knob: the round-1 kernel
  if (use_butterfly) {
    GP_KERNEL("gp_render_bwd_kernel", stream, gp_render_bwd_butterfly_kernel<<<gx * gy, GP_TILE_PIX, 0, stream>>>(
        im.ranges, point_list, W, H, gx, bg, g.xy_depth, g.conic_opacity, g.rgb, colors_precomp, im.final_T,
        im.n_contrib, dL_dpix, g.hdr, gp_list_packed(P), grec));
  } else {
    static bool attr_set[GP_STATIC_MAX_DEVICES_ATTR] = {};
    int dev = 0;
    GP_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= GP_STATIC_MAX_DEVICES_ATTR || !attr_set[dev]) {
      GP_CUDA_TRY(cudaFuncSetAttribute(gp_render_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GP_BW_SMEM));
      if (dev >= 0 && dev < GP_STATIC_MAX_DEVICES_ATTR) attr_set[dev] = true;
    }
    GP_KERNEL("gp_render_bwd_kernel", stream, gp_render_bwd_kernel<<<gx * gy, GP_TILE_PIX, GP_BW_SMEM, stream>>>(
        im.ranges, point_list, W, H, gx, bg, g.xy_depth, g.conic_opacity, g.rgb, colors_precomp, im.final_T,
        im.n_contrib, dL_dpix, g.hdr, gp_list_packed(P), grec));
  }
  GP_LAUNCH_CHECK();
  const float* cov3D_ptr = cov3D_precomp ? cov3D_precomp : g.cov3D;
  // SH rows go through shared memory when they are the standard 16 coefficients and 16-byte aligned
  const int staged = (shs != nullptr && cfg->sh_coeffs == 16 &&
                      ((((size_t)shs) | ((size_t)dL_dsh) | ((size_t)cfg->sh_rest) | ((size_t)cfg->dL_dsh_rest)) & 15) == 0) ? 1 : 0;
  if (cfg->sh_rest && (!staged || !cfg->dL_dsh_rest)) return GP_ERR_UNSUPPORTED;  // split SH storage (gp_b200.h)
  if (cfg->dL_dsh_accumulate && !staged) return GP_ERR_UNSUPPORTED;               // only the staged rows are summed in place
  const size_t pb_smem = staged ? (size_t)8 * GP_SH_TILE * sizeof(float) : 0;
  auto kernel = cfg->sh_rest ? gp_preprocess_bwd_kernel<true> : gp_preprocess_bwd_kernel<false>;
  if (pb_smem > 48 * 1024)
    GP_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pb_smem));
  GP_KERNEL("gp_preprocess_bwd_kernel", stream, kernel<<<(P + 255) / 256, 256, pb_smem, stream>>>(
      P, cfg->sh_degree, cfg->sh_coeffs, means3D, radii, shs, cfg->sh_rest, g.clamped, has_sr ? scales : nullptr,
      has_sr ? rotations : nullptr, cfg->scale_modifier, cov3D_ptr, viewmatrix, projmatrix, focal_x, focal_y,
      cfg->tanfovx, cfg->tanfovy, campos, grec, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
      cfg->dL_dsh_rest, dL_dscales, dL_drots, staged, cfg->dL_dsh_accumulate));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of the backward
// ------------------------------------------------------------------------------------------------
//
// gp_raster_backward needs 48 bytes of scratch per Gaussian (the packed gradient accumulator that the
// blend backward sums into).  Some hosts of this library cannot provide it: a real-time render loop
// where cudaMalloc is forbidden (it synchronises the device and may fail under memory pressure),
// plugins embedded in an engine that owns the allocator and hands out no device memory, or bindings
// from languages whose runtime has no device allocator.
// For them the scratch lives in the library image itself: a fixed array in device global memory,
// present on every GPU that loads the library, from load until the process exits.
//
// GP_STATIC_MAX_GAUSSIANS bounds the call: 262,144 Gaussians x 48 B = 12 MiB per GPU.
#define GP_STATIC_MAX_GAUSSIANS 262144
static __device__ __align__(256) float gp_static_grad_scratch[(size_t)GP_STATIC_MAX_GAUSSIANS * GP_GREC];

#define GP_STATIC_MAX_DEVICES 64
// Device address of the array per device ordinal, filled once per device under a lock (two host threads may make
// their first call at the same time); the atomic flag lets later calls skip the lock.
static float* gp_static_scratch_ptr[GP_STATIC_MAX_DEVICES];
static std::atomic<bool> gp_static_scratch_ready[GP_STATIC_MAX_DEVICES];
static std::mutex gp_static_scratch_mu;

extern "C" int gp_raster_backward_static_max_gaussians(void) { return GP_STATIC_MAX_GAUSSIANS; }

// Same arguments and results as gp_raster_backward without `grad_ws`.
//
// * Limit: cfg->P <= GP_STATIC_MAX_GAUSSIANS, otherwise GP_ERR_LIMIT is returned and nothing is launched.
// * Allocates nothing: no cudaMalloc / cudaFree / allocating library call on this path.
// * The first call on each device resolves the device address of the scratch array
//   (cudaGetSymbolAddress, with that device current).  Make that first call outside any stream
//   capture; later calls only enqueue work on `stream`.
// * One array per GPU is shared by every call and keeps the previous call's contents: the part a
//   call uses is cleared on `stream` before the kernels run.  Two calls that can overlap on the same
//   GPU (different streams, or threads) would sum into the same memory and corrupt each other --
//   serialise them, or use gp_raster_backward with separate workspaces.
// * The resolved address is cached per device for the life of the process; host threads may race on the first call
//   (it is locked).  After cudaDeviceReset() the cache would be stale: do not reset a device this library has used.
extern "C" int gp_raster_backward_static(const gp_raster_cfg* cfg, const float* bg, const float* means3D,
                                         const float* shs, const float* colors_precomp, const float* scales,
                                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                         const float* projmatrix, const float* campos, const int* radii,
                                         const void* geom_ws, const void* binning_ws, long long capacity,
                                         const void* image_ws, const float* dL_dpix, float* dL_dmeans2D,
                                         float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                         float* dL_dsh, float* dL_dscales, float* dL_drots, gp_stream_t stream) {
  if (!cfg) return GP_ERR_BAD_ARG;
  if (cfg->P > GP_STATIC_MAX_GAUSSIANS) return GP_ERR_LIMIT;
  int dev = -1;
  GP_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= GP_STATIC_MAX_DEVICES) return GP_ERR_UNSUPPORTED;
  if (!gp_static_scratch_ready[dev].load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> lk(gp_static_scratch_mu);
    if (!gp_static_scratch_ready[dev].load(std::memory_order_relaxed)) {
      // In host code the array's name is only a handle for the runtime, not an address a kernel can use.
      float* p = nullptr;
      GP_CUDA_TRY(cudaGetSymbolAddress((void**)&p, gp_static_grad_scratch));
      if (!p) return GP_ERR_UNSUPPORTED;
      gp_static_scratch_ptr[dev] = p;
      gp_static_scratch_ready[dev].store(true, std::memory_order_release);
    }
  }
  // gp_raster_backward clears the first P records of the scratch on `stream` before accumulating.
  return gp_raster_backward(cfg, bg, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, viewmatrix,
                            projmatrix, campos, radii, geom_ws, binning_ws, capacity, image_ws, dL_dpix,
                            gp_static_scratch_ptr[dev], dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
                            dL_dsh, dL_dscales, dL_drots, stream);
}
