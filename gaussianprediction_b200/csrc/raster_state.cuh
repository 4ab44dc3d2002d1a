// Layout of the three rasterizer workspaces (caller-allocated, carved here).
//
// Plays the role of the reference's GeometryState / BinningState / ImageState
// (cuda_rasterizer/rasterizer_impl.h:29-65) but laid out for B200: per-Gaussian records are
// 16-byte float4 so that the render kernels gather them with single 128-bit loads, and the binning
// stage keeps two (key, value) u32 ping-pong pairs instead of 64-bit keys.
#pragma once
#include "../../include/gp_b200.h"
#include "common.cuh"
#include "radix_sort.cuh"

#define GP_SCAN_BLOCK 256  // sorted Gaussians handled by one block of the duplicate/scan kernels

struct GpGeom {
  int* hdr;                 // GP_HDR_WORDS int32
  float4* conic_opacity;    // [P] (conic.x, conic.y, conic.z, opacity)
  float4* xy_depth;         // [P] (pix x, pix y, view depth, unused)
  float4* rgb;              // [P] (r, g, b, unused)  -- only when colours come from SH
  float* cov3D;             // [P*6]
  uint8_t* clamped;         // [P] bit c set if channel c was clamped at 0
  uint32_t* tiles_touched;  // [P]
  uint32_t* dkeys[2];       // [P] depth bits (0xFFFFFFFF for culled), ping-pong
  uint32_t* dvals[2];       // [P] Gaussian index, ping-pong
  uint32_t* block_sums;     // [ceil(P/GP_SCAN_BLOCK)] tiles touched per block of sorted Gaussians (then exclusive)
  uint32_t* tile_hist;      // digit histograms + tickets of the TILE sort: kept here (not in the binning workspace) so
  uint32_t* tile_tickets;   // that a kernel of the preprocess stage can zero them before the binning buffers exist
  GpRadixTemp radix;
  size_t bytes;
};

struct GpBinning {
  uint32_t* tkeys[2];  // [capacity] tile id
  uint32_t* tvals[2];  // [capacity] Gaussian index
  GpRadixTemp radix;   // look-back status words of the tile sort
  size_t bytes;
};

struct GpImage {
  float* final_T;       // [H*W]
  uint32_t* n_contrib;  // [H*W]
  uint2* ranges;        // [tiles]
  size_t bytes;
};

static inline GpGeom gp_carve_geom(void* ws, int P) {
  GpCarver c(ws);
  GpGeom g;
  size_t n = (size_t)(P > 0 ? P : 1);
  g.hdr = c.take<int>(GP_HDR_WORDS);
  g.conic_opacity = c.take<float4>(n);
  g.xy_depth = c.take<float4>(n);
  g.rgb = c.take<float4>(n);
  g.cov3D = c.take<float>(n * 6);
  g.clamped = c.take<uint8_t>(n);
  g.tiles_touched = c.take<uint32_t>(n);
  g.dkeys[0] = c.take<uint32_t>(n);
  g.dkeys[1] = c.take<uint32_t>(n);
  g.dvals[0] = c.take<uint32_t>(n);
  g.dvals[1] = c.take<uint32_t>(n);
  g.block_sums = c.take<uint32_t>((n + GP_SCAN_BLOCK - 1) / GP_SCAN_BLOCK + 1);
  g.tile_hist = c.take<uint32_t>((size_t)GP_RS_MAX_PASSES * GP_RS_RADIX);
  g.tile_tickets = c.take<uint32_t>(GP_RS_MAX_PASSES);
  g.radix = gp_radix_carve(c, (long long)n);
  g.bytes = c.used();
  return g;
}

static inline GpBinning gp_carve_binning(void* ws, long long capacity) {
  GpCarver c(ws);
  GpBinning b;
  size_t n = (size_t)(capacity > 0 ? capacity : 1);
  b.tkeys[0] = c.take<uint32_t>(n);
  b.tkeys[1] = c.take<uint32_t>(n);
  b.tvals[0] = c.take<uint32_t>(n);
  b.tvals[1] = c.take<uint32_t>(n);
  b.radix = gp_radix_carve(c, (long long)n);
  b.bytes = c.used();
  return b;
}

static inline GpImage gp_carve_image(void* ws, int W, int H) {
  GpCarver c(ws);
  GpImage im;
  size_t npix = (size_t)W * H;
  size_t tiles = (size_t)((W + GP_TILE_X - 1) / GP_TILE_X) * ((H + GP_TILE_Y - 1) / GP_TILE_Y);
  im.final_T = c.take<float>(npix ? npix : 1);
  im.n_contrib = c.take<uint32_t>(npix ? npix : 1);
  im.ranges = c.take<uint2>(tiles ? tiles : 1);
  im.bytes = c.used();
  return im;
}

// Depth keys are full 32-bit float patterns -> 4 passes.  Tile ids need ceil(log2(tiles)) bits.
static inline int gp_tile_bits(int W, int H) {
  unsigned tiles = (unsigned)((W + GP_TILE_X - 1) / GP_TILE_X) * (unsigned)((H + GP_TILE_Y - 1) / GP_TILE_Y);
  int bits = 1;
  while (bits < 32 && (1u << bits) < tiles) bits++;
  return bits;
}
static inline int gp_tile_sort_sel(int W, int H) { return ((gp_tile_bits(W, H) + 7) / 8) & 1; }
