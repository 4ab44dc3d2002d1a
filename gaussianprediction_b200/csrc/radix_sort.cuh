// Stable LSD radix sort of (u32 key, u32 value) pairs, 8-bit digits, element count read on the DEVICE.
//
// Replaces cub::DeviceRadixSort::SortPairs at reference
// submodules/diff-gaussian-rasterization-w-depth/cuda_rasterizer/rasterizer_impl.cu:306-311.
// The reference sorts R (tile|depth) 64-bit keys in one 6-pass sort; here the same ordering is
// produced by two short sorts (see raster_fwd.cu): P depth keys, then R tile ids.  Because the
// count of tile instances R is only known on the device, every kernel takes `n_ptr` and uses a
// fixed grid (a multiple of the 148 SMs) so no host round trip is needed to size a launch.
//
// Layout: the n keys are cut into GV contiguous "warp chunks" (one per warp of the grid).  A pass is
//   hist    : each warp counts the digits of its chunk                 -> hist[digit][chunk]
//   scan    : one block per digit, exclusive scan over chunks          -> hist (in place), totals[digit]
//   scatter : each warp walks its chunk 32 keys at a time; ranks inside a step come from
//             __match_any_sync (stable: lower lane = earlier key), running offsets live in shared memory.
#pragma once
#include "common.cuh"

#define GP_RS_BLOCKS (148 * 2)
#define GP_RS_WARPS 8
#define GP_RS_THREADS (GP_RS_WARPS * 32)
#define GP_RS_GV (GP_RS_BLOCKS * GP_RS_WARPS)
#define GP_RS_RADIX 256

struct GpRadixTemp {
  uint32_t* hist;    // [256][GV]
  uint32_t* totals;  // [256]
};

static inline size_t gp_radix_temp_bytes() {
  return gp_align_up((size_t)GP_RS_RADIX * GP_RS_GV * 4, 256) + gp_align_up(GP_RS_RADIX * 4, 256);
}

#ifdef __CUDACC__
// Number of warp chunks in use for n keys: at least GP_RS_MIN_CHUNK keys per chunk, so that a small sort does not
// pay for 256 x GV histogram counters (the scan only touches the chunks in use).
#define GP_RS_MIN_CHUNK 256
__device__ __forceinline__ int gp_rs_num_chunks(int n) {
  int c = (n + GP_RS_MIN_CHUNK - 1) / GP_RS_MIN_CHUNK;
  return c < 1 ? 1 : (c > GP_RS_GV ? GP_RS_GV : c);
}
__device__ __forceinline__ void gp_rs_chunk(int n, int gw, int& start, int& end) {
  const int nch = gp_rs_num_chunks(n);
  int sub = (n + nch - 1) / nch;
  sub = (sub + 31) & ~31;
  long long s = (long long)gw * sub;
  long long e = s + sub;
  start = (int)(s < n ? s : n);
  end = (int)(e < n ? e : n);
  if (gw >= nch) start = end = n;
}

static __global__ void __launch_bounds__(GP_RS_THREADS)
gp_rs_hist_kernel(const uint32_t* __restrict__ keys, const int* __restrict__ n_ptr, int shift,
                  uint32_t* __restrict__ hist) {
  __shared__ uint32_t wh[GP_RS_WARPS][GP_RS_RADIX];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gw = blockIdx.x * GP_RS_WARPS + w;
  const int n = *n_ptr;
  if (gw >= gp_rs_num_chunks(n)) return;
  for (int d = lane; d < GP_RS_RADIX; d += 32) wh[w][d] = 0;
  __syncwarp();
  int start, end;
  gp_rs_chunk(n, gw, start, end);
  constexpr int U = 4;
  for (int i0 = start; i0 < end; i0 += 32 * U) {
    uint32_t k[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * 32 + lane;
      k[u] = i < end ? __ldg(keys + i) : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * 32 + lane;
      const bool valid = i < end;
      const uint32_t d = valid ? ((k[u] >> shift) & 0xFFu) : 0x100u;
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      if (valid && (peers & gp_lanemask_lt()) == 0) wh[w][d] += __popc(peers);
      __syncwarp();
    }
  }
  for (int d = lane; d < GP_RS_RADIX; d += 32) hist[(size_t)d * GP_RS_GV + gw] = wh[w][d];
}

// One block per digit: exclusive scan of hist[d][0..GV) in place; totals[d] = sum.
static __global__ void __launch_bounds__(256)
gp_rs_scan_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ totals, const int* __restrict__ n_ptr) {
  constexpr int PER = (GP_RS_GV + 255) / 256;
  __shared__ uint32_t warp_sums[8];
  uint32_t* row = hist + (size_t)blockIdx.x * GP_RS_GV;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int nch = gp_rs_num_chunks(*n_ptr);
  uint32_t v[PER];
  uint32_t local = 0;
#pragma unroll
  for (int j = 0; j < PER; j++) {
    const int i = t * PER + j;
    v[j] = i < nch ? row[i] : 0u;
    local += v[j];
  }
  uint32_t incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  if (lane == 31) warp_sums[w] = incl;
  __syncthreads();
  uint32_t wbase = 0;
  for (int j = 0; j < w; j++) wbase += warp_sums[j];
  uint32_t run = wbase + incl - local;
#pragma unroll
  for (int j = 0; j < PER; j++) {
    const int i = t * PER + j;
    if (i < nch) row[i] = run;
    run += v[j];
  }
  if (t == 255) totals[blockIdx.x] = run;
}

static __global__ void __launch_bounds__(GP_RS_THREADS)
gp_rs_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                     const int* __restrict__ n_ptr, int shift, const uint32_t* __restrict__ hist,
                     const uint32_t* __restrict__ totals) {
  __shared__ uint32_t off[GP_RS_WARPS][GP_RS_RADIX];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gw = blockIdx.x * GP_RS_WARPS + w;
  const int n = *n_ptr;
  if (gw >= gp_rs_num_chunks(n)) return;
  // digit bases: exclusive scan of totals[256]; every warp does it redundantly (8 digits per lane).
  {
    uint32_t tv[8];
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      tv[j] = totals[lane * 8 + j];
      s += tv[j];
    }
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    uint32_t run = incl - s;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int d = lane * 8 + j;
      off[w][d] = run + hist[(size_t)d * GP_RS_GV + gw];
      run += tv[j];
    }
  }
  __syncwarp();
  int start, end;
  gp_rs_chunk(n, gw, start, end);
  constexpr int U = 4;
  for (int i0 = start; i0 < end; i0 += 32 * U) {
    uint32_t k[U], v[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * 32 + lane;
      const bool valid = i < end;
      k[u] = valid ? __ldg(keys_in + i) : 0u;
      v[u] = valid ? __ldg(vals_in + i) : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * 32 + lane;
      const bool valid = i < end;
      const uint32_t d = valid ? ((k[u] >> shift) & 0xFFu) : 0x100u;
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      const unsigned before = peers & gp_lanemask_lt();
      uint32_t pos = 0;
      if (valid) pos = off[w][d] + __popc(before);
      __syncwarp();
      if (valid && before == 0) off[w][d] += __popc(peers);
      __syncwarp();
      if (valid) {
        keys_out[pos] = k[u];
        vals_out[pos] = v[u];
      }
    }
  }
}

// Sort bits [begin_bit, end_bit) of keys.  keys[0]/vals[0] hold the input; buffers ping-pong.
// Returns the index (0/1) of the buffer pair holding the sorted result.
static inline int gp_radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], const int* n_ptr, int begin_bit,
                                      int end_bit, const GpRadixTemp& tmp, cudaStream_t stream) {
  int sel = 0;
  for (int shift = begin_bit; shift < end_bit; shift += 8) {
    GP_KERNEL("gp_rs_hist_kernel", stream, gp_rs_hist_kernel<<<GP_RS_BLOCKS, GP_RS_THREADS, 0, stream>>>(keys[sel], n_ptr, shift, tmp.hist));
    GP_KERNEL("gp_rs_scan_kernel", stream, gp_rs_scan_kernel<<<GP_RS_RADIX, 256, 0, stream>>>(tmp.hist, tmp.totals, n_ptr));
    GP_KERNEL("gp_rs_scatter_kernel", stream, gp_rs_scatter_kernel<<<GP_RS_BLOCKS, GP_RS_THREADS, 0, stream>>>(keys[sel], vals[sel], keys[sel ^ 1], vals[sel ^ 1],
                                                                    n_ptr, shift, tmp.hist, tmp.totals));
    sel ^= 1;
  }
  return sel;
}
#endif
