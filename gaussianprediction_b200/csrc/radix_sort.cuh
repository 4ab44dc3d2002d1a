// Stable LSD radix sort of (u32 key, u32 value) pairs, 8-bit digits, element count read on the DEVICE.
//
// Replaces cub::DeviceRadixSort::SortPairs at reference
// submodules/diff-gaussian-rasterization-w-depth/cuda_rasterizer/rasterizer_impl.cu:306-311.
// The reference sorts R (tile|depth) 64-bit keys in one 6-pass sort; here the same ordering is
// produced by two short sorts (see raster_fwd.cu): P depth keys, then R tile ids.  Because the
// count of tile instances R is only known on the device, every kernel takes `n_ptr` and uses a
// fixed grid (a multiple of the 148 SMs) so no host round trip is needed to size a launch.
//
// One launch computes the digit histograms of ALL passes (they do not depend on the order of the
// keys); then each pass is ONE kernel in the "onesweep" style: a block takes the next 1024-key tile
// (atomic ticket, so lower tiles are always already running), ranks its keys stably
// (warp = 128 consecutive keys walked 32 at a time, rank inside a step from __match_any_sync, lower
// lane = earlier key), publishes its per-digit counts, finds the counts of all earlier tiles by
// decoupled look-back over per-tile status words (2 flag bits + 30-bit count), and scatters.
// A 4-pass sort of the depth keys is 6 launches, the tile sort 4, instead of 3 launches per pass.
#pragma once
#include "common.cuh"

#define GP_RS_BLOCKS (148 * 4)
#define GP_RS_WARPS 8
#define GP_RS_THREADS (GP_RS_WARPS * 32)
#define GP_RS_RADIX 256
#define GP_RS_STEPS 4                                // 32-key steps per warp per tile
#define GP_RS_TILE (GP_RS_WARPS * GP_RS_STEPS * 32)  // 1024 keys
#define GP_RS_MAX_PASSES 4
#define GP_RS_WINDOW 8  // look-back: status words fetched per round trip

#define GP_RS_FLAG_AGG 0x40000000u
#define GP_RS_FLAG_PREFIX 0x80000000u
#define GP_RS_COUNT_MASK 0x3FFFFFFFu

struct GpRadixTemp {
  uint32_t* hist;     // [MAX_PASSES][256] digit counts
  uint32_t* tickets;  // [MAX_PASSES]
  uint32_t* status;   // [MAX_PASSES][tiles_max][256]
  int tiles_max;
};

static inline int gp_radix_tiles(long long n_max) { return (int)((n_max + GP_RS_TILE - 1) / GP_RS_TILE) + 1; }

// carve the temporaries for sorts of up to n_max elements out of a GpCarver
static inline GpRadixTemp gp_radix_carve(GpCarver& c, long long n_max) {
  GpRadixTemp t;
  t.tiles_max = gp_radix_tiles(n_max);
  t.hist = c.take<uint32_t>((size_t)GP_RS_MAX_PASSES * GP_RS_RADIX);
  t.tickets = c.take<uint32_t>(GP_RS_MAX_PASSES);
  t.status = c.take<uint32_t>((size_t)GP_RS_MAX_PASSES * t.tiles_max * GP_RS_RADIX);
  return t;
}

#ifdef __CUDACC__
// hist and tickets must be zero before gp_rs_prepare_kernel accumulates into them.  Callers that launch some kernel of
// their own right before the sort let one block of it do this (gp_rs_reset_by_block) and pass reset = false.
__device__ __forceinline__ void gp_rs_reset_by_block(const GpRadixTemp& t) {
  for (int i = threadIdx.x; i < GP_RS_MAX_PASSES * GP_RS_RADIX; i += blockDim.x) t.hist[i] = 0u;
  if (threadIdx.x < GP_RS_MAX_PASSES) t.tickets[threadIdx.x] = 0u;
}
static __global__ void gp_rs_reset_kernel(GpRadixTemp t) { gp_rs_reset_by_block(t); }

// For a kernel that PRODUCES the keys of a sort: clear the look-back status words the passes will use (n keys, spread
// over the whole grid) ...
__device__ __forceinline__ void gp_rs_clear_status(const GpRadixTemp& t, int n, int passes) {
  const int ntiles = (n + GP_RS_TILE - 1) / GP_RS_TILE;
  const long long words = (long long)ntiles * GP_RS_RADIX;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int p = 0; p < passes; p++) {
    uint32_t* st = t.status + (size_t)p * t.tiles_max * GP_RS_RADIX;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += stride) st[i] = 0u;
  }
}
// ... and add a block's digit counts (sh[pass][digit], built in shared memory while the keys were written) to the
// global histograms, which some earlier kernel must have zeroed.
__device__ __forceinline__ void gp_rs_flush_histogram(const GpRadixTemp& t, const uint32_t (*sh)[GP_RS_RADIX], int passes) {
  for (int e = threadIdx.x; e < passes * GP_RS_RADIX; e += blockDim.x) {
    const uint32_t v = sh[e / GP_RS_RADIX][e % GP_RS_RADIX];
    if (v) atomicAdd(&t.hist[e], v);
  }
}

// Digit histograms of every pass + reset of the status words the passes will use.
static __global__ void __launch_bounds__(GP_RS_THREADS)
gp_rs_prepare_kernel(const uint32_t* __restrict__ keys, const int* __restrict__ n_ptr, int begin_bit, int passes,
                     GpRadixTemp t) {
  __shared__ uint32_t sh[GP_RS_MAX_PASSES][GP_RS_RADIX];
  const int n = *n_ptr;
  const int tid = threadIdx.x;
  for (int p = 0; p < passes; p++) sh[p][tid] = 0;
  __syncthreads();
  const int stride = gridDim.x * GP_RS_THREADS;
  bool any = false;
  for (int i = blockIdx.x * GP_RS_THREADS + tid; i < n; i += stride) {
    const uint32_t k = __ldg(keys + i);
    for (int p = 0; p < passes; p++) atomicAdd(&sh[p][(k >> (begin_bit + 8 * p)) & 0xFFu], 1u);
    any = true;
  }
  if (__syncthreads_or(any)) {
    for (int p = 0; p < passes; p++) {
      const uint32_t v = sh[p][tid];
      if (v) atomicAdd(&t.hist[p * GP_RS_RADIX + tid], v);
    }
  }
  // status words of the tiles in use
  const int ntiles = (n + GP_RS_TILE - 1) / GP_RS_TILE;
  const long long words = (long long)ntiles * GP_RS_RADIX;
  for (int p = 0; p < passes; p++) {
    uint32_t* st = t.status + (size_t)p * t.tiles_max * GP_RS_RADIX;
    for (long long i = (long long)blockIdx.x * GP_RS_THREADS + tid; i < words; i += stride) st[i] = 0u;
  }
}

// STEPS 32-key steps per warp per tile: tile = 8 warps x STEPS x 32 keys.  Small sorts use 1024-key tiles (enough
// tiles to occupy the SMs), large ones 4096-key tiles (fewer tickets / look-backs / barriers per key, and longer runs
// of equal digits per tile: the keys of a tile are first put in digit order in shared memory, then written out by
// consecutive threads, so a run of m keys with the same digit is one contiguous 4m-byte store).
template <int STEPS>
static __global__ void __launch_bounds__(GP_RS_THREADS)
gp_rs_pass_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                  uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, const int* __restrict__ n_ptr,
                  int shift, int nbits, int pass, GpRadixTemp t) {
  constexpr int TILE = GP_RS_WARPS * STEPS * 32;
  __shared__ uint32_t wcnt[GP_RS_WARPS][GP_RS_RADIX];
  __shared__ uint32_t gshift[GP_RS_RADIX];   // (global position of the tile's first key of digit d) - (its slot in the tile)
  __shared__ uint32_t tstart[GP_RS_RADIX];   // slot of the tile's first key of digit d
  __shared__ uint32_t digit_base[GP_RS_RADIX];
  __shared__ uint32_t warp_tot[GP_RS_WARPS];
  __shared__ uint32_t skv[2 * TILE];
  uint32_t* const sk = skv;
  uint32_t* const sv = skv + TILE;
  __shared__ int s_tile;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int n = *n_ptr;
  const int ntiles = (n + TILE - 1) / TILE;
  if ((int)blockIdx.x >= ntiles) return;  // tickets are handed out in order: this block could only draw a tile >= ntiles
  volatile uint32_t* status = t.status + (size_t)pass * t.tiles_max * GP_RS_RADIX;

  // exclusive scan over the 256 digits of one value per thread (thread = digit)
  auto digit_scan = [&](uint32_t v) -> uint32_t {
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    __syncthreads();                       // warp_tot may still be read from the previous use
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    uint32_t wb = 0;
    for (int j = 0; j < w; j++) wb += warp_tot[j];
    return wb + incl - v;
  };
  // this pass's digit histogram -> first output position of every digit
  digit_base[tid] = digit_scan(t.hist[pass * GP_RS_RADIX + tid]);

  for (;;) {
    __syncthreads();
    if (tid == 0) s_tile = (int)atomicAdd(&t.tickets[pass], 1u);
#pragma unroll
    for (int d = lane; d < GP_RS_RADIX; d += 32) wcnt[w][d] = 0u;
    __syncthreads();
    const int tile = s_tile;
    if (tile >= ntiles) break;
    const int base = tile * TILE + w * (STEPS * 32);
    uint32_t k[STEPS], v[STEPS], local[STEPS];
#pragma unroll
    for (int u = 0; u < STEPS; u++) {
      const int i = base + u * 32 + lane;
      k[u] = i < n ? __ldg(keys_in + i) : 0u;
      v[u] = i < n ? __ldg(vals_in + i) : 0u;
    }
#pragma unroll
    for (int u = 0; u < STEPS; u++) {
      const int i = base + u * 32 + lane;
      const bool valid = i < n;
      const uint32_t d = valid ? ((k[u] >> shift) & 0xFFu) : 0x100u;
      // lanes with the same digit find each other through 8 ballots (one per digit bit): constant cost, whereas
      // match.any takes time that grows with the number of distinct digits in the warp (~30 on a low byte)
      unsigned peers = __ballot_sync(0xffffffffu, valid);
      // (bits of the digit above the sort's last bit are zero in every key: no ballot for them)
#pragma unroll
      for (int b = 0; b < 8; b++) {
        if (b < nbits) {
          const bool bit = (d >> b) & 1u;
          const unsigned m = __ballot_sync(0xffffffffu, bit);
          peers &= bit ? m : ~m;
        }
      }
      const unsigned before = peers & gp_lanemask_lt();
      local[u] = valid ? wcnt[w][d] + __popc(before) : 0u;
      __syncwarp();
      if (valid && before == 0) {
        wcnt[w][d] += __popc(peers);
      }
      __syncwarp();
    }
    __syncthreads();
    // thread `tid` owns digit `tid`: exclusive prefix over the 8 warps, tile total, look-back
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < GP_RS_WARPS; ww++) {
      const uint32_t c = wcnt[ww][tid];
      wcnt[ww][tid] = run;
      run += c;
    }
    status[(size_t)tile * GP_RS_RADIX + tid] = GP_RS_FLAG_AGG | run;
    const uint32_t slot0 = digit_scan(run);  // slot of this digit's first key in the tile (overlaps the look-back latency)
    // look-back: walk the earlier tiles until one has published an inclusive prefix.  The status words of a window of
    // GP_RS_WINDOW tiles are fetched together (independent loads: one L2 round trip instead of one per tile); an entry
    // that is not published yet is polled on its own.  Early in a pass all tiles run at once and nothing but
    // aggregates exists, so the walk is long: at 100k keys the serial version cost ~1/3 of the pass.
    uint32_t excl = 0;
    for (int t0 = tile - 1; t0 >= 0; t0 -= GP_RS_WINDOW) {
      uint32_t sw[GP_RS_WINDOW];
#pragma unroll
      for (int j = 0; j < GP_RS_WINDOW; j++) sw[j] = t0 - j >= 0 ? status[(size_t)(t0 - j) * GP_RS_RADIX + tid] : GP_RS_FLAG_PREFIX;
      bool done = false;
#pragma unroll
      for (int j = 0; j < GP_RS_WINDOW; j++) {
        if (done) continue;
        uint32_t s = sw[j];
        unsigned polls = 0;
        while ((s & (GP_RS_FLAG_AGG | GP_RS_FLAG_PREFIX)) == 0u) {
          s = status[(size_t)(t0 - j) * GP_RS_RADIX + tid];
          // an earlier tile never published: after ~2^30 polls (minutes -- far beyond any time-slicing or debugger pause)
          // fail loudly instead of hanging the device forever
          if (++polls > (1u << 30)) __trap();
        }
        excl += s & GP_RS_COUNT_MASK;
        if (s & GP_RS_FLAG_PREFIX) done = true;
      }
      if (done) break;
    }
    status[(size_t)tile * GP_RS_RADIX + tid] = GP_RS_FLAG_PREFIX | (excl + run);
    tstart[tid] = slot0;
    gshift[tid] = digit_base[tid] + excl - slot0;
    __syncthreads();
    // keys into digit order inside the tile
#pragma unroll
    for (int u = 0; u < STEPS; u++) {
      const int i = base + u * 32 + lane;
      if (i < n) {
        const uint32_t d = (k[u] >> shift) & 0xFFu;
        const uint32_t slot = tstart[d] + wcnt[w][d] + local[u];
        sk[slot] = k[u];
        sv[slot] = v[u];
      }
    }
    __syncthreads();
    const int nt = min(TILE, n - tile * TILE);
#pragma unroll
    for (int u = 0; u < STEPS * 32 / (GP_RS_THREADS / GP_RS_WARPS); u++) {
      const int slot = u * GP_RS_THREADS + tid;
      if (slot < nt) {
        const uint32_t key = sk[slot];
        const uint32_t pos = gshift[(key >> shift) & 0xFFu] + (uint32_t)slot;
        keys_out[pos] = key;
        vals_out[pos] = sv[slot];
      }
    }
  }
}

#define GP_RS_BIG_SORT (1 << 19)  // from this many keys (upper bound known to the host) on: 4096-key tiles

// Sort bits [begin_bit, end_bit) of keys (at most 32 bits = 4 passes); the keys must be ZERO above end_bit (the last
// pass ranks on the bits below end_bit only, but buckets by the whole byte).  keys[0]/vals[0] hold the input; buffers
// ping-pong.  Returns the index (0/1) of the buffer pair holding the sorted result.
static inline int gp_radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], const int* n_ptr, int begin_bit,
                                      int end_bit, const GpRadixTemp& tmp, cudaStream_t stream, bool reset = true,
                                      bool prepared = false) {
  // prepared: the kernel that produced the keys has already built the digit histograms and cleared the status words
  // (gp_rs_block_histogram / gp_rs_clear_status), so no separate pass over the keys is needed
  const int passes = (end_bit - begin_bit + 7) / 8;
  if (reset) GP_KERNEL("gp_rs_reset_kernel", stream, gp_rs_reset_kernel<<<1, 256, 0, stream>>>(tmp));
  if (!prepared)
    GP_KERNEL("gp_rs_prepare_kernel", stream,
              gp_rs_prepare_kernel<<<GP_RS_BLOCKS, GP_RS_THREADS, 0, stream>>>(keys[0], n_ptr, begin_bit, passes, tmp));
  int sel = 0;
  for (int p = 0; p < passes; p++) {
    auto kernel = (long long)tmp.tiles_max * GP_RS_TILE >= GP_RS_BIG_SORT ? gp_rs_pass_kernel<16> : gp_rs_pass_kernel<GP_RS_STEPS>;
    GP_KERNEL("gp_rs_pass_kernel", stream,
              kernel<<<GP_RS_BLOCKS, GP_RS_THREADS, 0, stream>>>(keys[sel], vals[sel], keys[sel ^ 1], vals[sel ^ 1], n_ptr,
                                                                  begin_bit + 8 * p, min(8, end_bit - begin_bit - 8 * p), p, tmp));
    sel ^= 1;
  }
  return sel;
}
#endif
