// Multiresolution hash-grid encoding of the blend-weight network, forward and backward, for sm_100a.
//
// Reference call replaced: `weights = self.weights_model(self.get_xyz.detach())` (scene/gaussian_model.py:257), a
// tinycudann `NetworkWithInputEncoding` built at :370-392 -- Grid/Hash encoding, L = 16 levels, F = 4 features per
// level, T = 2^19 entries per level, base resolution 16, per-level scale b = exp(ln(2048/16)/15), linear interpolation --
// followed by a 64-wide two-hidden-layer ReLU MLP without biases (that part runs through csrc/linear_tc.cu).
// tinycudann is not vendored in the reference tree (scripts/utils/env.sh:7 installs an unpinned fork), so this follows
// the PUBLISHED algorithm (Mueller et al., "Instant Neural Graphics Primitives", 2022, and the tiny-cuda-nn grid
// encoding): per level l
//     scale_l = base * b^l - 1,  resolution_l = ceil(scale_l) + 1,
//     pos = x * scale_l + 0.5,  cell = floor(pos),  w = pos - cell,
//     entry(corner) = dense index (x + y R + z R^2) while R^3 fits the level's table, else the spatial hash
//                     x ^ y * 2654435761 ^ z * 805459861, both modulo the level's entry count,
//     feature_l = sum over the 8 corners of prod_d (w_d or 1 - w_d) * table_l[entry]
// Parity: unpinned (no reference source or vectors); checked against oracle/hashgrid_oracle.py.
//
// Layout: the table is float4 per entry (F = 4), levels back to back (offsets in entries).  One WARP works on one level
// of 32 consecutive points: with the points in Morton order (the same permutation the neighbour search uses) the 32
// lanes touch neighbouring cells, so the coarse and middle levels are served by L1/L2.  Backward: each lane adds its
// 8 weighted gradients with one red.global.add.v4.f32 per corner; when the whole warp hits the SAME entry (coarse levels,
// Morton-ordered points) the 32 contributions are summed with a butterfly first and one lane adds.
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_HG_MAX_LEVELS 32

struct GpHashLevels {
  float scale[GP_HG_MAX_LEVELS];
  unsigned resolution[GP_HG_MAX_LEVELS];
  unsigned offset[GP_HG_MAX_LEVELS + 1];   // in entries
  int n_levels;
};

__device__ __forceinline__ unsigned gp_hg_index(unsigned x, unsigned y, unsigned z, unsigned res, unsigned size) {
  // dense while the running stride stays within the table (tiny-cuda-nn grid_index), else the spatial hash
  unsigned long long stride = 1;
  unsigned idx = 0;
  bool dense = true;
  idx += x * (unsigned)stride;
  stride *= res;
  if (stride > size) dense = false;
  if (dense) {
    idx += y * (unsigned)stride;
    stride *= res;
    if (stride > size) dense = false;
  }
  if (dense) {
    idx += z * (unsigned)stride;
    stride *= res;
  }
  if (!dense || size < stride) idx = x ^ (y * 2654435761u) ^ (z * 805459861u);
  return idx % size;
}

__device__ __forceinline__ void gp_hg_cell(const float* __restrict__ xyz, int p, float scale, unsigned (&c)[3], float (&w)[3]) {
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float pos = fmaf(scale, xyz[3 * (size_t)p + d], 0.5f);
    const float f = floorf(pos);
    c[d] = (unsigned)(int)f;
    w[d] = pos - f;
  }
}

__global__ void __launch_bounds__(256)
gp_hashgrid_fwd_kernel(int N, const float* __restrict__ xyz, const int* __restrict__ order, const float4* __restrict__ table,
                       GpHashLevels lv, float* __restrict__ enc) {
  const int L = lv.n_levels;
  const int warps_total = gridDim.x * (blockDim.x >> 5);
  const int lane = threadIdx.x & 31;
  const int n_chunks = (N + 31) / 32;
  for (long long unit = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < (long long)n_chunks * L; unit += warps_total) {
    const int chunk = (int)(unit / L), l = (int)(unit % L);
    const int slot = chunk * 32 + lane;
    if (slot >= N) continue;
    const int p = order ? order[slot] : slot;
    unsigned c[3];
    float w[3];
    gp_hg_cell(xyz, p, lv.scale[l], c, w);
    const unsigned res = lv.resolution[l], size = lv.offset[l + 1] - lv.offset[l];
    const float4* tl = table + lv.offset[l];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const unsigned dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
      const float wt = (dx ? w[0] : 1.f - w[0]) * (dy ? w[1] : 1.f - w[1]) * (dz ? w[2] : 1.f - w[2]);
      const float4 f = __ldg(tl + gp_hg_index(c[0] + dx, c[1] + dy, c[2] + dz, res, size));
      acc.x = fmaf(wt, f.x, acc.x);
      acc.y = fmaf(wt, f.y, acc.y);
      acc.z = fmaf(wt, f.z, acc.z);
      acc.w = fmaf(wt, f.w, acc.w);
    }
    *(float4*)(enc + (size_t)p * (4 * L) + 4 * l) = acc;
  }
}

__global__ void __launch_bounds__(256)
gp_hashgrid_bwd_kernel(int N, const float* __restrict__ xyz, const int* __restrict__ order, GpHashLevels lv,
                       const float* __restrict__ d_enc, float* __restrict__ d_table) {
  const int L = lv.n_levels;
  const int warps_total = gridDim.x * (blockDim.x >> 5);
  const int lane = threadIdx.x & 31;
  const int n_chunks = (N + 31) / 32;
  for (long long unit = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < (long long)n_chunks * L; unit += warps_total) {
    const int chunk = (int)(unit / L), l = (int)(unit % L);
    const int slot = chunk * 32 + lane;
    const bool valid = slot < N;
    const int p = valid ? (order ? order[slot] : slot) : 0;
    unsigned c[3];
    float w[3];
    gp_hg_cell(xyz, p, lv.scale[l], c, w);
    const unsigned res = lv.resolution[l], size = lv.offset[l + 1] - lv.offset[l];
    float* tl = d_table + 4 * (size_t)lv.offset[l];
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) g = __ldg((const float4*)(d_enc + (size_t)p * (4 * L) + 4 * l));
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const unsigned dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
      const float wt = valid ? (dx ? w[0] : 1.f - w[0]) * (dy ? w[1] : 1.f - w[1]) * (dz ? w[2] : 1.f - w[2]) : 0.f;
      const unsigned idx = valid ? gp_hg_index(c[0] + dx, c[1] + dy, c[2] + dz, res, size) : 0xFFFFFFFFu;
      float4 v = make_float4(wt * g.x, wt * g.y, wt * g.z, wt * g.w);
      // merge the lanes that hit the same entry: the lowest lane of each group adds the group's sum
      const unsigned peers = __match_any_sync(0xffffffffu, idx);
      const int leader = __ffs(peers) - 1;
      if (peers == 0xffffffffu) {
        // the whole warp sits on one entry (coarse levels, Morton-ordered points): plain butterfly
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
          v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
          v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
          v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
        }
      }
      // Partial groups are NOT merged: walking a group with shuffles costs more than the few extra atomics it saves
      // (measured at 2 M points: 4.1 ms with the walk, see profiles/r2_hashgrid_bench.json).
      const bool writer = (peers == 0xffffffffu) ? (lane == leader) : true;
      if (valid && writer)
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(tl + 4 * (size_t)idx), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
  }
}

static int gp_hg_levels(int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size, GpHashLevels* lv) {
  if (n_levels < 1 || n_levels > GP_HG_MAX_LEVELS || base_resolution < 1 || log2_hashmap_size < 1 || log2_hashmap_size > 28) return GP_ERR_BAD_ARG;
  lv->n_levels = n_levels;
  unsigned off = 0;
  const unsigned cap = 1u << log2_hashmap_size;
  for (int l = 0; l < n_levels; l++) {
    // evaluated in double and rounded once, so that every implementation (this host code, the CPU oracle) gets the
    // same float: at the finest levels one ulp of the scale moves the interpolation weights by ~1e-4
    const float scale = (float)(exp2((double)l * log2((double)per_level_scale)) * (double)base_resolution - 1.0);
    const unsigned res = (unsigned)ceilf(scale) + 1u;
    unsigned long long dense = (unsigned long long)res * res * res;
    dense = (dense + 7ull) / 8ull * 8ull;                        // tiny-cuda-nn aligns a level to 8 entries
    const unsigned size = (unsigned)(dense < cap ? dense : cap);
    lv->scale[l] = scale;
    lv->resolution[l] = res;
    lv->offset[l] = off;
    off += size;
  }
  lv->offset[n_levels] = off;
  return GP_OK;
}

extern "C" long long gp_hashgrid_entries(int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size) {
  GpHashLevels lv;
  if (gp_hg_levels(n_levels, base_resolution, per_level_scale, log2_hashmap_size, &lv) != GP_OK) return -1;
  return (long long)lv.offset[n_levels];
}

extern "C" int gp_hashgrid_forward(int N, int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size,
                                   const float* xyz, const int* order, const float* table, float* enc, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GpHashLevels lv;
  const int rc = gp_hg_levels(n_levels, base_resolution, per_level_scale, log2_hashmap_size, &lv);
  if (rc != GP_OK) return rc;
  if (N < 0 || (N > 0 && (!xyz || !table || !enc)) || (((size_t)table) & 15) || (((size_t)enc) & 15)) return GP_ERR_BAD_ARG;
  if (N == 0) return GP_OK;
  GP_KERNEL("gp_hashgrid_fwd_kernel", stream, gp_hashgrid_fwd_kernel<<<148 * 8, 256, 0, stream>>>(N, xyz, order, (const float4*)table, lv, enc));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// d_table must be zero (or hold a running sum) on entry: gradients are ADDED.
extern "C" int gp_hashgrid_backward(int N, int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size,
                                    const float* xyz, const int* order, const float* d_enc, float* d_table, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GpHashLevels lv;
  const int rc = gp_hg_levels(n_levels, base_resolution, per_level_scale, log2_hashmap_size, &lv);
  if (rc != GP_OK) return rc;
  if (N < 0 || (N > 0 && (!xyz || !d_enc || !d_table)) || (((size_t)d_table) & 15) || (((size_t)d_enc) & 15)) return GP_ERR_BAD_ARG;
  if (N == 0) return GP_OK;
  GP_KERNEL("gp_hashgrid_bwd_kernel", stream, gp_hashgrid_bwd_kernel<<<148 * 8, 256, 0, stream>>>(N, xyz, order, lv, d_enc, d_table));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
