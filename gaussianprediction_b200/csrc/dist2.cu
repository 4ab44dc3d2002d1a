// Mean squared distance to the 3 nearest neighbours of every point (exact), on a uniform hash grid.
//
// Reference path replaced: simple_knn._C.distCUDA2 (submodules/simple-knn/spatial.cu:15-26) ->
// SimpleKNN::knn (simple_knn.cu:185-221): two CUB reductions with blocking D2H copies, Morton codes,
// a CUB radix sort, 1024-point boxes and a scan over every box that cannot be rejected.
// Here: one bounding-box reduction, a counting sort of the points into grid cells, and a ring search
// per point that stops as soon as the third-best distance is inside the searched block of cells.
// No host synchronisation, no allocation (workspace from the caller).
//
// Parity: the result is a function of the SET of the three smallest squared distances, evaluated as
// the reference does -- d = other - self; dist = fma(dz,dz, fma(dx,dx, dy*dy)) (nvcc's contraction of
// simple_knn.cu:141-142) -- and summed in ascending order then divided by 3.0f (:182), so it is
// bit-identical whatever the search order.
#include <float.h>

#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_D2_MAXRES 128

struct GpDist2Ws {
  int* bbox;        // 6 ordered-int encodings: min xyz, max xyz
  int* grid_dims;   // gx, gy, gz, ncells (written on device)
  float* grid_f;    // minx, miny, minz, inv_cell, cell
  uint32_t* cell_of;    // [P]
  uint32_t* cnt;        // [MAXCELLS] counts, then exclusive offsets
  uint32_t* fill;       // [MAXCELLS]
  uint32_t* blk;        // [MAXCELLS/1024 + 1]
  float4* sorted;       // [P] (x, y, z, original index bits)
  size_t bytes;
};

static inline GpDist2Ws gp_carve_dist2(void* ws, int P) {
  GpCarver c(ws);
  GpDist2Ws w;
  const size_t n = (size_t)(P > 0 ? P : 1);
  const size_t cells = (size_t)GP_D2_MAXRES * GP_D2_MAXRES * GP_D2_MAXRES;
  w.bbox = c.take<int>(8);
  w.grid_dims = c.take<int>(4);
  w.grid_f = c.take<float>(8);
  w.cell_of = c.take<uint32_t>(n);
  w.cnt = c.take<uint32_t>(cells + 1);
  w.fill = c.take<uint32_t>(cells + 1);
  w.blk = c.take<uint32_t>(cells / 1024 + 2);
  w.sorted = c.take<float4>(n);
  w.bytes = c.used();
  return w;
}

__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void gp_d2_init_kernel(GpDist2Ws w) {
  if (threadIdx.x < 3) w.bbox[threadIdx.x] = 0x7fffffff;
  else if (threadIdx.x < 6) w.bbox[threadIdx.x] = (int)0x80000000;
}

__global__ void __launch_bounds__(256) gp_d2_bbox_kernel(int P, const float* __restrict__ pts, GpDist2Ws w) {
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const float v = pts[(size_t)i * 3 + a];
      mn[a] = fminf(mn[a], v);
      mx[a] = fmaxf(mx[a], v);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      atomicMin(&w.bbox[a], f2ord(mn[a]));
      atomicMax(&w.bbox[3 + a], f2ord(mx[a]));
    }
  }
}

__global__ void gp_d2_grid_kernel(int P, GpDist2Ws w) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float mn[3], ext[3];
  for (int a = 0; a < 3; a++) {
    mn[a] = ord2f(w.bbox[a]);
    ext[a] = fmaxf(ord2f(w.bbox[3 + a]) - mn[a], 0.f);
  }
  // aim at ~2 points per cell assuming the points fill the box; clamp each axis to MAXRES cells
  const float emax = fmaxf(fmaxf(ext[0], ext[1]), ext[2]);
  float cell;
  if (!(emax > 0.f)) {
    cell = 1.f;
  } else {
    const float e0 = fmaxf(ext[0], emax * 1e-3f), e1 = fmaxf(ext[1], emax * 1e-3f), e2 = fmaxf(ext[2], emax * 1e-3f);
    cell = cbrtf(e0 * e1 * e2 / fmaxf((float)P * 0.5f, 1.f));
    cell = fmaxf(cell, emax / (float)GP_D2_MAXRES * 1.0001f);
  }
  int g[3];
  for (int a = 0; a < 3; a++) {
    int r = (int)floorf(ext[a] / cell) + 1;
    g[a] = min(max(r, 1), GP_D2_MAXRES);
  }
  w.grid_dims[0] = g[0];
  w.grid_dims[1] = g[1];
  w.grid_dims[2] = g[2];
  w.grid_dims[3] = g[0] * g[1] * g[2];
  w.grid_f[0] = mn[0];
  w.grid_f[1] = mn[1];
  w.grid_f[2] = mn[2];
  w.grid_f[3] = 1.0f / cell;
  w.grid_f[4] = cell;
}

__device__ __forceinline__ int cell_coord(float v, float mn, float inv, int g) {
  int c = (int)floorf((v - mn) * inv);
  return min(max(c, 0), g - 1);
}

__global__ void gp_d2_zero_kernel(GpDist2Ws w) {
  const int n = w.grid_dims[3];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) {
    w.cnt[i] = 0;
    w.fill[i] = 0;
  }
}

__global__ void __launch_bounds__(256) gp_d2_count_kernel(int P, const float* __restrict__ pts, GpDist2Ws w) {
  const int gx = w.grid_dims[0], gy = w.grid_dims[1], gz = w.grid_dims[2];
  const float mx = w.grid_f[0], my = w.grid_f[1], mz = w.grid_f[2], inv = w.grid_f[3];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    const int cx = cell_coord(pts[(size_t)i * 3], mx, inv, gx);
    const int cy = cell_coord(pts[(size_t)i * 3 + 1], my, inv, gy);
    const int cz = cell_coord(pts[(size_t)i * 3 + 2], mz, inv, gz);
    const uint32_t c = (uint32_t)((cz * gy + cy) * gx + cx);
    w.cell_of[i] = c;
    atomicAdd(&w.cnt[c], 1u);
  }
}

// exclusive scan of cnt[0..ncells] in three steps (1024 per block)
__global__ void __launch_bounds__(1024) gp_d2_scan1_kernel(GpDist2Ws w) {
  __shared__ uint32_t ws[32];
  const int n = w.grid_dims[3] + 1;
  const int nblk = (n + 1023) / 1024;
  for (int b = blockIdx.x; b < nblk; b += gridDim.x) {
    const int i = b * 1024 + threadIdx.x;
    uint32_t v = i < n ? w.cnt[i] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t t = 0;
      for (int j = 0; j < 32; j++) t += ws[j];
      w.blk[b] = t;
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(1024) gp_d2_scan2_kernel(GpDist2Ws w) {
  __shared__ uint32_t ws[32];
  __shared__ uint32_t carry;
  const int n = w.grid_dims[3] + 1;
  const int nblk = (n + 1023) / 1024;
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < nblk ? w.blk[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) ws[wp] = incl;
    __syncthreads();
    uint32_t wb = 0;
    for (int j = 0; j < wp; j++) wb += ws[j];
    const uint32_t c = carry;
    if (i < nblk) w.blk[i] = c + wb + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = c + wb + incl;
    __syncthreads();
  }
}
__global__ void __launch_bounds__(1024) gp_d2_scan3_kernel(GpDist2Ws w) {
  __shared__ uint32_t ws[32];
  const int n = w.grid_dims[3] + 1;
  const int nblk = (n + 1023) / 1024;
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  for (int b = blockIdx.x; b < nblk; b += gridDim.x) {
    const int i = b * 1024 + threadIdx.x;
    const uint32_t v = i < n ? w.cnt[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) ws[wp] = incl;
    __syncthreads();
    uint32_t wb = 0;
    for (int j = 0; j < wp; j++) wb += ws[j];
    if (i < n) w.cnt[i] = w.blk[b] + wb + incl - v;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) gp_d2_scatter_kernel(int P, const float* __restrict__ pts, GpDist2Ws w) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    const uint32_t c = w.cell_of[i];
    const uint32_t pos = w.cnt[c] + atomicAdd(&w.fill[c], 1u);
    w.sorted[pos] = make_float4(pts[(size_t)i * 3], pts[(size_t)i * 3 + 1], pts[(size_t)i * 3 + 2], __int_as_float(i));
  }
}

__device__ __forceinline__ void kbest3(float dist, float* best) {
#pragma unroll
  for (int j = 0; j < 3; j++) {
    if (best[j] > dist) {
      const float t = best[j];
      best[j] = dist;
      dist = t;
    }
  }
}

__global__ void __launch_bounds__(128) gp_d2_query_kernel(int P, GpDist2Ws w, float* __restrict__ out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;  // position in cell order: neighbours share cache lines
  if (s >= P) return;
  const int gx = w.grid_dims[0], gy = w.grid_dims[1], gz = w.grid_dims[2];
  const float mnx = w.grid_f[0], mny = w.grid_f[1], mnz = w.grid_f[2], inv = w.grid_f[3], cell = w.grid_f[4];
  const float4 me = w.sorted[s];
  const int cx = cell_coord(me.x, mnx, inv, gx), cy = cell_coord(me.y, mny, inv, gy), cz = cell_coord(me.z, mnz, inv, gz);
  float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  const int maxring = max(max(gx, gy), gz);
  for (int ring = 0; ring <= maxring; ring++) {
    const int x0 = max(cx - ring, 0), x1 = min(cx + ring, gx - 1);
    const int y0 = max(cy - ring, 0), y1 = min(cy + ring, gy - 1);
    const int z0 = max(cz - ring, 0), z1 = min(cz + ring, gz - 1);
    auto visit = [&](int x, int y, int z) {
      const uint32_t c = (uint32_t)((z * gy + y) * gx + x);
      const uint32_t b = w.cnt[c], e = w.cnt[c + 1];
      for (uint32_t k = b; k < e; k++) {
        if ((int)k == s) continue;
        const float4 o = w.sorted[k];
        const float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
        const float dist = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
        kbest3(dist, best);
      }
    };
    // visit only the cells at Chebyshev distance == ring from the centre cell
    for (int z = z0; z <= z1; z++) {
      const bool zshell = (z == cz - ring) || (z == cz + ring);
      for (int y = y0; y <= y1; y++) {
        if (zshell || (y == cy - ring) || (y == cy + ring)) {
          for (int x = x0; x <= x1; x++) visit(x, y, z);
        } else {
          if (cx - ring >= 0) visit(cx - ring, y, z);
          if (cx + ring <= gx - 1) visit(cx + ring, y, z);
        }
      }
    }
    // distance from the point to the nearest face of the searched block that still has cells behind it
    float bound = FLT_MAX;
    if (cx - ring > 0) bound = fminf(bound, me.x - (mnx + (float)(cx - ring) * cell));
    if (cx + ring < gx - 1) bound = fminf(bound, (mnx + (float)(cx + ring + 1) * cell) - me.x);
    if (cy - ring > 0) bound = fminf(bound, me.y - (mny + (float)(cy - ring) * cell));
    if (cy + ring < gy - 1) bound = fminf(bound, (mny + (float)(cy + ring + 1) * cell) - me.y);
    if (cz - ring > 0) bound = fminf(bound, me.z - (mnz + (float)(cz - ring) * cell));
    if (cz + ring < gz - 1) bound = fminf(bound, (mnz + (float)(cz + ring + 1) * cell) - me.z);
    if (bound == FLT_MAX) break;  // whole grid searched
    bound -= cell * 1e-3f;        // safety margin for the rounding of the cell assignment
    if (bound > 0.f && best[2] <= bound * bound) break;
  }
  out[__float_as_int(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

extern "C" size_t gp_dist2_ws_bytes(int P) { return gp_carve_dist2(nullptr, P).bytes; }

extern "C" int gp_dist2_knn3(int P, const float* points, float* mean_dist2, void* ws, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return GP_ERR_BAD_ARG;
  if (P == 0) return GP_OK;
  if (!points || !mean_dist2 || !ws) return GP_ERR_BAD_ARG;
  GpDist2Ws w = gp_carve_dist2(ws, P);
  const int grid = 148 * 4;
  GP_KERNEL("gp_d2_init_kernel", stream, gp_d2_init_kernel<<<1, 32, 0, stream>>>(w));
  GP_KERNEL("gp_d2_bbox_kernel", stream, gp_d2_bbox_kernel<<<grid, 256, 0, stream>>>(P, points, w));
  GP_KERNEL("gp_d2_grid_kernel", stream, gp_d2_grid_kernel<<<1, 32, 0, stream>>>(P, w));
  GP_KERNEL("gp_d2_zero_kernel", stream, gp_d2_zero_kernel<<<grid, 256, 0, stream>>>(w));
  GP_KERNEL("gp_d2_count_kernel", stream, gp_d2_count_kernel<<<grid, 256, 0, stream>>>(P, points, w));
  GP_KERNEL("gp_d2_scan1_kernel", stream, gp_d2_scan1_kernel<<<grid, 1024, 0, stream>>>(w));
  GP_KERNEL("gp_d2_scan2_kernel", stream, gp_d2_scan2_kernel<<<1, 1024, 0, stream>>>(w));
  GP_KERNEL("gp_d2_scan3_kernel", stream, gp_d2_scan3_kernel<<<grid, 1024, 0, stream>>>(w));
  GP_KERNEL("gp_d2_scatter_kernel", stream, gp_d2_scatter_kernel<<<grid, 256, 0, stream>>>(P, points, w));
  GP_KERNEL("gp_d2_query_kernel", stream, gp_d2_query_kernel<<<(P + 127) / 128, 128, 0, stream>>>(P, w, mean_dist2));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
