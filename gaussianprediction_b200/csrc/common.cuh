// Shared helpers for the gaussianprediction_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GP_TILE_X 16
#define GP_TILE_Y 16
#define GP_TILE_PIX (GP_TILE_X * GP_TILE_Y)
#define GP_MAX_INSTANCES (1ll << 30)   // sort keys per frame (and Gaussians): the radix sort packs 30-bit counts

// Error codes returned across the C ABI (0 = ok, >0 = cudaError_t, <0 = ours).
#define GP_OK 0
#define GP_ERR_BAD_ARG (-1)
#define GP_ERR_WORKSPACE_TOO_SMALL (-2)
#define GP_ERR_UNSUPPORTED (-3)
#define GP_ERR_LIMIT (-4)

#define GP_CUDA_TRY(expr)                        \
  do {                                           \
    cudaError_t _e = (expr);                     \
    if (_e != cudaSuccess) return (int)_e;       \
  } while (0)

#define GP_LAUNCH_CHECK() GP_CUDA_TRY(cudaGetLastError())

static inline size_t gp_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-provided workspace.  256-byte alignment keeps every
// sub-array eligible for 128-bit vector access and TMA bulk copies.
struct GpCarver {
  char* base;
  size_t off;
  __host__ explicit GpCarver(void* p) : base((char*)p), off(0) {}
  template <typename T>
  __host__ T* take(size_t count) {
    off = gp_align_up(off, 256);
    T* p = base ? (T*)(base + off) : (T*)nullptr;
    off += count * sizeof(T);
    return p;
  }
  __host__ size_t used() const { return gp_align_up(off, 256); }
};

// Launch accounting and optional per-kernel timing (api_misc.cu).  GP_KERNEL(name, stream, launch) counts the
// launch and, when profiling is enabled, brackets it with CUDA events on the launching stream.
void gp_prof_begin(const char* name, cudaStream_t stream);
void gp_prof_end(cudaStream_t stream);
#define GP_KERNEL(name, stream, ...) \
  do {                               \
    gp_prof_begin(name, stream);     \
    __VA_ARGS__;                     \
    gp_prof_end(stream);             \
  } while (0)

#ifdef __CUDACC__
__device__ __forceinline__ unsigned gp_lanemask_lt() {
  unsigned m;
  asm volatile("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
__device__ __forceinline__ int gp_lane() { return threadIdx.x & 31; }
#endif
