// Dense layer on the 5th-generation tensor cores (tcgen05 + TMEM), fp32 in / fp32 out with error-compensated TF32.
//
// Reference path replaced: the per-Gaussian MLPs of GaussianModel.forward -- the first-stage deformation
// (scene/gaussian_model.py:240-244,265) and the step-opacity pass (:291-298): `Deformable_Field`
// (scene/deformable_field.py:112-127), [N,104] -> 4 x (Linear 256 + ReLU) -> Linear 8 over ALL N Gaussians, which the
// reference runs as cuBLAS SGEMMs + elementwise kernels.  It is the one large dense contraction of the path
// (N x 226 k multiply-adds per frame), so it belongs on the tensor cores; the parity bar is the fp32 result of the
// torch modules, so plain TF32 (10-bit mantissa, ~1e-3) is not enough:
//
//   x = hi(x) + lo(x),  hi = x with the low 13 mantissa bits cleared (what the TF32 datapath reads anyway),
//                       lo = x - hi (exact in fp32; its own truncation costs 2^-22 relative to x)
//   x w  ~=  hi(x) hi(w) + hi(x) lo(w) + lo(x) hi(w)        (the dropped lo lo term is 2^-22)
//
// i.e. three tcgen05.mma per K step, accumulated in fp32 in tensor memory.
//
// A CTA walks 128-row tiles of C = act(A W^T + b) (persistent grid, two CTAs per SM):
//   * A [M,K] row major and W [Nout,K] row major (nn.Linear layout) are both "K-major" operands.  K is walked in chunks
//     of 16; all 256 threads load a chunk of A and W with 128-bit loads, split it into hi / lo and store it in the
//     canonical no-swizzle K-major core-matrix layout: 16-byte K slabs, inside a slab the rows are 16 bytes apart
//     (8 rows = one 8 x 16 B core matrix), slab stride = rows * 16 + 16 bytes (the +16 spreads the slabs over the
//     banks).  Descriptor: start, LBO = slab stride, SBO = 128 B, version 1, no swizzle.
//   * two smem stages: while the tensor core works on chunk c the loads of chunk c + 1 are already in flight (registers);
//     `tcgen05.commit` to an mbarrier frees a stage.  TMEM columns, barriers and their phases carry over from tile to
//     tile, and the first chunk of the next tile is loaded while this tile's accumulator is read back.
//   * the accumulator (128 lanes x Nout columns fp32) lives in TMEM; after the last chunk four warps read it back with
//     tcgen05.ld (32 lanes x 16 columns per instruction), add the bias, apply ReLU and write C.
// An optional mask [M,Nout] (the saved output of the ReLU below) zeroes C where mask <= 0: with W passed transposed the
// same kernel is the backward of a layer, dZ_below = (dZ W) * [h > 0].
#include "../../include/gp_b200.h"
#include "common.cuh"

#define GP_LT_THREADS 256
#define GP_LT_ROWS 128
#define GP_LT_KC 16                                    // K chunk per stage (floats)
#define GP_LT_SLABS (GP_LT_KC / 4)                      // 16-byte K slabs per chunk

__device__ __forceinline__ uint32_t gp_lt_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t gp_lt_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;                              // descriptor version (Blackwell)
  return d;                                            // base offset 0, layout type 0 = no swizzle
}

__device__ __forceinline__ void gp_lt_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(gp_lt_smem_u32(bar)), "r"(count));
}

// Bounded wait: a barrier that never completes (a bug) traps after ~seconds instead of hanging the GPU.
__device__ __forceinline__ void gp_lt_mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = gp_lt_smem_u32(bar);
  uint32_t done = 0;
  for (unsigned long long spins = 0; !done; spins++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(a), "r"(parity)
        : "memory");
    if (spins > (1ull << 26)) __trap();
  }
}

__device__ __forceinline__ void gp_lt_mma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void gp_lt_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(gp_lt_smem_u32(bar)) : "memory");
}

// split 4 floats into hi (low 13 mantissa bits cleared) and lo = x - hi
__device__ __forceinline__ void gp_lt_split(const float4 v, float4& hi, float4& lo) {
  hi.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u);
  hi.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
  hi.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u);
  hi.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
  lo.x = v.x - hi.x;
  lo.y = v.y - hi.y;
  lo.z = v.z - hi.z;
  lo.w = v.w - hi.w;
}

// NP = Nout padded to a multiple of 16 (UMMA N); TMEM columns = max(32, pow2 >= NP)
template <int NP>
__global__ void __launch_bounds__(GP_LT_THREADS, 2)
gp_linear_tc_kernel(int M, int K, int Nout, const float* __restrict__ A, int lda, const float* __restrict__ mask, int ldm,
                    const float* __restrict__ W, const float* __restrict__ bias, int relu, float* __restrict__ C, int ldc) {
  constexpr int TMEM_COLS = NP <= 32 ? 32 : NP <= 64 ? 64 : NP <= 128 ? 128 : 256;
  constexpr uint32_t A_SLAB = GP_LT_ROWS * 16 + 16;          // bytes between K slabs of A
  constexpr uint32_t B_SLAB = NP * 16 + 16;
  constexpr uint32_t A_BYTES = GP_LT_SLABS * A_SLAB;          // one of hi / lo
  constexpr uint32_t B_BYTES = GP_LT_SLABS * B_SLAB;
  constexpr uint32_t STAGE = 2 * A_BYTES + 2 * B_BYTES;       // A hi | A lo | B hi | B lo
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar_free[2];               // stage s may be overwritten (its MMAs are done)
  __shared__ __align__(8) uint64_t bar_done;                  // the whole accumulation is done
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int n_tiles = (M + GP_LT_ROWS - 1) / GP_LT_ROWS;

  if (tid == 0) {
    gp_lt_mbar_init(&bar_free[0], 1);
    gp_lt_mbar_init(&bar_free[1], 1);
    gp_lt_mbar_init(&bar_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {   // one warp allocates the accumulator columns (and frees them at the end)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(gp_lt_smem_u32(&s_tmem)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;

  // instruction descriptor: D fp32, A / B tf32, both K-major, N = NP, M = 128
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NP >> 3) << 17) | ((uint32_t)(GP_LT_ROWS >> 4) << 24);
  const uint32_t sbase = gp_lt_smem_u32(smem);
  const int n_chunks = (K + GP_LT_KC - 1) / GP_LT_KC;
  const bool vec_a = ((lda & 3) == 0) && ((((size_t)A) & 15) == 0);
  const bool vec_w = ((K & 3) == 0) && ((((size_t)W) & 15) == 0);

  // Global loads of chunk c + 1 are issued (into registers) before the MMAs of chunk c are launched, so their latency
  // hides behind the tensor-core work and the barrier wait; element e = r * 8 + j (row, 16-byte K slab), j fastest.
  constexpr int A_IT = GP_LT_ROWS * GP_LT_SLABS / GP_LT_THREADS;   // 4
  constexpr int W_IT = (NP * GP_LT_SLABS + GP_LT_THREADS - 1) / GP_LT_THREADS;
  float4 ra[A_IT], rw[W_IT];
  auto load_chunk = [&](int row0, int c) {
    const int k0 = c * GP_LT_KC;
#pragma unroll
    for (int it = 0; it < A_IT; it++) {
      const int e = tid + it * GP_LT_THREADS;
      const int r = e / GP_LT_SLABS, j = e % GP_LT_SLABS;
      const int gr = row0 + r, k = k0 + 4 * j;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gr < M && k < K) {
        if (vec_a && k + 3 < K) {
          v = __ldg((const float4*)(A + (size_t)gr * lda + k));
        } else {
          float t[4];
          for (int u = 0; u < 4; u++) t[u] = (k + u < K) ? A[(size_t)gr * lda + k + u] : 0.f;
          v = make_float4(t[0], t[1], t[2], t[3]);
        }
      }
      ra[it] = v;
    }
#pragma unroll
    for (int it = 0; it < W_IT; it++) {
      const int e = tid + it * GP_LT_THREADS;
      const int r = e / GP_LT_SLABS, j = e % GP_LT_SLABS;
      const int k = k0 + 4 * j;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < Nout && k < K) {
        if (vec_w && k + 3 < K) {
          v = __ldg((const float4*)(W + (size_t)r * K + k));
        } else {
          float t[4];
          for (int u = 0; u < 4; u++) t[u] = (k + u < K) ? W[(size_t)r * K + k + u] : 0.f;
          v = make_float4(t[0], t[1], t[2], t[3]);
        }
      }
      rw[it] = v;
    }
  };
  // Persistent: a CTA walks row tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the TMEM columns, the barriers and their
  // phases carry over (g counts the chunks this CTA has issued, over all its tiles), and the first chunk of the NEXT
  // tile is already being loaded while this tile's accumulator is read back.
  int g = 0, tiles_done = 0;
  if ((int)blockIdx.x < n_tiles) load_chunk(blockIdx.x * GP_LT_ROWS, 0);
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, tiles_done++) {
  const int row0 = tile * GP_LT_ROWS;
  for (int c = 0; c < n_chunks; c++, g++) {
    const int s = g & 1;
    if (g >= 2) gp_lt_mbar_wait(&bar_free[s], ((g >> 1) - 1) & 1);   // the MMAs of two chunks ago have read this stage
    unsigned char* st = smem + (size_t)s * STAGE;
    const int k0 = c * GP_LT_KC;
#pragma unroll
    for (int it = 0; it < A_IT; it++) {
      const int e = tid + it * GP_LT_THREADS;
      const int r = e / GP_LT_SLABS, j = e % GP_LT_SLABS;
      float4 hi, lo;
      gp_lt_split(ra[it], hi, lo);
      *(float4*)(st + j * A_SLAB + r * 16) = hi;
      *(float4*)(st + A_BYTES + j * A_SLAB + r * 16) = lo;
    }
#pragma unroll
    for (int it = 0; it < W_IT; it++) {
      const int e = tid + it * GP_LT_THREADS;
      const int r = e / GP_LT_SLABS, j = e % GP_LT_SLABS;
      if (r < NP) {
        float4 hi, lo;
        gp_lt_split(rw[it], hi, lo);
        *(float4*)(st + 2 * A_BYTES + j * B_SLAB + r * 16) = hi;
        *(float4*)(st + 2 * A_BYTES + B_BYTES + j * B_SLAB + r * 16) = lo;
      }
    }
    if (c + 1 < n_chunks) load_chunk(row0, c + 1);
    else if (tile + (int)gridDim.x < n_tiles) load_chunk((tile + gridDim.x) * GP_LT_ROWS, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core
    __syncthreads();
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a_hi = sbase + s * STAGE, a_lo = a_hi + A_BYTES, b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
      const int ksteps = min(GP_LT_KC / 8, (K - k0 + 7) / 8);       // >= 1; the rest of the chunk is zero padding
      for (int t = 0; t < ksteps; t++) {      // one UMMA K step = 8 tf32 = two 16-byte slabs
        const uint64_t dah = gp_lt_desc(a_hi + 2 * t * A_SLAB, A_SLAB, 128), dal = gp_lt_desc(a_lo + 2 * t * A_SLAB, A_SLAB, 128);
        const uint64_t dbh = gp_lt_desc(b_hi + 2 * t * B_SLAB, B_SLAB, 128), dbl = gp_lt_desc(b_lo + 2 * t * B_SLAB, B_SLAB, 128);
        gp_lt_mma_tf32(tmem, dal, dbh, idesc, (c | t) ? 1u : 0u);   // small terms first
        gp_lt_mma_tf32(tmem, dah, dbl, idesc, 1u);
        gp_lt_mma_tf32(tmem, dah, dbh, idesc, 1u);
      }
      gp_lt_commit(&bar_free[s]);                 // arrives when the MMAs above have finished reading the stage
      if (c == n_chunks - 1) gp_lt_commit(&bar_done);
    }
  }

  // ---- epilogue: warps 0..3 own TMEM lanes 32 w .. 32 w + 31 = rows of the tile ----
  gp_lt_mbar_wait(&bar_done, tiles_done & 1);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp < 4) {
    const int r = row0 + warp * 32 + (tid & 31);
    for (int n0 = 0; n0 < NP; n0 += 16) {
      uint32_t v[16];
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)n0;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(taddr)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (r < M) {
        float o[16];
#pragma unroll
        for (int u = 0; u < 16; u++) {
          const int n = n0 + u;
          float x = __uint_as_float(v[u]) + ((bias != nullptr && n < Nout) ? __ldg(bias + n) : 0.f);
          x = relu ? fmaxf(x, 0.f) : x;
          if (mask != nullptr && n < Nout && !(__ldg(mask + (size_t)r * ldm + n) > 0.f)) x = 0.f;
          o[u] = x;
        }
        float* dst = C + (size_t)r * ldc + n0;
        if (n0 + 16 <= Nout && ((ldc & 3) == 0) && ((((size_t)C) & 15) == 0)) {
#pragma unroll
          for (int u = 0; u < 16; u += 4) *(float4*)(dst + u) = make_float4(o[u], o[u + 1], o[u + 2], o[u + 3]);
        } else {
#pragma unroll
          for (int u = 0; u < 16; u++)
            if (n0 + u < Nout) dst[u] = o[u];
        }
      }
    }
  }
  // the accumulator has been read: the next tile's first MMA (accumulate = 0) may overwrite it
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }  // tiles
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
  }
}

template <int NP>
static int gp_linear_launch(int M, int K, int Nout, const float* A, int lda, const float* mask, int ldm, const float* W,
                            const float* bias, int relu, float* C, int ldc, cudaStream_t stream) {
  constexpr size_t A_BYTES = (size_t)GP_LT_SLABS * (GP_LT_ROWS * 16 + 16), B_BYTES = (size_t)GP_LT_SLABS * (NP * 16 + 16);
  const size_t smem = 2 * (2 * A_BYTES + 2 * B_BYTES);
  auto kern = gp_linear_tc_kernel<NP>;
  GP_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int n_tiles = (M + GP_LT_ROWS - 1) / GP_LT_ROWS;
  int dev = 0, sms = 148;
  GP_CUDA_TRY(cudaGetDevice(&dev));
  GP_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = n_tiles < 2 * sms ? n_tiles : 2 * sms;          // persistent: two CTAs per SM
  GP_KERNEL("gp_linear_tc_kernel", stream, kern<<<grid, GP_LT_THREADS, smem, stream>>>(
      M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_linear_forward(int M, int K, int Nout, const float* A, int lda, const float* mask, int ldm, const float* W,
                                 const float* bias, int relu, float* C, int ldc, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (M < 0 || K < 1 || Nout < 1 || Nout > 256 || !A || !W || !C || lda < K || ldc < Nout || (mask && ldm < Nout)) return GP_ERR_BAD_ARG;
  if (M == 0) return GP_OK;
  if (Nout <= 16) return gp_linear_launch<16>(M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc, stream);
  if (Nout <= 32) return gp_linear_launch<32>(M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc, stream);
  if (Nout <= 64) return gp_linear_launch<64>(M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc, stream);
  if (Nout <= 128) return gp_linear_launch<128>(M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc, stream);
  return gp_linear_launch<256>(M, K, Nout, A, lda, mask, ldm, W, bias, relu, C, ldc, stream);
}
