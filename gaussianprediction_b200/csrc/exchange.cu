// Gradient exchange of the data-parallel frame loop over NVLink 5 / NVSwitch peer memory (sm_100a).
//
// The reference is single-GPU; the seam is its --batch accumulation (train.py:113-133): the gradients of the step's
// (camera, time) samples are SUMMED.  With one process per GPU every rank holds the sum of its own samples in one flat
// buffer (parallel.FlatGradAllReduce) and the step needs the sum over the ranks in every copy.
//
// One kernel does the whole all-reduce in place on a SYMMETRIC allocation (the same virtual layout on every rank,
// mapped into every process: torch.distributed._symmetric_memory supplies the allocation and the address exchange):
//   1. flag barrier: block b of every rank tells block b of every peer that it has started -- its predecessors on the
//      stream, the kernels that wrote the gradients, are therefore complete;
//   2. rank r owns slice r of the range.  With an NVSwitch multicast mapping, one `multimem.ld_reduce.add.v4.f32` per
//      16 bytes makes the SWITCH read the word from all ranks and add them (NVLS), and one `multimem.st.v4.f32` writes
//      the sum back to all ranks: every byte crosses this GPU's links once in and once out, instead of 2 (n - 1) / n
//      times for a ring.  Without multicast (peer mapping only) the owner loads the word from every peer in rank
//      order, adds, and stores it to every peer;
//   3. flag barrier: all slices are written everywhere.
// No host synchronisation, no NCCL launch: the kernel is stream-ordered (capturable), uses a few dozen CTAs and runs
// next to the backward kernels of the following stage.  Each owner computes a word once and broadcasts it, so all
// replicas receive bit-identical sums.
#include "../../include/gp_b200.h"
#include "common.cuh"
#include <limits.h>

#define GP_AR_MAX_WORLD 16
#define GP_AR_THREADS 512
#define GP_AR_FLAG_WORDS (GP_AR_MAX_WORLD + 1)   // per block: one arrival word per peer + this block's epoch

struct GpPeers {
  char* p[GP_AR_MAX_WORLD];
};

__device__ __forceinline__ void gp_st_release_sys(uint32_t* a, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t gp_ld_acquire_sys(const uint32_t* a) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a) : "memory");
  return v;
}

// Block b of this rank <-> block b of every peer.  Epochs only grow, so no flag is ever reset and a fast peer that is
// already one call ahead cannot be mistaken for the current one (the comparison is on the signed difference).
__device__ __forceinline__ void gp_peer_barrier(const GpPeers& peers, size_t flags_off, int rank, int world, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < (unsigned)world) {
    const int q = threadIdx.x;
    __threadfence_system();
    uint32_t* theirs = (uint32_t*)(peers.p[q] + flags_off) + (size_t)blockIdx.x * GP_AR_FLAG_WORDS + rank;
    gp_st_release_sys(theirs, epoch);
    const uint32_t* mine = (const uint32_t*)(peers.p[rank] + flags_off) + (size_t)blockIdx.x * GP_AR_FLAG_WORDS + q;
    // bounded: a peer that never arrives (a rank that died, mismatched calls) ends in a trap after ~half a minute
    // instead of a kernel that spins for ever
    unsigned long long polls = 0;
    while ((int)(gp_ld_acquire_sys(mine) - epoch) < 0) {
      if (++polls > (1ull << 26)) __trap();
    }
  }
  __syncthreads();
}

// This rank's slice of an n4 x 16-byte range: sum over the ranks, result to every rank.
template <bool MULTICAST>
__device__ __forceinline__ void gp_sum_slice(const GpPeers& peers, char* mc_base, int rank, int world, size_t byte_off, size_t n4) {
  const size_t per = (n4 + world - 1) / world;
  const size_t lo = min(n4, per * rank), hi = min(n4, lo + per);
  const size_t stride = (size_t)gridDim.x * GP_AR_THREADS;
  if (MULTICAST) {
    char* base = mc_base + byte_off;
    size_t i = lo + (size_t)blockIdx.x * GP_AR_THREADS + threadIdx.x;
    // four independent 16-byte reductions in flight per thread: the round trip through the switch is microseconds
    for (; i + 3 * stride < hi; i += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; u++)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                     : "l"(base + (i + u * stride) * 16)
                     : "memory");
#pragma unroll
      for (int u = 0; u < 4; u++)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(base + (i + u * stride) * 16),
                     "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                     : "memory");
    }
    for (; i < hi; i += stride) {
      float4 v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                   : "l"(base + i * 16)
                   : "memory");
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(base + i * 16), "f"(v.x), "f"(v.y),
                   "f"(v.z), "f"(v.w)
                   : "memory");
    }
  } else {
    for (size_t i = lo + (size_t)blockIdx.x * GP_AR_THREADS + threadIdx.x; i < hi; i += stride) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int q = 0; q < world; q++) {   // rank order: every owner adds in the same order
        float4 v;   // bypass L1: the peers wrote this memory
        asm volatile("ld.global.relaxed.sys.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "l"(peers.p[q] + byte_off + i * 16)
                     : "memory");
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      for (int q = 0; q < world; q++) *(float4*)(peers.p[q] + byte_off + i * 16) = acc;
    }
  }
}

// Element-wise MAX of int32 (the radii of the step's samples: max over samples and ranks, train.py:122; visibility is
// radii > 0).  4-byte granules (the integer forms of multimem.ld_reduce are scalar).
template <bool MULTICAST>
__device__ __forceinline__ void gp_max_slice(const GpPeers& peers, char* mc_base, int rank, int world, size_t byte_off, size_t n) {
  const size_t per = (n + world - 1) / world;
  const size_t lo = min(n, per * rank), hi = min(n, lo + per);
  const size_t stride = (size_t)gridDim.x * GP_AR_THREADS;
  for (size_t i = lo + (size_t)blockIdx.x * GP_AR_THREADS + threadIdx.x; i < hi; i += stride) {
    if (MULTICAST) {
      int v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.max.s32 %0, [%1];" : "=r"(v) : "l"(mc_base + byte_off + i * 4) : "memory");
      asm volatile("multimem.st.relaxed.sys.global.b32 [%0], %1;" ::"l"(mc_base + byte_off + i * 4), "r"(v) : "memory");
    } else {
      int m = INT_MIN;
      for (int q = 0; q < world; q++) {
        int v;
        asm volatile("ld.global.relaxed.sys.s32 %0, [%1];" : "=r"(v) : "l"(peers.p[q] + byte_off + i * 4) : "memory");
        m = max(m, v);
      }
      for (int q = 0; q < world; q++) *(int*)(peers.p[q] + byte_off + i * 4) = m;
    }
  }
}

template <bool MULTICAST>
__global__ void __launch_bounds__(GP_AR_THREADS)
gp_allreduce_sum_kernel(GpPeers peers, char* mc_base, int rank, int world, size_t byte_off, size_t n4, size_t flags_off) {
  __shared__ uint32_t s_epoch;
  uint32_t* my_epoch = (uint32_t*)(peers.p[rank] + flags_off) + (size_t)blockIdx.x * GP_AR_FLAG_WORDS + GP_AR_MAX_WORLD;
  if (threadIdx.x == 0) s_epoch = *my_epoch;
  __syncthreads();
  const uint32_t e = s_epoch;
  gp_peer_barrier(peers, flags_off, rank, world, e + 1);
  gp_sum_slice<MULTICAST>(peers, mc_base, rank, world, byte_off, n4);
  gp_peer_barrier(peers, flags_off, rank, world, e + 2);
  if (threadIdx.x == 0) *my_epoch = e + 2;
}

template <bool MULTICAST>
__global__ void __launch_bounds__(GP_AR_THREADS)
gp_allreduce_max_kernel(GpPeers peers, char* mc_base, int rank, int world, size_t byte_off, size_t n, size_t flags_off) {
  __shared__ uint32_t s_epoch;
  uint32_t* my_epoch = (uint32_t*)(peers.p[rank] + flags_off) + (size_t)blockIdx.x * GP_AR_FLAG_WORDS + GP_AR_MAX_WORLD;
  if (threadIdx.x == 0) s_epoch = *my_epoch;
  __syncthreads();
  const uint32_t e = s_epoch;
  gp_peer_barrier(peers, flags_off, rank, world, e + 1);
  gp_max_slice<MULTICAST>(peers, mc_base, rank, world, byte_off, n);
  gp_peer_barrier(peers, flags_off, rank, world, e + 2);
  if (threadIdx.x == 0) *my_epoch = e + 2;
}

// Several ranges in ONE resident kernel.  A kernel launched while the GPU is full waits for free SM slots; kernels with
// persistent blocks (the neighbour search) hold theirs for their whole run, so an exchange launched after them starts
// late however its stream is prioritised.  This kernel is launched ONCE, when the last gradients of the step are final,
// takes its few CTAs, and then works through the ranges in the order the next step needs them while that step's
// kernels fill the rest of the machine.  After each range (all ranks have written it: one flag barrier per range) the
// last CTA to finish publishes `call_id` in sync[GP_XS_MAX + s]; the consumer stream waits for that word with
// gp_stream_wait_word just in front of the first reader.  sync[s] (s < GP_XS_MAX) are CTA counters (zero between calls).
#define GP_XS_MAX 8
struct GpSegments {
  size_t byte_off[GP_XS_MAX];
  size_t count[GP_XS_MAX];   // floats (sum, multiple of 4) or ints (max)
  int op[GP_XS_MAX];         // 0: sum f32, 1: max s32
  int n;
};

template <bool MULTICAST>
__global__ void __launch_bounds__(GP_AR_THREADS)
gp_exchange_segments_kernel(GpPeers peers, char* mc_base, int rank, int world, size_t flags_off, GpSegments segs,
                            unsigned* sync, unsigned call_id) {
  __shared__ uint32_t s_epoch;
  uint32_t* my_epoch = (uint32_t*)(peers.p[rank] + flags_off) + (size_t)blockIdx.x * GP_AR_FLAG_WORDS + GP_AR_MAX_WORLD;
  if (threadIdx.x == 0) s_epoch = *my_epoch;
  __syncthreads();
  const uint32_t e = s_epoch;
  gp_peer_barrier(peers, flags_off, rank, world, e + 1);   // every rank's gradients are final (all ranges)
  for (int s = 0; s < segs.n; s++) {
    if (segs.op[s] == 0)
      gp_sum_slice<MULTICAST>(peers, mc_base, rank, world, segs.byte_off[s], segs.count[s] / 4);
    else
      gp_max_slice<MULTICAST>(peers, mc_base, rank, world, segs.byte_off[s], segs.count[s]);
    gp_peer_barrier(peers, flags_off, rank, world, e + 2 + s);   // every rank's slice of this range is written everywhere
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(&sync[s], 1u) == gridDim.x - 1) {
        sync[s] = 0u;
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(sync + GP_XS_MAX + s), "r"(call_id) : "memory");
      }
    }
  }
  if (threadIdx.x == 0) *my_epoch = e + 1 + segs.n;
}

// Stream-ordered wait for a word to reach a value (signed distance: ids only grow).  One thread; sleeps between polls;
// gives up with a trap after ~1 minute instead of hanging the device if the producer never comes.
__global__ void gp_wait_word_kernel(const unsigned* word, unsigned value) {
  unsigned v;
  unsigned long long polls = 0;
  for (;;) {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
    if ((int)(v - value) >= 0) break;
    __nanosleep(500);
    if (++polls > (1ull << 27)) __trap();
  }
}

extern "C" size_t gp_allreduce_flag_bytes(int blocks) {
  return gp_align_up((size_t)(blocks > 0 ? blocks : 1) * GP_AR_FLAG_WORDS * sizeof(uint32_t), 256);
}

extern "C" int gp_allreduce_sum_f32(int rank, int world, void* multicast_base, void* const* peer_bases, size_t byte_offset,
                                    size_t n_floats, size_t flags_byte_offset, int blocks, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world < 1 || world > GP_AR_MAX_WORLD || rank < 0 || rank >= world || !peer_bases || blocks < 1) return GP_ERR_BAD_ARG;
  if ((byte_offset & 15) || (n_floats & 3) || (flags_byte_offset & 15)) return GP_ERR_BAD_ARG;   // 16-byte granules
  if (n_floats == 0) return GP_OK;
  GpPeers peers;
  for (int q = 0; q < GP_AR_MAX_WORLD; q++) peers.p[q] = q < world ? (char*)peer_bases[q] : nullptr;
  for (int q = 0; q < world; q++)
    if (!peers.p[q]) return GP_ERR_BAD_ARG;
  const size_t n4 = n_floats / 4;
  if (multicast_base) {
    GP_KERNEL("gp_allreduce_sum_kernel", stream, gp_allreduce_sum_kernel<true><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, (char*)multicast_base, rank, world, byte_offset, n4, flags_byte_offset));
  } else {
    GP_KERNEL("gp_allreduce_sum_kernel", stream, gp_allreduce_sum_kernel<false><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, nullptr, rank, world, byte_offset, n4, flags_byte_offset));
  }
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_allreduce_max_s32(int rank, int world, void* multicast_base, void* const* peer_bases, size_t byte_offset,
                                    size_t n_ints, size_t flags_byte_offset, int blocks, gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world < 1 || world > GP_AR_MAX_WORLD || rank < 0 || rank >= world || !peer_bases || blocks < 1) return GP_ERR_BAD_ARG;
  if ((byte_offset & 3) || (flags_byte_offset & 15)) return GP_ERR_BAD_ARG;
  if (n_ints == 0) return GP_OK;
  GpPeers peers;
  for (int q = 0; q < GP_AR_MAX_WORLD; q++) peers.p[q] = q < world ? (char*)peer_bases[q] : nullptr;
  for (int q = 0; q < world; q++)
    if (!peers.p[q]) return GP_ERR_BAD_ARG;
  if (multicast_base) {
    GP_KERNEL("gp_allreduce_max_kernel", stream, gp_allreduce_max_kernel<true><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, (char*)multicast_base, rank, world, byte_offset, n_ints, flags_byte_offset));
  } else {
    GP_KERNEL("gp_allreduce_max_kernel", stream, gp_allreduce_max_kernel<false><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, nullptr, rank, world, byte_offset, n_ints, flags_byte_offset));
  }
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_exchange_segments(int rank, int world, void* multicast_base, void* const* peer_bases,
                                    size_t flags_byte_offset, int blocks, int n_segments, const size_t* byte_offsets,
                                    const size_t* counts, const int* ops, unsigned* sync_words, unsigned call_id,
                                    gp_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world < 1 || world > GP_AR_MAX_WORLD || rank < 0 || rank >= world || !peer_bases || blocks < 1) return GP_ERR_BAD_ARG;
  if (n_segments < 1 || n_segments > GP_XS_MAX || !byte_offsets || !counts || !ops || !sync_words) return GP_ERR_BAD_ARG;
  if (flags_byte_offset & 15) return GP_ERR_BAD_ARG;
  GpSegments segs;
  segs.n = n_segments;
  for (int s = 0; s < GP_XS_MAX; s++) {
    segs.byte_off[s] = s < n_segments ? byte_offsets[s] : 0;
    segs.count[s] = s < n_segments ? counts[s] : 0;
    segs.op[s] = s < n_segments ? ops[s] : 0;
    if (s < n_segments) {
      if (ops[s] == 0 && ((byte_offsets[s] & 15) || (counts[s] & 3))) return GP_ERR_BAD_ARG;
      if (ops[s] == 1 && (byte_offsets[s] & 3)) return GP_ERR_BAD_ARG;
      if (ops[s] != 0 && ops[s] != 1) return GP_ERR_BAD_ARG;
    }
  }
  GpPeers peers;
  for (int q = 0; q < GP_AR_MAX_WORLD; q++) peers.p[q] = q < world ? (char*)peer_bases[q] : nullptr;
  for (int q = 0; q < world; q++)
    if (!peers.p[q]) return GP_ERR_BAD_ARG;
  if (multicast_base) {
    GP_KERNEL("gp_exchange_segments_kernel", stream, gp_exchange_segments_kernel<true><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, (char*)multicast_base, rank, world, flags_byte_offset, segs, sync_words, call_id));
  } else {
    GP_KERNEL("gp_exchange_segments_kernel", stream, gp_exchange_segments_kernel<false><<<blocks, GP_AR_THREADS, 0, stream>>>(
        peers, nullptr, rank, world, flags_byte_offset, segs, sync_words, call_id));
  }
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------------------------------------
// Two ranks: the copy engines move the bytes.  A rank pulls its peer's range into local staging with cudaMemcpyAsync (no
// SM involved: measured 733 GB/s for 544 MiB, and the neighbour search next to it runs at full speed, whereas it
// crawls next to the 148-CTA peer-load kernel, tools/ce_copy_probe.py) and adds it in place right in front of the first
// reader.  The kernels below are the pieces: a barrier between the ranks on one flag slot (before the copy: the peer's
// gradients are final; before the in-place add: the peer has finished reading this rank's range), the copy, the add.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) gp_barrier_kernel(GpPeers peers, int rank, int world, size_t flags_off, int slot) {
  uint32_t* my_epoch = (uint32_t*)(peers.p[rank] + flags_off) + (size_t)slot * GP_AR_FLAG_WORDS + GP_AR_MAX_WORLD;
  const uint32_t e = *my_epoch + 1;
  __syncwarp();
  if (threadIdx.x < (unsigned)world) {
    const int q = threadIdx.x;
    __threadfence_system();
    gp_st_release_sys((uint32_t*)(peers.p[q] + flags_off) + (size_t)slot * GP_AR_FLAG_WORDS + rank, e);
    const uint32_t* mine = (const uint32_t*)(peers.p[rank] + flags_off) + (size_t)slot * GP_AR_FLAG_WORDS + q;
    unsigned long long polls = 0;
    while ((int)(gp_ld_acquire_sys(mine) - e) < 0) {
      if (++polls > (1ull << 26)) __trap();     // a peer that never arrives: fail loudly (~half a minute)
    }
  }
  __syncwarp();
  if (threadIdx.x == 0) *my_epoch = e;
}

__global__ void __launch_bounds__(256) gp_add_inplace_kernel(float4* __restrict__ dst, const float4* __restrict__ src, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
    float4 a = dst[i];
    const float4 b = src[i];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    dst[i] = a;
  }
}

__global__ void __launch_bounds__(256) gp_max_inplace_kernel(int* __restrict__ dst, const int* __restrict__ src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) dst[i] = max(dst[i], src[i]);
}

extern "C" int gp_peer_barrier_launch(int rank, int world, void* const* peer_bases, size_t flags_byte_offset, int slot,
                                      gp_stream_t stream_) {
  if (world < 1 || world > GP_AR_MAX_WORLD || rank < 0 || rank >= world || !peer_bases || slot < 0) return GP_ERR_BAD_ARG;
  if (flags_byte_offset & 15) return GP_ERR_BAD_ARG;
  GpPeers peers;
  for (int q = 0; q < GP_AR_MAX_WORLD; q++) peers.p[q] = q < world ? (char*)peer_bases[q] : nullptr;
  for (int q = 0; q < world; q++)
    if (!peers.p[q]) return GP_ERR_BAD_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  GP_KERNEL("gp_barrier_kernel", stream, gp_barrier_kernel<<<1, 32, 0, stream>>>(peers, rank, world, flags_byte_offset, slot));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_copy_async(void* dst, const void* src, size_t bytes, gp_stream_t stream_) {
  if (bytes == 0) return GP_OK;
  if (!dst || !src) return GP_ERR_BAD_ARG;
  GP_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
  return GP_OK;
}

extern "C" int gp_add_inplace_f32(float* dst, const float* src, size_t n_floats, gp_stream_t stream_) {
  if (n_floats == 0) return GP_OK;
  if (!dst || !src || (n_floats & 3) || (((size_t)dst | (size_t)src) & 15)) return GP_ERR_BAD_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  GP_KERNEL("gp_add_inplace_kernel", stream, gp_add_inplace_kernel<<<148 * 8, 256, 0, stream>>>((float4*)dst, (const float4*)src, n_floats / 4));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_max_inplace_s32(int* dst, const int* src, size_t n, gp_stream_t stream_) {
  if (n == 0) return GP_OK;
  if (!dst || !src) return GP_ERR_BAD_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  GP_KERNEL("gp_max_inplace_kernel", stream, gp_max_inplace_kernel<<<148 * 4, 256, 0, stream>>>(dst, src, n));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_exchange_sync_words(void) { return 2 * GP_XS_MAX; }

extern "C" int gp_stream_wait_word(const unsigned* word, unsigned value, gp_stream_t stream_) {
  if (!word) return GP_ERR_BAD_ARG;
  gp_wait_word_kernel<<<1, 1, 0, (cudaStream_t)stream_>>>(word, value);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
