// The keypoint-motion MLP of the deformation stage, forward and backward, as three small kernels.
//
// Stands in for the torch ops of reference scene/gaussian_model.py:180-189 (get_motion_delta: repeat t, cat, df_model)
// on top of scene/deformable_field.py:63-72 (positional_encoding) and :74-127 (Deformable_Field: d x (Linear+ReLU),
// then Linear) for the keypoint call sites (:255-265).  With ~512 keypoints the torch path is ~70 launches of
// 3-10 us each per iteration (0.26 ms of a 1.18 ms train iteration on B200, profiles/r1_keypoint_mlp.md); the work
// itself is ~1 GFLOP, so it is launch/latency bound and is fused here:
//
//   forward    : a block owns 4 keypoints; 512 threads = 256 hidden units x 2 groups that split every chunk's summation
//                range.  Builds the input row [feature | PE(xyz) | PE(t)] in shared memory, walks the layers with the
//                activations ping-ponging between two shared-memory rows, keeps the post-ReLU activations for the
//                backward, normalises the quaternion part of the output if asked.  The weights stream through a
//                3-stage cp.async ring of 64 KB chunks that runs ahead across layer boundaries.
//   bwd (data) : same tiling and ring walked backwards, ReLU masks from the saved activations; writes
//                dL/d(pre-activation) of every layer, the feature gradient and the xyz gradient through the encoding.
//   bwd (weights): all layers in one launch, 32x64 output tiles of dW_l = dZ_l^T H_{l-1} summed over the keypoints
//                in a fixed order (no atomics: bit-reproducible), bias gradient alongside, 3-stage cp.async ring.
// Every block re-reads all the weights (872 KB, L2 hits); what bounds the kernels is one SM's shared-memory / issue
// rate on the consume side, not L2 or HBM bandwidth (profiles/r1_keypoint_mlp.md).
//
// fp32 FMA throughout like the reference's fp32 cuBLAS GEMMs (no TF32); sums run in index order.
#include "../../include/gp_b200.h"
#include "common.cuh"

#define MLP_ROWS 4
#define MLP_MAXW 256
#define MLP_MAXL 9  // hidden layers + the output layer
#define MLP_THREADS 256

struct MlpDims {
  int rows, F, xf, tf, depth, w, out, din, norm_q;
};
struct MlpPtrs {
  const float* W[MLP_MAXL];
  const float* b[MLP_MAXL];
  float* gW[MLP_MAXL];
  float* gb[MLP_MAXL];
};
struct MlpTiles {
  int first[MLP_MAXL + 1];
};

static int mlp_dims(const gp_mlp_cfg* c, MlpDims* d) {
  if (!c) return GP_ERR_BAD_ARG;
  d->rows = c->rows; d->F = c->feature_dim; d->xf = c->xyz_freq; d->tf = c->time_freq;
  d->depth = c->depth; d->w = c->width; d->out = c->out_dim; d->norm_q = c->normalize_q != 0;
  d->din = c->feature_dim + 6 * c->xyz_freq + 2 * c->time_freq;
  if (c->rows < 0 || c->feature_dim < 0 || c->xyz_freq < 0 || c->time_freq < 0) return GP_ERR_BAD_ARG;
  if (c->xyz_freq > 24 || c->time_freq > 24) return GP_ERR_LIMIT;  // 2^f must stay exact in the encoding
  if (c->depth < 1 || c->depth > MLP_MAXL - 1) return GP_ERR_LIMIT;
  if (c->width < 4 || c->width > MLP_MAXW || (c->width & 3)) return GP_ERR_LIMIT;
  if (d->din < 4 || d->din > MLP_MAXW || (d->din & 3)) return GP_ERR_LIMIT;
  if (c->out_dim != 7 && c->out_dim != 8) return GP_ERR_LIMIT;  // (xyz, quaternion[, opacity]): get_motion_delta's columns
  return GP_OK;
}

// ------------------------------------------------------------------------------------------------ async copies

// Ampere-style asynchronous global->shared copies: the weight chunks of the next stages are in flight while the current
// one is being consumed, with no staging registers.  (A weight matrix is re-read by every block, so it comes from L2.)
__device__ __forceinline__ void mlp_cp16(float* smem, const float* gmem, bool valid = true) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 16 : 0;  // 0 source bytes: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(n) : "memory");
}
__device__ __forceinline__ void mlp_cp16s(unsigned smem_addr, const float* gmem) {  // shared address already converted
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void mlp_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void mlp_cp_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// The network's output row is (delta_xyz[3], delta_q[4][, delta_o]) (scene/gaussian_model.py:189); the three tensors
// cross the ABI separately so that no slicing copies are needed on either side.
struct MlpOut {
  float* xyz;  // [rows,3]
  float* q;    // [rows,4]
  float* o;    // [rows,1] or NULL
};
struct MlpOutGrad {
  const float* xyz;
  const float* q;
  const float* o;
};
#define MLP_NORM_EPS 1e-12f  // torch.nn.functional.normalize default eps

#define MLP_CH 64     // weight columns (forward) / rows (backward) per chunk
#define MLP_STAGES 3  // chunks resident in shared memory
#define MLP_SPLIT 2   // thread groups sharing one chunk: group h takes half of the chunk's summation range
#define MLP_BIG (MLP_MAXW * MLP_SPLIT)  // 512 threads: thread = (group h, unit u)
#define MLP_STAGE_FLOATS (MLP_CH * MLP_MAXW)
#define MLP_STAGE_BYTES (MLP_STAGE_FLOATS * 4)
// activations ping-pong | weight ring | output-layer weights | partial sums of the groups
#define MLP_SMEM_FLOATS (2 * MLP_ROWS * MLP_MAXW + MLP_STAGES * MLP_STAGE_FLOATS + 8 * MLP_MAXW + MLP_SPLIT * MLP_ROWS * MLP_MAXW)
#define MLP_SMEM_BYTES (MLP_SMEM_FLOATS * sizeof(float))

extern __shared__ __align__(16) float mlp_smem[];

// ------------------------------------------------------------------------------------------------ forward

// column c of the network input of keypoint `row` (scene/gaussian_model.py:188: cat([feature, xyz_embed, t]))
__device__ __forceinline__ float mlp_input(const MlpDims& d, const float* __restrict__ xyz, const float* __restrict__ feat,
                                           float t, int row, int c) {
  if (c < d.F) return __ldg(feat + (size_t)row * d.F + c);
  c -= d.F;
  float x;
  int f;
  if (c < 6 * d.xf) {
    const int dim = c / (2 * d.xf);
    f = (c >> 1) % d.xf;
    x = __ldg(xyz + (size_t)row * 3 + dim);
  } else {
    c -= 6 * d.xf;
    f = c >> 1;
    x = t;
  }
  const float u = x * (float)(1u << f);  // scene/deformable_field.py:66-67: positions * 2^f
  return (c & 1) ? cosf(u) : sinf(u);    // :69: [sin, cos] interleaved
}

// Hidden unit u needs row u of W_l ([out][in]).  A chunk is 64 input columns of all rows; in shared memory row j keeps
// its sixteen 16-byte granules at positions g ^ (j & 7), so that the 128-bit reads of 8 consecutive threads (one
// granule index, 8 consecutive rows) fall on 8 different bank groups.  One block = 4 keypoints x 256 units x 2 groups;
// group h accumulates granules 8h..8h+7 of every chunk and the partial sums meet in shared memory at the end of a
// layer.  (History, profiles/r1_keypoint_mlp.md: 256 threads and 32-column chunks sat at 2 warps per scheduler and 41%
// issue utilisation; 1024 threads were issue-bound on per-chunk bookkeeping, 110 of 155 instructions per chunk.)
static __global__ void __launch_bounds__(MLP_BIG)
gp_kpmlp_fwd_kernel(MlpDims d, MlpPtrs p, const float* __restrict__ xyz, const float* __restrict__ feat,
                    const float* __restrict__ t_ptr, float* __restrict__ x0, float* __restrict__ acts,
                    float* __restrict__ raw_out, MlpOut out) {
  float(*xs)[MLP_ROWS][MLP_MAXW] = reinterpret_cast<float(*)[MLP_ROWS][MLP_MAXW]>(mlp_smem);
  float* stages = mlp_smem + 2 * MLP_ROWS * MLP_MAXW;
  float* wlast = stages + MLP_STAGES * MLP_STAGE_FLOATS;  // [out][w]
  float(*red)[MLP_ROWS][MLP_MAXW] = reinterpret_cast<float(*)[MLP_ROWS][MLP_MAXW]>(wlast + 8 * MLP_MAXW);
  const int tid = threadIdx.x, h = tid >> 8, u = tid & (MLP_MAXW - 1);
  const int r0 = blockIdx.x * MLP_ROWS;
  const int n0 = (d.din + MLP_CH - 1) / MLP_CH, nw = (d.w + MLP_CH - 1) / MLP_CH;
  const int total = n0 + (d.depth - 1) * nw;
  // The stream of chunks: layer 0 (n0 chunks), then layers 1..depth-1 (nw each).  Blocks walk the chunks of a layer in
  // rotated orders (start = blockIdx mod n) so that the blocks are not all pulling the same 64 KB at the same moment.
  // Two cursors (copy-issue and consume) step through the stream; everything per-thread about a copy is hoisted.
  const int rot0 = blockIdx.x % n0, rotw = blockIdx.x % nw;
  const unsigned stages_s = (unsigned)__cvta_generic_to_shared(stages);
  const int jb = tid >> 4, g = tid & 15;  // this thread copies granule g of rows jb, jb+32, ..., jb+224
  const unsigned dst_thread = (unsigned)((jb * 16 + (g ^ (jb & 7))) * 16);
  int il = 0, ikc = 0, istage = 0;
  auto issue = [&]() {
    if (il < d.depth) {
      const int n = il == 0 ? n0 : nw, K = il == 0 ? d.din : d.w;
      int kq = ikc + (il == 0 ? rot0 : rotw);
      if (kq >= n) kq -= n;
      const int k = kq * MLP_CH + 4 * g;
      if (k < K) {
        const float* src = p.W[il] + (unsigned)(jb * K + k);
        const unsigned dst = stages_s + istage * MLP_STAGE_BYTES + dst_thread;
        const unsigned step = 32u * K;
#pragma unroll
        for (int i = 0; i < 8; i++)
          if (jb + 32 * i < d.w) mlp_cp16s(dst + i * (32 * 16 * 16), src + i * step);
      }
      if (++ikc == n) { ikc = 0; il++; }
    }
    if (++istage == MLP_STAGES) istage = 0;
    mlp_cp_commit();
  };
  // the output layer's weights ride along with the first chunk
  for (int idx = tid; idx < (d.out * d.w) >> 2; idx += MLP_BIG) mlp_cp16(wlast + idx * 4, p.W[d.depth] + idx * 4);
  for (int c = 0; c < MLP_STAGES - 1; c++) issue();
  const float t = (d.tf > 0 && t_ptr) ? __ldg(t_ptr) : 0.f;
  for (int i = tid; i < MLP_ROWS * d.din; i += MLP_BIG) {
    const int r = i / d.din, c = i - r * d.din, row = r0 + r;
    float v = 0.f;
    if (row < d.rows) {
      v = mlp_input(d, xyz, feat, t, row, c);
      x0[(size_t)row * d.din + c] = v;
    }
    xs[0][r][c] = v;
  }
  int cur = 0, cl = 0, ckc = 0, cstage = 0;
  float acc[MLP_ROWS];
  const int sw = u & 7;
  constexpr int GPT = MLP_CH / 4 / MLP_SPLIT;  // granules per thread per chunk
  for (int c = 0; c < total; c++) {
    mlp_cp_wait<MLP_STAGES - 2>();
    __syncthreads();  // chunk c is visible to all; every thread is done with chunk c-1, whose stage is refilled next
    issue();
    const int n = cl == 0 ? n0 : nw, K = cl == 0 ? d.din : d.w;
    int kq = ckc + (cl == 0 ? rot0 : rotw);
    if (kq >= n) kq -= n;
    const int k0 = kq * MLP_CH;
    if (ckc == 0) {
#pragma unroll
      for (int r = 0; r < MLP_ROWS; r++) acc[r] = 0.f;
    }
    if (u < d.w) {
      const float* st = stages + cstage * MLP_STAGE_FLOATS + u * MLP_CH + h * (GPT * 4);
      const float* xr = &xs[cur][0][k0 + h * (GPT * 4)];
      auto granule = [&](int q) {  // granule h*GPT + q of the chunk; the swizzle only touches the low 3 bits
        const float4 wv = *reinterpret_cast<const float4*>(st + ((q ^ sw) << 2));
#pragma unroll
        for (int r = 0; r < MLP_ROWS; r++) {
          const float4 xv = *reinterpret_cast<const float4*>(xr + r * MLP_MAXW + 4 * q);
          acc[r] = fmaf(xv.x, wv.x, acc[r]);
          acc[r] = fmaf(xv.y, wv.y, acc[r]);
          acc[r] = fmaf(xv.z, wv.z, acc[r]);
          acc[r] = fmaf(xv.w, wv.w, acc[r]);
        }
      };
      const int left = (K - k0 - h * (GPT * 4)) >> 2;  // valid granules of this thread's range
      if (left >= GPT) {
#pragma unroll
        for (int q = 0; q < GPT; q++) granule(q);
      } else {
        for (int q = 0; q < left; q++) granule(q);
      }
    }
    if (ckc == n - 1) {
      // layer complete: the groups' partial sums meet; thread (h, u) finishes rows 2h, 2h+1 of unit u:
      // bias, ReLU, keep for the backward, hand to the next layer
#pragma unroll
      for (int r = 0; r < MLP_ROWS; r++) red[h][r][u] = acc[r];
      __syncthreads();
      if (u < d.w) {
        const float b = __ldg(p.b[cl] + u);
#pragma unroll
        for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) {
          const int r = h * (MLP_ROWS / MLP_SPLIT) + rr;
          float v = b;
#pragma unroll
          for (int hh = 0; hh < MLP_SPLIT; hh++) v += red[hh][r][u];
          v = fmaxf(v, 0.f);
          xs[cur ^ 1][r][u] = v;
          if (r0 + r < d.rows) acts[((size_t)cl * d.rows + r0 + r) * d.w + u] = v;
        }
      }
      ckc = 0; cl++; cur ^= 1;
    } else {
      ckc++;
    }
    if (++cstage == MLP_STAGES) cstage = 0;
  }
  mlp_cp_wait<0>();
  __syncthreads();
  // output layer: (row, output, k-slice) = (4, 8, 8) threads, slices summed by shuffles
  if (tid < 256) {
    const int r = tid >> 6, j = (tid >> 3) & 7, s = tid & 7;
    float part = 0.f;
    if (j < d.out)
      for (int k = s; k < d.w; k += 8) part = fmaf(xs[cur][r][k], wlast[j * d.w + k], part);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    part += __shfl_xor_sync(0xffffffffu, part, 4);
    if (s == 0) red[0][r][j] = j < d.out ? part + __ldg(p.b[d.depth] + j) : 0.f;
  }
  __syncthreads();
  if (tid < MLP_ROWS * 8) {
    // thread (row, column): keep the raw row for the backward, normalise the quaternion if asked
    // (scene/gaussian_model.py:266-267: rotation_activation = torch.nn.functional.normalize), write the three tensors
    const int r = tid >> 3, j = tid & 7, row = r0 + r;
    if (row < d.rows) {
      const float* o = &red[0][r][0];
      float v = o[j];
      raw_out[(size_t)row * 8 + j] = v;
      if (d.norm_q && j >= 3 && j < 7) {
        const float n = sqrtf(o[3] * o[3] + o[4] * o[4] + o[5] * o[5] + o[6] * o[6]);
        v = v / fmaxf(n, MLP_NORM_EPS);
      }
      if (j < 3) out.xyz[(size_t)row * 3 + j] = v;
      else if (j < 7) out.q[(size_t)row * 4 + (j - 3)] = v;
      else if (j < d.out && out.o) out.o[row] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward: data

// Input unit u of layer l needs column u of W_l: a chunk is 64 whole rows of W_l, read from shared memory with
// consecutive threads on consecutive words; group h takes rows 32h..32h+31 of the chunk.
static __global__ void __launch_bounds__(MLP_BIG)
gp_kpmlp_bwd_data_kernel(MlpDims d, MlpPtrs p, const float* __restrict__ xyz, const float* __restrict__ t_ptr,
                         const float* __restrict__ acts, const float* __restrict__ raw_out, MlpOutGrad gout,
                         float* __restrict__ dout, float* __restrict__ dz, float* __restrict__ g_xyz,
                         float* __restrict__ g_feat) {
  float(*gs)[MLP_ROWS][MLP_MAXW] = reinterpret_cast<float(*)[MLP_ROWS][MLP_MAXW]>(mlp_smem);
  float* stages = mlp_smem + 2 * MLP_ROWS * MLP_MAXW;
  float(*red)[MLP_ROWS][MLP_MAXW] =
      reinterpret_cast<float(*)[MLP_ROWS][MLP_MAXW]>(stages + MLP_STAGES * MLP_STAGE_FLOATS + 8 * MLP_MAXW);
  __shared__ float go[MLP_ROWS][8];
  const int tid = threadIdx.x, h = tid >> 8, u = tid & (MLP_MAXW - 1);
  const int r0 = blockIdx.x * MLP_ROWS;
  const int nw = (d.w + MLP_CH - 1) / MLP_CH;
  const int total = d.depth * nw;
  // chunks: layers depth-1 .. 0, nw row-blocks of 64 rows each, rotated per block like the forward.  In shared memory
  // the rows keep a fixed stride of MLP_MAXW floats whatever the layer's row length.
  const int rotw = blockIdx.x % nw;
  const unsigned stages_s = (unsigned)__cvta_generic_to_shared(stages);
  const int lane = tid & 31, wp = tid >> 5;  // warp wp copies rows wp, wp+16, wp+32, wp+48 of the chunk
  int il = d.depth - 1, ijc = 0, istage = 0;
  auto issue = [&]() {
    if (il >= 0) {
      const int K = il == 0 ? d.din : d.w;
      int jq = ijc + rotw;
      if (jq >= nw) jq -= nw;
      const int j = jq * MLP_CH + wp;
      const float* src = p.W[il] + (unsigned)(j * K + 4 * lane);
      const unsigned dst = stages_s + istage * MLP_STAGE_BYTES + (unsigned)(wp * MLP_MAXW + 4 * lane) * 4;
      const unsigned step = 16u * K;
#pragma unroll
      for (int i = 0; i < MLP_CH / 16; i++) {
        if (j + 16 * i < d.w) {
          if (4 * lane < K) mlp_cp16s(dst + i * (16 * MLP_MAXW * 4), src + i * step);
          if (4 * lane + 128 < K) mlp_cp16s(dst + i * (16 * MLP_MAXW * 4) + 512, src + i * step + 128);
        }
      }
      if (++ijc == nw) { ijc = 0; il--; }
    }
    if (++istage == MLP_STAGES) istage = 0;
    mlp_cp_commit();
  };
  for (int c = 0; c < MLP_STAGES - 1; c++) issue();
  if (tid < MLP_ROWS * 8) {
    const int r = tid >> 3, j = tid & 7;
    // dL/d(raw output row): the incoming gradients of the three tensors, the quaternion part taken back through the
    // normalisation (x / max(|x|, eps)): g/n - x (x.g)/n^3 above eps, g/eps below.  Kept in `dout` ([rows][8], zero
    // padded) for the weight-gradient kernel.
    const int row = r0 + r;
    float v = 0.f;
    if (row < d.rows) {
      if (j < 3) v = gout.xyz ? __ldg(gout.xyz + (size_t)row * 3 + j) : 0.f;
      else if (j < 7) {
        v = gout.q ? __ldg(gout.q + (size_t)row * 4 + (j - 3)) : 0.f;
        if (d.norm_q && gout.q) {
          const float* x = raw_out + (size_t)row * 8 + 3;
          const float* gq = gout.q + (size_t)row * 4;
          const float x0 = __ldg(x), x1 = __ldg(x + 1), x2 = __ldg(x + 2), x3 = __ldg(x + 3);
          const float n = sqrtf(x0 * x0 + x1 * x1 + x2 * x2 + x3 * x3);
          if (n > MLP_NORM_EPS) {
            const float dot = x0 * __ldg(gq) + x1 * __ldg(gq + 1) + x2 * __ldg(gq + 2) + x3 * __ldg(gq + 3);
            const float inv = 1.f / n;
            v = v * inv - __ldg(x + (j - 3)) * (dot * inv * inv * inv);
          } else {
            v = v / MLP_NORM_EPS;
          }
        }
      } else if (j < d.out) v = gout.o ? __ldg(gout.o + row) : 0.f;
      dout[(size_t)row * 8 + j] = v;
    }
    go[r][j] = v;
  }
  __syncthreads();
  if (u < d.w) {  // thread (h, u): rows 2h, 2h+1 through the output Linear, then the ReLU of the last hidden layer
    float a0[MLP_ROWS / MLP_SPLIT];
#pragma unroll
    for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) a0[rr] = 0.f;
    for (int j = 0; j < d.out; j++) {
      const float wv = __ldg(p.W[d.depth] + (size_t)j * d.w + u);
#pragma unroll
      for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) a0[rr] = fmaf(go[h * (MLP_ROWS / MLP_SPLIT) + rr][j], wv, a0[rr]);
    }
#pragma unroll
    for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) {
      const int r = h * (MLP_ROWS / MLP_SPLIT) + rr, row = r0 + r;
      const size_t a = ((size_t)(d.depth - 1) * d.rows + row) * d.w + u;
      const float v = (row < d.rows && __ldg(acts + a) > 0.f) ? a0[rr] : 0.f;
      gs[0][r][u] = v;
      if (row < d.rows) dz[a] = v;
    }
  }
  int cur = 0, cl = d.depth - 1, cjc = 0, cstage = 0;
  float acc[MLP_ROWS];
  float hprev[MLP_ROWS / MLP_SPLIT];  // activations (rows 2h, 2h+1; unit u) of layer cl-1: the ReLU mask
  constexpr int RPT = MLP_CH / MLP_SPLIT;  // chunk rows per thread
  for (int c = 0; c < total; c++) {
    mlp_cp_wait<MLP_STAGES - 2>();
    __syncthreads();
    issue();
    // gradient wrt the input of layer cl: sum_j gs[r][j] W[j][k]; for cl > 0 masked by the ReLU of layer cl-1
    const int K = cl == 0 ? d.din : d.w;
    int jq = cjc + rotw;
    if (jq >= nw) jq -= nw;
    const int j0 = jq * MLP_CH + h * RPT;  // first row of this thread's range
    if (cjc == 0) {
#pragma unroll
      for (int r = 0; r < MLP_ROWS; r++) acc[r] = 0.f;
#pragma unroll
      for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) {
        const int row = r0 + h * (MLP_ROWS / MLP_SPLIT) + rr;
        hprev[rr] = (cl > 0 && u < d.w && row < d.rows) ? __ldg(acts + ((size_t)(cl - 1) * d.rows + row) * d.w + u) : 0.f;
      }
    }
    if (u < K) {
      const float* st = stages + cstage * MLP_STAGE_FLOATS + h * (RPT * MLP_MAXW) + u;
      const float* gr = &gs[cur][0][j0];
      auto quad = [&](int jj) {
        const float w0 = st[(jj + 0) * MLP_MAXW], w1 = st[(jj + 1) * MLP_MAXW];
        const float w2 = st[(jj + 2) * MLP_MAXW], w3 = st[(jj + 3) * MLP_MAXW];
#pragma unroll
        for (int r = 0; r < MLP_ROWS; r++) {
          const float4 gv = *reinterpret_cast<const float4*>(gr + r * MLP_MAXW + jj);
          acc[r] = fmaf(gv.x, w0, acc[r]);
          acc[r] = fmaf(gv.y, w1, acc[r]);
          acc[r] = fmaf(gv.z, w2, acc[r]);
          acc[r] = fmaf(gv.w, w3, acc[r]);
        }
      };
      const int left = d.w - j0;
      if (left >= RPT) {
#pragma unroll
        for (int jj = 0; jj < RPT; jj += 4) quad(jj);
      } else {
        for (int jj = 0; jj < left; jj += 4) quad(jj);
      }
    }
    if (cjc == nw - 1) {
#pragma unroll
      for (int r = 0; r < MLP_ROWS; r++) red[h][r][u] = acc[r];
      __syncthreads();
      if (u < K) {
#pragma unroll
        for (int rr = 0; rr < MLP_ROWS / MLP_SPLIT; rr++) {
          const int r = h * (MLP_ROWS / MLP_SPLIT) + rr;
          float v = 0.f;
#pragma unroll
          for (int hh = 0; hh < MLP_SPLIT; hh++) v += red[hh][r][u];
          if (cl > 0) {
            v = hprev[rr] > 0.f ? v : 0.f;
            if (r0 + r < d.rows) dz[((size_t)(cl - 1) * d.rows + r0 + r) * d.w + u] = v;
          }
          gs[cur ^ 1][r][u] = v;
        }
      }
      cjc = 0; cl--; cur ^= 1;
    } else {
      cjc++;
    }
    if (++cstage == MLP_STAGES) cstage = 0;
  }
  mlp_cp_wait<0>();
  __syncthreads();
  // gs[cur] now holds dL/d(input row).  Features pass through; xyz goes back through the encoding.
  if (g_feat) {
    for (int i = tid; i < MLP_ROWS * d.F; i += MLP_BIG) {
      const int r = i / d.F, c = i - r * d.F;
      if (r0 + r < d.rows) g_feat[(size_t)(r0 + r) * d.F + c] = gs[cur][r][c];
    }
  }
  if (g_xyz && tid < MLP_ROWS * 3) {
    const int r = tid / 3, dim = tid - r * 3, row = r0 + r;
    if (row < d.rows) {
      const float x = __ldg(xyz + (size_t)row * 3 + dim);
      float a = 0.f;
      for (int f = 0; f < d.xf; f++) {
        const float sc = (float)(1u << f), uu = x * sc;
        const int c = d.F + (dim * d.xf + f) * 2;
        a += sc * (gs[cur][r][c] * cosf(uu) - gs[cur][r][c + 1] * sinf(uu));
      }
      g_xyz[(size_t)row * 3 + dim] = a;
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward: weights

#define MLP_TJ 32
#define MLP_TK 64
#define MLP_RC 32
#define MLP_WSTAGES 3
#define MLP_WTHREADS 512  // two groups of 256: group h takes keypoints 16h..16h+15 of every 32-keypoint chunk

static __global__ void __launch_bounds__(MLP_WTHREADS)
gp_kpmlp_bwd_weights_kernel(MlpDims d, MlpPtrs p, MlpTiles tl, const float* __restrict__ x0,
                            const float* __restrict__ acts, const float* __restrict__ dz,
                            const float* __restrict__ dout) {
  __shared__ __align__(16) float a_s[MLP_WSTAGES][MLP_RC][MLP_TJ];
  __shared__ __align__(16) float b_s[MLP_WSTAGES][MLP_RC][MLP_TK];
  int l = 0;
  while (l < d.depth && (int)blockIdx.x >= tl.first[l + 1]) l++;
  const int J = l < d.depth ? d.w : d.out, K = l == 0 ? d.din : d.w;
  const float* __restrict__ A = l < d.depth ? dz + (size_t)l * d.rows * d.w : dout;  // [rows][ldA], zero beyond J
  const int ldA = l < d.depth ? d.w : 8;
  const float* __restrict__ B = l == 0 ? x0 : acts + (size_t)(l - 1) * d.rows * d.w;    // [rows][K]
  const int tiles_k = (K + MLP_TK - 1) / MLP_TK;
  const int local = blockIdx.x - tl.first[l];
  const int tk = local % tiles_k;
  const int j0 = (local / tiles_k) * MLP_TJ, k0 = tk * MLP_TK;
  const int tid = threadIdx.x, h = tid >> 8, ty = (tid >> 4) & 15, tx = tid & 15;
  const int nchunks = (d.rows + MLP_RC - 1) / MLP_RC;
  int ic = blockIdx.x % nchunks, istage = 0, issued = 0;  // rotated start: see the forward
  auto issue = [&]() {
    if (issued < nchunks) {
      const int rc = ic * MLP_RC;
      if (tid < 256) {
        const int rr = tid >> 3, c4 = (tid & 7) * 4, row = rc + rr, j = j0 + c4;
        const bool ok = row < d.rows && j < ldA;
        mlp_cp16(&a_s[istage][rr][c4], ok ? A + (size_t)row * ldA + j : A, ok);
      }
      {
        const int rr = tid >> 4, c4 = (tid & 15) * 4, row = rc + rr, k = k0 + c4;
        const bool ok = row < d.rows && k < K;
        mlp_cp16(&b_s[istage][rr][c4], ok ? B + (size_t)row * K + k : B, ok);
      }
      issued++;
      if (++ic == nchunks) ic = 0;
    }
    if (++istage == MLP_WSTAGES) istage = 0;
    mlp_cp_commit();
  };
  for (int c = 0; c < MLP_WSTAGES - 1; c++) issue();
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  float bacc[2] = {0.f, 0.f};
  int cstage = 0;
  for (int c = 0; c < nchunks; c++) {
    mlp_cp_wait<MLP_WSTAGES - 2>();
    __syncthreads();
    issue();
#pragma unroll
    for (int q = 0; q < MLP_RC / 2; q++) {
      const int rr = h * (MLP_RC / 2) + q;
      const float2 a = *reinterpret_cast<const float2*>(&a_s[cstage][rr][2 * ty]);
      const float4 b = *reinterpret_cast<const float4*>(&b_s[cstage][rr][4 * tx]);
      acc[0][0] = fmaf(a.x, b.x, acc[0][0]); acc[0][1] = fmaf(a.x, b.y, acc[0][1]);
      acc[0][2] = fmaf(a.x, b.z, acc[0][2]); acc[0][3] = fmaf(a.x, b.w, acc[0][3]);
      acc[1][0] = fmaf(a.y, b.x, acc[1][0]); acc[1][1] = fmaf(a.y, b.y, acc[1][1]);
      acc[1][2] = fmaf(a.y, b.z, acc[1][2]); acc[1][3] = fmaf(a.y, b.w, acc[1][3]);
      if (tk == 0) { bacc[0] += a.x; bacc[1] += a.y; }
    }
    if (++cstage == MLP_WSTAGES) cstage = 0;
  }
  mlp_cp_wait<0>();
  __syncthreads();
  // group 1 hands its partial sums to group 0 through the (now idle) b_s ring
  float* red = &b_s[0][0][0];  // [256][10]
  if (h == 1) {
    float* r = red + (tid & 255) * 10;
#pragma unroll
    for (int jj = 0; jj < 2; jj++) {
#pragma unroll
      for (int i = 0; i < 4; i++) r[jj * 4 + i] = acc[jj][i];
      r[8 + jj] = bacc[jj];
    }
  }
  __syncthreads();
  if (h == 0) {
    const float* r = red + tid * 10;
#pragma unroll
    for (int jj = 0; jj < 2; jj++) {
      const int j = j0 + 2 * ty + jj, k = k0 + 4 * tx;
      if (j < J) {
        if (k < K)
          *reinterpret_cast<float4*>(p.gW[l] + (size_t)j * K + k) =
              make_float4(acc[jj][0] + r[jj * 4 + 0], acc[jj][1] + r[jj * 4 + 1], acc[jj][2] + r[jj * 4 + 2],
                          acc[jj][3] + r[jj * 4 + 3]);
        if (tk == 0 && tx == 0) p.gb[l][j] = bacc[jj] + r[8 + jj];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ entry points

extern "C" int gp_kpmlp_sizes(const gp_mlp_cfg* cfg, size_t* saved_bytes, size_t* bwd_scratch_bytes) {
  MlpDims d;
  int rc = mlp_dims(cfg, &d);
  if (rc) return rc;
  const size_t rows = (size_t)(d.rows > 0 ? d.rows : 1);
  if (saved_bytes) *saved_bytes = gp_align_up(rows * ((size_t)d.din + (size_t)d.depth * d.w + 8) * sizeof(float), 256);
  if (bwd_scratch_bytes) *bwd_scratch_bytes = gp_align_up(rows * ((size_t)d.depth * d.w + 8) * sizeof(float), 256);
  return GP_OK;
}

extern "C" int gp_kpmlp_forward(const gp_mlp_cfg* cfg, const float* kp_xyz, const float* kp_feat, const float* t,
                                const float* const* weights, const float* const* biases, void* saved, float* delta_xyz,
                                float* delta_q, float* delta_o, gp_stream_t stream_) {
  MlpDims d;
  int rc = mlp_dims(cfg, &d);
  if (rc) return rc;
  if (!weights || !biases || !saved || !delta_xyz || !delta_q) return GP_ERR_BAD_ARG;
  if ((d.xf > 0 && !kp_xyz) || (d.F > 0 && !kp_feat) || (d.tf > 0 && !t)) return GP_ERR_BAD_ARG;
  if (d.rows == 0) return GP_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  MlpPtrs p = {};
  for (int l = 0; l <= d.depth; l++) {
    if (!weights[l] || !biases[l]) return GP_ERR_BAD_ARG;
    p.W[l] = weights[l];
    p.b[l] = biases[l];
  }
  float* x0 = (float*)saved;
  float* acts = x0 + (size_t)d.rows * d.din;
  float* raw_out = acts + (size_t)d.depth * d.rows * d.w;
  MlpOut out = {delta_xyz, delta_q, delta_o};
  GP_CUDA_TRY(cudaFuncSetAttribute(gp_kpmlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM_BYTES));
  const int blocks = (d.rows + MLP_ROWS - 1) / MLP_ROWS;
  GP_KERNEL("gp_kpmlp_fwd_kernel", stream,
            gp_kpmlp_fwd_kernel<<<blocks, MLP_BIG, MLP_SMEM_BYTES, stream>>>(d, p, kp_xyz, kp_feat, t, x0, acts, raw_out,
                                                                              out));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_kpmlp_backward(const gp_mlp_cfg* cfg, const float* kp_xyz, const float* t,
                                 const float* const* weights, const void* saved, const float* dL_ddelta_xyz,
                                 const float* dL_ddelta_q, const float* dL_ddelta_o, void* bwd_scratch, float* dL_dxyz, float* dL_dfeat, float* const* dL_dweights,
                                 float* const* dL_dbiases, gp_stream_t stream_) {
  MlpDims d;
  int rc = mlp_dims(cfg, &d);
  if (rc) return rc;
  if (!weights || !saved || !bwd_scratch || !dL_dweights || !dL_dbiases) return GP_ERR_BAD_ARG;
  if (d.xf > 0 && dL_dxyz && !kp_xyz) return GP_ERR_BAD_ARG;
  if (d.rows == 0) return GP_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  MlpPtrs p = {};
  for (int l = 0; l <= d.depth; l++) {
    if (!weights[l] || !dL_dweights[l] || !dL_dbiases[l]) return GP_ERR_BAD_ARG;
    if (((size_t)dL_dweights[l]) & 15) return GP_ERR_BAD_ARG;   // the weight gradients are written with 128-bit stores
    p.W[l] = weights[l];
    p.gW[l] = dL_dweights[l];
    p.gb[l] = dL_dbiases[l];
  }
  const float* x0 = (const float*)saved;
  const float* acts = x0 + (size_t)d.rows * d.din;
  const float* raw_out = acts + (size_t)d.depth * d.rows * d.w;
  float* dz = (float*)bwd_scratch;
  float* dout = dz + (size_t)d.depth * d.rows * d.w;
  MlpOutGrad gout = {dL_ddelta_xyz, dL_ddelta_q, dL_ddelta_o};
  const int blocks = (d.rows + MLP_ROWS - 1) / MLP_ROWS;
  GP_CUDA_TRY(cudaFuncSetAttribute(gp_kpmlp_bwd_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM_BYTES));
  GP_KERNEL("gp_kpmlp_bwd_data_kernel", stream,
            gp_kpmlp_bwd_data_kernel<<<blocks, MLP_BIG, MLP_SMEM_BYTES, stream>>>(d, p, kp_xyz, t, acts, raw_out, gout,
                                                                                   dout, dz, dL_dxyz, dL_dfeat));
  GP_LAUNCH_CHECK();
  MlpTiles tl = {};
  for (int l = 0; l <= d.depth; l++) {
    const int J = l < d.depth ? d.w : d.out, K = l == 0 ? d.din : d.w;
    tl.first[l + 1] = tl.first[l] + ((J + MLP_TJ - 1) / MLP_TJ) * ((K + MLP_TK - 1) / MLP_TK);
  }
  GP_KERNEL("gp_kpmlp_bwd_weights_kernel", stream,
            gp_kpmlp_bwd_weights_kernel<<<tl.first[d.depth + 1], MLP_WTHREADS, 0, stream>>>(d, p, tl, x0, acts, dz,
                                                                                            dout));
  GP_LAUNCH_CHECK();
  return GP_OK;
}
