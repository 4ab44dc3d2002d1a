/*
 * gp_b200.h -- C ABI of libgp_b200.so, the B200 (sm_100a) hot path of GaussianPrediction:
 * keypoint-driven deformation followed by depth-aware tile rasterization, forward + backward.
 *
 * The reference has no C ABI for this path: it is reached through pybind11 torch-extension
 * modules.  Each entry point below names the reference binding it stands in for (paths relative to
 * the reference root).  Conventions: plain device pointers and sizes, caller-owned outputs and
 * workspaces, an explicit cudaStream_t, no exceptions; the return value is 0 on success, a positive
 * cudaError_t, or a negative GP_ERR_* code.  Nothing here allocates device memory unless stated.
 */
#ifndef GP_B200_H
#define GP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* gp_stream_t; /* == cudaStream_t */

#define GP_OK 0
#define GP_ERR_BAD_ARG (-1)
#define GP_ERR_WORKSPACE_TOO_SMALL (-2)
#define GP_ERR_UNSUPPORTED (-3)
#define GP_ERR_LIMIT (-4)

/* Version of this ABI (bumped on any signature change). */
int gp_abi_version(void);
/* Static text for an error code returned by any entry point. */
const char* gp_error_string(int code);

/* Measurement hooks (no reference counterpart; the reference only brackets a whole iteration with CUDA events,
 * train.py:62-63,77,135).  gp_launch_count: kernels launched by this library since load.  gp_profile_enable(1)
 * makes every launch record a CUDA event pair on its stream; gp_profile_report writes one line
 * "<kernel> <launches> <total ms>" per kernel into buf (synchronises on the recorded events). */
long long gp_launch_count(void);
int gp_profile_enable(int on);
int gp_profile_report(char* buf, size_t buf_bytes);

/* FP32 FMA-pipe micro-benchmark (no reference counterpart): launches a kernel of dependent FFMA chains (8 per thread,
 * 8 blocks of 256 threads per SM, `iters` x 64 FFMA per thread) on `stream` and returns the floating-point operations
 * it performs in *flops; the caller times it with events.  bench.py uses it as the measured peak of the FP32-bound
 * kernels' roofline. */
int gp_fma_peak(int iters, float* out, double* flops, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Rasterizer.  Stands in for diff_gaussian_rasterization._C.rasterize_gaussians /
 * rasterize_gaussians_backward / mark_visible
 * (submodules/diff-gaussian-rasterization-w-depth/ext.cpp:15-19, rasterize_points.cu:35-218,
 *  cuda_rasterizer/rasterizer_impl.cu:141-439).
 * ---------------------------------------------------------------------------------------------- */

typedef struct {
  int P;              /* number of Gaussians */
  int width, height;  /* image size in pixels */
  int sh_degree;      /* active SH degree D (0..3) */
  int sh_coeffs;      /* M: coefficients stored per Gaussian (0 when colours are precomputed) */
  float tanfovx, tanfovy;
  float scale_modifier;
  int prefiltered;
  /* Optional split SH storage.  The reference keeps coefficient 0 (_features_dc [P,1,3]) and the rest
   * (_features_rest [P,15,3]) as two parameters and concatenates them every frame (scene/gaussian_model.py:155-158).
   * With sh_rest non-NULL the `shs` argument of gp_raster_preprocess / gp_raster_backward holds only coefficient 0 and
   * sh_rest the other 15 (sh_coeffs must be 16, arrays 16-byte aligned; GP_ERR_UNSUPPORTED otherwise); the backward
   * then writes dL_dsh as [P,1,3] and dL_dsh_rest as [P,15,3].  NULL: `shs` is the concatenated [P,sh_coeffs,3]. */
  const float* sh_rest;
  float* dL_dsh_rest;
  /* Non-zero: gp_raster_backward ADDS the SH gradients to what dL_dsh (and dL_dsh_rest) hold instead of overwriting
   * them -- the running sum over the samples of a gradient-accumulation step (the reference's batch loop,
   * train.py:99-124, sums them through autograd's .grad accumulation).  Needs the staged SH path (sh_coeffs = 16,
   * 16-byte aligned arrays); GP_ERR_UNSUPPORTED otherwise.  All other gradients are written as usual. */
  int dL_dsh_accumulate;
} gp_raster_cfg;

/* Header words at the start of the geometry workspace (int32 each); readable by the host. */
#define GP_HDR_NUM_RENDERED 0 /* R = sum of tiles touched (may exceed the binning capacity) */
#define GP_HDR_OVERFLOW 1     /* 1 if R > capacity: the frame was NOT rendered (outputs = the empty image), retry with more */
#define GP_HDR_NUM_POINTS 2   /* P */
#define GP_HDR_NUM_SORTED 3   /* min(R, capacity) */
#define GP_HDR_SCAN_TICKET 4  /* internal: blocks of the tiles-touched scan that have finished */
#define GP_HDR_WORDS 64

/* Workspace sizes in bytes for P Gaussians, a width x height image and room for `capacity` tile
 * instances.  (The reference grows three torch byte tensors through callbacks instead,
 * rasterize_points.cu:27-33,72-79.)  P and capacity must be below 2^30 (GP_ERR_LIMIT): the tile sort keeps 30-bit
 * counts in its look-back words. */
int gp_raster_sizes(int P, int width, int height, long long capacity, size_t* geom_bytes, size_t* binning_bytes,
                    size_t* image_bytes);

/* Stage 1 of the forward: per-Gaussian preprocess, depth sort and tile-count scan.
 * Reference: FORWARD::preprocess + cub InclusiveSum, rasterizer_impl.cu:250-280.
 * Writes radii[P] and the geometry workspace; leaves R in header word GP_HDR_NUM_RENDERED.
 * shs / colors_precomp: exactly one non-NULL.  scales+rotations / cov3D_precomp: exactly one set. */
int gp_raster_preprocess(const gp_raster_cfg* cfg, const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, const float* rotations,
                         const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                         const float* campos, void* geom_ws, int* radii, gp_stream_t stream);

/* Blocking read of R after gp_raster_preprocess (the reference's own D2H sync, rasterizer_impl.cu:283-284). */
int gp_raster_num_rendered(const void* geom_ws, gp_stream_t stream, int* num_rendered);

/* Stage 2 of the forward: duplicate-with-keys, tile sort, tile ranges, alpha/median-depth blending.
 * Reference: rasterizer_impl.cu:286-339, forward.cu:261-398.  `capacity` is the number of tile
 * instances binning_ws was sized for; if R > capacity nothing is rendered, GP_HDR_OVERFLOW is set and the outputs
 * hold the EMPTY image (background colour, depth 15, tidx -1; the backward then yields zero gradients) -- never
 * uninitialised memory.  No host sync is needed to launch this stage. */
int gp_raster_render(const gp_raster_cfg* cfg, const float* bg, const float* colors_precomp, void* geom_ws,
                     void* binning_ws, long long capacity, void* image_ws, const int* radii, float* out_color,
                     float* out_depth, int* out_tidx, gp_stream_t stream);

/* Backward of the two stages above.  Reference: Rasterizer::backward, rasterizer_impl.cu:346-439,
 * backward.cu:144-557.  `capacity` must be the value given to gp_raster_render for this frame.  All
 * dL_* outputs are fully written (no pre-zeroing required); 12 floats per Gaussian of scratch are
 * taken from `grad_ws` (size: gp_raster_backward_ws_bytes).  Outputs that do not apply may be NULL:
 * dL_dsh (colours precomputed), dL_dcolors (SH given), dL_dscales/dL_drots (cov3D precomputed),
 * dL_dcov3D (scales/rotations given). */
size_t gp_raster_backward_ws_bytes(int P);
int gp_raster_backward(const gp_raster_cfg* cfg, const float* bg, const float* means3D, const float* shs,
                       const float* colors_precomp, const float* scales, const float* rotations,
                       const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                       const float* campos, const int* radii, const void* geom_ws, const void* binning_ws,
                       long long capacity, const void* image_ws, const float* dL_dpix, void* grad_ws, float* dL_dmeans2D /*[P,3]*/,
                       float* dL_dcolors /*[P,3]*/, float* dL_dopacity /*[P,1]*/, float* dL_dmeans3D /*[P,3]*/,
                       float* dL_dcov3D /*[P,6]*/, float* dL_dsh /*[P,M,3]*/, float* dL_dscales /*[P,3]*/,
                       float* dL_drots /*[P,4]*/, gp_stream_t stream);

/* Backward for callers that cannot allocate device memory.  No counterpart in the reference (its backward
 * always allocates through torch, rasterize_points.cu:153-161).  Identical to gp_raster_backward, except that the
 * 48 B/Gaussian scratch is a fixed array inside the library (12 MiB on every GPU that loads it) instead of
 * `grad_ws`.  Meant for hosts where cudaMalloc is off limits: real-time render loops, plugins inside an engine
 * that owns the device allocator, language bindings without one.
 *   - cfg->P must be <= gp_raster_backward_static_max_gaussians() (262,144); larger -> GP_ERR_LIMIT, nothing runs.
 *   - calls no cudaMalloc / cudaFree / allocating routine.
 *   - the first call on a device looks up the array's device address; do it before any stream capture.
 *   - the array is shared by all calls on a GPU and is cleared (the part in use) at the start of each call;
 *     calls that may overlap on one GPU race on it and must be serialised by the caller. */
int gp_raster_backward_static_max_gaussians(void);
int gp_raster_backward_static(const gp_raster_cfg* cfg, const float* bg, const float* means3D, const float* shs,
                              const float* colors_precomp, const float* scales, const float* rotations,
                              const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                              const float* campos, const int* radii, const void* geom_ws, const void* binning_ws,
                              long long capacity, const void* image_ws, const float* dL_dpix,
                              float* dL_dmeans2D /*[P,3]*/, float* dL_dcolors /*[P,3]*/, float* dL_dopacity /*[P,1]*/,
                              float* dL_dmeans3D /*[P,3]*/, float* dL_dcov3D /*[P,6]*/, float* dL_dsh /*[P,M,3]*/,
                              float* dL_dscales /*[P,3]*/, float* dL_drots /*[P,4]*/, gp_stream_t stream);

/* present[i] = (view-space z > 0.2).  Reference: checkFrustum, rasterizer_impl.cu:54-66,141-153. */
int gp_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                    gp_stream_t stream);

/* Test/introspection: unpack the internal state into the reference's array shapes so parity tests can
 * compare them bit for bit with the reference's geom/binning/image buffers.  Any output may be NULL. */
int gp_raster_export_state(const gp_raster_cfg* cfg, const void* geom_ws, const void* image_ws, float* depths /*[P]*/,
                           float* means2D /*[P,2]*/, float* cov3D /*[P,6]*/, float* conic_opacity /*[P,4]*/,
                           float* rgb /*[P,3]*/, uint32_t* tiles_touched /*[P]*/, uint8_t* clamped /*[P,3]*/,
                           uint32_t* ranges /*[tiles,2]*/, float* final_T /*[H*W]*/, uint32_t* n_contrib /*[H*W]*/,
                           gp_stream_t stream);
/* Sorted instance list: point_list[R] (Gaussian index) and the reference's 64-bit keys
 * (tile << 32 | depth bits, rasterizer_impl.cu:101-104) rebuilt from tile id and depth. */
int gp_raster_export_binning(const gp_raster_cfg* cfg, const void* geom_ws, const void* binning_ws, long long capacity,
                             uint32_t* point_list, uint64_t* point_list_keys, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Keypoint search.  Stands in for frnn.frnn_grid_points as get_nearest_mask calls it
 * (scene/gaussian_model.py:110-125 -> submodules/FRNN/frnn/frnn.py:210-347 -> frnn._C.insert_points_cuda,
 * prefix_sum_cuda, counting_sort_cuda, find_nbrs_cuda; frnn/csrc/ext.cpp:10-27).
 * ---------------------------------------------------------------------------------------------- */

/* K nearest points2 (within radius r) of every points1 row, both row-major [n, D].  idxs int64 [N,K] and
 * dists [N,K] (squared), ascending, padded with -1 like the reference (find_nbrs.cu:621-622).  K <= 32. */
int gp_knn_points(int N, int Kp, int D, int K, float r, const float* points1, const float* points2, long long* idxs,
                  float* dists, gp_stream_t stream);

/* Same search on the un-concatenated model tensors: query = [xyz | motion_feature * feature_amplify]
 * (the torch.cat of scene/gaussian_model.py:115-116 is folded into the kernel; F = 0 selects the "3D" mode of
 * :112-113).  Either index output may be NULL; `order` and `hint` may be NULL.
 * `hint` (int32 [N,K], e.g. the idx32 output of the previous frame; may alias idx32) only speeds the search up: the
 * distances to any K distinct keypoints bound the K-th nearest distance, which lets a warp drop every keypoint whose
 * distance from the warp's bounding box exceeds that bound.  Invalid hint rows are ignored; results never depend on it. */
int gp_knn_keypoints(int N, int Kp, int F, int K, float r, float feature_amplify, const float* xyz,
                     const float* motion_feature, const float* kp_xyz, const float* kp_feature, const int* order,
                     const int* hint, long long* idx64, int* idx32, float* dists, gp_stream_t stream);

/* Optional acceleration hint for gp_knn_keypoints: `order` = a permutation of 0..N-1 in Morton order of xyz, so that
 * the 32 queries of a warp are spatial neighbours and the kernel's exact warp-wide early exit skips most keypoints.
 * The search result does not depend on `order` (NULL = identity); a stale order only costs speed.  (The reference's
 * FRNN path sorts the queries into grid cells for the same reason, frnn.py:137-170 -- with one cell here.) */
size_t gp_spatial_order_ws_bytes(int N);
int gp_spatial_order(int N, const float* xyz, int* order /*[N]*/, void* ws, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Keypoint blending.  Stands in for the torch ops of GaussianModel.forward's keypoint stage
 * (scene/gaussian_model.py:214-229 fill_nearest, :272-273 dense products, :285-287, :304, :314-315;
 * quat_mul utils/camera_utils.py:158-170) and of render_kpts (eval.py:140-141).
 * ---------------------------------------------------------------------------------------------- */

/* logits [N,2K] (translation weights then rotation weights, softmax-ed inside), idx (int32 or int64, [N,K]),
 * per-keypoint deltas [Kp,3] / [Kp,4], canonical xyz [N,3], raw rotation [N,4], log-scale [N,3], opacity
 * logit [N,1].  Outputs xyz_t [N,3], q_t [N,4] (normalised), s_t [N,3] = exp, o_t [N,1] = sigmoid (the last
 * two optional).  K <= 16, Kp <= 3200.  rot, q_t (and g_q_t, g_rot in the backward) are moved as 128-bit rows: 16-byte
 * aligned, GP_ERR_BAD_ARG otherwise. */
int gp_blend_forward(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                     const float* kp_delta_xyz, const float* kp_delta_q, const float* xyz, const float* rot,
                     const float* scaling, const float* opacity, float* xyz_t, float* q_t, float* s_t, float* o_t,
                     gp_stream_t stream);

/* Backward of gp_blend_forward (what torch autograd computes through the reference's dense ops).
 * g_kp_delta_* are zeroed here and accumulated; other outputs are written; any g_* input may be NULL (= 0). */
int gp_blend_backward(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                      const float* kp_delta_xyz, const float* kp_delta_q, const float* rot, const float* s_t,
                      const float* o_t, const float* g_xyz_t, const float* g_q_t, const float* g_s_t,
                      const float* g_o_t, float* g_logits, float* g_xyz, float* g_rot, float* g_scaling,
                      float* g_opacity, float* g_kp_delta_xyz, float* g_kp_delta_q, gp_stream_t stream);

/* Same, but the per-Gaussian outputs (g_logits, g_xyz, g_rot, g_scaling, g_opacity) are ADDED to what the arrays hold:
 * the running sum over the samples of a gradient-accumulation step (train.py:99-124), with no scratch copy and no
 * separate add pass.  The keypoint outputs are zeroed and summed as in gp_blend_backward. */
int gp_blend_backward_accumulate(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                                 const float* kp_delta_xyz, const float* kp_delta_q, const float* rot, const float* s_t,
                                 const float* o_t, const float* g_xyz_t, const float* g_q_t, const float* g_s_t,
                                 const float* g_o_t, float* g_logits, float* g_xyz, float* g_rot, float* g_scaling,
                                 float* g_opacity, float* g_kp_delta_xyz, float* g_kp_delta_q, gp_stream_t stream);

/* Dense [N,Kp] weight matrices for API parity: weights_xyz / weights_r (return_weights=True,
 * scene/gaussian_model.py:302-303) and weights_sum (:263).  Any output may be NULL. */
int gp_blend_dense_weights(int N, int Kp, int K, const float* logits, const int* idx32, const long long* idx64,
                           float* w_xyz, float* w_r, float* w_sum, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Initial scale estimate.  Stands in for simple_knn._C.distCUDA2 (submodules/simple-knn/ext.cpp:15-17,
 * spatial.cu:15-26, simple_knn.cu:185-221): mean of the 3 smallest squared distances to other points.
 * ---------------------------------------------------------------------------------------------- */
size_t gp_dist2_ws_bytes(int P);
int gp_dist2_knn3(int P, const float* points /*[P,3]*/, float* mean_dist2 /*[P]*/, void* ws, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Image loss (the step right after the rasterizer).  Stands in for the torch ops of
 * utils/loss_utils.py:54-100 (l1_loss, ssim/_ssim with the 11x11 sigma-1.5 window, zero padding) as train.py:105-108
 * combines them, and for their autograd backward.  Images are [C,H,W] float32 planes (any leading dims folded into C).
 * ---------------------------------------------------------------------------------------------- */

/* sums[0] = mean |img - gt|, sums[1] = mean SSIM map (device floats).  dmaps: scratch of 3*C*H*W floats that the
 * backward reads (NULL if no backward will follow); ws: gp_image_loss_ws_bytes bytes. */
size_t gp_image_loss_ws_bytes(int C, int H, int W);
int gp_image_loss_forward(int C, int H, int W, const float* img, const float* gt, float* dmaps, void* ws, float* sums,
                          gp_stream_t stream);
/* grad_img = (*grad_scale) * d/d img [ w_l1 * mean|img-gt| + w_ssim * mean SSIM ];  grad_scale: device scalar or NULL (=1).
 * train.py:108 is w_l1 = 1 - lambda_dssim, w_ssim = -lambda_dssim. */
int gp_image_loss_backward(int C, int H, int W, const float* img, const float* gt, const float* dmaps, float w_l1,
                           float w_ssim, const float* grad_scale, float* grad_img, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Keypoint-motion MLP, fused.  Stands in for the torch ops of scene/gaussian_model.py:180-189 (get_motion_delta:
 * repeat t, cat([feature, xyz_embed, t]), df_model) with scene/deformable_field.py:63-72 (positional_encoding of the
 * keypoint positions and of t) and :74-127 (Deformable_Field: depth x (Linear+ReLU), then Linear), and for their
 * autograd backward, at the keypoint call site (:255-265).  Input row = [feature(F) | PE(xyz)(6*xyz_freq) |
 * PE(t)(2*time_freq)].  weights[l] / biases[l], l = 0..depth: device pointers in nn.Linear layout ([out][in] row-major);
 * the arrays of pointers themselves live on the host.  Limits (GP_ERR_LIMIT otherwise): width and input row length
 * multiples of 4 and <= 256, depth <= 8, out_dim 7 or 8.
 * ---------------------------------------------------------------------------------------------- */
typedef struct gp_mlp_cfg {
  int rows;        /* number of keypoints */
  int feature_dim; /* F */
  int xyz_freq;    /* frequencies of the position encoding (0: no such columns) */
  int time_freq;   /* frequencies of the time encoding (0: no such columns) */
  int depth;       /* hidden Linear+ReLU layers */
  int width;       /* hidden width */
  int out_dim;     /* columns of the last Linear: 7 = (delta_xyz, delta_q); 8 with step opacity (+ delta_o) */
  int normalize_q; /* 1: delta_q = normalize(delta_q) as with --norm_rotation (scene/gaussian_model.py:266-267) */
} gp_mlp_cfg;

/* saved: activations the backward needs (forward writes, backward reads); bwd_scratch holds no state between calls */
int gp_kpmlp_sizes(const gp_mlp_cfg* cfg, size_t* saved_bytes, size_t* bwd_scratch_bytes);
/* (delta_xyz [rows,3], delta_q [rows,4], delta_o [rows,1] or NULL) = the three tensors get_motion_delta returns
 * (scene/gaussian_model.py:189), as separate contiguous arrays.  t: device scalar. */
int gp_kpmlp_forward(const gp_mlp_cfg* cfg, const float* kp_xyz /*[rows,3]*/, const float* kp_feat /*[rows,F]*/,
                     const float* t, const float* const* weights, const float* const* biases, void* saved,
                     float* delta_xyz, float* delta_q, float* delta_o, gp_stream_t stream);
/* Every requested gradient is fully written (no accumulation).  dL_dxyz / dL_dfeat may be NULL; a NULL incoming
 * gradient (dL_ddelta_*) counts as zeros. */
int gp_kpmlp_backward(const gp_mlp_cfg* cfg, const float* kp_xyz, const float* t, const float* const* weights,
                      const void* saved, const float* dL_ddelta_xyz, const float* dL_ddelta_q,
                      const float* dL_ddelta_o, void* bwd_scratch,
                      float* dL_dxyz /*[rows,3]*/, float* dL_dfeat /*[rows,F]*/, float* const* dL_dweights,
                      float* const* dL_dbiases, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * After the backward.  gp_adam_step stands in for `gaussians.optimizer.step()` (train.py:192-194):
 * torch.optim.Adam without weight decay / amsgrad as scene/gaussian_model.py:452 builds it, one launch for every tensor
 * of the step.  `tensors` is a HOST array; per tensor its own lr and its 1-based step count (the bias corrections are
 * computed from it in double on the host, as torch does).  exp_avg / exp_avg_sq are the optimizer state of that name.
 * gp_densification_stats stands in for train.py:166-167 / scene/gaussian_model.py:756-760: for every Gaussian with
 * radii > 0: max_radii2D = max(max_radii2D, radii); g = |viewspace_grad.xy|; xyz_gradient_accum += g; denom += 1;
 * xyz_gradient_accum_max = max(., g).  max_radii2D and xyz_gradient_accum_max may be NULL.
 * ---------------------------------------------------------------------------------------------- */
typedef struct gp_adam_tensor {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  long long n; /* elements */
  float lr;
  int step;    /* 1 for the first update */
} gp_adam_tensor;
int gp_adam_step(int count, const gp_adam_tensor* tensors, double beta1, double beta2, double eps, gp_stream_t stream);
int gp_densification_stats(int P, const float* viewspace_grad /*[P,3]*/, const int* radii /*[P]*/,
                           float* max_radii2D /*[P]*/, float* xyz_gradient_accum /*[P,1]*/, float* denom /*[P,1]*/,
                           float* xyz_gradient_accum_max /*[P,1]*/, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Multiresolution hash-grid encoding of the blend-weight network.  Stands in for the encoding half of the tinycudann
 * `NetworkWithInputEncoding` the reference builds at scene/gaussian_model.py:370-392 and calls every keypoint-stage
 * frame at :257 (`weights_model(get_xyz.detach())`): Grid/Hash, linear interpolation; the reference's values are
 * n_levels 16, base_resolution 16, per_level_scale exp(ln(2048/16)/15), log2_hashmap_size 19, 4 features per level
 * (fixed here).  tinycudann is not part of the reference tree: this follows the published algorithm (parity unpinned,
 * oracle/hashgrid_oracle.py).  table: float[entries][4], entries = gp_hashgrid_entries(...); enc: float[N][4 * n_levels];
 * order: optional permutation of the points (e.g. gp_spatial_order) that only changes which thread serves which point.
 * gp_hashgrid_backward ADDS dL/dtable into d_table (zero it first); the points get no gradient (the reference detaches). */
long long gp_hashgrid_entries(int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size);
int gp_hashgrid_forward(int N, int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size,
                        const float* xyz, const int* order, const float* table, float* enc, gp_stream_t stream);
int gp_hashgrid_backward(int N, int n_levels, int base_resolution, float per_level_scale, int log2_hashmap_size,
                         const float* xyz, const int* order, const float* d_enc, float* d_table, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Dense layer on the tensor cores: C[M,Nout] = act(A[M,K] W[Nout,K]^T + bias), fp32 in and out, error-compensated
 * TF32 (three tcgen05.mma per K step, fp32 accumulation in tensor memory; ~1e-6 relative, csrc/linear_tc.cu).
 * Stands in for the cuBLAS SGEMM + elementwise kernels behind the per-Gaussian `Deformable_Field` calls of
 * GaussianModel.forward -- first-stage deformation (scene/gaussian_model.py:240-244,265) and step-opacity pass
 * (:291-298; scene/deformable_field.py:112-127): [N,104] -> 4 x (Linear 256 + ReLU) -> Linear 8 over all N Gaussians.
 * W is in nn.Linear layout.  relu != 0 applies max(0, .).  mask (optional, [M,Nout], leading dimension ldm) zeroes C
 * where mask <= 0: with W passed transposed the same call is the backward of a layer through the ReLU below it,
 * dZ_below = (dZ W) * [h > 0].  Nout <= 256; lda >= K, ldc >= Nout (elements). */
int gp_linear_forward(int M, int K, int Nout, const float* A, int lda, const float* mask, int ldm, const float* W,
                      const float* bias, int relu, float* C, int ldc, gp_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Gradient exchange of the data-parallel frame loop.  No reference counterpart: the reference is single-GPU; the
 * seam is its --batch accumulation (train.py:113-133), which SUMS the gradients of the step's samples.
 *
 * gp_allreduce_sum_f32 sums `n_floats` floats at `byte_offset` of a SYMMETRIC allocation over `world` ranks, in
 * place, in one kernel on `stream` (csrc/exchange.cu): peer_bases[q] is rank q's copy of the allocation as mapped into
 * this process (host array of `world` device addresses), multicast_base the NVSwitch multicast mapping of the same
 * allocation or NULL (then plain peer loads/stores are used).  The allocation also holds gp_allreduce_flag_bytes(blocks)
 * bytes of zero-initialised flags at `flags_byte_offset`; calls that may overlap in time must use different flag
 * regions (calls issued one after the other on one stream may share one).  Every rank must issue the same call (same sizes, same `blocks`, <= the number of co-resident CTAs).
 * byte_offset, flags_byte_offset: multiples of 16; n_floats: multiple of 4; world <= 16. */
size_t gp_allreduce_flag_bytes(int blocks);
int gp_allreduce_sum_f32(int rank, int world, void* multicast_base, void* const* peer_bases, size_t byte_offset,
                         size_t n_floats, size_t flags_byte_offset, int blocks, gp_stream_t stream);
/* Same protocol, element-wise MAX of int32: the radii of the step's samples ("radii = max", "visibility = or" =
 * radii > 0, train.py:122-123).  byte_offset: multiple of 4. */
int gp_allreduce_max_s32(int rank, int world, void* multicast_base, void* const* peer_bases, size_t byte_offset,
                         size_t n_ints, size_t flags_byte_offset, int blocks, gp_stream_t stream);
/* Up to 8 ranges of the same allocation in ONE kernel that stays resident and works through them in order (ops[s]: 0 =
 * sum of counts[s] floats, 1 = max of counts[s] int32).  Launched when the step's last gradients are final, it holds
 * its few CTAs while the next step's kernels fill the rest of the GPU -- an exchange launched later would wait behind
 * kernels with persistent blocks.  After range s is complete on every rank the kernel stores `call_id` (ids must grow
 * from call to call) to sync_words[8 + s]; sync_words: gp_exchange_sync_words() zero-initialised device words owned by
 * the caller.  gp_stream_wait_word makes `stream` wait (a one-thread kernel) until *word >= value: put it in front of
 * the first reader of the range. */
int gp_exchange_sync_words(void);
int gp_exchange_segments(int rank, int world, void* multicast_base, void* const* peer_bases, size_t flags_byte_offset,
                         int blocks, int n_segments, const size_t* byte_offsets, const size_t* counts, const int* ops,
                         unsigned* sync_words, unsigned call_id, gp_stream_t stream);
int gp_stream_wait_word(const unsigned* word, unsigned value, gp_stream_t stream);

/* Two ranks: the pieces of an exchange whose bytes are moved by the copy engines instead of SMs.
 * gp_peer_barrier_launch: a one-warp kernel on `stream` that meets the same call of every peer on flag slot `slot` of
 * the symmetric allocation (slots are the per-CTA flag groups of gp_allreduce_flag_bytes; calls on one slot must come
 * in the same order on every rank and never overlap in time).  gp_copy_async: cudaMemcpyAsync device-to-device (a
 * peer's mapped buffer as source pulls it over NVLink).  gp_add_inplace_f32 / gp_max_inplace_s32: dst op= src. */
int gp_peer_barrier_launch(int rank, int world, void* const* peer_bases, size_t flags_byte_offset, int slot,
                           gp_stream_t stream);
int gp_copy_async(void* dst, const void* src, size_t bytes, gp_stream_t stream);
int gp_add_inplace_f32(float* dst, const float* src, size_t n_floats, gp_stream_t stream);
int gp_max_inplace_s32(int* dst, const int* src, size_t n, gp_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GP_B200_H */
