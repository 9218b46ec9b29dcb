/*
 * splat_b200.h -- C ABI of the B200-native (sm_100a) differentiable Gaussian-splat rasterizer.
 *
 * This is the drop-in boundary.  Every entry point takes plain device
 * pointers, sizes and a cudaStream_t (passed as void*); there are no torch types.  The caller
 * owns every buffer; the library never allocates or frees device memory (the only exception
 * is the optional host-buffer session at the bottom, which exists for hosts that have no
 * device allocator of their own).
 *
 * Reference interfaces replaced (paths relative to /root/reference; HX = hessian-diff-gaussian-
 * rasterization-w-depth, DG = diff-gaussian-rasterization-w-depth.git):
 *   splat_forward_geometry + splat_forward_render  <-  CudaRasterizer::Rasterizer::forward
 *        (HX/cuda_rasterizer/rasterizer.h:36-60, rasterizer_impl.cu:198-339), bound to Python
 *        as _C.rasterize_gaussians (HX/ext.cpp:15, HX/rasterize_points.cu:35-115)
 *   splat_backward                                 <-  CudaRasterizer::Rasterizer::backward
 *        (HX/cuda_rasterizer/rasterizer.h:62-92, rasterizer_impl.cu:343-433; DG variant
 *        DG/cuda_rasterizer/rasterizer_impl.cu:343-435), bound as
 *        _C.rasterize_gaussians_backward (HX/ext.cpp:16, HX/rasterize_points.cu:117-196)
 *   splat_mark_visible                             <-  CudaRasterizer::Rasterizer::markVisible
 *        (HX/cuda_rasterizer/rasterizer.h:28-34), bound as _C.mark_visible (HX/ext.cpp:17)
 *   splat_*_state_bytes                            <-  required<GeometryState/ImageState/
 *        BinningState>() (HX/cuda_rasterizer/rasterizer_impl.h:66-72)
 *
 * Conventions kept from the reference: matrices are passed "transposed" (element (r,c) of the
 * math matrix is m[c*4+r], HX/cuda_rasterizer/auxiliary.h:58-77); quaternions are (w,x,y,z) and
 * used un-normalised; colours/outputs are planar [3,H,W]; radii are int32, 0 == culled.
 */
#ifndef SPLAT_B200_H_
#define SPLAT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPLAT_ABI_VERSION 7   /* v6 (round 2): splat_peer_reduce_adam, splat_peer_adam_shard_floats, splat_set_gather, splat_slam_track_loss_noalloc, splat_debug_exp_mismatches; v7: splat_peer_set_tuning */

/* Error codes.  Positive values >= 1000 are 1000 + cudaError_t of a failed runtime call. */
enum {
  SPLAT_OK = 0,
  SPLAT_ERR_BAD_ARG = 1,          /* null/mis-sized argument; see splat_last_error()            */
  SPLAT_ERR_SCRATCH_TOO_SMALL = 2,/* binning capacity exceeded (splat_forward only)             */
  SPLAT_ERR_UNSUPPORTED = 3,      /* e.g. sh_degree > 3                                         */
  SPLAT_ERR_CUDA_BASE = 1000
};

/* Camera / raster settings.  Mirrors GaussianRasterizationSettings
 * (HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py:140-151) field by field, plus the
 * near-plane constant that distinguishes the two reference builds
 * (HX/cuda_rasterizer/auxiliary.h:154 uses 0.001f, DG/cuda_rasterizer/auxiliary.h:154 uses 0.2f). */
typedef struct SplatCamera {
  int32_t image_height;
  int32_t image_width;
  float tanfovx;
  float tanfovy;
  const float* bg;          /* device, 3 floats                                   */
  float scale_modifier;
  const float* viewmatrix;  /* device, 16 floats, transposed layout               */
  const float* projmatrix;  /* device, 16 floats, transposed layout               */
  int32_t sh_degree;        /* 0..3 (only read when shs != NULL)                  */
  const float* campos;      /* device, 3 floats (only read when shs != NULL)      */
  int32_t prefiltered;      /* kept for API parity; a culled point is never fatal */
  float near_plane;         /* 0.2f == DG semantics, 0.001f == HX semantics       */
  int32_t depth_sil;        /* 0: 3 feature channels (the reference's pass).  1: fused RGB + depth + silhouette:
                             * out_color / dL_dout_color have SIX planes, planes 3..5 render the features
                             * (z, 1, z*z), z = view-space depth -- what scripts/splatam.py:257-259 obtains from a
                             * second rasterizer pass with colors_precomp = get_depth_and_silhouette()
                             * (utils/slam_helpers.py:124-139); their gradient reaches dL_dmeans3D through z.
                             * power must be 1. */
} SplatCamera;

/* ---- scratch sizing (all sizes in bytes, all buffers must be 256-byte aligned) ------------- */
size_t splat_geom_state_bytes(int32_t P);
size_t splat_image_state_bytes(int32_t width, int32_t height);
size_t splat_binning_state_bytes(int32_t width, int32_t height, int64_t R_capacity);

/* ---- forward, phase 1: per-Gaussian geometry + per-tile counts + tile ranges ----------------
 * Replaces preprocessCUDA + InclusiveSum + the blocking num_rendered read
 * (HX/cuda_rasterizer/rasterizer_impl.cu:248-282).  Fully asynchronous.  On return (in stream
 * order) geom_state holds one 48-byte record per visible Gaussian (pixel mean, conic, opacity, depth,
 * colour -- SH-evaluated if shs is given), image_state holds the bit-exact tile ranges, radii[P] is
 * written, and *num_rendered_dev holds R (a uint32 in device memory OR in mapped pinned host memory: the
 * Python host passes a pinned word so that reading R back needs no copy-engine transfer).
 * Exactly one of (scales && rotations) or cov3D_precomp must be non-NULL; exactly one of
 * colors_precomp or shs must be non-NULL (same rule as HX/.../__init__.py:174-178). */
int splat_forward_geometry(const SplatCamera* cam, int32_t P,
                           const float* means3D,        /* [P,3] */
                           const float* scales,         /* [P,3] or NULL */
                           const float* rotations,      /* [P,4] or NULL */
                           const float* opacities,      /* [P]   */
                           const float* cov3D_precomp,  /* [P,6] or NULL */
                           const float* shs,            /* [P,M,3] or NULL */
                           int32_t M,
                           const float* colors_precomp, /* [P,3] or NULL */
                           void* geom_state, void* image_state,
                           int32_t* radii,              /* [P] out */
                           uint32_t* num_rendered_dev,  /* device scalar out */
                           void* stream);

/* ---- forward, phase 2: scatter instances to tiles, per-tile depth sort, blend ---------------
 * Replaces duplicateWithKeys + DeviceRadixSort + identifyTileRanges + renderCUDA
 * (HX/cuda_rasterizer/rasterizer_impl.cu:290-337).  The sort also writes the sorted 48-byte instance
 * records that the forward and backward blend kernels stream with the TMA unit.  R_capacity is the number of (tile,Gaussian)
 * instances binning_state was sized for.  If the device-side R exceeds it, every kernel of this
 * phase exits without touching memory out of bounds and the outputs are undefined; the caller
 * learns R from *num_rendered_dev (phase 1) and simply calls phase 2 again with a larger state.
 * out_depth may be NULL. */
int splat_forward_render(const SplatCamera* cam, int32_t P,
                         const float* colors_precomp,   /* unused (colours were captured by phase 1); kept for symmetry */
                         const int32_t* radii,          /* [P] as written by phase 1 */
                         const void* geom_state, void* image_state, void* binning_state,
                         int64_t R_capacity,
                         float* out_color,              /* [3,H,W] */
                         float* out_depth,              /* [1,H,W] median depth, default 15.0 */
                         void* stream);

/* ---- backward -------------------------------------------------------------------------------
 * power == 1 with fused == 0 reproduces DG (BACKWARD::render + BACKWARD::preprocess);
 * fused != 0 reproduces HX's renderCUDAFused semantics: every per-(pixel,Gaussian) gradient
 * term is raised to `power` before accumulation (HX/cuda_rasterizer/backward.cu:1093-1137);
 * power == 2 is the fork's diagonal Gauss-Newton ("Hessian") accumulation.
 * `R` is the R_capacity that was passed to splat_forward_render (it fixes the state layout).
 * Every dL_* output that is passed is fully written by the call (no pre-zeroing needed).  dL_dconic is a
 * [P,4] output (x,y,_,w as in the reference's [P,2,2]).  dL_dsh may be NULL when shs is.
 * With power == 1 and no SHs every gradient array except dL_dmeans3D is OPTIONAL (NULL = not needed): it is then
 * neither computed nor written, and with the fused depth/silhouette render a call without dL_dcolors and
 * dL_dopacity skips those sums in the blend backward altogether (pose-only tracking, where the reference computes
 * them for an optimizer with zero learning rates, configs/replica/splatam.py:71-79).
 * `scratch` is a 16-byte aligned device buffer of at least splat_backward_scratch_bytes(P, power)
 * bytes (the packed per-Gaussian accumulator rows the blend-backward kernel reduces into); the
 * call zeroes it itself.  colors_precomp is accepted for signature parity with the reference
 * (HX/rasterize_points.h:43-65) but the colours of the forward call are used.
 * out_color (optional) only CONDITIONS the power-1 blend backward: dL/dalpha of a pair is invariant under a constant
 * shift of all feature vectors of a pixel (sum of blend weights + final T == 1), so the kernel measures features
 * relative to the pixel's rendered value, which keeps its one running suffix sum small where neighbouring splats
 * have similar colours.  Any finite values give the same result up to float rounding; NULL means no shift. */
size_t splat_backward_scratch_bytes(int32_t P, int32_t power);
int splat_backward(const SplatCamera* cam, int32_t P, int64_t R,
                   const float* means3D, const float* scales, const float* rotations,
                   const float* cov3D_precomp, const float* shs, int32_t M,
                   const float* colors_precomp, const int32_t* radii,
                   const void* geom_state, const void* image_state, const void* binning_state,
                   const float* dL_dout_color,          /* [3,H,W] ([6,H,W] with depth_sil) */
                   const float* out_color,              /* the forward's rendered image, same shape, or NULL (see below) */
                   int32_t power, int32_t fused,
                   float* dL_dmeans3D,  /* [P,3] */
                   float* dL_dmeans2D,  /* [P,3] (z always 0) */
                   float* dL_dcolors,   /* [P,3] */
                   float* dL_dopacity,  /* [P]   */
                   float* dL_dscales,   /* [P,3] */
                   float* dL_drotations,/* [P,4] */
                   float* dL_dcov3D,    /* [P,6] */
                   float* dL_dconic,    /* [P,4] */
                   float* dL_dsh,       /* [P,M,3] or NULL */
                   void* scratch, size_t scratch_bytes,
                   void* stream);

/* ---- markVisible (HX/cuda_rasterizer/rasterizer_impl.cu:54-66,141-153) ---------------------- */
int splat_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                       const float* projmatrix, float near_plane, uint8_t* present, void* stream);

/* ---- SplaTAM glue around the rasterizer (ABI v4) -------------------------
 * The reference runs these steps as ~45 small PyTorch kernels + autograd nodes per optimisation
 * iteration; here each is one HBM-bound kernel (plus a one-block finalize).  All reductions are
 * two-stage and deterministic.
 *
 * splat_rendervar_forward  <-  transform_to_frame + transformed_params2rendervar
 *      (utils/slam_helpers.py:253-304, :123-139; build_rotation utils/slam_external.py:25-42):
 *      out_means3D = R(q) p + t with q = normalize(cam_unnorm_rots[0,:,time_idx]), t = cam_trans[0,:,time_idx];
 *      out_rotations = normalize(unnorm_rotations)                           (scale_dim == 1, isotropic)
 *                    = normalize(quat_mult(q, normalize(unnorm_rotations)))  (scale_dim == 3);
 *      out_opacities = sigmoid(logit_opacities); out_scales[P,3] = exp(log_scales) (tiled when scale_dim == 1).
 *      colors_precomp is params['rgb_colors'] itself (no kernel).  cam_* are the full [1,4,F] / [1,3,F] tensors.
 * splat_rendervar_backward <-  the autograd graph of the above.  dL_d{means3D_cam, rotations, opacities, scales}
 *      are the rasterizer's gradients (any may be NULL == zero).  gaussians_grad / camera_grad are
 *      transform_to_frame's flags: with gaussians_grad the four Gaussian-side gradients are written
 *      ([P,3], [P,4], [P,1], [P,scale_dim]); with camera_grad the pose gradient -- the sum over ALL Gaussians of
 *      dL/dp' and dL/dp' (x) p, chained through build_rotation and both normalisations -- is written into column
 *      time_idx of dL_dcam_unnorm_rots [1,4,F] / dL_dcam_trans [1,3,F] (other columns untouched).
 *      NOTE the reference detaches means3D AND unnorm_rotations when gaussians_grad is false, but opacity,
 *      scale and colour gradients still flow (slam_helpers.py:282-288 vs :133-136): dL_dlogit_opacities and
 *      dL_dlog_scales are written whenever their pointers are non-NULL. */
int splat_rendervar_forward(int32_t P, int32_t scale_dim,
                            const float* means3D, const float* unnorm_rotations, const float* logit_opacities,
                            const float* log_scales, const float* cam_unnorm_rots, const float* cam_trans,
                            int32_t n_frames, int32_t time_idx,
                            float* out_means3D, float* out_rotations, float* out_opacities, float* out_scales,
                            void* stream);
size_t splat_rendervar_scratch_bytes(int32_t P);
int splat_rendervar_backward(int32_t P, int32_t scale_dim,
                             const float* means3D, const float* unnorm_rotations, const float* logit_opacities,
                             const float* log_scales, const float* cam_unnorm_rots, const float* cam_trans,
                             int32_t n_frames, int32_t time_idx,
                             const float* dL_dmeans3D_cam, const float* dL_drotations, const float* dL_dopacities,
                             const float* dL_dscales, int32_t gaussians_grad, int32_t camera_grad,
                             float* dL_dmeans3D, float* dL_dunnorm_rotations, float* dL_dlogit_opacities,
                             float* dL_dlog_scales, float* dL_dcam_unnorm_rots, float* dL_dcam_trans,
                             void* scratch, size_t scratch_bytes, void* stream);

/* splat_slam_loss  <-  the image-space part of get_loss (scripts/splatam.py:262-296,342-346) with
 *      use_l1 = True and ignore_outlier_depth_loss = False (every shipped config): loss AND its gradient
 *      w.r.t. the two rendered images in one call.
 *      im [3,H,W] / depth_sil [3,H,W] = (depth, silhouette, depth^2) are the two Renderer outputs (or planes
 *      0-2 / 3-5 of the fused render).  mask = gt_depth > 0 & !isnan(depth) & !isnan(depth^2 - depth*depth)
 *      [& silhouette > sil_thres when tracking && use_sil_for_loss].
 *      tracking != 0: losses = (w_im * sum_mask |gt_im - im|, w_depth * sum_mask |gt_depth - depth|, total, count)
 *      tracking == 0: losses = (w_im * (0.8 mean|im - gt_im| + 0.2 (1 - SSIM_11x11(im, gt_im))),
 *                               w_depth * mean_mask |gt_depth - depth|, total, count)
 *      (calc_ssim, utils/slam_external.py:54-97).  `losses` is a device float[4].  dL_dim / dL_ddepth_sil are fully
 *      written (planes 1 and 2 of dL_ddepth_sil are zero: silhouette only enters the mask, uncertainty is detached). */
size_t splat_slam_loss_scratch_bytes(int32_t width, int32_t height, int32_t tracking);
int splat_slam_loss(int32_t width, int32_t height, int32_t tracking, int32_t use_sil_for_loss, float sil_thres,
                    float w_im, float w_depth,
                    const float* im, const float* depth_sil, const float* gt_im, const float* gt_depth,
                    float* losses, float* dL_dim, float* dL_ddepth_sil,
                    void* scratch, size_t scratch_bytes, void* stream);

/* splat_slam_track_loss_noalloc: NO COUNTERPART IN THE REFERENCE -- added for
 *      callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses that
 *      forbid cudaMalloc).  Same kernels, arguments and results as splat_slam_loss(tracking = 1, ...), but the scratch
 *      (per-block partial sums + stats, 19,200 bytes) is a `static __device__` array inside the library
 *      (csrc/noalloc.cu) instead of a caller buffer: the call allocates and frees nothing.
 *      Limit: width * height <= SPLAT_NOALLOC_MAX_PIXELS; larger images return SPLAT_ERR_SCRATCH_TOO_SMALL (use
 *      splat_slam_loss with your own scratch).  Every GPU that loads the library holds the array until exit.  All calls
 *      share it: calls that OVERLAP on the GPU (two streams, two host threads) race on it -- serialise them.  Nothing in
 *      it needs to start at zero (each call writes every row it reads), so back-to-back calls need no clearing.  The
 *      first call on each device looks the array's device address up (cudaGetSymbolAddress, no allocation). */
#define SPLAT_NOALLOC_MAX_PIXELS (4096 * 4096)
int splat_slam_track_loss_noalloc(int32_t width, int32_t height, int32_t use_sil_for_loss, float sil_thres,
                                  float w_im, float w_depth,
                                  const float* im, const float* depth_sil, const float* gt_im, const float* gt_depth,
                                  float* losses, float* dL_dim, float* dL_ddepth_sil, void* stream);

/* splat_adam_step  <-  the per-iteration optimizer.step() of SplaTAM (scripts/splatam.py:724-725, :889-890):
 *      torch.optim.Adam over one param group per tensor (scripts/splatam.py:166-172; betas (0.9, 0.999), eps 1e-8 while
 *      tracking / 1e-15 while mapping, no weight decay, no amsgrad), all tensors in ONE launch.  Arithmetic as torch/optim/adam.py::_single_tensor_adam:
 *        exp_avg += (1 - beta1) (grad - exp_avg);  exp_avg_sq = beta2 exp_avg_sq + (1 - beta2) grad^2;
 *        param  -= (lr / bias_correction1) * exp_avg / (sqrt(exp_avg_sq) / bias_correction2_sqrt + eps)
 *      with bias_correction1 = 1 - beta1^step and bias_correction2_sqrt = sqrt(1 - beta2^step) computed by the
 *      library from each tensor's `step` (a tensor without a gradient this step is simply left out, as torch does).
 *      `tensors` is a HOST array of `count` <= 16 descriptors (copied into the launch). */
typedef struct SplatAdamTensor {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;   /* device, n floats each */
  int64_t n;
  float lr;
  int32_t step;            /* number of THIS update for this tensor (>= 1): bias corrections use beta^step */
} SplatAdamTensor;
/* replay_counter == NULL: bias corrections are formed on the host from `step`.  replay_counter != NULL (a device
 * int32, zero-initialised by the caller) makes the call CUDA-graph capturable: a one-block prologue kernel forms the
 * bias corrections for step + *replay_counter on the device (into replay_scratch, >= 32 device floats) and then
 * increments the counter, so every replay of a captured step advances the optimizer by one step. */
int splat_adam_step(const SplatAdamTensor* tensors, int32_t count, double beta1, double beta2, double eps,
                    int32_t* replay_counter, float* replay_scratch, void* stream);

/* Capture guard.  Inside a CUDA graph the host cannot read num_rendered to re-size binning_state (phase 2 then runs
 * with a capacity fixed at capture time and simply exits if R exceeds it).  This one-thread kernel ORs
 * (R > R_capacity) into flags[0] and keeps max(R) in flags[1] (device or mapped pinned uint32[2]) so that the host can
 * tell, after any number of replays, whether every one of them rendered. */
int splat_forward_guard(const void* image_state, int32_t width, int32_t height, int64_t R_capacity,
                        uint32_t* flags, void* stream);

/* ---- per-iteration bookkeeping ---------------------------------------------------------------------------
 * splat_seen_update: `seen = radius > 0; max_2D_radius[seen] = max(radius[seen], max_2D_radius[seen])`
 * (scripts/splatam.py:348-350) in one pass over the radii the forward wrote; seen (P bytes, torch.bool) may be NULL.
 * splat_best_pose_update: the tracking loop's candidate pose (scripts/splatam.py:726-731) on the device:
 * if (*loss < *min_loss) { *min_loss = *loss; best_rot[4] / best_tran[3] = column time_idx of cam_unnorm_rots
 * [1,4,n_frames] / cam_trans [1,3,n_frames] } -- no host read of the loss, capturable in a CUDA graph. */
int splat_seen_update(int32_t P, const int32_t* radii, float* max_2D_radius, uint8_t* seen, void* stream);
/* x[0..n) *= *scale (device scalar), skipped on the device when *scale == 1: applies the incoming gradient of the loss
 * to the gradient image the loss kernel already formed (loss.backward() passes exactly 1). */
int splat_scale_inplace(float* x, int64_t n, const float* scale, void* stream);
int splat_best_pose_update(const float* loss, float* min_loss, const float* cam_unnorm_rots, const float* cam_trans,
                           int32_t n_frames, int32_t time_idx, float* best_rot, float* best_tran, void* stream);

/* ---- per-frame map surgery -----------------------------------------------------
 * Replaces the torch boolean-index / torch.cat surgery of utils/slam_external.py:123-195 (remove_points,
 * cat_params_to_optimizer, prune_gaussians) and scripts/splatam.py:73-123,354-426 (get_pointcloud,
 * initialize_new_params, add_new_gaussians).  All pointers are device pointers unless noted.
 *
 * splat_prune_flags: keep[i] = !( sigmoid(logit_opacities[i]) < opacity_threshold
 *                                 || (big_threshold >= 0 && max_j exp(log_scales[i,j]) > big_threshold) )
 *   -- the `to_remove` mask of prune_gaussians (:180-186), big_threshold = 0.1 * scene_radius or < 0 for "not yet". */
int splat_prune_flags(int32_t P, int32_t scale_dim, const float* logit_opacities, const float* log_scales,
                      float opacity_threshold, float big_threshold, uint8_t* keep, void* stream);

/* splat_compact_rows: dst = src[keep] for `count` (<= 24) float32 row tensors at once (parameters, Adam exp_avg /
 * exp_avg_sq, per-Gaussian variables), order preserving, out of place (src != dst).  keep: n bytes, 4-byte aligned,
 * non-zero = keep.  *n_kept_out (device or mapped pinned host uint32, may be NULL) receives the number of rows kept;
 * scratch >= splat_compact_scratch_bytes(n), 16-byte aligned. */
typedef struct SplatRowTensor {
  const float* src;
  float* dst;
  int32_t width;      /* floats per row */
} SplatRowTensor;
size_t splat_compact_scratch_bytes(int64_t n);
int splat_compact_rows(int64_t n, const uint8_t* keep, const SplatRowTensor* tensors, int32_t count, void* scratch,
                       size_t scratch_bytes, uint32_t* n_kept_out, void* stream);

/* splat_new_gaussian_mask: the pixels add_new_gaussians spawns Gaussians from (scripts/splatam.py:390-408):
 *   depth_error = |gt - render_depth| * (gt > 0);   median = torch.median(depth_error)  (exact lower median, NaN
 *   propagating);   pre = (silhouette < sil_thres) | ((render_depth > gt) & (depth_error > 50 * median));
 *   flags = pre & (gt > 0).
 * depth_sil: [3,H,W] silhouette render (depth, silhouette, depth^2), gt_depth: [H,W], flags: W*H bytes (4-byte
 * aligned).  counts_out (device or mapped pinned uint32[2], may be NULL) = { sum(pre), sum(flags) } -- the reference
 * appends only if sum(pre) > 0.  scratch >= splat_spawn_scratch_bytes(W,H); it also carries the scan that
 * splat_spawn_gaussians consumes, so pass the SAME scratch, untouched, to that call. */
size_t splat_spawn_scratch_bytes(int32_t width, int32_t height);
int splat_new_gaussian_mask(int32_t width, int32_t height, const float* depth_sil, const float* gt_depth,
                            float sil_thres, uint8_t* flags, void* scratch, size_t scratch_bytes,
                            uint32_t* counts_out, void* stream);

/* splat_spawn_gaussians: get_pointcloud (:73-123, transform_pts, projective mean-square distance) +
 * initialize_new_params (:354-381) for the flagged pixels, in raster order, written to rows P_old.. of the six
 * destination tensors (which must have room for P_old + counts_out[1] rows): means3D = c2w * (xx*z, yy*z, z, 1) with
 * c2w the inverse of the frame's pose (cam_unnorm_rots [1,4,n_frames], cam_trans [1,3,n_frames], column time_idx),
 * rgb = im[:, pixel], rotation (1,0,0,0), logit opacity 0, log_scales = log(sqrt((z / ((fx+fy)/2))^2)) x scale_dim,
 * timestep = time_idx (timestep may be NULL). */
int splat_spawn_gaussians(int32_t width, int32_t height, const uint8_t* flags, const void* scratch, const float* im,
                          const float* gt_depth, float fx, float fy, float cx, float cy, const float* cam_unnorm_rots,
                          const float* cam_trans, int32_t n_frames, int32_t time_idx, int32_t scale_dim, int64_t P_old,
                          float* means3D, float* rgb_colors, float* unnorm_rotations, float* logit_opacities,
                          float* log_scales, float* timestep, void* stream);

/* ---- multi-GPU exchange (new; the reference is single-GPU, configs/replica/splatam.py:8) -------
 * Sum of one flat float32 bucket over the `world` ranks of one NVSwitch node, in place, on `stream`.
 * buffers[j] / flag_pads[j] are the device addresses, valid in THIS process, of rank j's bucket and of
 * rank j's flag pad (>= 256 bytes, zero-initialised once) -- e.g. the buffer_ptrs of two
 * torch.distributed._symmetric_memory rendezvous handles.  `epoch` must be the same on every rank and
 * increase by one per call.  n_floats % 4 == 0, buffers 16-byte aligned.  Rank r reduces slice r over
 * peer loads and stores the sums into every rank's bucket; all ranks end up with bit-identical sums.
 * multicast_buffer != 0: the NVSwitch multicast address of the same bucket; the sums are then formed inside the
 * switch (multimem.ld_reduce / multimem.st) and each link carries the bucket once instead of 2(N-1)/N times. */
int splat_peer_allreduce_sum(const uint64_t* buffers, const uint64_t* flag_pads, uint64_t multicast_buffer,
                             int32_t rank, int32_t world, uint32_t epoch, int64_t n_floats, void* stream);

/* Fused exchange + optimizer step of the keyframe-sharded mapping iteration (scripts/splatam.py:889's optimizer.step()
 * after a gradient sum over ranks): ONE kernel per rank does reduce-scatter -> Adam -> all-gather over peer memory.
 * Rank r sums slice r of the gradient bucket (multimem.ld_reduce in the NVSwitch when multicast addresses are given, else
 * peer loads in rank order), applies Adam (arithmetic of splat_adam_step, shared code) to slice r of the PARAMETER
 * bucket with its own 1/N shard of the moments, and stores the updated parameters into every replica.  grad_buffers[j] /
 * param_buffers[j] are rank j's two buckets (same layout, n_floats each, symmetric memory); segment k covers the elements
 * [segment_ends[k-1], segment_ends[k]) and has learning rate segment_lrs[k] (one segment per parameter tensor).  `step`
 * (>= 1, same on every rank) is the Adam step number; the shards hold splat_peer_adam_shard_floats(n_floats, world)
 * floats each and must start zeroed.  `epoch` as for splat_peer_allreduce_sum (shared counter: every call of either
 * function advances it by one). */
size_t splat_peer_adam_shard_floats(int64_t n_floats, int32_t world);
/* Measurement knobs of the two in-switch (multicast) kernels, no reference counterpart: CTAs per launch (0 = the measured
 * default) and multimem loads in flight per thread of the plain all-reduce (1, 2, 4 or 8; anything else = the default).  Results do
 * not depend on either.  Process-wide; set the same values on every rank. */
void splat_peer_set_tuning(int32_t blocks, int32_t unroll);
int splat_peer_reduce_adam(const uint64_t* grad_buffers, const uint64_t* param_buffers, const uint64_t* flag_pads,
                           uint64_t grad_multicast, uint64_t param_multicast, int32_t rank, int32_t world,
                           uint32_t epoch, int64_t n_floats, const int64_t* segment_ends, const float* segment_lrs,
                           int32_t n_segments, int32_t step, double beta1, double beta2, double eps,
                           float* exp_avg_shard, float* exp_avg_sq_shard, void* stream);

/* ---- introspection used by the parity tests (read-only views into the states) ---------------
 * ranges: uint2[T] exactly as identifyTileRanges leaves them (untouched tiles are (0,0));
 * point_list: uint32[R] sorted Gaussian ids; n_contrib uint32[N]; final_T float[N]. */
const void* splat_state_ranges(const void* image_state, int32_t width, int32_t height);
const void* splat_state_n_contrib(const void* image_state, int32_t width, int32_t height);
const void* splat_state_final_T(const void* image_state, int32_t width, int32_t height);
const void* splat_state_point_list(const void* binning_state, int32_t width, int32_t height,
                                   int64_t R_capacity);
/* float[P,12] per-Gaussian records (x, y, conic A, B | conic C, opacity, view depth, - | r, g, b, -);
 * only rows with radii > 0 are defined.  float[R,12] sorted instance records: same layout with the
 * Gaussian id (uint32 bits) in slot 7 -- the byte stream the blend kernels' TMA ring consumes. */
const void* splat_state_gaussian_records(const void* geom_state, int32_t P);
const void* splat_state_instance_records(const void* binning_state, int32_t width, int32_t height,
                                         int64_t R_capacity);

/* ---- measurement helpers (used by bench.py; not on the product path) -----------------------
 * splat_profile_enable(1) makes every kernel launch of this library record a cudaEvent pair on
 * its stream; splat_profile_read() synchronises, returns the summed milliseconds and launch
 * counts per kernel id (order given by splat_profile_kernel_name(k), k = 0.. until it returns "") and resets. */
int splat_profile_enable(int on);
int splat_profile_read(float* ms_sum, int32_t* launches, int32_t n);
/* idle milliseconds on the stream before each kernel id (previous profiled launch's end -> its start),
 * for the pairs consumed by the last splat_profile_read(). */
int splat_profile_read_gaps(float* gap_ms_sum, int32_t n);
const char* splat_profile_kernel_name(int kernel_id);
int splat_debug_count_blended(const SplatCamera* cam, int32_t P, const float* colors, const void* geom_state,
                              void* image_state, const void* binning_state, int64_t R_capacity,
                              float* scratch_color, uint32_t* blended, void* stream);
int splat_debug_fp32_peak(float* scratch_dev, float* tflops_out, void* stream);
/* How the blend kernels get a tile's records (process-wide; returns the previous setting; also SPLAT_B200_GATHER=0/1):
 *   1  gather: three 16-byte cp.async per instance straight from the Gaussians' own 48-byte records (L2 resident),
 *      addressed through the sorted id list -- no emit_records pass, the instance-record area of binning_state unused;
 *   0  stream: emit_records re-writes the records in list order and the blend kernels stream them with the TMA unit.
 * Results are identical; which one is the default is a measured choice. */
int splat_set_gather(int on);
/* Parity probe for the blend kernels' packed exponential (csrc/packed.cuh): evaluates it and the device library's
 * expf() -- what the reference build calls (HX/cuda_rasterizer/forward.cu:349) -- on the `count` consecutive float bit
 * patterns starting at first_bits, and ADDS the number of bit-level disagreements to *mismatches_dev (device uint64). */
int splat_debug_exp_mismatches(uint32_t first_bits, uint64_t count, unsigned long long* mismatches_dev, void* stream);

/* ---- misc ------------------------------------------------------------------------------------ */
int splat_abi_version(void);
const char* splat_strerror(int code);
const char* splat_last_error(void);      /* thread-local detail string of the last failure */
int splat_kernel_launch_count(void);     /* kernels launched by this library in this process */

#ifdef __cplusplus
}
#endif
#endif /* SPLAT_B200_H_ */
