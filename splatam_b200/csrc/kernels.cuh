// kernels.cuh -- argument blocks and host-side launchers of every kernel in libsplat_b200.
#pragma once
#include "common.cuh"

namespace splat {

struct PreFwdArgs {
  int P, W, H, gx, gy, M, sh_degree;
  float tanx, tany, fx, fy, mod, near_plane;
  const float* means; const float* scales; const float* rots; const float* opac;
  const float* cov3D_pre; const float* shs; const float* colors_pre;
  const float* V; const float* Pm; const float* campos;
  GeomView g; uint32_t* tile_count; int* radii;
};

struct PreBwdArgs {
  int P, M, sh_degree;
  float tanx, tany, fx, fy, mod;
  const float* means; const float* scales; const float* rots; const float* cov3D_pre;
  const float* shs; const uint8_t* clamped; const float* campos;
  const float* V; const float* Pm; const int* radii;
  int W, H;
  const float4* grec;           // [P*3] forward geometry records (conic + opacity)
  const float* acc;             // [P,12] raw blend-backward sums (see render_bwd.cu)
  int depth_sil;                // fused RGB+depth+silhouette: acc[9], acc[10] = sums for the z and z*z channels
  float* dL_dmean2D;  // [P,3] out
  float* dL_dconic;   // [P,4] out
  float* dL_dcolor;   // [P,3] out (and SH backward input)
  float* dL_dopacity; // [P] out
  float* dL_dmeans; float* dL_dcov3D; float* dL_dscales; float* dL_drots; float* dL_dsh;
};

struct HessArgs {
  int P;
  float tanx, tany, fx, fy, mod;
  const float* means; const float* scales; const float* rots; const float* cov3D_pre;
  int W, H;
  const float* V; const float* Pm; const int* radii; const float* acc; const float4* grec;
  float* dL_dmeans; float* dL_dmean2D; float* dL_dcolors; float* dL_dopacity;
  float* dL_dscales; float* dL_drots; float* dL_dcov3D; float* dL_dconic;
};

struct RenderFwdArgs {
  int W, H, gx;
  int depth_sil;          // 1: six output planes, planes 3..5 = features (z, 1, z*z)
  const uint32_t* order;  // blockIdx.x -> tile (ImageView::tile_order)
  const uint2* ranges; const float4* inst;
  const float* bg;
  float* final_T; uint32_t* n_contrib; float* out_color; float* out_depth;
  const ImageHeader* hdr; uint32_t Rcap;
  uint32_t* blended;  // debug: per-pixel count of blended pairs (NULL in production)
  int zero;           // always 0: an operand ptxas cannot fold (packed.cuh ExpConsts)
  const float4* grec; const uint32_t* plist;   // gather mode: per-Gaussian records + the sorted id list (inst is not written)
  int P;              // number of Gaussians (bounds checks)
};

struct RenderBwdArgs {
  int W, H, gx;
  const uint32_t* order;
  const uint2* ranges; const float4* inst; const float* bg;
  const float* final_T; const uint32_t* n_contrib; const float* dL_dpix;
  const float* out_color;   // optional: the forward's image, reference level of the per-pixel feature sums (MODE 0/2)
  float* acc;   // MODE 0: [P,12] raw sums.  MODE 1: [P,20] moment accumulator.  Pre-zeroed.
  const ImageHeader* hdr; uint32_t Rcap;   // R > Rcap (a graph replay that overflowed its binning capacity): nothing to walk
  int zero;     // always 0 (packed.cuh ExpConsts)
  const float4* grec; const uint32_t* plist;   // gather mode (see RenderFwdArgs)
  int P;        // number of Gaussians (bounds checks)
};

// SplaTAM glue (slam_ops.cu).  Render-variable builder: forward outputs o_*, backward inputs g_* (any may be NULL ==
// zero), Gaussian-side gradients d_* and the pose gradient columns d_cam_*[:, :, t].
struct RvArgs {
  int P, aniso, n_frames, t;
  const float* means; const float* unnorm_rots; const float* logit_op; const float* log_scales;
  const float* cam_rots; const float* cam_trans;       // [1,4,F], [1,3,F]
  float* o_means; float* o_rots; float* o_op; float* o_scales;
  const float* g_means; const float* g_rots; const float* g_op; const float* g_scales;
  float* d_means; float* d_unnorm_rots; float* d_logit; float* d_logscales;
  float* d_cam_rots; float* d_cam_trans;
};
struct LossArgs {
  int W, H, tracking, use_sil;
  float sil_thres, w_im, w_depth;
  float gauss[11];                                      // utils/slam_external.py:54-56
  const float* im; const float* depth_sil; const float* gt_im; const float* gt_depth;
  float* losses; float* g_im; float* g_depth_sil;
  float* dmaps;                                         // [3 maps][3 ch][H][W] mapping loss only
};
struct SplatAdamTensorDev {   // == SplatAdamTensor of the C ABI + a vector-access flag
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t n; float lr, bias_correction1, bias_correction2_sqrt; int vec;
};
typedef SplatAdamTensorDev SplatAdamTensorK;
#define ADAM_MAX_TENSORS 16
struct AdamArgs {
  int count;
  float beta2, w1, w2, eps;   // w = 1 - beta, rounded from double
  double beta1d, beta2d;
  int step[ADAM_MAX_TENSORS];
  int* replay_counter;        // capturable mode: effective step = step[i] + *replay_counter (then incremented)
  float* dev_bias;            // [2 * ADAM_MAX_TENSORS]: (bias_correction1, bias_correction2_sqrt) per tensor, device-formed
  int blocks[ADAM_MAX_TENSORS];
  SplatAdamTensorK t[ADAM_MAX_TENSORS];
};
int launch_adam(AdamArgs& a, cudaStream_t st);
int launch_forward_guard(const ImageHeader* hdr, uint32_t Rcap, uint32_t* flags, cudaStream_t st);
int rendervar_bwd_blocks(int P);
int launch_rendervar_fwd(const RvArgs& a, cudaStream_t st);
int launch_rendervar_bwd(const RvArgs& a, bool gg, bool cg, float* partial, cudaStream_t st);
int slam_loss_blocks(int W, int H, int tracking);
int launch_slam_loss(const LossArgs& a, float* partial, float* stats, cudaStream_t st);

int launch_preprocess_fwd(const PreFwdArgs& a, cudaStream_t st);
int launch_preprocess_bwd(const PreBwdArgs& a, cudaStream_t st);
int launch_hessian_finalize(const HessArgs& a, cudaStream_t st);
int launch_mark_visible(int P, const float* means, const float* V, float near_plane, uint8_t* present, cudaStream_t st);
int launch_tile_scan(int T, const ImageView& iv, uint32_t* num_rendered_out, uint32_t sort_cap, cudaStream_t st);
int launch_scatter(int P, int gx, int gy, const int* radii, const GeomView& g, const ImageView& iv,
                   const BinView& b, uint32_t Rcap, cudaStream_t st);
int launch_tile_sort(int T, const ImageView& iv, const GeomView& g, const BinView& b, uint32_t Rcap, cudaStream_t st);
int sort_smem_cap();
int launch_render_fwd(const RenderFwdArgs& a, int T, cudaStream_t st);
int launch_render_bwd(const RenderBwdArgs& a, int T, int mode, cudaStream_t st);
extern int g_disable_cull;
extern int g_gather;   // 1: the blend kernels gather the Gaussians' own records by id (no emit_records pass)
int launch_exp_check(uint32_t first, unsigned long long count, unsigned long long* out, cudaStream_t st);

}  // namespace splat
