// slam_ops.cu -- the per-iteration glue SplaTAM runs in PyTorch on either side of the rasterizer
// (SURVEY.md 8f-2 / 8f-3), as five HBM-bound kernels:
//   rendervar_fwd / rendervar_bwd (+ pose_finalize)  <-  transform_to_frame + transformed_params2rendervar
//        (utils/slam_helpers.py:253-304, :123-139; build_rotation utils/slam_external.py:25-42) and the
//        autograd graph torch builds for them (cat, matmul, normalize, quat_mult, sigmoid, exp, tile);
//        the camera-pose gradient (sum over all Gaussians of dL/dp' and dL/dp' (x) p) is reduced in-kernel;
//   track_loss                                        <-  get_loss(tracking=True), scripts/splatam.py:262-285
//   map_loss_fwd / map_loss_bwd (+ loss_finalize)     <-  get_loss(mapping=True): masked L1 depth mean +
//        0.8 L1 + 0.2 (1 - SSIM) colour (scripts/splatam.py:279-296, utils/slam_external.py:54-97), producing
//        dL/d(im, depth_sil) directly -- what torch autograd computes through five 11x11 convolutions.
// All reductions are two-stage and deterministic (per-block partials, then one block).
#include "kernels.cuh"

namespace splat {

// ------------------------------------------------------------------------------------------------ pose
struct Pose {
  float q1[4];   // F.normalize(cam_unnorm_rots[..., t])  (eps 1e-12)        slam_helpers.py:268
  float q2[4];   // q1 / |q1|                                                 slam_external.py:26-27
  float n0, n1;  // |u|, |q1|
  float R[9];    // row-major rotation of q2                                  slam_external.py:33-41
  float t[3];
};

__device__ __forceinline__ Pose load_pose(const float* cam_rots, const float* cam_trans, int F, int t) {
  Pose p;
  float u[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) u[k] = cam_rots[k * F + t];
  p.n0 = sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2] + u[3] * u[3]);
  const float d0 = fmaxf(p.n0, 1e-12f);
#pragma unroll
  for (int k = 0; k < 4; ++k) p.q1[k] = u[k] / d0;
  p.n1 = sqrtf(p.q1[0] * p.q1[0] + p.q1[1] * p.q1[1] + p.q1[2] * p.q1[2] + p.q1[3] * p.q1[3]);
#pragma unroll
  for (int k = 0; k < 4; ++k) p.q2[k] = p.q1[k] / p.n1;
  const float r = p.q2[0], x = p.q2[1], y = p.q2[2], z = p.q2[3];
  p.R[0] = 1.f - 2.f * (y * y + z * z); p.R[1] = 2.f * (x * y - r * z);       p.R[2] = 2.f * (x * z + r * y);
  p.R[3] = 2.f * (x * y + r * z);       p.R[4] = 1.f - 2.f * (x * x + z * z); p.R[5] = 2.f * (y * z - r * x);
  p.R[6] = 2.f * (x * z - r * y);       p.R[7] = 2.f * (y * z + r * x);       p.R[8] = 1.f - 2.f * (x * x + y * y);
#pragma unroll
  for (int k = 0; k < 3; ++k) p.t[k] = cam_trans[k * F + t];
  return p;
}

// F.normalize of a 4-vector: v / max(|v|, 1e-12); returns |v|
__device__ __forceinline__ float normalize4(const float4 v, float4& o) {
  const float n = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
  const float d = fmaxf(n, 1e-12f);
  o = make_float4(v.x / d, v.y / d, v.z / d, v.w / d);
  return n;
}
// backward of o = v / max(|v|, eps) given o, |v|
__device__ __forceinline__ float4 normalize4_bwd(const float4 o, const float n, const float4 g) {
  if (n < 1e-12f) return make_float4(g.x / 1e-12f, g.y / 1e-12f, g.z / 1e-12f, g.w / 1e-12f);
  const float dot = o.x * g.x + o.y * g.y + o.z * g.z + o.w * g.w;
  const float inv = 1.f / n;
  return make_float4((g.x - o.x * dot) * inv, (g.y - o.y * dot) * inv, (g.z - o.z * dot) * inv, (g.w - o.w * dot) * inv);
}
// m = c (x) n, Hamilton product in (w,x,y,z) order (slam_helpers.py:21-28)
__device__ __forceinline__ float4 quat_mul(const float* c, const float4 n) {
  return make_float4(c[0] * n.x - c[1] * n.y - c[2] * n.z - c[3] * n.w,
                     c[0] * n.y + c[1] * n.x + c[2] * n.w - c[3] * n.z,
                     c[0] * n.z - c[1] * n.w + c[2] * n.x + c[3] * n.y,
                     c[0] * n.w + c[1] * n.z - c[2] * n.y + c[3] * n.x);
}

// ------------------------------------------------------------------------------------------------ builder fwd
__global__ void __launch_bounds__(256) rendervar_fwd_kernel(const RvArgs a) {
  pdl_wait();
  pdl_trigger();
  __shared__ Pose ps;   // R(q), t: built by one thread per CTA (two square roots + eight divides), broadcast from shared memory
  if (threadIdx.x == 0) ps = load_pose(a.cam_rots, a.cam_trans, a.n_frames, a.t);
  __syncthreads();
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= a.P) return;
  const float px = a.means[3 * i], py = a.means[3 * i + 1], pz = a.means[3 * i + 2];
  const float4 q = reinterpret_cast<const float4*>(a.unnorm_rots)[i];
  const float lo = a.logit_op[i];
  float ls[3];
  if (a.aniso) { ls[0] = a.log_scales[3 * i]; ls[1] = a.log_scales[3 * i + 1]; ls[2] = a.log_scales[3 * i + 2]; }
  else { ls[0] = ls[1] = ls[2] = a.log_scales[i]; }
  a.o_means[3 * i + 0] = ps.R[0] * px + ps.R[1] * py + ps.R[2] * pz + ps.t[0];
  a.o_means[3 * i + 1] = ps.R[3] * px + ps.R[4] * py + ps.R[5] * pz + ps.t[1];
  a.o_means[3 * i + 2] = ps.R[6] * px + ps.R[7] * py + ps.R[8] * pz + ps.t[2];
  float4 n, o;
  normalize4(q, n);
  if (a.aniso) { normalize4(quat_mul(ps.q1, n), o); } else { o = n; }
  reinterpret_cast<float4*>(a.o_rots)[i] = o;
  a.o_op[i] = 1.f / (1.f + expf(-lo));
#pragma unroll
  for (int k = 0; k < 3; ++k) a.o_scales[3 * i + k] = expf(ls[k]);
}

// ------------------------------------------------------------------------------------------------ builder bwd
#define RV_SUMS 16   // 0..2 sum dL/dp' | 3..11 sum dL/dp' (x) p (row-major) | 12..15 sum dL/dq1 through quat_mult

template <bool GG, bool CG, bool ANISO>
__global__ void __launch_bounds__(256) rendervar_bwd_kernel(const RvArgs a, float* __restrict__ partial) {
  pdl_wait();
  pdl_trigger();
  __shared__ Pose ps;
  if (threadIdx.x == 0) ps = load_pose(a.cam_rots, a.cam_trans, a.n_frames, a.t);
  __syncthreads();
  float s[RV_SUMS];
#pragma unroll
  for (int k = 0; k < RV_SUMS; ++k) s[k] = 0.f;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < a.P; i += gridDim.x * 256) {
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (a.g_means) { gx = a.g_means[3 * i]; gy = a.g_means[3 * i + 1]; gz = a.g_means[3 * i + 2]; }
    if (CG) {
      const float px = a.means[3 * i], py = a.means[3 * i + 1], pz = a.means[3 * i + 2];
      s[0] += gx; s[1] += gy; s[2] += gz;
      s[3] += gx * px; s[4] += gx * py; s[5] += gx * pz;
      s[6] += gy * px; s[7] += gy * py; s[8] += gy * pz;
      s[9] += gz * px; s[10] += gz * py; s[11] += gz * pz;
    }
    if (GG) {
      a.d_means[3 * i + 0] = ps.R[0] * gx + ps.R[3] * gy + ps.R[6] * gz;
      a.d_means[3 * i + 1] = ps.R[1] * gx + ps.R[4] * gy + ps.R[7] * gz;
      a.d_means[3 * i + 2] = ps.R[2] * gx + ps.R[5] * gy + ps.R[8] * gz;
    }
    if (a.d_logit) {   // opacity / scale gradients flow even when gaussians_grad is false (slam_helpers.py:133-136)
      const float sg = 1.f / (1.f + expf(-a.logit_op[i]));
      a.d_logit[i] = a.g_op ? a.g_op[i] * sg * (1.f - sg) : 0.f;
    }
    if (a.d_logscales) {
      float g3[3] = {0.f, 0.f, 0.f};
      if (a.g_scales) { g3[0] = a.g_scales[3 * i]; g3[1] = a.g_scales[3 * i + 1]; g3[2] = a.g_scales[3 * i + 2]; }
      if (ANISO) {
#pragma unroll
        for (int k = 0; k < 3; ++k) a.d_logscales[3 * i + k] = g3[k] * expf(a.log_scales[3 * i + k]);
      } else {
        a.d_logscales[i] = (g3[0] + g3[1] + g3[2]) * expf(a.log_scales[i]);
      }
    }
    if (GG || (CG && ANISO)) {
      const float4 q = reinterpret_cast<const float4*>(a.unnorm_rots)[i];
      float4 g = a.g_rots ? reinterpret_cast<const float4*>(a.g_rots)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 n;
      const float nq = normalize4(q, n);
      if (ANISO) {
        const float4 m = quat_mul(ps.q1, n);
        float4 o;
        const float nm = normalize4(m, o);
        const float4 dm = normalize4_bwd(o, nm, g);
        const float* c = ps.q1;
        if (CG) {
          s[12] += n.x * dm.x + n.y * dm.y + n.z * dm.z + n.w * dm.w;
          s[13] += -n.y * dm.x + n.x * dm.y - n.w * dm.z + n.z * dm.w;
          s[14] += -n.z * dm.x + n.w * dm.y + n.x * dm.z - n.y * dm.w;
          s[15] += -n.w * dm.x - n.z * dm.y + n.y * dm.z + n.x * dm.w;
        }
        g = make_float4(c[0] * dm.x + c[1] * dm.y + c[2] * dm.z + c[3] * dm.w,
                        -c[1] * dm.x + c[0] * dm.y + c[3] * dm.z - c[2] * dm.w,
                        -c[2] * dm.x - c[3] * dm.y + c[0] * dm.z + c[1] * dm.w,
                        -c[3] * dm.x + c[2] * dm.y - c[1] * dm.z + c[0] * dm.w);
      }
      if (GG) reinterpret_cast<float4*>(a.d_unnorm_rots)[i] = normalize4_bwd(n, nq, g);
    }
  }
  if (CG) {
    __shared__ float red[8][RV_SUMS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < RV_SUMS; ++k) {
      float v = s[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[w][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < RV_SUMS) {
      float v = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) v += red[j][threadIdx.x];
      partial[blockIdx.x * RV_SUMS + threadIdx.x] = v;
    }
  }
}

// One block: sums the per-block partials (double), then runs the 7-scalar chain
// dR, dt, dq1 -> cam_unnorm_rots[:, :, t], cam_trans[:, :, t].
__global__ void __launch_bounds__(512) pose_finalize_kernel(const RvArgs a, const float* __restrict__ partial, int nblk) {
  pdl_wait();
  pdl_trigger();
  __shared__ double S[RV_SUMS];
  const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;   // 16 warps, one per sum
  double v = 0.0;
  float p4[4];
  for (int b0 = lane; b0 < nblk; b0 += 128) {   // four independent loads in flight per lane (single block: latency-bound)
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int b = b0 + 32 * u; p4[u] = b < nblk ? partial[b * RV_SUMS + k] : 0.f; }
    v += ((double)p4[0] + (double)p4[1]) + ((double)p4[2] + (double)p4[3]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) S[k] = v;
  __syncthreads();
  if (threadIdx.x != 0) return;
  const Pose ps = load_pose(a.cam_rots, a.cam_trans, a.n_frames, a.t);
  const double r = ps.q2[0], x = ps.q2[1], y = ps.q2[2], z = ps.q2[3];
  const double* G = &S[3];   // G[3*row+col] = dL/dR[row][col]
  double d2[4];
  d2[0] = 2.0 * (-z * G[1] + y * G[2] + z * G[3] - x * G[5] - y * G[6] + x * G[7]);
  d2[1] = 2.0 * (y * G[1] + z * G[2] + y * G[3] - 2.0 * x * G[4] - r * G[5] + z * G[6] + r * G[7] - 2.0 * x * G[8]);
  d2[2] = 2.0 * (-2.0 * y * G[0] + x * G[1] + r * G[2] + x * G[3] + z * G[5] - r * G[6] + z * G[7] - 2.0 * y * G[8]);
  d2[3] = 2.0 * (-2.0 * z * G[0] - r * G[1] + x * G[2] + r * G[3] - 2.0 * z * G[4] + y * G[5] + x * G[6] + y * G[7]);
  // q2 = q1 / |q1|
  double dot = 0.0;
  for (int j = 0; j < 4; ++j) dot += (double)ps.q2[j] * d2[j];
  double d1[4];
  for (int j = 0; j < 4; ++j) d1[j] = (d2[j] - (double)ps.q2[j] * dot) / (double)ps.n1 + S[12 + j];
  // q1 = u / max(|u|, eps)
  double du[4];
  if (ps.n0 < 1e-12f) {
    for (int j = 0; j < 4; ++j) du[j] = d1[j] / 1e-12;
  } else {
    dot = 0.0;
    for (int j = 0; j < 4; ++j) dot += (double)ps.q1[j] * d1[j];
    for (int j = 0; j < 4; ++j) du[j] = (d1[j] - (double)ps.q1[j] * dot) / (double)ps.n0;
  }
  for (int j = 0; j < 4; ++j) a.d_cam_rots[j * a.n_frames + a.t] = (float)du[j];
  for (int j = 0; j < 3; ++j) a.d_cam_trans[j * a.n_frames + a.t] = (float)S[j];
}

int rendervar_bwd_blocks(int P) {
  int nb = (P + 255) / 256;
  return nb < 1 ? 1 : (nb > 148 * 4 ? 148 * 4 : nb);   // grid-stride loop; few partial rows for the finalize kernel
}

int launch_rendervar_fwd(const RvArgs& a, cudaStream_t st) {
  ProfScope prof_(K_RV_FWD, st);
  if (a.P <= 0) return SPLAT_OK;
  SPLAT_CUDA(launch_k(rendervar_fwd_kernel, dim3((a.P + 255) / 256), dim3(256), 0, st, a));
  SPLAT_LAUNCH_CHECK("rendervar_fwd_kernel");
  return SPLAT_OK;
}

int launch_rendervar_bwd(const RvArgs& a, bool gg, bool cg, float* partial, cudaStream_t st) {
  ProfScope prof_(K_RV_BWD, st);
  if (!gg && !cg && !a.d_logit && !a.d_logscales) return SPLAT_OK;
  const int nb = rendervar_bwd_blocks(a.P);
  if (a.P > 0) {
#define RV_CASE(G, C, A) SPLAT_CUDA(launch_k(rendervar_bwd_kernel<G, C, A>, dim3(nb), dim3(256), 0, st, a, partial))
    const int key = (gg ? 4 : 0) | (cg ? 2 : 0) | (a.aniso ? 1 : 0);
    switch (key) {
      case 0: RV_CASE(false, false, false); break;
      case 1: RV_CASE(false, false, true); break;
      case 2: RV_CASE(false, true, false); break;
      case 3: RV_CASE(false, true, true); break;
      case 4: RV_CASE(true, false, false); break;
      case 5: RV_CASE(true, false, true); break;
      case 6: RV_CASE(true, true, false); break;
      default: RV_CASE(true, true, true); break;
    }
#undef RV_CASE
    SPLAT_LAUNCH_CHECK("rendervar_bwd_kernel");
  }
  if (cg) {
    SPLAT_CUDA(launch_k(pose_finalize_kernel, dim3(1), dim3(512), 0, st, a, (const float*)partial, a.P > 0 ? nb : 0));
    SPLAT_LAUNCH_CHECK("pose_finalize_kernel");
  }
  return SPLAT_OK;
}

// ================================================================================================ losses
#define LOSS_SUMS 4   // tracking: (sum |dim|, sum |ddepth|, mask count, -)   mapping: (ssim sum, L1 sum, depth sum, count)

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

__device__ __forceinline__ void block_partials(float (&s)[LOSS_SUMS], float* __restrict__ partial, int block_id) {
  __shared__ float red[8][LOSS_SUMS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < LOSS_SUMS; ++k) {
    float v = s[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[w][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < LOSS_SUMS) {
    float v = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) v += red[j][threadIdx.x];
    partial[block_id * LOSS_SUMS + threadIdx.x] = v;
  }
}

// mask of scripts/splatam.py:268-280 (ignore_outlier_depth_loss == False)
__device__ __forceinline__ bool loss_mask(float gt_d, float d, float d2, float sil, bool use_sil, float thres) {
  const float unc = d2 - d * d;
  bool m = (gt_d > 0.f) && !isnan(d) && !isnan(unc);
  if (use_sil) m = m && (sil > thres);
  return m;
}

// get_loss(tracking=True): sums over the mask, gradient written in the same pass.
__global__ void __launch_bounds__(256) track_loss_kernel(const LossArgs a, float* __restrict__ partial) {
  pdl_wait();
  pdl_trigger();
  const int N = a.W * a.H;
  float s[LOSS_SUMS] = {0.f, 0.f, 0.f, 0.f};
  for (int p = blockIdx.x * 256 + threadIdx.x; p < N; p += gridDim.x * 256) {
    const float d = a.depth_sil[p], sil = a.depth_sil[N + p], d2 = a.depth_sil[2 * N + p];
    const float gd = a.gt_depth[p];
    const bool m = loss_mask(gd, d, d2, sil, a.use_sil, a.sil_thres);
    const bool cm = a.use_sil ? m : true;   // colour mask only when use_sil_for_loss (splatam.py:288-293)
    const float dd = gd - d;
    s[1] += m ? fabsf(dd) : 0.f;
    s[2] += m ? 1.f : 0.f;
    a.g_depth_sil[p] = m ? -a.w_depth * sgn(dd) : 0.f;
    a.g_depth_sil[N + p] = 0.f;
    a.g_depth_sil[2 * N + p] = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float di = a.gt_im[c * N + p] - a.im[c * N + p];
      s[0] += cm ? fabsf(di) : 0.f;
      a.g_im[c * N + p] = cm ? -a.w_im * sgn(di) : 0.f;
    }
  }
  block_partials(s, partial, blockIdx.x);
}

// ---- mapping loss: 11x11 Gaussian-window SSIM, separable, 32x32 output tile per CTA and channel.
// Both passes are register-blocked: a thread produces 4 adjacent outputs from 14 inputs (horizontal pass: four
// 128-bit shared loads per image row segment; vertical pass: 14 conflict-free column loads per quantity), so an
// output costs 3.5 shared-memory words per quantity instead of 11.
#define ST 32
#define SHALO 5
#define SW (ST + 2 * SHALO)   // 42 rows / columns incl. halo
#define SWP 44                // padded input row: 16-byte aligned segments, columns up to 28 + 15
#define HZP 36                // padded row of the horizontal-pass results

__device__ __forceinline__ void lds16(const float* row, float (&u)[16]) {
  const float4* r4 = reinterpret_cast<const float4*>(row);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float4 q = r4[t];
    u[4 * t] = q.x; u[4 * t + 1] = q.y; u[4 * t + 2] = q.z; u[4 * t + 3] = q.w;
  }
}

__global__ void __launch_bounds__(256) map_loss_fwd_kernel(const LossArgs a, float* __restrict__ partial) {
  pdl_wait();
  pdl_trigger();
  __shared__ __align__(16) float sx[SW][SWP], sy[SW][SWP];
  __shared__ __align__(16) float hz[5][SW][HZP];
  const int W = a.W, H = a.H, N = W * H, c = blockIdx.z;
  const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST;
  const float* X = a.im + (size_t)c * N;
  const float* Y = a.gt_im + (size_t)c * N;
  for (int k = threadIdx.x; k < SW * SWP; k += 256) {
    const int ly = k / SWP, lx = k - ly * SWP, gx = x0 + lx - SHALO, gy = y0 + ly - SHALO;
    const bool in = lx < SW && gx >= 0 && gx < W && gy >= 0 && gy < H;
    sx[ly][lx] = in ? X[gy * W + gx] : 0.f;
    sy[ly][lx] = in ? Y[gy * W + gx] : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < SW * 8; k += 256) {
    const int ly = k >> 3, lx0 = (k & 7) * 4;
    float u[16], v[16];
    lds16(&sx[ly][lx0], u);
    lds16(&sy[ly][lx0], v);
    float o[5][4];
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int i = 0; i < 4; ++i) o[q][i] = 0.f;
#pragma unroll
    for (int j = 0; j < 14; ++j) {
      const float uu = u[j] * u[j], vv = v[j] * v[j], uv = u[j] * v[j];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = j - i;
        if (t >= 0 && t < 11) {
          const float g = a.gauss[t];
          o[0][i] += g * u[j]; o[1][i] += g * v[j]; o[2][i] += g * uu; o[3][i] += g * vv; o[4][i] += g * uv;
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 5; ++q)
      *reinterpret_cast<float4*>(&hz[q][ly][lx0]) = make_float4(o[q][0], o[q][1], o[q][2], o[q][3]);
  }
  __syncthreads();
  float s[LOSS_SUMS] = {0.f, 0.f, 0.f, 0.f};
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  {
    const int lx = threadIdx.x & 31, ly0 = (threadIdx.x >> 5) * 4;
    float o[5][4];
#pragma unroll
    for (int q = 0; q < 5; ++q)
#pragma unroll
      for (int i = 0; i < 4; ++i) o[q][i] = 0.f;
#pragma unroll
    for (int j = 0; j < 14; ++j) {
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        const float h = hz[q][ly0 + j][lx];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int t = j - i;
          if (t >= 0 && t < 11) o[q][i] += a.gauss[t] * h;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int ly = ly0 + i, gx = x0 + lx, gy = y0 + ly;
      if (gx >= W || gy >= H) continue;
      const float m1 = o[0][i], m2 = o[1][i], e11 = o[2][i], e22 = o[3][i], e12 = o[4][i];
      const float s1 = e11 - m1 * m1, s2 = e22 - m2 * m2, s12 = e12 - m1 * m2;
      const float A = 2.f * m1 * m2 + C1, B = 2.f * s12 + C2, Cc = m1 * m1 + m2 * m2 + C1, D = s1 + s2 + C2;
      const float inv = 1.f / (Cc * D);
      const float m = A * B * inv;
      // partial derivatives of the map value w.r.t. the three window means that depend on the rendered image:
      // conv(x) [with sigma1, sigma12 depending on it], conv(x*x), conv(x*y)
      const float d_xy = 2.f * A * inv;
      const float d_xx = -m / D;
      const float d_mu = 2.f * m2 * B * inv - 2.f * m1 * m / Cc - m2 * d_xy - 2.f * m1 * d_xx;
      const size_t p = (size_t)gy * W + gx;
      a.dmaps[(size_t)(0 * 3 + c) * N + p] = d_mu;
      a.dmaps[(size_t)(1 * 3 + c) * N + p] = d_xx;
      a.dmaps[(size_t)(2 * 3 + c) * N + p] = d_xy;
      s[0] += m;
      s[1] += fabsf(sx[ly + SHALO][lx + SHALO] - sy[ly + SHALO][lx + SHALO]);
      if (c == 0) {
        const float d = a.depth_sil[p], sil = a.depth_sil[N + p], d2 = a.depth_sil[2 * N + p], gd = a.gt_depth[p];
        const bool mk = loss_mask(gd, d, d2, sil, false, 0.f);
        s[2] += mk ? fabsf(gd - d) : 0.f;
        s[3] += mk ? 1.f : 0.f;
      }
    }
  }
  block_partials(s, partial, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x);
}

__global__ void __launch_bounds__(256) map_loss_bwd_kernel(const LossArgs a, const float* __restrict__ stats) {
  pdl_wait();
  pdl_trigger();
  __shared__ __align__(16) float sm[3][SW][SWP];
  __shared__ __align__(16) float hz[3][SW][HZP];
  const int W = a.W, H = a.H, N = W * H, c = blockIdx.z;
  const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST;
  for (int k = threadIdx.x; k < SW * SWP; k += 256) {
    const int ly = k / SWP, lx = k - ly * SWP, gx = x0 + lx - SHALO, gy = y0 + ly - SHALO;
    const bool in = lx < SW && gx >= 0 && gx < W && gy >= 0 && gy < H;
#pragma unroll
    for (int q = 0; q < 3; ++q) sm[q][ly][lx] = in ? a.dmaps[(size_t)(q * 3 + c) * N + (size_t)gy * W + gx] : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < SW * 8; k += 256) {
    const int ly = k >> 3, lx0 = (k & 7) * 4;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float u[16];
      lds16(&sm[q][ly][lx0], u);
      float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < 14; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int t = j - i;
          if (t >= 0 && t < 11) o[i] += a.gauss[t] * u[j];
        }
      *reinterpret_cast<float4*>(&hz[q][ly][lx0]) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
  __syncthreads();
  const float inv3N = 1.f / (3.f * (float)N);
  const float inv_cnt = 1.f / stats[3];
  const int lx = threadIdx.x & 31, ly0 = (threadIdx.x >> 5) * 4;
  float o[3][4];
#pragma unroll
  for (int q = 0; q < 3; ++q)
#pragma unroll
    for (int i = 0; i < 4; ++i) o[q][i] = 0.f;
#pragma unroll
  for (int j = 0; j < 14; ++j)
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const float h = hz[q][ly0 + j][lx];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = j - i;
        if (t >= 0 && t < 11) o[q][i] += a.gauss[t] * h;
      }
    }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gx = x0 + lx, gy = y0 + ly0 + i;
    if (gx >= W || gy >= H) continue;
    const size_t p = (size_t)gy * W + gx;
    const float x = a.im[(size_t)c * N + p], y = a.gt_im[(size_t)c * N + p];
    const float dssim = o[0][i] + 2.f * x * o[1][i] + y * o[2][i];
    a.g_im[(size_t)c * N + p] = a.w_im * inv3N * (0.8f * sgn(x - y) - 0.2f * dssim);
    if (c == 0) {
      const float d = a.depth_sil[p], sil = a.depth_sil[N + p], d2 = a.depth_sil[2 * N + p], gd = a.gt_depth[p];
      const bool mk = loss_mask(gd, d, d2, sil, false, 0.f);
      a.g_depth_sil[p] = mk ? -a.w_depth * sgn(gd - d) * inv_cnt : 0.f;
      a.g_depth_sil[N + p] = 0.f;
      a.g_depth_sil[2 * N + p] = 0.f;
    }
  }
}

// One block: partials -> stats[0..3] (raw sums, double accumulation) and the weighted losses
// losses = (w_im * im, w_depth * depth, total, mask count)   (scripts/splatam.py:342-346)
__global__ void __launch_bounds__(256) loss_finalize_kernel(const LossArgs a, const float* __restrict__ partial, int nblk,
                                                            float* __restrict__ stats) {
  pdl_wait();
  pdl_trigger();
  __shared__ double S[8][LOSS_SUMS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double v[LOSS_SUMS] = {0.0, 0.0, 0.0, 0.0};
  const float4* p4 = reinterpret_cast<const float4*>(partial);   // one 16-byte row per block
  for (int b0 = threadIdx.x; b0 < nblk; b0 += 1024) {            // four independent row loads in flight per thread
    float4 r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int b = b0 + 256 * u; r[u] = b < nblk ? p4[b] : make_float4(0.f, 0.f, 0.f, 0.f); }
#pragma unroll
    for (int u = 0; u < 4; ++u) { v[0] += (double)r[u].x; v[1] += (double)r[u].y; v[2] += (double)r[u].z; v[3] += (double)r[u].w; }
  }
#pragma unroll
  for (int k = 0; k < LOSS_SUMS; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if (lane == 0) S[w][k] = v[k];
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double t[LOSS_SUMS];
  for (int k = 0; k < LOSS_SUMS; ++k) { t[k] = 0.0; for (int j = 0; j < 8; ++j) t[k] += S[j][k]; stats[k] = (float)t[k]; }
  double l_im, l_d, cnt;
  if (a.tracking) {
    l_im = t[0]; l_d = t[1]; cnt = t[2];
  } else {
    const double n3 = 3.0 * (double)a.W * (double)a.H;
    l_im = 0.8 * (t[1] / n3) + 0.2 * (1.0 - t[0] / n3);
    l_d = t[2] / t[3];   // mean over an empty mask is NaN in torch as well
    cnt = t[3];
  }
  const float wi = a.w_im * (float)l_im, wd = a.w_depth * (float)l_d;
  a.losses[0] = wi; a.losses[1] = wd; a.losses[2] = wi + wd; a.losses[3] = (float)cnt;
}

int slam_loss_blocks(int W, int H, int tracking) {
  if (tracking) {
    int nb = (W * H + 255) / 256;
    return nb > 148 * 8 ? 148 * 8 : (nb < 1 ? 1 : nb);
  }
  return ((W + ST - 1) / ST) * ((H + ST - 1) / ST) * 3;
}

int launch_slam_loss(const LossArgs& a, float* partial, float* stats, cudaStream_t st) {
  const int nb = slam_loss_blocks(a.W, a.H, a.tracking);
  if (a.tracking) {
    ProfScope prof_(K_LOSS, st);
    SPLAT_CUDA(launch_k(track_loss_kernel, dim3(nb), dim3(256), 0, st, a, partial));
    SPLAT_LAUNCH_CHECK("track_loss_kernel");
    SPLAT_CUDA(launch_k(loss_finalize_kernel, dim3(1), dim3(256), 0, st, a, (const float*)partial, nb, stats));
    SPLAT_LAUNCH_CHECK("loss_finalize_kernel");
    return SPLAT_OK;
  }
  ProfScope prof_(K_LOSS, st);
  dim3 grid((a.W + ST - 1) / ST, (a.H + ST - 1) / ST, 3);
  SPLAT_CUDA(launch_k(map_loss_fwd_kernel, grid, dim3(256), 0, st, a, partial));
  SPLAT_LAUNCH_CHECK("map_loss_fwd_kernel");
  SPLAT_CUDA(launch_k(loss_finalize_kernel, dim3(1), dim3(256), 0, st, a, (const float*)partial, nb, stats));
  SPLAT_LAUNCH_CHECK("loss_finalize_kernel");
  SPLAT_CUDA(launch_k(map_loss_bwd_kernel, grid, dim3(256), 0, st, a, (const float*)stats));
  SPLAT_LAUNCH_CHECK("map_loss_bwd_kernel");
  return SPLAT_OK;
}

// ================================================================================================ Adam
// One launch for every parameter tensor of the optimizer (SplaTAM: 5 Gaussian tensors + 2 pose tensors,
// scripts/splatam.py:166-172 builds one param group per tensor, each with its own lr): torch.optim.Adam's foreach path
// runs ~8 multi-tensor kernels and several hundred microseconds of host code per step.  Arithmetic follows
// torch/optim/adam.py::_single_tensor_adam (lerp, mul+addcmul, sqrt/bias2_sqrt + eps, addcdiv).
// HBM roofline: 16 B read + 12 B written per element.
// capturable mode: bias corrections for step[i] + *replay_counter, formed on the device in double; one block
__global__ void adam_prep_kernel(const AdamArgs a) {
  pdl_wait();
  pdl_trigger();
  const int i = threadIdx.x;
  const int replays = *a.replay_counter;
  if (i < a.count) {
    const double t = (double)(a.step[i] + replays);
    a.dev_bias[2 * i] = (float)(1.0 - pow(a.beta1d, t));
    a.dev_bias[2 * i + 1] = (float)sqrt(1.0 - pow(a.beta2d, t));
  }
  __syncthreads();
  if (i == 0) *a.replay_counter = replays + 1;
}

__global__ void forward_guard_kernel(const ImageHeader* hdr, uint32_t Rcap, uint32_t* flags) {
  pdl_wait();
  pdl_trigger();
  const uint32_t R = hdr->num_rendered;
  if (R > Rcap) flags[0] = 1u;
  if (R > flags[1]) flags[1] = R;
}

__global__ void __launch_bounds__(256) adam_kernel(const AdamArgs a) {
  pdl_wait();
  pdl_trigger();
  int t = 0;
  int b = (int)blockIdx.x;
#pragma unroll 1
  while (t < a.count - 1 && b >= a.blocks[t]) { b -= a.blocks[t]; ++t; }
  const SplatAdamTensorK& T = a.t[t];
  const float bc1 = a.dev_bias ? a.dev_bias[2 * t] : T.bias_correction1;
  const float bc2s = a.dev_bias ? a.dev_bias[2 * t + 1] : T.bias_correction2_sqrt;
  const float lr1 = T.lr / bc1, inv_b2 = 1.f / bc2s;
  const float w1 = a.w1, w2 = a.w2;
  const int64_t base = (int64_t)b * 1024 + threadIdx.x * 4;
  if (base >= T.n) return;
  if (T.vec && base + 4 <= T.n) {
    const float4 g = *reinterpret_cast<const float4*>(T.grad + base);
    float4 m = *reinterpret_cast<float4*>(T.exp_avg + base);
    float4 v = *reinterpret_cast<float4*>(T.exp_avg_sq + base);
    float4 p = *reinterpret_cast<float4*>(T.param + base);
    adam_update(g.x, m.x, v.x, p.x, w1, w2, a.beta2, lr1, inv_b2, a.eps);
    adam_update(g.y, m.y, v.y, p.y, w1, w2, a.beta2, lr1, inv_b2, a.eps);
    adam_update(g.z, m.z, v.z, p.z, w1, w2, a.beta2, lr1, inv_b2, a.eps);
    adam_update(g.w, m.w, v.w, p.w, w1, w2, a.beta2, lr1, inv_b2, a.eps);
    *reinterpret_cast<float4*>(T.exp_avg + base) = m;
    *reinterpret_cast<float4*>(T.exp_avg_sq + base) = v;
    *reinterpret_cast<float4*>(T.param + base) = p;
  } else {
    for (int64_t i = base; i < T.n && i < base + 4; ++i) {
      float m = T.exp_avg[i], v = T.exp_avg_sq[i], pp = T.param[i];
      adam_update(T.grad[i], m, v, pp, w1, w2, a.beta2, lr1, inv_b2, a.eps);
      T.exp_avg[i] = m; T.exp_avg_sq[i] = v;
      T.param[i] = pp;
    }
  }
}

int launch_adam(AdamArgs& a, cudaStream_t st) {
  ProfScope prof_(K_ADAM, st);
  int total = 0;
  for (int i = 0; i < a.count; ++i) {
    a.blocks[i] = (int)((a.t[i].n + 1023) / 1024);
    total += a.blocks[i];
  }
  if (total == 0) return SPLAT_OK;
  if (a.replay_counter) {
    SPLAT_CUDA(launch_k(adam_prep_kernel, dim3(1), dim3(32), 0, st, a));
    SPLAT_LAUNCH_CHECK("adam_prep_kernel");
  }
  SPLAT_CUDA(launch_k(adam_kernel, dim3(total), dim3(256), 0, st, a));
  SPLAT_LAUNCH_CHECK("adam_kernel");
  return SPLAT_OK;
}

int launch_forward_guard(const ImageHeader* hdr, uint32_t Rcap, uint32_t* flags, cudaStream_t st) {
  SPLAT_CUDA(launch_k(forward_guard_kernel, dim3(1), dim3(1), 0, st, hdr, Rcap, flags));
  SPLAT_LAUNCH_CHECK("forward_guard_kernel");
  return SPLAT_OK;
}

}  // namespace splat

// ------------------------------------------------------------------------------------- per-iteration bookkeeping
// Two pieces of get_loss / the tracking loop that the reference spreads over ~10 tiny torch kernels per iteration
// (and, for the candidate pose, a blocking host read of the loss): scripts/splatam.py:348-350 and :726-731.
namespace splat {

__global__ void __launch_bounds__(256) seen_update_kernel(int P, const int* __restrict__ radii,
                                                          float* __restrict__ max_2D_radius, uint8_t* __restrict__ seen) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  const bool s = r > 0;
  if (seen) seen[i] = s ? 1 : 0;
  if (s) {                                       // max_2D_radius[seen] = max(radius[seen], max_2D_radius[seen])
    const float m = max_2D_radius[i], rf = (float)r;
    if (!(m >= rf)) max_2D_radius[i] = (m != m) ? m : rf;     // torch.max propagates NaN
  }
}

__global__ void best_pose_kernel(const float* __restrict__ loss, float* min_loss, const float* __restrict__ cam_unnorm_rots,
                                 const float* __restrict__ cam_trans, int n_frames, int time_idx, float* best_rot,
                                 float* best_tran) {
  __shared__ int better;
  if (threadIdx.x == 0) {
    const float l = *loss;
    better = l < *min_loss;
    if (better) *min_loss = l;
  }
  __syncthreads();
  if (!better) return;
  if (threadIdx.x < 4) best_rot[threadIdx.x] = cam_unnorm_rots[threadIdx.x * n_frames + time_idx];
  else if (threadIdx.x < 7) best_tran[threadIdx.x - 4] = cam_trans[(threadIdx.x - 4) * n_frames + time_idx];
}

// grad *= *scale, skipped entirely when *scale == 1 (loss.backward() on the loss itself): every block reads the
// scalar first and exits before touching the gradient image.
__global__ void __launch_bounds__(256) scale_inplace_kernel(float4* __restrict__ x4, float* __restrict__ x, int64_t n,
                                                            const float* __restrict__ scale) {
  const float s = *scale;
  if (s == 1.0f) return;
  const int64_t n4 = x4 ? n / 4 : 0;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    float4 v = x4[i];
    v.x *= s; v.y *= s; v.z *= s; v.w *= s;
    x4[i] = v;
  }
  for (int64_t i = n4 * 4 + (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) x[i] *= s;
}

}  // namespace splat

extern "C" {

int splat_scale_inplace(float* x, int64_t n, const float* scale, void* stream) {
  using namespace splat;
  if (n < 0) return fail(SPLAT_ERR_BAD_ARG, "n < 0");
  if (n == 0) return SPLAT_OK;
  if (!x || !scale) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  const bool vec = (((uintptr_t)x) & 15) == 0;
  const int64_t work = vec ? (n + 3) / 4 : n;
  int blocks = (int)((work + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  scale_inplace_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(vec ? reinterpret_cast<float4*>(x) : nullptr, x, n, scale);
  SPLAT_LAUNCH_CHECK("scale_inplace_kernel");
  return SPLAT_OK;
}

int splat_seen_update(int32_t P, const int32_t* radii, float* max_2D_radius, uint8_t* seen, void* stream) {
  using namespace splat;
  if (P < 0) return fail(SPLAT_ERR_BAD_ARG, "P < 0");
  if (P == 0) return SPLAT_OK;
  if (!radii || !max_2D_radius) return fail(SPLAT_ERR_BAD_ARG, "radii / max_2D_radius is NULL");
  seen_update_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, radii, max_2D_radius, seen);
  SPLAT_LAUNCH_CHECK("seen_update_kernel");
  return SPLAT_OK;
}

int splat_best_pose_update(const float* loss, float* min_loss, const float* cam_unnorm_rots, const float* cam_trans,
                           int32_t n_frames, int32_t time_idx, float* best_rot, float* best_tran, void* stream) {
  using namespace splat;
  if (!loss || !min_loss || !cam_unnorm_rots || !cam_trans || !best_rot || !best_tran)
    return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  if (n_frames <= 0 || time_idx < 0 || time_idx >= n_frames) return fail(SPLAT_ERR_BAD_ARG, "time_idx outside [0, n_frames)");
  best_pose_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(loss, min_loss, cam_unnorm_rots, cam_trans, n_frames, time_idx,
                                                      best_rot, best_tran);
  SPLAT_LAUNCH_CHECK("best_pose_kernel");
  return SPLAT_OK;
}

}  // extern "C"
