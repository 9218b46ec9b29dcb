// api.cu -- the extern "C" boundary declared in include/splat_b200.h.
// Argument validation, state carving and kernel sequencing only; no device allocation.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "kernels.cuh"

namespace splat {

std::atomic<int> g_launch_count{0};
int g_pdl = 1;
int g_prof_on = 0;
struct ProfPair { cudaEvent_t a, b, prev; };   // prev = end event of the launch before this one (idle-gap accounting)
static cudaEvent_t g_prof_last_end = nullptr;
static float g_prof_gap[K_COUNT];
static std::vector<ProfPair> g_prof[K_COUNT];
static std::vector<cudaEvent_t> g_prof_free;
static cudaEvent_t prof_new_event() {
  if (!g_prof_free.empty()) { cudaEvent_t e = g_prof_free.back(); g_prof_free.pop_back(); return e; }
  cudaEvent_t e; cudaEventCreate(&e); return e;
}
void prof_record(int id, cudaStream_t st, bool start) {
  if (start) {
    ProfPair p; p.a = prof_new_event(); p.b = prof_new_event(); p.prev = g_prof_last_end;
    cudaEventRecord(p.a, st); g_prof[id].push_back(p);
  } else if (!g_prof[id].empty()) {
    cudaEventRecord(g_prof[id].back().b, st);
    g_prof_last_end = g_prof[id].back().b;
  }
}
static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return SPLAT_OK;
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return SPLAT_ERR_CUDA_BASE + (int)e;
}

static int check_cam(const SplatCamera* cam) {
  if (!cam) return fail(SPLAT_ERR_BAD_ARG, "camera is NULL");
  if (cam->image_width <= 0 || cam->image_height <= 0)
    return fail(SPLAT_ERR_BAD_ARG, "bad image size %dx%d", cam->image_width, cam->image_height);
  if (!cam->viewmatrix || !cam->projmatrix || !cam->bg)
    return fail(SPLAT_ERR_BAD_ARG, "camera viewmatrix/projmatrix/bg must be device pointers");
  return SPLAT_OK;
}

static void read_env_once() {
  static bool done = false;
  if (done) return;
  done = true;
  const char* e = getenv("SPLAT_B200_NO_CULL");
  if (e && e[0] == '1') g_disable_cull = 1;
  e = getenv("SPLAT_B200_PDL");      // "0": plain stream-ordered launches (no programmatic dependent launch)
  if (e) g_pdl = e[0] != '0';
  e = getenv("SPLAT_B200_GATHER");   // "0": blend kernels stream the re-written instance records (emit_records + TMA)
  if (e) g_gather = e[0] != '0';
}

static const bool g_env_read_at_load = (read_env_once(), true);   // switches apply to every entry point from the first call

}  // namespace splat

using namespace splat;

extern "C" {

size_t splat_geom_state_bytes(int32_t P) { return geom_bytes(P); }
size_t splat_image_state_bytes(int32_t W, int32_t H) { return image_bytes(W, H); }
size_t splat_binning_state_bytes(int32_t, int32_t, int64_t R) { return bin_bytes(R); }

int splat_forward_geometry(const SplatCamera* cam, int32_t P, const float* means3D, const float* scales,
                           const float* rotations, const float* opacities, const float* cov3D_precomp,
                           const float* shs, int32_t M, const float* colors_precomp, void* geom_state,
                           void* image_state, int32_t* radii, uint32_t* num_rendered_dev, void* stream) {
  read_env_once();
  if (int rc = check_cam(cam)) return rc;
  if (P < 0) return fail(SPLAT_ERR_BAD_ARG, "P < 0");
  if (!geom_state || !image_state || !num_rendered_dev) return fail(SPLAT_ERR_BAD_ARG, "state buffers must be non-NULL");
  if (P > 0) {
    if (!means3D || !opacities || !radii) return fail(SPLAT_ERR_BAD_ARG, "means3D/opacities/radii must be non-NULL");
    const bool sr = scales && rotations;
    if ((sr && cov3D_precomp) || (!sr && !cov3D_precomp) || ((scales != nullptr) != (rotations != nullptr)))
      return fail(SPLAT_ERR_BAD_ARG, "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    if ((shs && colors_precomp) || (!shs && !colors_precomp))
      return fail(SPLAT_ERR_BAD_ARG, "Please provide excatly one of either SHs or precomputed colors!");
    if (rotations && ((uintptr_t)rotations & 15)) return fail(SPLAT_ERR_BAD_ARG, "rotations must be 16-byte aligned");
    if (shs) {
      if (cam->sh_degree < 0 || cam->sh_degree > 3) return fail(SPLAT_ERR_UNSUPPORTED, "sh_degree must be 0..3");
      if (M < (cam->sh_degree + 1) * (cam->sh_degree + 1)) return fail(SPLAT_ERR_BAD_ARG, "M too small for sh_degree");
      if (!cam->campos) return fail(SPLAT_ERR_BAD_ARG, "campos required with SHs");
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int W = cam->image_width, H = cam->image_height;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, T = gx * gy;
  ImageView iv = image_view(image_state, W, H);
  GeomView gv = geom_view(geom_state, P);
  SPLAT_CUDA(cudaMemsetAsync(iv.tile_count, 0, (size_t)T * 4 * CSTRIDE, st));
  PreFwdArgs a;
  a.P = P; a.W = W; a.H = H; a.gx = gx; a.gy = gy; a.M = M; a.sh_degree = cam->sh_degree;
  a.tanx = cam->tanfovx; a.tany = cam->tanfovy;
  a.fx = W / (2.0f * cam->tanfovx); a.fy = H / (2.0f * cam->tanfovy);
  a.mod = cam->scale_modifier; a.near_plane = cam->near_plane;
  a.means = means3D; a.scales = scales; a.rots = rotations; a.opac = opacities;
  a.cov3D_pre = cov3D_precomp; a.shs = shs; a.colors_pre = colors_precomp;
  a.V = cam->viewmatrix; a.Pm = cam->projmatrix; a.campos = cam->campos;
  a.g = gv; a.tile_count = iv.tile_count; a.radii = radii;
  if (int rc = launch_preprocess_fwd(a, st)) return rc;
  return launch_tile_scan(T, iv, num_rendered_dev, (uint32_t)sort_smem_cap(), st);
}

int splat_forward_render(const SplatCamera* cam, int32_t P, const float* colors_precomp, const int32_t* radii,
                         const void* geom_state,
                         void* image_state, void* binning_state, int64_t R_capacity, float* out_color,
                         float* out_depth, void* stream) {
  read_env_once();
  (void)colors_precomp;  // already folded into the geometry records by phase 1
  if (int rc = check_cam(cam)) return rc;
  if (!geom_state || !image_state || !binning_state || !out_color) return fail(SPLAT_ERR_BAD_ARG, "NULL buffer");
  if (R_capacity < 0 || R_capacity > 0xffffffffll) return fail(SPLAT_ERR_BAD_ARG, "R_capacity out of range");
  cudaStream_t st = (cudaStream_t)stream;
  const int W = cam->image_width, H = cam->image_height;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, T = gx * gy;
  ImageView iv = image_view(image_state, W, H);
  GeomView gv = geom_view(const_cast<void*>(geom_state), P);
  BinView bv = bin_view(binning_state, R_capacity);
  const uint32_t Rcap = (uint32_t)R_capacity;
  if (P > 0 && !radii) return fail(SPLAT_ERR_BAD_ARG, "radii is NULL");
  if (int rc = launch_scatter(P, gx, gy, radii, gv, iv, bv, Rcap, st)) return rc;
  if (int rc = launch_tile_sort(T, iv, gv, bv, Rcap, st)) return rc;
  RenderFwdArgs r;
  r.W = W; r.H = H; r.gx = gx;
  r.ranges = iv.ranges; r.inst = bv.inst; r.order = iv.tile_order;
  r.bg = cam->bg;
  r.final_T = iv.final_T; r.n_contrib = iv.n_contrib; r.out_color = out_color; r.out_depth = out_depth;
  r.hdr = iv.hdr; r.Rcap = Rcap; r.blended = nullptr; r.depth_sil = cam->depth_sil ? 1 : 0; r.zero = 0;
  r.grec = gv.grec; r.plist = bv.point_list; r.P = P;
  return launch_render_fwd(r, T, st);
}


size_t splat_backward_scratch_bytes(int32_t P, int32_t power) {
  return align_up((size_t)(P > 0 ? P : 1) * (power == 1 ? 12 : 20) * sizeof(float));
}

int splat_backward(const SplatCamera* cam, int32_t P, int64_t R, const float* means3D, const float* scales,
                   const float* rotations, const float* cov3D_precomp, const float* shs, int32_t M,
                   const float* colors_precomp, const int32_t* radii, const void* geom_state,
                   const void* image_state, const void* binning_state, const float* dL_dout_color,
                   const float* out_color, int32_t power, int32_t fused, float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors,
                   float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                   float* dL_dconic, float* dL_dsh, void* scratch, size_t scratch_bytes, void* stream) {
  read_env_once();
  (void)colors_precomp;  // the colours travel inside the instance records written by the forward pass
  if (int rc = check_cam(cam)) return rc;
  if (P < 0 || R < 0) return fail(SPLAT_ERR_BAD_ARG, "P or R negative");
  if (P == 0) return SPLAT_OK;
  if (!means3D || !radii || !geom_state || !image_state || !binning_state || !dL_dout_color)
    return fail(SPLAT_ERR_BAD_ARG, "NULL input buffer");
  if (!dL_dmeans3D) return fail(SPLAT_ERR_BAD_ARG, "dL_dmeans3D must be non-NULL");
  const bool all_out = dL_dmeans2D && dL_dcolors && dL_dopacity && dL_dscales && dL_drotations && dL_dcov3D && dL_dconic;
  if (!all_out && (fused && power != 1))
    return fail(SPLAT_ERR_BAD_ARG, "the power-2 backward writes every gradient array: NULL gradient buffer");
  if (!all_out && shs) return fail(SPLAT_ERR_BAD_ARG, "the SH backward needs every gradient array");
  if (shs && !dL_dsh) return fail(SPLAT_ERR_BAD_ARG, "dL_dsh required with SHs");
  const bool sr = scales && rotations;
  if (!sr && !cov3D_precomp) return fail(SPLAT_ERR_BAD_ARG, "need scales+rotations or cov3D_precomp");
  if ((rotations && ((uintptr_t)rotations & 15)) || ((uintptr_t)dL_drotations & 15) || ((uintptr_t)dL_dconic & 15))
    return fail(SPLAT_ERR_BAD_ARG, "rotations, dL_drotations and dL_dconic must be 16-byte aligned");
  const bool hess = fused && power != 1;
  if (hess && power != 2)
    return fail(SPLAT_ERR_UNSUPPORTED, "backward_power must be 1 or 2 (the reference only ever uses those, ros_handler.py:884)");
  if (hess && shs) return fail(SPLAT_ERR_UNSUPPORTED, "SH + Hessian backward is undefined in the reference (backward.cu:1067)");
  if (hess && cam->depth_sil) return fail(SPLAT_ERR_UNSUPPORTED, "the fused depth/silhouette channels support backward_power 1 only");
  const size_t need = splat_backward_scratch_bytes(P, hess ? 2 : 1);
  if (!scratch || ((uintptr_t)scratch & 15)) return fail(SPLAT_ERR_BAD_ARG, "scratch must be a 16-byte aligned device buffer");
  if (scratch_bytes < need)
    return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "backward scratch needs %zu bytes, got %zu", need, scratch_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  const int W = cam->image_width, H = cam->image_height;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, T = gx * gy;
  ImageView iv = image_view(const_cast<void*>(image_state), W, H);
  GeomView gv = geom_view(const_cast<void*>(geom_state), P);
  BinView bv = bin_view(const_cast<void*>(binning_state), R);
  const float fx = W / (2.0f * cam->tanfovx), fy = H / (2.0f * cam->tanfovy);
  float* acc = reinterpret_cast<float*>(scratch);
  SPLAT_CUDA(cudaMemsetAsync(acc, 0, need, st));

  RenderBwdArgs r;
  r.W = W; r.H = H; r.gx = gx; r.order = iv.tile_order; r.ranges = iv.ranges; r.inst = bv.inst; r.bg = cam->bg;
  r.final_T = iv.final_T; r.n_contrib = iv.n_contrib; r.dL_dpix = dL_dout_color; r.out_color = out_color;
  r.acc = acc;
  r.hdr = iv.hdr; r.Rcap = (uint32_t)(R > 0xffffffffll ? 0xffffffffll : R); r.zero = 0;
  r.grec = gv.grec; r.plist = bv.point_list; r.P = P;
  if (!hess) {
    // callers that need neither colour nor opacity gradients (pose-only tracking) get the blend backward without those sums
    const int mode = cam->depth_sil ? ((!dL_dcolors && !dL_dopacity) ? 3 : 2) : 0;
    if (R > 0) if (int rc = launch_render_bwd(r, T, mode, st)) return rc;
    PreBwdArgs b;
    b.depth_sil = cam->depth_sil ? 1 : 0;
    b.P = P; b.M = M; b.sh_degree = cam->sh_degree; b.tanx = cam->tanfovx; b.tany = cam->tanfovy;
    b.fx = fx; b.fy = fy; b.mod = cam->scale_modifier;
    b.means = means3D; b.scales = sr ? scales : nullptr; b.rots = sr ? rotations : nullptr;
    b.cov3D_pre = cov3D_precomp; b.shs = shs; b.clamped = gv.clamped; b.campos = cam->campos;
    b.V = cam->viewmatrix; b.Pm = cam->projmatrix; b.radii = radii;
    b.W = W; b.H = H; b.grec = gv.grec; b.acc = acc;
    b.dL_dmean2D = dL_dmeans2D; b.dL_dconic = dL_dconic; b.dL_dcolor = dL_dcolors; b.dL_dopacity = dL_dopacity;
    b.dL_dmeans = dL_dmeans3D; b.dL_dcov3D = dL_dcov3D; b.dL_dscales = dL_dscales; b.dL_drots = dL_drotations;
    b.dL_dsh = dL_dsh;
    return launch_preprocess_bwd(b, st);
  }
  if (R > 0) if (int rc = launch_render_bwd(r, T, 1, st)) return rc;
  HessArgs h;
  h.P = P; h.tanx = cam->tanfovx; h.tany = cam->tanfovy; h.fx = fx; h.fy = fy; h.mod = cam->scale_modifier;
  h.W = W; h.H = H;
  h.means = means3D; h.scales = sr ? scales : nullptr; h.rots = sr ? rotations : nullptr; h.cov3D_pre = cov3D_precomp;
  h.V = cam->viewmatrix; h.Pm = cam->projmatrix; h.radii = radii; h.acc = acc; h.grec = gv.grec;
  h.dL_dmeans = dL_dmeans3D; h.dL_dmean2D = dL_dmeans2D; h.dL_dcolors = dL_dcolors; h.dL_dopacity = dL_dopacity;
  h.dL_dscales = dL_dscales; h.dL_drots = dL_drotations; h.dL_dcov3D = dL_dcov3D; h.dL_dconic = dL_dconic;
  return launch_hessian_finalize(h, st);
}

int splat_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       float near_plane, uint8_t* present, void* stream) {
  (void)projmatrix;  // the reference computes p_proj but only tests view-space z (auxiliary.h:150-154)
  if (P < 0) return fail(SPLAT_ERR_BAD_ARG, "P < 0");
  if (P == 0) return SPLAT_OK;
  if (!means3D || !viewmatrix || !present) return fail(SPLAT_ERR_BAD_ARG, "NULL buffer");
  return launch_mark_visible(P, means3D, viewmatrix, near_plane, present, (cudaStream_t)stream);
}

const void* splat_state_ranges(const void* s, int32_t W, int32_t H) { return image_view(const_cast<void*>(s), W, H).ranges; }
const void* splat_state_n_contrib(const void* s, int32_t W, int32_t H) { return image_view(const_cast<void*>(s), W, H).n_contrib; }
const void* splat_state_final_T(const void* s, int32_t W, int32_t H) { return image_view(const_cast<void*>(s), W, H).final_T; }
const void* splat_state_point_list(const void* s, int32_t, int32_t, int64_t R) { return bin_view(const_cast<void*>(s), R).point_list; }
const void* splat_state_gaussian_records(const void* s, int32_t P) { return geom_view(const_cast<void*>(s), P).grec; }
const void* splat_state_instance_records(const void* s, int32_t, int32_t, int64_t R) { return bin_view(const_cast<void*>(s), R).inst; }

// Debug/measurement helper: re-runs the forward blend over an existing (geom, image, binning) state
// WITHOUT culling and writes the number of blended pairs per pixel (F_alg of the roofline model).
int splat_debug_count_blended(const SplatCamera* cam, int32_t P, const float* colors, const void* geom_state,
                              void* image_state, const void* binning_state, int64_t R_capacity,
                              float* scratch_color, uint32_t* blended, void* stream) {
  if (int rc = check_cam(cam)) return rc;
  if (!colors || !geom_state || !image_state || !binning_state || !scratch_color || !blended)
    return fail(SPLAT_ERR_BAD_ARG, "NULL buffer");
  const int W = cam->image_width, H = cam->image_height;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
  ImageView iv = image_view(image_state, W, H);
  GeomView gv = geom_view(const_cast<void*>(geom_state), P);
  BinView bv = bin_view(const_cast<void*>(binning_state), R_capacity);
  RenderFwdArgs r;
  (void)colors;
  r.W = W; r.H = H; r.gx = gx; r.order = iv.tile_order; r.ranges = iv.ranges; r.inst = bv.inst; r.bg = cam->bg;
  r.final_T = iv.final_T; r.n_contrib = iv.n_contrib; r.out_color = scratch_color; r.out_depth = nullptr;
  r.hdr = iv.hdr; r.Rcap = (uint32_t)R_capacity; r.blended = blended; r.depth_sil = 0; r.zero = 0;
  r.grec = gv.grec; r.plist = bv.point_list; r.P = P;
  return launch_render_fwd(r, gx * gy, (cudaStream_t)stream);
}

__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f,
        a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float b = 1.0000001f, c = 1e-7f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      a0 = fmaf(a0, b, c); a1 = fmaf(a1, b, c); a2 = fmaf(a2, b, c); a3 = fmaf(a3, b, c);
      a4 = fmaf(a4, b, c); a5 = fmaf(a5, b, c); a6 = fmaf(a6, b, c); a7 = fmaf(a7, b, c);
    }
  }
  if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 123.456f) out[0] = a0;
}

// Non-tensor FP32 peak of this device (FFMA microbenchmark, 2 FLOP per FMA): the roofline
// denominator of the blend kernels.  MEASURED_PEAKS.json carries HBM and bf16-tensor peaks only.
int splat_debug_fp32_peak(float* scratch_dev, float* tflops_out, void* stream) {
  if (!scratch_dev || !tflops_out) return fail(SPLAT_ERR_BAD_ARG, "NULL buffer");
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 0;
  SPLAT_CUDA(cudaGetDevice(&dev));
  SPLAT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int iters = 20000, blocks = sms * 8;
  cudaEvent_t e0, e1;
  SPLAT_CUDA(cudaEventCreate(&e0)); SPLAT_CUDA(cudaEventCreate(&e1));
  float best = 0.f;
  for (int rep = 0; rep < 4; ++rep) {
    SPLAT_CUDA(cudaEventRecord(e0, st));
    fp32_peak_kernel<<<blocks, 256, 0, st>>>(scratch_dev, iters);
    SPLAT_CUDA(cudaEventRecord(e1, st));
    SPLAT_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SPLAT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double flop = 2.0 * 64.0 * iters * 256.0 * blocks;
    const float tf = (float)(flop / (ms * 1e-3) / 1e12);
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *tflops_out = best;
  return SPLAT_OK;
}

int splat_debug_exp_mismatches(uint32_t first_bits, uint64_t count, unsigned long long* mismatches_dev, void* stream) {
  if (!mismatches_dev) return fail(SPLAT_ERR_BAD_ARG, "mismatches_dev is NULL");
  if (count > (1ull << 32)) return fail(SPLAT_ERR_BAD_ARG, "count exceeds the 2^32 float bit patterns");
  return launch_exp_check(first_bits, count, mismatches_dev, (cudaStream_t)stream);
}

int splat_set_gather(int on) {
  read_env_once();
  const int prev = g_gather;
  g_gather = on ? 1 : 0;
  return prev;
}

int splat_profile_enable(int on) { g_prof_on = on ? 1 : 0; return SPLAT_OK; }

// Synchronises the device, sums the event-pair durations recorded since the last read and resets.
int splat_profile_read(float* ms_sum, int32_t* launches, int32_t n) {
  if (!ms_sum || !launches || n < K_COUNT) return fail(SPLAT_ERR_BAD_ARG, "need arrays of %d", (int)K_COUNT);
  SPLAT_CUDA(cudaDeviceSynchronize());
  for (int k = 0; k < K_COUNT; ++k) {  // gaps first: they reference other pairs' end events
    float gap = 0.f;
    for (auto& p : g_prof[k]) {
      float ms = 0.f;
      if (p.prev && cudaEventElapsedTime(&ms, p.prev, p.a) == cudaSuccess) gap += ms;
    }
    g_prof_gap[k] = gap;
  }
  for (int k = 0; k < K_COUNT; ++k) {
    float tot = 0.f;
    for (auto& p : g_prof[k]) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) tot += ms;
      g_prof_free.push_back(p.a); g_prof_free.push_back(p.b);
    }
    ms_sum[k] = tot; launches[k] = (int32_t)g_prof[k].size();
    g_prof[k].clear();
  }
  g_prof_last_end = nullptr;
  cudaGetLastError();
  return SPLAT_OK;
}
// Idle time on the stream BEFORE each kernel (previous profiled launch's end -> this launch's start),
// summed per kernel id over the pairs consumed by the last splat_profile_read.  Includes whatever
// foreign work (torch kernels, copies) ran in between.
int splat_profile_read_gaps(float* gap_ms_sum, int32_t n) {
  if (!gap_ms_sum || n < K_COUNT) return fail(SPLAT_ERR_BAD_ARG, "need an array of %d", (int)K_COUNT);
  for (int k = 0; k < K_COUNT; ++k) gap_ms_sum[k] = g_prof_gap[k];
  return SPLAT_OK;
}
const char* splat_profile_kernel_name(int k) {
  static const char* names[K_COUNT] = {"preprocess_fwd", "tile_scan", "scatter", "tile_sort", "render_fwd",
                                       "render_bwd", "preprocess_bwd", "hessian_finalize", "emit_records",
                                       "rendervar_fwd", "rendervar_bwd", "slam_loss", "adam", "map_surgery"};
  return (k >= 0 && k < K_COUNT) ? names[k] : "";
}

// ---- SplaTAM glue -------------------------------------------------------------
static int check_rv(int32_t P, int32_t scale_dim, const float* means3D, const float* unnorm_rotations,
                    const float* logit_opacities, const float* log_scales, const float* cam_unnorm_rots,
                    const float* cam_trans, int32_t n_frames, int32_t time_idx) {
  if (P < 0) return fail(SPLAT_ERR_BAD_ARG, "P < 0");
  if (scale_dim != 1 && scale_dim != 3) return fail(SPLAT_ERR_BAD_ARG, "log_scales must have 1 or 3 columns, got %d", scale_dim);
  if (!cam_unnorm_rots || !cam_trans) return fail(SPLAT_ERR_BAD_ARG, "cam_unnorm_rots/cam_trans must be non-NULL");
  if (n_frames <= 0 || time_idx < 0 || time_idx >= n_frames)
    return fail(SPLAT_ERR_BAD_ARG, "time_idx %d out of range for %d frames", time_idx, n_frames);
  if (P > 0 && (!means3D || !unnorm_rotations || !logit_opacities || !log_scales))
    return fail(SPLAT_ERR_BAD_ARG, "means3D/unnorm_rotations/logit_opacities/log_scales must be non-NULL");
  if ((uintptr_t)unnorm_rotations & 15) return fail(SPLAT_ERR_BAD_ARG, "unnorm_rotations must be 16-byte aligned");
  return SPLAT_OK;
}

int splat_rendervar_forward(int32_t P, int32_t scale_dim, const float* means3D, const float* unnorm_rotations,
                            const float* logit_opacities, const float* log_scales, const float* cam_unnorm_rots,
                            const float* cam_trans, int32_t n_frames, int32_t time_idx, float* out_means3D,
                            float* out_rotations, float* out_opacities, float* out_scales, void* stream) {
  if (int rc = check_rv(P, scale_dim, means3D, unnorm_rotations, logit_opacities, log_scales, cam_unnorm_rots,
                        cam_trans, n_frames, time_idx)) return rc;
  if (P > 0 && (!out_means3D || !out_rotations || !out_opacities || !out_scales))
    return fail(SPLAT_ERR_BAD_ARG, "rendervar outputs must be non-NULL");
  if ((uintptr_t)out_rotations & 15) return fail(SPLAT_ERR_BAD_ARG, "out_rotations must be 16-byte aligned");
  RvArgs a = {};
  a.P = P; a.aniso = scale_dim == 3; a.n_frames = n_frames; a.t = time_idx;
  a.means = means3D; a.unnorm_rots = unnorm_rotations; a.logit_op = logit_opacities; a.log_scales = log_scales;
  a.cam_rots = cam_unnorm_rots; a.cam_trans = cam_trans;
  a.o_means = out_means3D; a.o_rots = out_rotations; a.o_op = out_opacities; a.o_scales = out_scales;
  return launch_rendervar_fwd(a, (cudaStream_t)stream);
}

size_t splat_rendervar_scratch_bytes(int32_t P) { return align_up((size_t)rendervar_bwd_blocks(P) * 16 * sizeof(float)); }

int splat_rendervar_backward(int32_t P, int32_t scale_dim, const float* means3D, const float* unnorm_rotations,
                             const float* logit_opacities, const float* log_scales, const float* cam_unnorm_rots,
                             const float* cam_trans, int32_t n_frames, int32_t time_idx,
                             const float* dL_dmeans3D_cam, const float* dL_drotations, const float* dL_dopacities,
                             const float* dL_dscales, int32_t gaussians_grad, int32_t camera_grad,
                             float* dL_dmeans3D, float* dL_dunnorm_rotations, float* dL_dlogit_opacities,
                             float* dL_dlog_scales, float* dL_dcam_unnorm_rots, float* dL_dcam_trans,
                             void* scratch, size_t scratch_bytes, void* stream) {
  if (int rc = check_rv(P, scale_dim, means3D, unnorm_rotations, logit_opacities, log_scales, cam_unnorm_rots,
                        cam_trans, n_frames, time_idx)) return rc;
  if (gaussians_grad && P > 0 && (!dL_dmeans3D || !dL_dunnorm_rotations))
    return fail(SPLAT_ERR_BAD_ARG, "gaussians_grad needs dL_dmeans3D and dL_dunnorm_rotations");
  if (camera_grad && (!dL_dcam_unnorm_rots || !dL_dcam_trans))
    return fail(SPLAT_ERR_BAD_ARG, "camera_grad needs dL_dcam_unnorm_rots and dL_dcam_trans");
  if (camera_grad && (!scratch || scratch_bytes < splat_rendervar_scratch_bytes(P)))
    return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "rendervar scratch needs %zu bytes", splat_rendervar_scratch_bytes(P));
  if (((uintptr_t)dL_drotations & 15) || ((uintptr_t)dL_dunnorm_rotations & 15))
    return fail(SPLAT_ERR_BAD_ARG, "rotation gradients must be 16-byte aligned");
  RvArgs a = {};
  a.P = P; a.aniso = scale_dim == 3; a.n_frames = n_frames; a.t = time_idx;
  a.means = means3D; a.unnorm_rots = unnorm_rotations; a.logit_op = logit_opacities; a.log_scales = log_scales;
  a.cam_rots = cam_unnorm_rots; a.cam_trans = cam_trans;
  a.g_means = dL_dmeans3D_cam; a.g_rots = dL_drotations; a.g_op = dL_dopacities; a.g_scales = dL_dscales;
  a.d_means = dL_dmeans3D; a.d_unnorm_rots = dL_dunnorm_rotations; a.d_logit = dL_dlogit_opacities;
  a.d_logscales = dL_dlog_scales; a.d_cam_rots = dL_dcam_unnorm_rots; a.d_cam_trans = dL_dcam_trans;
  return launch_rendervar_bwd(a, gaussians_grad != 0, camera_grad != 0, (float*)scratch, (cudaStream_t)stream);
}

// scratch = [partials: blocks x 4 floats][stats: 4 floats (256 B)][mapping only: 9 derivative planes]
static size_t loss_partial_bytes(int32_t W, int32_t H) {
  const int nb_t = slam_loss_blocks(W, H, 1), nb_m = slam_loss_blocks(W, H, 0);
  return align_up((size_t)(nb_t > nb_m ? nb_t : nb_m) * 4 * sizeof(float));
}
size_t splat_slam_loss_scratch_bytes(int32_t W, int32_t H, int32_t tracking) {
  if (W <= 0 || H <= 0) return 0;
  return loss_partial_bytes(W, H) + 256 + (tracking ? 0 : align_up((size_t)9 * W * H * sizeof(float)));
}

int splat_slam_loss(int32_t W, int32_t H, int32_t tracking, int32_t use_sil_for_loss, float sil_thres, float w_im,
                    float w_depth, const float* im, const float* depth_sil, const float* gt_im, const float* gt_depth,
                    float* losses, float* dL_dim, float* dL_ddepth_sil, void* scratch, size_t scratch_bytes,
                    void* stream) {
  if (W <= 0 || H <= 0) return fail(SPLAT_ERR_BAD_ARG, "bad image size %dx%d", W, H);
  if (!im || !depth_sil || !gt_im || !gt_depth || !losses || !dL_dim || !dL_ddepth_sil)
    return fail(SPLAT_ERR_BAD_ARG, "image / loss / gradient pointers must be non-NULL");
  if ((uintptr_t)scratch & 15) return fail(SPLAT_ERR_BAD_ARG, "loss scratch must be 16-byte aligned");
  if (!scratch || scratch_bytes < splat_slam_loss_scratch_bytes(W, H, tracking))
    return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "loss scratch needs %zu bytes", splat_slam_loss_scratch_bytes(W, H, tracking));
  LossArgs a = {};
  a.W = W; a.H = H; a.tracking = tracking != 0; a.use_sil = (tracking != 0) && (use_sil_for_loss != 0);
  a.sil_thres = sil_thres; a.w_im = w_im; a.w_depth = w_depth;
  {  // gaussian(11, 1.5) exactly as the reference builds it: float32 tensor of exp(...) divided by its float32 sum
    float g[11], sum = 0.f;
    for (int x = 0; x < 11; ++x) { g[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5)); sum += g[x]; }
    for (int x = 0; x < 11; ++x) a.gauss[x] = g[x] / sum;
  }
  a.im = im; a.depth_sil = depth_sil; a.gt_im = gt_im; a.gt_depth = gt_depth;
  a.losses = losses; a.g_im = dL_dim; a.g_depth_sil = dL_ddepth_sil;
  char* c = (char*)scratch;
  float* partial = (float*)c; c += loss_partial_bytes(W, H);
  float* stats = (float*)c;   c += 256;
  a.dmaps = tracking ? nullptr : (float*)c;
  return launch_slam_loss(a, partial, stats, (cudaStream_t)stream);
}

int splat_adam_step(const SplatAdamTensor* tensors, int32_t count, double beta1, double beta2, double eps,
                    int32_t* replay_counter, float* replay_scratch, void* stream) {
  if (count < 0 || count > ADAM_MAX_TENSORS) return fail(SPLAT_ERR_BAD_ARG, "count must be 0..%d", ADAM_MAX_TENSORS);
  if (count == 0) return SPLAT_OK;
  if (!tensors) return fail(SPLAT_ERR_BAD_ARG, "tensors is NULL");
  if (replay_counter && !replay_scratch) return fail(SPLAT_ERR_BAD_ARG, "replay_counter needs replay_scratch (32 device floats)");
  AdamArgs a = {};
  // (1 - beta) is formed in double like torch's Python scalars: 1 - 0.999f is off by 1.3e-5 relative
  a.count = 0; a.beta2 = (float)beta2; a.w1 = (float)(1.0 - beta1); a.w2 = (float)(1.0 - beta2); a.eps = (float)eps;
  a.beta1d = beta1; a.beta2d = beta2;
  a.replay_counter = replay_counter; a.dev_bias = replay_counter ? replay_scratch : nullptr;
  for (int i = 0; i < count; ++i) {
    const SplatAdamTensor& t = tensors[i];
    if (t.n < 0) return fail(SPLAT_ERR_BAD_ARG, "tensor %d: n < 0", i);
    if (t.n == 0) continue;
    if (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq) return fail(SPLAT_ERR_BAD_ARG, "tensor %d: NULL pointer", i);
    if (t.step < 1) return fail(SPLAT_ERR_BAD_ARG, "tensor %d: step must be >= 1", i);
    a.step[a.count] = t.step;
    SplatAdamTensorK& k = a.t[a.count++];
    k.param = t.param; k.grad = t.grad; k.exp_avg = t.exp_avg; k.exp_avg_sq = t.exp_avg_sq; k.n = t.n;
    k.lr = t.lr;
    k.bias_correction1 = (float)(1.0 - pow(beta1, (double)t.step));
    k.bias_correction2_sqrt = (float)sqrt(1.0 - pow(beta2, (double)t.step));
    k.vec = (((uintptr_t)t.param | (uintptr_t)t.grad | (uintptr_t)t.exp_avg | (uintptr_t)t.exp_avg_sq) & 15) == 0;
  }
  return launch_adam(a, (cudaStream_t)stream);
}

int splat_forward_guard(const void* image_state, int32_t W, int32_t H, int64_t R_capacity, uint32_t* flags, void* stream) {
  if (!image_state || !flags || W <= 0 || H <= 0 || R_capacity < 0) return fail(SPLAT_ERR_BAD_ARG, "bad guard arguments");
  ImageView iv = image_view(const_cast<void*>(image_state), W, H);
  return launch_forward_guard(iv.hdr, (uint32_t)(R_capacity > 0xffffffffll ? 0xffffffffll : R_capacity), flags,
                              (cudaStream_t)stream);
}

int splat_abi_version(void) { return SPLAT_ABI_VERSION; }
const char* splat_last_error(void) { return g_err; }
int splat_kernel_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }
const char* splat_strerror(int code) {
  switch (code) {
    case SPLAT_OK: return "ok";
    case SPLAT_ERR_BAD_ARG: return "bad argument";
    case SPLAT_ERR_SCRATCH_TOO_SMALL: return "scratch buffer too small";
    case SPLAT_ERR_UNSUPPORTED: return "unsupported configuration";
    default: return code >= SPLAT_ERR_CUDA_BASE ? "CUDA runtime error" : "unknown error";
  }
}

}  // extern "C"
