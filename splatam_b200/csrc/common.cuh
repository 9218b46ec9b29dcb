// common.cuh -- shared device helpers, state layouts and launch bookkeeping for libsplat_b200.
//
// sm_100a only.  The math helpers restate the per-Gaussian projection algebra of the reference
// (HX/cuda_rasterizer/forward.cu:74-152, auxiliary.h:41-77,139-163) with a small column-major
// mat3 type of our own (the reference uses glm; element [c][r] == column c, row r, SURVEY.md
// Appendix A).  Expression trees are kept in the reference's association order on purpose:
// radii / tile rects must come out bit-identical, and nvcc's FMA contraction follows the tree.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include "../../include/splat_b200.h"

#define TILE_X 16
#define TILE_Y 16
#define TILE_PIX 256
#define NUM_CH 3

namespace splat {

// ---------------------------------------------------------------- launch bookkeeping
// Process-wide state of the library, all of it measurement / configuration, none of it on the data path: the launch
// counter (atomic: entry points may be called from several host threads), the per-launch event profiler (off unless
// splat_profile_enable(1); NOT thread-safe, a bench-only facility), and the two switches read once from the environment
// (SPLAT_B200_NO_CULL, SPLAT_B200_GATHER / splat_set_gather).  Everything else an entry point touches is caller-owned.
extern std::atomic<int> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

// ---------------------------------------------------------------- per-kernel event profiling
enum KernelId { K_PRE_FWD = 0, K_TILE_SCAN, K_SCATTER, K_TILE_SORT, K_RENDER_FWD, K_RENDER_BWD, K_PRE_BWD,
                K_HESS_FIN, K_EMIT, K_RV_FWD, K_RV_BWD, K_LOSS, K_ADAM, K_SURGERY, K_COUNT };
extern int g_prof_on;
void prof_record(int id, cudaStream_t st, bool start);
struct ProfScope {  // cudaEvent pair around one launch, only when splat_profile_enable(1) was called
  int id; cudaStream_t st;
  ProfScope(int id_, cudaStream_t st_) : id(id_), st(st_) { if (g_prof_on) prof_record(id, st, true); }
  ~ProfScope() { if (g_prof_on) prof_record(id, st, false); }
};

// ---------------------------------------------------------------- programmatic dependent launch
// Every kernel of the library is launched with the programmatic-stream-serialization attribute (unless SPLAT_B200_PDL=0):
// its CTAs may be scheduled, and run up to pdl_wait(), while the previous kernel of the stream is still draining, so the
// ~2.6 us launch gap per kernel boundary (8 per render iteration, ~20 per fused SLAM iteration) overlaps the predecessor's
// tail.  pdl_wait() (griddepcontrol.wait) is the FIRST statement of every such kernel: nothing a predecessor wrote is read
// before it.  Works inside CUDA-graph capture (the edge is captured as a programmatic dependency).
extern int g_pdl;
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Early trigger: lets the NEXT kernel's CTAs be scheduled before this grid has finished (they still block in their own
// pdl_wait() until it has).  Measured at C3 (r02): 1050 it/s with the trigger (right after the wait in the short kernels,
// after the main loop in the blend kernels) vs 1052 without -- waiting CTAs take slots the draining grid could use -- so it
// is compiled out; the implicit trigger at grid completion keeps the launch-latency overlap (1052 vs 1034 it/s without PDL).
#ifndef SPLAT_PDL_TRIGGER
#define SPLAT_PDL_TRIGGER 0
#endif
__device__ __forceinline__ void pdl_trigger() {
#if SPLAT_PDL_TRIGGER
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- state layouts
__host__ __device__ inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Per-Gaussian forward geometry, one 48-byte record per Gaussian in the SAME layout as the sorted
// instance records below (the tile sort copies it and patches the id in):
//   q0 = (x, y, conic A, conic B)   q1 = (conic C, opacity, view depth, -)   q2 = (r, g, b, -)
// Only records of visible Gaussians (radii > 0) are written / ever read.
#define INST_F4 3
#define INST_BYTES 48

struct GeomView {
  float4* grec;           // [P * INST_F4]
  uint8_t* clamped;       // [P,3] SH clamp flags (SH path only)
};
inline size_t geom_bytes(int P) {
  size_t p = (size_t)(P > 0 ? P : 1);
  return align_up(p * INST_BYTES) + align_up(p * 3);
}
inline GeomView geom_view(void* base, int P) {
  size_t p = (size_t)(P > 0 ? P : 1);
  char* c = (char*)base;
  GeomView v;
  v.grec = (float4*)c;           c += align_up(p * INST_BYTES);
  v.clamped = (uint8_t*)c;
  return v;
}

struct ImageHeader {          // first 256 bytes of the image state
  uint32_t num_rendered;      // R, written by tile_scan
  uint32_t max_tile_count;    // longest tile list
  uint32_t big_tile_count;    // number of tiles over the smem sort capacity
  uint32_t pad[61];
};
// Per-tile atomic counters live CSTRIDE words (256 bytes) apart: the L2 atomic unit serialises
// same-sector traffic, and ~2.5M atomics onto T packed words (T/8 sectors) made preprocess and
// scatter atomic-bound (ncu r01a); one counter per 256-byte block removes the hot sectors.
#define CSTRIDE 64
struct ImageView {
  ImageHeader* hdr;
  uint32_t* tile_count;   // [T*CSTRIDE] instances per tile (atomics in preprocess), word t*CSTRIDE
  uint32_t* tile_cursor;  // [T*CSTRIDE] scatter cursors (start .. end), word t*CSTRIDE
  uint2* ranges;          // [T] (start,end) -- bit-exact vs identifyTileRanges
  float* final_T;         // [N]
  uint32_t* n_contrib;    // [N]
  uint32_t* tile_order;   // [T] tiles by descending list length (64 length classes): launch order of the blend kernels
};
inline int num_tiles(int W, int H) { return ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y); }
inline size_t image_bytes(int W, int H) {
  size_t T = (size_t)num_tiles(W, H), N = (size_t)W * H;
  return 256 + align_up(T * 4 * CSTRIDE) * 2 + align_up(T * 8) + align_up(N * 4) * 2 + align_up(T * 4);
}
inline ImageView image_view(void* base, int W, int H) {
  size_t T = (size_t)num_tiles(W, H), N = (size_t)W * H;
  char* c = (char*)base;
  ImageView v;
  v.hdr = (ImageHeader*)c;     c += 256;
  v.tile_count = (uint32_t*)c; c += align_up(T * 4 * CSTRIDE);
  v.tile_cursor = (uint32_t*)c;c += align_up(T * 4 * CSTRIDE);
  v.ranges = (uint2*)c;        c += align_up(T * 8);
  v.final_T = (float*)c;       c += align_up(N * 4);
  v.n_contrib = (uint32_t*)c;  c += align_up(N * 4);
  v.tile_order = (uint32_t*)c;
  return v;
}

// One sorted (tile, Gaussian) instance as the blend kernels consume it: 48 bytes, written once by
// the tile sort in list order so that a tile's list is ONE contiguous, 16-byte aligned byte range
// that the TMA unit can stream into shared memory (cp.async.bulk).
//   q0 = (x, y, conic A, conic B)   q1 = (conic C, opacity, depth, gaussian id bits)   q2 = (r, g, b, 0)
struct BinView {
  uint64_t* keys;        // [Rcap] (depth_bits << 32) | gaussian id, bucketed by tile (unsorted in-bucket)
  uint64_t* keys_alt;    // [Rcap] ping-pong buffer for lists that do not fit in shared memory
  uint32_t* point_list;  // [Rcap] sorted gaussian ids
  float4* inst;          // [Rcap * INST_F4] sorted instance records
};
inline size_t bin_bytes(int64_t Rcap) {
  size_t r = (size_t)(Rcap > 0 ? Rcap : 1);
  return align_up(r * 8) * 2 + align_up(r * 4) + align_up(r * INST_BYTES);
}
inline BinView bin_view(void* base, int64_t Rcap) {
  size_t r = (size_t)(Rcap > 0 ? Rcap : 1);
  char* c = (char*)base;
  BinView v;
  v.keys = (uint64_t*)c;      c += align_up(r * 8);
  v.keys_alt = (uint64_t*)c;  c += align_up(r * 8);
  v.point_list = (uint32_t*)c; c += align_up(r * 4);
  v.inst = (float4*)c;
  return v;
}

// ---------------------------------------------------------------- bounds checks (debug build)
// -DSPLAT_BOUNDS_CHECK: every index the blend kernels, the scatter and the tile sort form (Gaussian ids, list positions,
// ring / transpose-tile / bin addresses, accumulator rows) is checked on the device and a violation traps -- the
// parity suite is run against that build (scripts/gpu_bounds_check.sh, profiles/r02_bounds_check.txt).  It stands in for
// compute-sanitizer, which is closed on this pool.  Compiled out (zero instructions) in the product build.
#ifdef SPLAT_BOUNDS_CHECK
#define SPLAT_CHECK(cond) do { if (!(cond)) { printf("SPLAT_CHECK failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define SPLAT_CHECK(cond) do {} while (0)
#endif

// ---------------------------------------------------------------- small math
struct Mat3 {  // column-major: m[c][r]
  float m[3][3];
};
// a0*b0 + a1*b1 + a2*b2 with the roundings of the reference build: nvcc contracts glm's and
// auxiliary.h's three-term sums to fma(a2,b2, fma(a0,b0, rn(a1*b1))) (SASS of preprocessCUDA<3>,
// forward.cu:74-152).  Spelled with _rn intrinsics so no other contraction can be chosen here:
// conics then match the reference bit for bit, and with them every alpha >= 1/255 decision.
__device__ __forceinline__ float dot3_ref(float a0, float b0, float a1, float b1, float a2, float b2) {
  return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}
// R = A * B (math product), R[c][r] = sum_k A[k][r] * B[c][k].
__device__ __forceinline__ Mat3 mat_mul(const Mat3& A, const Mat3& B) {
  Mat3 R;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r)
      R.m[c][r] = dot3_ref(A.m[0][r], B.m[c][0], A.m[1][r], B.m[c][1], A.m[2][r], B.m[c][2]);
  return R;
}
__device__ __forceinline__ Mat3 mat_transpose(const Mat3& A) {
  Mat3 R;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[r][c];
  return R;
}

// point (x,y,z,1) through a "transposed" 4x4: rows 0..2 (HX auxiliary.h:58-66)
__device__ __forceinline__ float3 xform_point_4x3(const float3 p, const float* __restrict__ m) {
  float3 t;
  t.x = __fadd_rn(dot3_ref(m[0], p.x, m[4], p.y, m[8], p.z), m[12]);
  t.y = __fadd_rn(dot3_ref(m[1], p.x, m[5], p.y, m[9], p.z), m[13]);
  t.z = __fadd_rn(dot3_ref(m[2], p.x, m[6], p.y, m[10], p.z), m[14]);
  return t;
}
__device__ __forceinline__ float4 xform_point_4x4(const float3 p, const float* __restrict__ m) {
  float4 t;
  t.x = __fadd_rn(dot3_ref(m[0], p.x, m[4], p.y, m[8], p.z), m[12]);
  t.y = __fadd_rn(dot3_ref(m[1], p.x, m[5], p.y, m[9], p.z), m[13]);
  t.z = __fadd_rn(dot3_ref(m[2], p.x, m[6], p.y, m[10], p.z), m[14]);
  t.w = __fadd_rn(dot3_ref(m[3], p.x, m[7], p.y, m[11], p.z), m[15]);
  return t;
}

// NDC -> pixel; the reference evaluates this in double (auxiliary.h:41-44) and we must too.
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle touched by a splat of integer radius r centred at p (auxiliary.h:46-56).
__device__ __forceinline__ void tile_rect(const float2 p, int r, int gx, int gy, int& x0, int& y0,
                                          int& x1, int& y1) {
  x0 = min(gx, max(0, (int)((p.x - r) / TILE_X)));
  y0 = min(gy, max(0, (int)((p.y - r) / TILE_Y)));
  x1 = min(gx, max(0, (int)((p.x + r + TILE_X - 1) / TILE_X)));
  y1 = min(gy, max(0, (int)((p.y + r + TILE_Y - 1) / TILE_Y)));
}

// Sigma = (S R)^T (S R) from scale*mod and UN-normalised quaternion (w,x,y,z)
// (forward.cu:118-152).  Output: upper triangle xx,xy,xz,yy,yz,zz.
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 s, float mod, const float4 q,
                                                     float* cov) {
  const float sx = __fmul_rn(s.x, mod), sy = __fmul_rn(s.y, mod), sz = __fmul_rn(s.z, mod);
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  // Rotation entries with the reference build's roundings: the shared product of each +/- pair is
  // rounded, the other one is fused; y*y + z*z is two rounded products.
  const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
  const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
  const float two_xz_p = 2.f * __fmaf_rn(r, y, xz), two_xz_m = 2.f * __fmaf_rn(-r, y, xz);
  const float two_yz_m = 2.f * __fmaf_rn(y, z, -rx), two_yz_p = 2.f * __fmaf_rn(y, z, rx);
  const float two_xy_m = 2.f * __fmaf_rn(x, y, -rz), two_xy_p = 2.f * __fmaf_rn(x, y, rz);
  const float d0 = __fadd_rn(1.f, -2.f * __fadd_rn(yy, zz));
  const float d1 = __fadd_rn(1.f, -2.f * __fmaf_rn(x, x, zz));
  const float d2 = __fadd_rn(1.f, -2.f * __fmaf_rn(x, x, yy));
  // M = S * R: M[c][k] = s_k * R[c][k] (one rounding each)
  Mat3 Mx = {{{__fmul_rn(sx, d0), __fmul_rn(sy, two_xy_m), __fmul_rn(sz, two_xz_p)},
              {__fmul_rn(sx, two_xy_p), __fmul_rn(sy, d1), __fmul_rn(sz, two_yz_m)},
              {__fmul_rn(sx, two_xz_m), __fmul_rn(sy, two_yz_p), __fmul_rn(sz, d2)}}};
  Mat3 Sg = mat_mul(mat_transpose(Mx), Mx);
  cov[0] = Sg.m[0][0]; cov[1] = Sg.m[0][1]; cov[2] = Sg.m[0][2];
  cov[3] = Sg.m[1][1]; cov[4] = Sg.m[1][2]; cov[5] = Sg.m[2][2];
}

// EWA 2D covariance (forward.cu:74-113): returns (a, b, c) with the +0.3 dilation applied.
__device__ __forceinline__ float3 cov2d_ewa(const float3 mean, float fx, float fy, float tanx,
                                            float tany, const float* cov3D,
                                            const float* __restrict__ V) {
  float3 t = xform_point_4x3(mean, V);
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), t.z);
  t.y = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), t.z);
  const float tz2 = __fmul_rn(t.z, t.z);
  Mat3 J = {{{fx / t.z, 0.f, -__fmul_rn(fx, t.x) / tz2},
             {0.f, fy / t.z, -__fmul_rn(fy, t.y) / tz2},
             {0.f, 0.f, 0.f}}};
  Mat3 Wm = {{{V[0], V[4], V[8]}, {V[1], V[5], V[9]}, {V[2], V[6], V[10]}}};
  Mat3 Tm = mat_mul(Wm, J);
  Mat3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
  Mat3 c = mat_mul(mat_mul(mat_transpose(Tm), mat_transpose(Vrk)), Tm);
  return make_float3(__fadd_rn(c.m[0][0], 0.3f), c.m[0][1], __fadd_rn(c.m[1][1], 0.3f));
}

// ---------------------------------------------------------------- Adam element update
// torch/optim/adam.py::_single_tensor_adam for one element (no weight decay, no amsgrad); w1 = 1 - beta1, w2 = 1 - beta2,
// lr1 = lr / bias_correction1, inv_b2 = 1 / sqrt(bias_correction2).  ONE definition for splat_adam_step's kernel and for the
// fused reduce + Adam + broadcast kernel of the multi-GPU mapping step, so both round identically.
__device__ __forceinline__ void adam_update(float g, float& m, float& v, float& p, float w1, float w2, float beta2,
                                            float lr1, float inv_b2, float eps) {
  m = m + w1 * (g - m);
  v = v * beta2 + w2 * (g * g);
  p = p - lr1 * (m / (sqrtf(v) * inv_b2 + eps));
}

// ---------------------------------------------------------------- warp helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

}  // namespace splat

// ---------------------------------------------------------------- error plumbing (api.cu)
namespace splat {
int fail(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int fail(int code, const char* fmt, ...);   // formats splat_last_error(), returns code
}  // namespace splat
#define SPLAT_CUDA(x)                                           \
  do {                                                          \
    int _rc = splat::check_cuda((x), #x);                       \
    if (_rc) return _rc;                                        \
  } while (0)
#define SPLAT_LAUNCH_CHECK(name)                                \
  do {                                                          \
    int _rc = splat::check_cuda(cudaGetLastError(), name);      \
    if (_rc) return _rc;                                        \
    splat::count_launch();                                      \
  } while (0)
