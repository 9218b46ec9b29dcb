// render.cuh -- pieces shared by the forward and backward blend kernels.
//
// Tile = 16x16 pixels = one CTA of CTA_WARPS = 4 warps (same tile semantics as the reference,
// HX/cuda_rasterizer/config.h:15-17, so tile-rect culling is identical), but the thread->pixel map is ours: warp w owns
// the 8x8 block (w&1, w>>1) of the tile and every lane owns TWO pixels of it -- column lane&7, rows lane>>3 and
// (lane>>3)+4 -- so that
//   * one lane-parallel footprint test + ballot rejects a Gaussian for 64 pixels at once,
//   * one record read, one ballot pop and (in the backward) one transposed reduction serve 64 pixels, and
//   * the per-pixel arithmetic of the pair is issued as packed FP32x2 instructions (packed.cuh): dx and every
//     per-Gaussian scalar are shared by the lane's two pixels and enter as broadcast operands.
// Round 1 ran one pixel per lane on 8x4 sub-tiles with scalar FP32: 49 warp-instructions per tested warp-pair in
// the backward against ~7 algorithmic FMAs (profiles/r01e_ncu_full_summary.txt); both blend kernels are bound by
// instruction issue, not by the FP32 pipe, and packing halves the issue slots per pixel.
#pragma once
#include "common.cuh"
#include "packed.cuh"

namespace splat {

// Instance ring shared by both blend kernels: RING_NS stages of RING_CH records (48 B each), filled by
// the TMA unit and consumed by the CTA's CTA_WARPS warps (one 8x8 block each).  There is no producer
// warp: a stage is released through a shared-memory counter, and the warp whose release is the
// last issues the refill of that stage itself (one elected lane, ~10 instructions per 6 KB) -- a
// polling producer lane cost 6.6% of the backward kernel's issue slots (profiles/r01c).
#define RING_CH 128
#ifndef RING_NS
#define RING_NS 3
#endif
#define CTA_WARPS 4
#define RING_THREADS (CTA_WARPS * 32)

struct SubTile {
  int px, pya, pyb;  // this lane's two pixels: (px, pya) and (px, pyb = pya + 4)
  float fx0, fx1;    // block x extent [fx0, fx1] in pixel coordinates
  float fy0, fy1;
};

__device__ __forceinline__ SubTile make_subtile(int tile_x, int tile_y) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  SubTile s;
  const int sx0 = tile_x * TILE_X + (wid & 1) * 8;
  const int sy0 = tile_y * TILE_Y + (wid >> 1) * 8;
  s.px = sx0 + (lane & 7);
  s.pya = sy0 + (lane >> 3);
  s.pyb = s.pya + 4;
  s.fx0 = (float)sx0; s.fx1 = (float)(sx0 + 7);
  s.fy0 = (float)sy0; s.fy1 = (float)(sy0 + 7);
  return s;
}

// power = -0.5 (A dx^2 + C dy^2) - B dx dy of one Gaussian at the lane's two pixels, with exactly the roundings of
// the reference build (SASS of renderCUDA<3>, forward.cu:341-343 as nvcc 12.9 contracts it):
//   fma( fma(dx, rn(A dx), rn(dy rn(C dy))), -0.5, -rn(dy rn(B dx)) ).
// dx is shared by the two pixels (same column); 3 scalar + 6 packed instructions for the pair.
__device__ __forceinline__ f2 pair_power(float dx, f2 dy2, float A, float B, float C) {
  const float Adx = __fmul_rn(A, dx), nBdx = -__fmul_rn(B, dx);
  const f2 t = mul2(dy2, mul2(bc(C), dy2));
  const f2 s = fma2(bc(dx), bc(Adx), t);
  return fma2(s, bc(-0.5f), mul2(dy2, bc(nBdx)));
}

// Conservative screen-space bounding box of the set of pixels at which the reference's tests
//   power <= 0  &&  min(0.99, o*exp(power)) >= 1/255        (forward.cu:341-350)
// can pass, as (xmin, xmax, ymin, ymax).  The level set power >= -tau, tau = ln(255 o), is the
// ellipse d^T Q d <= 2 tau whose half extents are sqrt(2 tau C/det), sqrt(2 tau A/det).  Margins:
//  * tau is padded by the worst-case float error of `power` over the tile (cancellation between
//    the three quadratic terms), of expf and of the product;
//  * ill-conditioned conics (det tiny vs A*C, where the rounded conic may not even be positive
//    definite) are never culled: their box is infinite.
// A pair outside the box is guaranteed to be skipped by the reference, so skipping it here is
// exact (n_contrib, T and colours are unchanged).
__device__ __forceinline__ float4 cull_box(const float2 xy, const float4 co, float tile_cx, float tile_cy) {
  const float A = co.x, B = co.y, C = co.z, o = co.w;
  const float inf = __int_as_float(0x7f800000);
  if (o < 1.0f / 255.0f) return make_float4(inf, -inf, inf, -inf);  // can never reach 1/255
  const float det = A * C - B * B;
  if (!(A > 0.f && C > 0.f && det > 1e-3f * A * C)) return make_float4(-inf, inf, -inf, inf);
  const float Dx = fabsf(xy.x - tile_cx) + 8.5f, Dy = fabsf(xy.y - tile_cy) + 8.5f;
  const float delta = 1e-6f * (A * Dx * Dx + C * Dy * Dy + 2.f * fabsf(B) * Dx * Dy);
  const float tau = __logf(255.0f * o) * 1.00001f + 1e-4f + delta;
  const float k = 2.004f * tau / det;
  const float ex = sqrtf(k * C) + 0.01f, ey = sqrtf(k * A) + 0.01f;
  return make_float4(xy.x - ex, xy.x + ex, xy.y - ey, xy.y + ey);
}

__device__ __forceinline__ bool box_hits(const float4 bb, const SubTile& s) {
  return bb.y >= s.fx0 && bb.x <= s.fx1 && bb.w >= s.fy0 && bb.z <= s.fy1;
}

// Exact footprint test of one Gaussian against one warp block (8x8 pixels): the bounding box above, then the
// true minimum of q(u) = A ux^2 + 2 B ux uy + C uy^2 (power = -q/2) over the sub-tile rectangle
// (expanded by 0.02 px) against the same padded threshold 2*tau.  For a convex quadratic whose
// centre lies outside the rectangle the minimiser sits on an edge that faces the centre, so at most
// two clamped 1-D minimisations are needed.  Everything it rejects fails the reference's
// alpha >= 1/255 test at every pixel of the sub-tile; it removes the bounding box's corner hits
// (~9% of hits for round splats, more for rotated thin ones).
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// EXACT adds the corner test.  Measured A/B on one box (scripts/gpu_ab_exact_cull.sh, C3): the blend BACKWARD gains 1.5%
// (a rejected pair saves its whole 75-flop gradient path), the blend FORWARD loses 2% (a rejected pair only saves ~20
// flops, less than the test costs) -- so the backward instantiates <true>, the forward <false>.  -DSPLAT_EXACT_CULL=0/1
// forces one setting for both.
template <bool EXACT_DEFAULT>
__device__ __forceinline__ bool footprint_hits(const float2 xy, const float4 co, float tile_cx, float tile_cy,
                                               const SubTile& s) {
  const float A = co.x, B = co.y, C = co.z, o = co.w;
  if (o < 1.0f / 255.0f) return false;  // can never reach 1/255
  const float det = A * C - B * B;
  if (!(A > 0.f && C > 0.f && det > 1e-3f * A * C)) return true;  // ill-conditioned: never culled
  const float Dx = fabsf(xy.x - tile_cx) + 8.5f, Dy = fabsf(xy.y - tile_cy) + 8.5f;
  const float delta = 1e-6f * (A * Dx * Dx + C * Dy * Dy + 2.f * fabsf(B) * Dx * Dy);
  const float tau = __logf(255.0f * o) * 1.00001f + 1e-4f + delta;
  // approximate division / square roots (MUFU, ~1e-6 relative): covered by the 0.2% pad on tau
  const float k = __fdividef(2.004f * tau, det);
  const float ex = sqrt_approx(k * C) + 0.01f, ey = sqrt_approx(k * A) + 0.01f;
  if (!(xy.x + ex >= s.fx0 && xy.x - ex <= s.fx1 && xy.y + ey >= s.fy0 && xy.y - ey <= s.fy1)) return false;
#ifdef SPLAT_EXACT_CULL
  constexpr bool EXACT = SPLAT_EXACT_CULL != 0;
#else
  constexpr bool EXACT = EXACT_DEFAULT;
#endif
  if (!EXACT) return true;
  const float ux0 = s.fx0 - 0.02f - xy.x, ux1 = s.fx1 + 0.02f - xy.x;
  const float uy0 = s.fy0 - 0.02f - xy.y, uy1 = s.fy1 + 0.02f - xy.y;
  const float fx = ux0 > 0.f ? ux0 : (ux1 < 0.f ? ux1 : 0.f);
  const float fy = uy0 > 0.f ? uy0 : (uy1 < 0.f ? uy1 : 0.f);
  if (fx == 0.f || fy == 0.f) return true;  // not in a corner region: the bounding box test is exact enough
  // corner region: true minimum of q over the two facing edges (clamped 1-D minimisations)
  const float y = fminf(uy1, fmaxf(uy0, __fdividef(-B * fx, C)));
  const float x = fminf(ux1, fmaxf(ux0, __fdividef(-B * fy, A)));
  const float qa = A * fx * fx + 2.f * B * fx * y + C * y * y;
  const float qb = A * x * x + 2.f * B * x * fy + C * fy * fy;
  return fminf(qa, qb) <= 2.004f * tau;
}

}  // namespace splat
