// preprocess.cu -- per-Gaussian forward geometry (+ per-tile instance counting), markVisible,
// and the per-Gaussian backward chain (conic -> cov2D -> cov3D -> scale/rotation, mean2D -> mean3D).
//
// Restates: preprocessCUDA<3> fwd (HX/cuda_rasterizer/forward.cu:155-256), in_frustum
// (auxiliary.h:139-163), checkFrustum (rasterizer_impl.cu:54-66), computeCov2DCUDA bwd
// (backward.cu:144-274), preprocessCUDA<3> bwd + computeCov3D bwd (backward.cu:412-530),
// computeColorFromSH fwd/bwd (forward.cu:20-71, backward.cu:20-139).
//
// B200 notes: this stage is HBM-bound (44 B in / 32 B out per Gaussian forward).  One thread per
// Gaussian; AoS float3 inputs are read through the read-only path (the three 4-byte loads of a
// warp cover 384 contiguous bytes, so every fetched sector is fully used).  Instead of the
// reference's tiles_touched[] + device-wide prefix scan over P, each visible Gaussian bumps the
// per-tile instance counters directly (RED.ADD on T addresses that live in L2); the scan then
// runs over T tiles, not P Gaussians (binning.cu).
#include "kernels.cuh"

namespace splat {

// Real SH basis up to degree 3 in the 3DGS sign convention (forward.cu:20-71): b[0..15].
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float* b) {
  b[0] = 0.28209479177387814f;
  if (deg > 0) {
    const float c1 = 0.4886025119029199f;
    b[1] = -c1 * y; b[2] = c1 * z; b[3] = -c1 * x;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = 1.0925484305920792f * xy;
      b[5] = -1.0925484305920792f * yz;
      b[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
      b[7] = -1.0925484305920792f * xz;
      b[8] = 0.5462742152960396f * (xx - yy);
      if (deg > 2) {
        b[9] = -0.5900435899266435f * y * (3.0f * xx - yy);
        b[10] = 2.890611442640554f * xy * z;
        b[11] = -0.4570457994644658f * y * (4.0f * zz - xx - yy);
        b[12] = 0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        b[13] = -0.4570457994644658f * x * (4.0f * zz - xx - yy);
        b[14] = 1.445305721320277f * z * (xx - yy);
        b[15] = -0.5900435899266435f * x * (xx - 3.0f * yy);
      }
    }
  }
}

// d b[k] / d(x,y,z) for the same basis (used by the SH backward, backward.cu:20-139).
__device__ __forceinline__ void sh_basis_grad(int deg, float x, float y, float z, float3* g) {
  g[0] = make_float3(0.f, 0.f, 0.f);
  if (deg > 0) {
    const float c1 = 0.4886025119029199f;
    g[1] = make_float3(0.f, -c1, 0.f);
    g[2] = make_float3(0.f, 0.f, c1);
    g[3] = make_float3(-c1, 0.f, 0.f);
    if (deg > 1) {
      const float k0 = 1.0925484305920792f, k2 = 0.31539156525252005f, k4 = 0.5462742152960396f;
      g[4] = make_float3(k0 * y, k0 * x, 0.f);
      g[5] = make_float3(0.f, -k0 * z, -k0 * y);
      g[6] = make_float3(-2.f * k2 * x, -2.f * k2 * y, 4.f * k2 * z);
      g[7] = make_float3(-k0 * z, 0.f, -k0 * x);
      g[8] = make_float3(2.f * k4 * x, -2.f * k4 * y, 0.f);
      if (deg > 2) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        const float a0 = -0.5900435899266435f, a1 = 2.890611442640554f, a2 = -0.4570457994644658f,
                    a3 = 0.3731763325901154f, a5 = 1.445305721320277f;
        g[9] = make_float3(a0 * 6.f * xy, a0 * 3.f * (xx - yy), 0.f);
        g[10] = make_float3(a1 * yz, a1 * xz, a1 * xy);
        g[11] = make_float3(a2 * -2.f * xy, a2 * (4.f * zz - xx - 3.f * yy), a2 * 8.f * yz);
        g[12] = make_float3(a3 * -6.f * xz, a3 * -6.f * yz, a3 * 3.f * (2.f * zz - xx - yy));
        g[13] = make_float3(a2 * (4.f * zz - 3.f * xx - yy), a2 * -2.f * xy, a2 * 8.f * xz);
        g[14] = make_float3(a5 * 2.f * xz, a5 * -2.f * yz, a5 * (xx - yy));
        g[15] = make_float3(a0 * 3.f * (xx - yy), a0 * -6.f * xy, 0.f);
      }
    }
  }
}


// Slope t = tan(theta), |t| <= 1, of the rotation that brings the pixel axes onto the splat's
// principal axes (theta = half the angle of (a-c, 2b), folded into [-45deg, 45deg]).  The
// power-2 backward accumulates its second moments in the frame d' = (dx + t*dy, dy - t*dx):
// there the quartic forms of hessian_finalize lose eps*aspect^2 instead of eps*aspect^4 for a thin
// splat.  Any t gives the same sums in exact arithmetic; t only conditions them.
__device__ __forceinline__ float frame_slope(float a, float b, float c) {
  const float D = a - c, b2 = 2.f * b;
  const float den = D + copysignf(sqrtf(D * D + b2 * b2), D);
  const float t = b2 / den;
  return (fabsf(t) <= 1.f) ? t : 0.f;  // den == 0 (isotropic) or overflow -> pixel axes
}

__global__ void __launch_bounds__(256) preprocess_fwd_kernel(const PreFwdArgs a) {
  pdl_wait();
  pdl_trigger();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  bool vis = false;
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
  if (idx < a.P) {
    int my_radius = 0;
    // every input of this Gaussian is requested up front (independent loads in flight together);
    // the few culled Gaussians cost their bytes, everyone else saves a dependent round trip
    const float3 p = make_float3(__ldg(a.means + 3 * idx), __ldg(a.means + 3 * idx + 1),
                                 __ldg(a.means + 3 * idx + 2));
    float3 s = make_float3(0.f, 0.f, 0.f);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    float cov3D[6];
    if (a.cov3D_pre) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov3D[k] = __ldg(a.cov3D_pre + 6 * idx + k);
    } else {
      s = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
      q = __ldg(reinterpret_cast<const float4*>(a.rots) + idx);
    }
    const float opacity = __ldg(a.opac + idx);
    float3 rgb = make_float3(0.f, 0.f, 0.f);
    if (a.colors_pre)
      rgb = make_float3(__ldg(a.colors_pre + 3 * idx), __ldg(a.colors_pre + 3 * idx + 1), __ldg(a.colors_pre + 3 * idx + 2));
    const float3 pv = xform_point_4x3(p, a.V);
    if (pv.z > a.near_plane) {  // in_frustum: reject p_view.z <= near
      const float4 ph = xform_point_4x4(p, a.Pm);
      const float pw = 1.0f / (ph.w + 0.0000001f);
      const float ppx = __fmul_rn(ph.x, pw), ppy = __fmul_rn(ph.y, pw);
      if (!a.cov3D_pre) cov3d_from_scale_rot(s, a.mod, q, cov3D);
      const float3 cov = cov2d_ewa(p, a.fx, a.fy, a.tanx, a.tany, cov3D, a.V);
      // roundings pinned to the reference build's (fma(a, c, -rn(b*b)), fma(mid, mid, -det))
      const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
      if (det != 0.0f) {
        const float det_inv = 1.f / det;
        const float3 conic = make_float3(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv), __fmul_rn(cov.x, det_inv));
        const float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
        const float disc = sqrtf(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
        const float lambda1 = __fadd_rn(mid, disc);
        const float lambda2 = __fadd_rn(mid, -disc);
        const float rad = ceilf(__fmul_rn(3.f, sqrtf(fmaxf(lambda1, lambda2))));
        const float2 pix = make_float2(ndc_to_pix(ppx, a.W), ndc_to_pix(ppy, a.H));
        tile_rect(pix, (int)rad, a.gx, a.gy, x0, y0, x1, y1);
        if ((x1 - x0) * (y1 - y0) != 0) {
          vis = true;
          my_radius = (int)rad;
          if (a.colors_pre == nullptr) {  // SH -> RGB (+0.5, clamp at 0 with mask)
            const float3 c = make_float3(__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2));
            float dx = p.x - c.x, dy = p.y - c.y, dz = p.z - c.z;
            const float inv = 1.0f / sqrtf(dx * dx + dy * dy + dz * dz);
            dx *= inv; dy *= inv; dz *= inv;
            float b[16];
            sh_basis(a.sh_degree, dx, dy, dz, b);
            const int nb = (a.sh_degree + 1) * (a.sh_degree + 1);
            const float* sh = a.shs + (size_t)idx * a.M * 3;
            float r = 0.f, gcol = 0.f, bl = 0.f;
            for (int k = 0; k < nb; ++k) {
              r += b[k] * __ldg(sh + 3 * k); gcol += b[k] * __ldg(sh + 3 * k + 1); bl += b[k] * __ldg(sh + 3 * k + 2);
            }
            r += 0.5f; gcol += 0.5f; bl += 0.5f;
            a.g.clamped[3 * idx + 0] = r < 0; a.g.clamped[3 * idx + 1] = gcol < 0; a.g.clamped[3 * idx + 2] = bl < 0;
            rgb = make_float3(fmaxf(r, 0.f), fmaxf(gcol, 0.f), fmaxf(bl, 0.f));
          }
          float4* rec = a.g.grec + (size_t)idx * INST_F4;
          rec[0] = make_float4(pix.x, pix.y, conic.x, conic.y);
          rec[1] = make_float4(conic.z, opacity, pv.z, __uint_as_float((uint32_t)idx));   // id: the row IS the blend kernels' record
          rec[2] = make_float4(rgb.x, rgb.y, rgb.z, frame_slope(cov.x, cov.y, cov.z));
        }
      }
    }
    a.radii[idx] = my_radius;
  }
  // Per-tile instance counting.  Small rects: the owning thread issues its few REDs; rects with
  // more than 32 tiles are spread over the warp so one huge splat does not serialise a lane.
  const int ntile = vis ? (x1 - x0) * (y1 - y0) : 0;
  const bool big = ntile > 32;
  if (vis && !big) {
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) atomicAdd(a.tile_count + (size_t)(y * a.gx + x) * CSTRIDE, 1u);
  }
  unsigned bigmask = __ballot_sync(0xffffffffu, big);
  while (bigmask) {
    const int src = __ffs(bigmask) - 1;
    bigmask &= bigmask - 1;
    const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
    const int bx1 = __shfl_sync(0xffffffffu, x1, src), by1 = __shfl_sync(0xffffffffu, y1, src);
    const int w = bx1 - bx0, n = w * (by1 - by0);
    for (int t = lane; t < n; t += 32) atomicAdd(a.tile_count + (size_t)((by0 + t / w) * a.gx + bx0 + t % w) * CSTRIDE, 1u);
  }
}

// ------------------------------------------------------------------------------------------
// Per-Gaussian backward chain.  GaussChain caches everything that depends only on the Gaussian
// (recomputed forward intermediates); apply() is the LINEAR map
//   (dL_dconic.x, .y, .w, dL_dmean2D.x, .y) -> (dL_dmean3D[3], dL_dcov3D[6], dL_dscale[3], dL_drot[4]).
// power==1 calls apply() once per Gaussian on the summed pair gradients (DG two-stage backward,
// and numerically what HX's per-pair fusion sums to); the Hessian finaliser calls it on unit
// vectors to obtain the Jacobian (hessian.cu).
struct GaussChain {
  Mat3 T, Vrk, W, Rm;      // T = W*J, cov3D as matrix, view 3x3, rotation matrix
  float a, b, c, denom, denom2inv;
  float tz, tx, ty, xmul, ymul, fx, fy;
  float3 s;                // mod * scale
  float4 q;
  float pj[3][2];          // d(ndc xy)/d(mean) rows
  const float* V;
  bool has_scale;
};

__device__ __forceinline__ void chain_setup(GaussChain& g, const float3 mean, const float* cov3D,
                                            bool has_scale, const float3 scale, float mod, const float4 q,
                                            float fx, float fy, float tanx, float tany,
                                            const float* __restrict__ V, const float* __restrict__ Pm) {
  // --- cov2D forward intermediates (backward.cu:159-199)
  float3 t = xform_point_4x3(mean, V);
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  g.xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  g.ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  Mat3 J = {{{fx / t.z, 0.f, -(fx * t.x) / (t.z * t.z)},
             {0.f, fy / t.z, -(fy * t.y) / (t.z * t.z)},
             {0.f, 0.f, 0.f}}};
  g.W = {{{V[0], V[4], V[8]}, {V[1], V[5], V[9]}, {V[2], V[6], V[10]}}};
  g.Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
  g.T = mat_mul(g.W, J);
  Mat3 c2 = mat_mul(mat_mul(mat_transpose(g.T), mat_transpose(g.Vrk)), g.T);
  g.a = c2.m[0][0] + 0.3f; g.b = c2.m[0][1]; g.c = c2.m[1][1] + 0.3f;
  g.denom = g.a * g.c - g.b * g.b;
  g.denom2inv = 1.0f / ((g.denom * g.denom) + 0.0000001f);
  g.tx = t.x; g.ty = t.y; g.tz = 1.f / t.z; g.fx = fx; g.fy = fy; g.V = V;
  // --- projection Jacobian (backward.cu:504-516)
  const float4 mh = xform_point_4x4(mean, Pm);
  const float mw = 1.0f / (mh.w + 0.0000001f);
  const float mul1 = (Pm[0] * mean.x + Pm[4] * mean.y + Pm[8] * mean.z + Pm[12]) * mw * mw;
  const float mul2 = (Pm[1] * mean.x + Pm[5] * mean.y + Pm[9] * mean.z + Pm[13]) * mw * mw;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    g.pj[k][0] = Pm[4 * k + 0] * mw - Pm[4 * k + 3] * mul1;
    g.pj[k][1] = Pm[4 * k + 1] * mw - Pm[4 * k + 3] * mul2;
  }
  // --- scale / rotation (backward.cu:412-432)
  g.has_scale = has_scale;
  if (has_scale) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    g.Rm = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    g.s = make_float3(mod * scale.x, mod * scale.y, mod * scale.z);
    g.q = q;
  }
}

__device__ __forceinline__ void chain_apply(const GaussChain& g, float dcx, float dcy, float dcw,
                                            float dmx, float dmy, float* dmean /*3*/, float* dcov /*6*/,
                                            float* dscale /*3*/, float* drot /*4*/) {
  const Mat3& T = g.T; const Mat3& Vrk = g.Vrk; const Mat3& W = g.W;
  const float a = g.a, b = g.b, c = g.c, denom = g.denom, d2 = g.denom2inv;
  float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
  if (d2 != 0.f) {
    dL_da = d2 * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcw);
    dL_dc = d2 * (-a * a * dcw + 2 * a * b * dcy + (denom - a * c) * dcx);
    dL_db = d2 * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcw);
    dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
    dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
    dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
    dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][1] * dL_dc;
    dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][2] * dL_dc;
    dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db + 2 * T.m[1][1] * T.m[1][2] * dL_dc;
  } else {
#pragma unroll
    for (int i = 0; i < 6; ++i) dcov[i] = 0.f;
  }
  // dL/dT (upper 2x3), dL/dJ, dL/dt, dL/dmean (cov part)
  const float r0x = T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2];
  const float r0y = T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2];
  const float r0z = T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2];
  const float r1x = T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2];
  const float r1y = T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2];
  const float r1z = T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2];
  const float dT00 = 2 * r0x * dL_da + r1x * dL_db;
  const float dT01 = 2 * r0y * dL_da + r1y * dL_db;
  const float dT02 = 2 * r0z * dL_da + r1z * dL_db;
  const float dT10 = 2 * r1x * dL_dc + r0x * dL_db;
  const float dT11 = 2 * r1y * dL_dc + r0y * dL_db;
  const float dT12 = 2 * r1z * dL_dc + r0z * dL_db;
  const float dJ00 = W.m[0][0] * dT00 + W.m[0][1] * dT01 + W.m[0][2] * dT02;
  const float dJ02 = W.m[2][0] * dT00 + W.m[2][1] * dT01 + W.m[2][2] * dT02;
  const float dJ11 = W.m[1][0] * dT10 + W.m[1][1] * dT11 + W.m[1][2] * dT12;
  const float dJ12 = W.m[2][0] * dT10 + W.m[2][1] * dT11 + W.m[2][2] * dT12;
  const float tz = g.tz, tz2 = tz * tz, tz3 = tz2 * tz;
  const float dtx = g.xmul * -g.fx * tz2 * dJ02;
  const float dty = g.ymul * -g.fy * tz2 * dJ12;
  const float dtz = -g.fx * tz2 * dJ00 - g.fy * tz2 * dJ11 + (2 * g.fx * g.tx) * tz3 * dJ02 + (2 * g.fy * g.ty) * tz3 * dJ12;
  const float* V = g.V;
  dmean[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
  dmean[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
  dmean[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
  // + projection part
  dmean[0] += g.pj[0][0] * dmx + g.pj[0][1] * dmy;
  dmean[1] += g.pj[1][0] * dmx + g.pj[1][1] * dmy;
  dmean[2] += g.pj[2][0] * dmx + g.pj[2][1] * dmy;
  // cov3D -> scale, rotation (backward.cu:433-475); note: no d(mod*scale)/dscale factor, as in the reference
  if (g.has_scale) {
    Mat3 S = {{{g.s.x, 0.f, 0.f}, {0.f, g.s.y, 0.f}, {0.f, 0.f, g.s.z}}};
    Mat3 Mx = mat_mul(S, g.Rm);
    Mat3 dSig = {{{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                  {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                  {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}}};
    Mat3 M2;
#pragma unroll
    for (int cc = 0; cc < 3; ++cc)
#pragma unroll
      for (int rr = 0; rr < 3; ++rr) M2.m[cc][rr] = 2.0f * Mx.m[cc][rr];
    Mat3 dM = mat_mul(M2, dSig);
    Mat3 Rt = mat_transpose(g.Rm);
    Mat3 dMt = mat_transpose(dM);
    dscale[0] = Rt.m[0][0] * dMt.m[0][0] + Rt.m[0][1] * dMt.m[0][1] + Rt.m[0][2] * dMt.m[0][2];
    dscale[1] = Rt.m[1][0] * dMt.m[1][0] + Rt.m[1][1] * dMt.m[1][1] + Rt.m[1][2] * dMt.m[1][2];
    dscale[2] = Rt.m[2][0] * dMt.m[2][0] + Rt.m[2][1] * dMt.m[2][1] + Rt.m[2][2] * dMt.m[2][2];
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) { dMt.m[0][rr] *= g.s.x; dMt.m[1][rr] *= g.s.y; dMt.m[2][rr] *= g.s.z; }
    const float r = g.q.x, x = g.q.y, y = g.q.z, z = g.q.w;
    drot[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
    drot[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) - 4 * x * (dMt.m[2][2] + dMt.m[1][1]);
    drot[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) - 4 * y * (dMt.m[2][2] + dMt.m[0][0]);
    drot[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) - 4 * z * (dMt.m[1][1] + dMt.m[0][0]);
  } else {
    dscale[0] = dscale[1] = dscale[2] = 0.f;
    drot[0] = drot[1] = drot[2] = drot[3] = 0.f;
  }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means, const float* __restrict__ V,
                                    float near_plane, uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  present[idx] = xform_point_4x3(p, V).z > near_plane ? 1 : 0;
}


// One thread per Gaussian: finalises the raw blend-backward sums (render_bwd.cu MODE 0) into
// dL_dcolors / dL_dopacity / dL_dmean2D / dL_dconic, then DG's computeCov2DCUDA + preprocessCUDA
// backward in one pass.  Every output row is written (zeros for culled Gaussians: their
// accumulator rows were never touched), so the caller need not pre-zero the outputs.
#ifndef PREBWD_CTAS
#define PREBWD_CTAS 4   // r02 A/B at C3: 62.1 us at 2 CTAs per SM (98 registers), 51.5 at 3, 49.4 at 4 (64 registers, 120 B of spills)
#endif
__global__ void __launch_bounds__(256, PREBWD_CTAS) preprocess_bwd_kernel(const PreBwdArgs a) {
  pdl_wait();
  pdl_trigger();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P) return;
  float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
  const bool vis = a.radii[idx] > 0;
  const int nb = a.shs ? (a.sh_degree + 1) * (a.sh_degree + 1) : 0;
  float3 dcol = make_float3(0.f, 0.f, 0.f);
  float dop = 0.f, dm2x = 0.f, dm2y = 0.f;
  float4 dcon = make_float4(0.f, 0.f, 0.f, 0.f);
  if (vis) {
    // raw sums: [0..2] colour, [3] u*dx, [4] u*dy, [5] u*dx*dx, [6] u*dx*dy, [7] u*dy*dy, [8] u
    const float4* row = reinterpret_cast<const float4*>(a.acc + (size_t)idx * 12);
    const float4 r0 = __ldg(row), r1 = __ldg(row + 1), r2 = __ldg(row + 2);
    const float4 g0 = __ldg(a.grec + (size_t)idx * INST_F4);
    const float2 g1 = __ldg(reinterpret_cast<const float2*>(a.grec + (size_t)idx * INST_F4 + 1));
    const float4 co = make_float4(g0.z, g0.w, g1.x, g1.y);   // conic A, B, C, opacity
    const float o = co.w;
    dcol = make_float3(r0.x, r0.y, r0.z);
    dop = r2.x;                                                     // sum G * dL/dalpha
    dm2x = -(co.x * r0.w + co.y * r1.x) * o * (0.5f * a.W);         // backward.cu:717-725
    dm2y = -(co.z * r1.x + co.y * r0.w) * o * (0.5f * a.H);
    dcon = make_float4(-0.5f * o * r1.y, -0.5f * o * r1.z, 0.f, -0.5f * o * r1.w);   // backward.cu:740-742
    const float3 mean = make_float3(__ldg(a.means + 3 * idx), __ldg(a.means + 3 * idx + 1),
                                    __ldg(a.means + 3 * idx + 2));
    float cov3D[6];
    float3 s = make_float3(0.f, 0.f, 0.f);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool has_scale = a.scales != nullptr;
    if (a.cov3D_pre) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov3D[k] = __ldg(a.cov3D_pre + 6 * idx + k);
    }
    if (has_scale) {
      s = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
      q = __ldg(reinterpret_cast<const float4*>(a.rots) + idx);
      if (!a.cov3D_pre) cov3d_from_scale_rot(s, a.mod, q, cov3D);
    }
    GaussChain g;
    // the cov3D -> scale / rotation leg is skipped when the caller wants neither gradient (pose-only tracking)
    const bool scale_leg = has_scale && (a.dL_dscales != nullptr || a.dL_drots != nullptr);
    chain_setup(g, mean, cov3D, scale_leg, s, a.mod, q, a.fx, a.fy, a.tanx, a.tany, a.V, a.Pm);
    chain_apply(g, dcon.x, dcon.y, dcon.w, dm2x, dm2y, dmean, dcov, dscale, drot);
    if (a.depth_sil) {
      // fused depth/silhouette channels: features (z, 1, z*z) with z = view depth = V[2]x + V[6]y + V[10]z + V[14]
      // (utils/slam_helpers.py:124-139 computes them in torch); dL/dz = sum(alpha T dL/dpix3) + 2 z sum(alpha T dL/dpix5)
      const float z = __ldg(reinterpret_cast<const float*>(a.grec + (size_t)idx * INST_F4 + 1) + 2);
      const float dz = r2.y + 2.f * z * r2.z;
      dmean[0] += dz * a.V[2]; dmean[1] += dz * a.V[6]; dmean[2] += dz * a.V[10];
    }
    if (a.shs) {  // SH backward (backward.cu:20-139): dL_dsh, and dL_dmean through the view direction
      const float3 c = make_float3(__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2));
      const float ox = mean.x - c.x, oy = mean.y - c.y, oz = mean.z - c.z;
      const float sum2 = ox * ox + oy * oy + oz * oz;
      const float inv = 1.0f / sqrtf(sum2);
      const float x = ox * inv, y = oy * inv, z = oz * inv;
      float3 dRGB = dcol;
      if (a.clamped[3 * idx + 0]) dRGB.x = 0.f;
      if (a.clamped[3 * idx + 1]) dRGB.y = 0.f;
      if (a.clamped[3 * idx + 2]) dRGB.z = 0.f;
      float b[16]; float3 gb[16];
      sh_basis(a.sh_degree, x, y, z, b);
      sh_basis_grad(a.sh_degree, x, y, z, gb);
      const float* sh = a.shs + (size_t)idx * a.M * 3;
      float* dsh = a.dL_dsh + (size_t)idx * a.M * 3;
      float ddx = 0.f, ddy = 0.f, ddz = 0.f;
      for (int k = 0; k < nb; ++k) {
        dsh[3 * k + 0] = b[k] * dRGB.x; dsh[3 * k + 1] = b[k] * dRGB.y; dsh[3 * k + 2] = b[k] * dRGB.z;
        const float w = __ldg(sh + 3 * k) * dRGB.x + __ldg(sh + 3 * k + 1) * dRGB.y + __ldg(sh + 3 * k + 2) * dRGB.z;
        ddx += gb[k].x * w; ddy += gb[k].y * w; ddz += gb[k].z * w;
      }
      for (int k = nb; k < a.M; ++k) { dsh[3 * k + 0] = 0.f; dsh[3 * k + 1] = 0.f; dsh[3 * k + 2] = 0.f; }
      // through dir = o / |o| (auxiliary.h:105-116)
      const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmean[0] += ((sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv32;
      dmean[1] += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * inv32;
      dmean[2] += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * inv32;
    }
  } else if (a.shs) {
    float* dsh = a.dL_dsh + (size_t)idx * a.M * 3;
    for (int k = 0; k < a.M * 3; ++k) dsh[k] = 0.f;
  }
  // every gradient array except dL_dmeans is optional (NULL = the caller does not need it)
  if (a.dL_dcolor) { a.dL_dcolor[3 * idx] = dcol.x; a.dL_dcolor[3 * idx + 1] = dcol.y; a.dL_dcolor[3 * idx + 2] = dcol.z; }
  if (a.dL_dopacity) a.dL_dopacity[idx] = dop;
  if (a.dL_dmean2D) { a.dL_dmean2D[3 * idx] = dm2x; a.dL_dmean2D[3 * idx + 1] = dm2y; a.dL_dmean2D[3 * idx + 2] = 0.f; }
  if (a.dL_dconic) reinterpret_cast<float4*>(a.dL_dconic)[idx] = dcon;
#pragma unroll
  for (int k = 0; k < 3; ++k) a.dL_dmeans[3 * idx + k] = dmean[k];
  if (a.dL_dcov3D) {
#pragma unroll
    for (int k = 0; k < 6; ++k) a.dL_dcov3D[6 * idx + k] = dcov[k];
  }
  if (a.dL_dscales) {
#pragma unroll
    for (int k = 0; k < 3; ++k) a.dL_dscales[3 * idx + k] = dscale[k];
  }
  if (a.dL_drots) reinterpret_cast<float4*>(a.dL_drots)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
}

int launch_preprocess_bwd(const PreBwdArgs& a, cudaStream_t st) {
  ProfScope prof_(K_PRE_BWD, st);
  if (a.P <= 0) return SPLAT_OK;
  SPLAT_CUDA(launch_k(preprocess_bwd_kernel, dim3((a.P + 255) / 256), dim3(256), 0, st, a));
  SPLAT_LAUNCH_CHECK("preprocess_bwd_kernel");
  return SPLAT_OK;
}

// ------------------------------------------------------------------------------------------
// Hessian (power == 2) finaliser.  render_bwd MODE 1 left, per Gaussian, the 19 sums
//   acc[0..2] = sum dcolor_ch^2, acc[3] = sum dopacity^2, acc[4..18] = upper triangle of
//   Mom = sum_pairs v v^T,  v = (dconic.x, dconic.y, dconic.w, dmean2D.x, dmean2D.y).
// HX accumulates, for every output k of the per-Gaussian chain, sum_pairs (J_k . v)^2
// (backward.cu:1041-1137 with grad_power == 2) which equals J_k^T Mom J_k.  J is obtained by
// pushing the five unit vectors through the same chain_apply() used for power == 1.

#ifndef HESSFIN_CTAS
#define HESSFIN_CTAS 3
#endif
__global__ void __launch_bounds__(128, HESSFIN_CTAS) hessian_finalize_kernel(const HessArgs a) {
  pdl_wait();
  pdl_trigger();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P) return;
  float out[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) out[k] = 0.f;
  float m[19];
#pragma unroll
  for (int k = 0; k < 19; ++k) m[k] = 0.f;
  if (a.radii[idx] > 0) {
#pragma unroll
    for (int k = 0; k < 19; ++k) m[k] = a.acc[(size_t)idx * 20 + k];
    const float3 mean = make_float3(__ldg(a.means + 3 * idx), __ldg(a.means + 3 * idx + 1), __ldg(a.means + 3 * idx + 2));
    float cov3D[6];
    float3 s = make_float3(0.f, 0.f, 0.f);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool has_scale = a.scales != nullptr;
    if (a.cov3D_pre) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov3D[k] = __ldg(a.cov3D_pre + 6 * idx + k);
    }
    if (has_scale) {
      s = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
      q = __ldg(reinterpret_cast<const float4*>(a.rots) + idx);
      if (!a.cov3D_pre) cov3d_from_scale_rot(s, a.mod, q, cov3D);
    }
    GaussChain g;
    // the cov3D -> scale / rotation leg is skipped when the caller wants neither gradient (pose-only tracking)
    const bool scale_leg = has_scale && (a.dL_dscales != nullptr || a.dL_drots != nullptr);
    chain_setup(g, mean, cov3D, scale_leg, s, a.mod, q, a.fx, a.fy, a.tanx, a.tany, a.V, a.Pm);
    float J[5][16];
#pragma unroll
    for (int p = 0; p < 5; ++p)
      chain_apply(g, p == 0 ? 1.f : 0.f, p == 1 ? 1.f : 0.f, p == 2 ? 1.f : 0.f, p == 3 ? 1.f : 0.f,
                  p == 4 ? 1.f : 0.f, &J[p][0], &J[p][3], &J[p][9], &J[p][12]);
    // The conic moments were taken in the frame d' = (dx + t*dy, dy - t*dx) (render_bwd MODE 1):
    // (dx^2, dx*dy, dy^2) = Qinv (dx'^2, dx'*dy', dy'^2) / (1+t^2)^2, so the conic rows of J become
    // J Qinv / (1+t^2)^2.
    {
      const float t = a.grec[(size_t)idx * INST_F4 + 2].w;
      const float n = 1.f + t * t, in2 = 1.f / (n * n), t2 = t * t;
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const float ja = J[0][k], jb = J[1][k], jc = J[2][k];
        J[0][k] = (ja + t * jb + t2 * jc) * in2;
        J[1][k] = (2.f * t * (jc - ja) + (1.f - t2) * jb) * in2;
        J[2][k] = (t2 * ja - t * jb + jc) * in2;
      }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float acc = 0.f;
      int t = 4;
#pragma unroll
      for (int p = 0; p < 5; ++p)
#pragma unroll
        for (int qq = p; qq < 5; ++qq) {
          const float w = J[p][k] * J[qq][k] * m[t++];
          acc += (p == qq) ? w : 2.f * w;
        }
      out[k] = acc;
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) a.dL_dmeans[3 * idx + k] = out[k];
#pragma unroll
  for (int k = 0; k < 6; ++k) a.dL_dcov3D[6 * idx + k] = out[3 + k];
#pragma unroll
  for (int k = 0; k < 3; ++k) a.dL_dscales[3 * idx + k] = out[9 + k];
  reinterpret_cast<float4*>(a.dL_drots)[idx] = make_float4(out[12], out[13], out[14], out[15]);
  a.dL_dcolors[3 * idx + 0] = m[0]; a.dL_dcolors[3 * idx + 1] = m[1]; a.dL_dcolors[3 * idx + 2] = m[2];
  a.dL_dopacity[idx] = m[3];
  // diagonal moments: (0,0)->4, (1,1)->9, (2,2)->13, (3,3)->16, (4,4)->18
  reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(m[4], m[9], 0.f, m[13]);
  a.dL_dmean2D[3 * idx + 0] = m[16]; a.dL_dmean2D[3 * idx + 1] = m[18]; a.dL_dmean2D[3 * idx + 2] = 0.f;
}

int launch_hessian_finalize(const HessArgs& a, cudaStream_t st) {
  ProfScope prof_(K_HESS_FIN, st);
  if (a.P <= 0) return SPLAT_OK;
  SPLAT_CUDA(launch_k(hessian_finalize_kernel, dim3((a.P + 127) / 128), dim3(128), 0, st, a));
  SPLAT_LAUNCH_CHECK("hessian_finalize_kernel");
  return SPLAT_OK;
}

int launch_preprocess_fwd(const PreFwdArgs& a, cudaStream_t st) {
  ProfScope prof_(K_PRE_FWD, st);
  if (a.P <= 0) return SPLAT_OK;
  SPLAT_CUDA(launch_k(preprocess_fwd_kernel, dim3((a.P + 255) / 256), dim3(256), 0, st, a));
  SPLAT_LAUNCH_CHECK("preprocess_fwd_kernel");
  return SPLAT_OK;
}

int launch_mark_visible(int P, const float* means, const float* V, float near_plane, uint8_t* present,
                        cudaStream_t st) {
  if (P <= 0) return SPLAT_OK;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means, V, near_plane, present);
  SPLAT_LAUNCH_CHECK("mark_visible_kernel");
  return SPLAT_OK;
}

}  // namespace splat
