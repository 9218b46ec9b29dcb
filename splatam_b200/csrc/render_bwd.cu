// render_bwd.cu -- back-to-front re-blend and per-Gaussian gradient accumulation.
// Restates renderCUDA<3> backward (HX/cuda_rasterizer/backward.cu:586-748; live in DG) and the
// per-pair head of renderCUDAFused<3> (backward.cu:850-1040; live in HX).
//
// Reference: every blended (pixel,Gaussian) pair issues 9 (DG) or 25 (HX) global float atomics.
// Here, per CTA (tile) and per staged batch of 256 Gaussians:
//   1. each warp walks only the Gaussians whose footprint can touch its 8x4 pixels (render.cuh);
//   2. the per-pair partials of up to GROUP Gaussians are reduced across the 32 lanes with a
//      transposing butterfly (31 shuffles for 32 values instead of 5 per value); afterwards lane L
//      holds the warp total of value L;
//   3. lanes add their totals into a per-CTA shared accumulator row of the staged Gaussian;
//   4. after the batch, thread j flushes row j with one RED.ADD.F32 per value.
// => global atomics per (tile,Gaussian) instance instead of per (pixel,Gaussian) pair.
//
// MODE 0 (VALS=9):  sums of  dcolor[3], dmean2D.xy, dconic.xyw, dopacity        -> power == 1
// MODE 1 (VALS=19): sums of  dcolor^2[3], dopacity^2, and the 15 second moments of
//                   v = (dconic.x, dconic.y, dconic.w, dmean2D.x, dmean2D.y)     -> power == 2;
//                   hessian.cu turns the moments into sum_pairs (J_k . v)^2 per output k, which
//                   is what HX accumulates pair by pair (backward.cu:1093-1137).
#include "render.cuh"
#include "kernels.cuh"

namespace splat {


// After the call lane L holds sum over lanes of v[L].
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = lane & off;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

template <int MODE> struct BwdTraits;
template <> struct BwdTraits<0> { static const int VALS = 9, GROUP = 3, ROW = 9; };
template <> struct BwdTraits<1> { static const int VALS = 19, GROUP = 1, ROW = 19; };

template <bool CULL, int MODE>
__global__ void __launch_bounds__(256) render_bwd_kernel(const RenderBwdArgs a) {
  constexpr int VALS = BwdTraits<MODE>::VALS, GROUP = BwdTraits<MODE>::GROUP, ROW = BwdTraits<MODE>::ROW;
  __shared__ uint32_t s_id[BATCH];
  __shared__ float2 s_xy[BATCH];
  __shared__ float4 s_co[BATCH];
  __shared__ float4 s_col[BATCH];
  __shared__ float4 s_bb[BATCH];
  __shared__ float s_acc[BATCH * ROW];
  __shared__ int s_max[8];

  const int tile = blockIdx.x;
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const SubTile st = make_subtile(tile_x, tile_y);
  const bool inside = st.px < a.W && st.py < a.H;
  const float pxf = (float)st.px, pyf = (float)st.py;
  const float tile_cx = tile_x * TILE_X + 7.5f, tile_cy = tile_y * TILE_Y + 7.5f;
  const uint2 range = a.ranges[tile];
  const int pix = st.py * a.W + st.px;
  const size_t HW = (size_t)a.W * a.H;

  const float T_final = inside ? a.final_T[pix] : 0.f;
  float T = T_final;
  const int last = inside ? (int)a.n_contrib[pix] : 0;
  float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
  if (inside) { dp0 = a.dL_dpix[pix]; dp1 = a.dL_dpix[HW + pix]; dp2 = a.dL_dpix[2 * HW + pix]; }
  const float bg_dot = __ldg(a.bg) * dp0 + __ldg(a.bg + 1) * dp1 + __ldg(a.bg + 2) * dp2;
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
  const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;

  // list positions >= max(n_contrib) over the tile are never touched by anyone
  const int warp_hi = __reduce_max_sync(0xffffffffu, last);
  if (lane == 0) s_max[wid] = warp_hi;
  __syncthreads();
  int hi = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) hi = max(hi, s_max[w]);

  for (int top = hi; top > 0; top -= BATCH) {
    const int lo = max(0, top - BATCH), cnt = top - lo;
    __syncthreads();  // previous batch fully flushed
    if (tid < cnt) {
      const uint32_t id = a.point_list[range.x + lo + tid];
      const float2 xy = a.xy[id];
      const float4 co = a.conic_opacity[id];
      s_id[tid] = id;
      s_xy[tid] = xy;
      s_co[tid] = co;
      s_col[tid] = make_float4(__ldg(a.colors + 3 * id), __ldg(a.colors + 3 * id + 1), __ldg(a.colors + 3 * id + 2), 0.f);
      if (CULL) s_bb[tid] = cull_box(xy, co, tile_cx, tile_cy);
    }
#pragma unroll
    for (int k = 0; k < ROW; ++k) s_acc[k * BATCH + tid] = 0.f;
    __syncthreads();

    if (warp_hi > lo) {
      for (int c = ((cnt - 1) >> 5) << 5; c >= 0; c -= 32) {
        const int j = c + lane;
        bool hit = j < cnt && (lo + j) < warp_hi;
        if (CULL) hit = hit && box_hits(s_bb[j], st);
        unsigned m = __ballot_sync(0xffffffffu, hit);
        while (m) {
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0.f;
          int jsel[GROUP];
          bool any = false;
#pragma unroll
          for (int gsl = 0; gsl < GROUP; ++gsl) {
            jsel[gsl] = -1;
            if (m) {
              const int b = 31 - __clz(m);
              m &= ~(1u << b);
              const int jj = c + b;
              jsel[gsl] = jj;
              const int pos = lo + jj;
              bool valid = pos < last;
              const float2 xy = s_xy[jj];
              const float4 co = s_co[jj];
              const float dx = xy.x - pxf, dy = xy.y - pyf;
              const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
              valid = valid && !(power > 0.0f);
              const float G = expf(power);
              const float alpha = fminf(0.99f, co.w * G);
              valid = valid && !(alpha < 1.0f / 255.0f);
              if (valid) {
                T = T / (1.f - alpha);
                const float dchannel_dcolor = alpha * T;
                const float4 col = s_col[jj];
                float dL_dalpha = 0.f;
                ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0; lc0 = col.x; dL_dalpha += (col.x - ar0) * dp0;
                ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1; lc1 = col.y; dL_dalpha += (col.y - ar1) * dp1;
                ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2; lc2 = col.z; dL_dalpha += (col.z - ar2) * dp2;
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                const float dL_dG = co.w * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * co.x - gdy * co.y;
                const float dG_ddely = -gdy * co.z - gdx * co.y;
                const float c0 = dchannel_dcolor * dp0, c1 = dchannel_dcolor * dp1, c2 = dchannel_dcolor * dp2;
                const float mx = dL_dG * dG_ddelx * ddelx_dx, my = dL_dG * dG_ddely * ddely_dy;
                const float qx = -0.5f * gdx * dx * dL_dG, qy = -0.5f * gdx * dy * dL_dG, qw = -0.5f * gdy * dy * dL_dG;
                const float op = G * dL_dalpha;
                if (MODE == 0) {
                  float* o = &v[gsl * VALS];
                  o[0] = c0; o[1] = c1; o[2] = c2; o[3] = mx; o[4] = my; o[5] = qx; o[6] = qy; o[7] = qw; o[8] = op;
                } else {
                  v[0] = c0 * c0; v[1] = c1 * c1; v[2] = c2 * c2; v[3] = op * op;
                  const float u[5] = {qx, qy, qw, mx, my};
                  int k = 4;
#pragma unroll
                  for (int p = 0; p < 5; ++p)
#pragma unroll
                    for (int q = p; q < 5; ++q) v[k++] = u[p] * u[q];
                }
              }
              any = any || __any_sync(0xffffffffu, valid);
            }
          }
          if (!any) continue;
          const float tot = transpose_reduce32(v, lane);
          const int gsl = lane / VALS, k = lane - gsl * VALS;
          if (gsl < GROUP) {
            int jj = jsel[0];
#pragma unroll
            for (int q = 1; q < GROUP; ++q) if (gsl == q) jj = jsel[q];
            if (jj >= 0 && tot != 0.f) atomicAdd(&s_acc[k * BATCH + jj], tot);
          }
        }
      }
    }
    __syncthreads();
    if (tid < cnt) {
      const uint32_t id = s_id[tid];
      float r[ROW];
      bool nz = false;
#pragma unroll
      for (int k = 0; k < ROW; ++k) { r[k] = s_acc[k * BATCH + tid]; nz = nz || r[k] != 0.f; }
      if (nz) {
        if (MODE == 0) {
          atomicAdd(a.dL_dcolors + 3 * id + 0, r[0]);
          atomicAdd(a.dL_dcolors + 3 * id + 1, r[1]);
          atomicAdd(a.dL_dcolors + 3 * id + 2, r[2]);
          atomicAdd(a.dL_dmean2D + 3 * id + 0, r[3]);
          atomicAdd(a.dL_dmean2D + 3 * id + 1, r[4]);
          atomicAdd(a.dL_dconic + 4 * id + 0, r[5]);
          atomicAdd(a.dL_dconic + 4 * id + 1, r[6]);
          atomicAdd(a.dL_dconic + 4 * id + 3, r[7]);
          atomicAdd(a.dL_dopacity + id, r[8]);
        } else {
#pragma unroll
          for (int k = 0; k < ROW; ++k) atomicAdd(a.acc + (size_t)id * 20 + k, r[k]);
        }
      }
    }
  }
}


int launch_render_bwd(const RenderBwdArgs& a, int T, int mode, cudaStream_t st) {
  ProfScope prof_(K_RENDER_BWD, st);
  if (T <= 0) return SPLAT_OK;
  if (mode == 0) {
    if (g_disable_cull) render_bwd_kernel<false, 0><<<T, 256, 0, st>>>(a);
    else render_bwd_kernel<true, 0><<<T, 256, 0, st>>>(a);
  } else {
    if (g_disable_cull) render_bwd_kernel<false, 1><<<T, 256, 0, st>>>(a);
    else render_bwd_kernel<true, 1><<<T, 256, 0, st>>>(a);
  }
  SPLAT_LAUNCH_CHECK("render_bwd_kernel");
  return SPLAT_OK;
}

}  // namespace splat
