// render_fwd.cu -- per-tile front-to-back alpha compositing (colour[3], median depth, final T,
// n_contrib).  Restates renderCUDA<3> forward (HX/cuda_rasterizer/forward.cu:261-393).
//
// Differences in HOW (results identical):
//  * per-warp 8x4 sub-tiles with an exact conservative footprint test (render.cuh): a warp only
//    evaluates Gaussians whose alpha>=1/255 ellipse can touch its 32 pixels (ballot + ffs walk);
//  * the Gaussian's colour and depth are staged in shared memory with the geometry (the
//    reference re-reads colour from global memory per pixel, forward.cu:361);
//  * warp-level early termination (__all_sync(done)) on top of the block-level vote.
// Per-pixel arithmetic keeps the reference's expression order and precise expf so the discrete
// tests (power>0, alpha<1/255, T<1e-4, median crossing) take the same branches.
#include "render.cuh"
#include "kernels.cuh"

namespace splat {


template <bool CULL, bool COUNT>
__global__ void __launch_bounds__(256) render_fwd_kernel(const RenderFwdArgs a) {
  if (a.hdr->num_rendered > a.Rcap) return;
  __shared__ float2 s_xy[BATCH];
  __shared__ float4 s_co[BATCH];
  __shared__ float4 s_cd[BATCH];  // r, g, b, depth
  __shared__ float4 s_bb[BATCH];

  const int tile = blockIdx.x;
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  const int tid = threadIdx.x, lane = tid & 31;
  const SubTile st = make_subtile(tile_x, tile_y);
  const bool inside = st.px < a.W && st.py < a.H;
  const float pxf = (float)st.px, pyf = (float)st.py;
  const float tile_cx = tile_x * TILE_X + 7.5f, tile_cy = tile_y * TILE_Y + 7.5f;

  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);

  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 15.0f;
  uint32_t last = 0, nblend = 0;
  bool done = !inside;
  bool warp_done = __all_sync(0xffffffffu, done);

  for (int base = 0; base < n; base += BATCH) {
    if (__syncthreads_count(done) == TILE_PIX) break;
    {
      const int i = base + tid;
      if (i < n) {
        const uint32_t id = a.point_list[range.x + i];
        const float2 xy = a.xy[id];
        const float4 co = a.conic_opacity[id];
        s_xy[tid] = xy;
        s_co[tid] = co;
        s_cd[tid] = make_float4(__ldg(a.colors + 3 * id), __ldg(a.colors + 3 * id + 1),
                                __ldg(a.colors + 3 * id + 2), a.depths[id]);
        if (CULL) s_bb[tid] = cull_box(xy, co, tile_cx, tile_cy);
      }
    }
    __syncthreads();
    if (warp_done) continue;
    const int cnt = min(BATCH, n - base);
    for (int c = 0; c < cnt; c += 32) {
      const int j = c + lane;
      bool hit = j < cnt;
      if (CULL) hit = hit && box_hits(s_bb[j], st);
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int jj = c + __ffs(m) - 1;
        m &= m - 1;
        if (done) continue;
        const float2 xy = s_xy[jj];
        const float4 co = s_co[jj];
        const float dx = xy.x - pxf, dy = xy.y - pyf;
        const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, co.w * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) { done = true; continue; }
        const float4 cd = s_cd[jj];
        C0 += cd.x * alpha * T;
        C1 += cd.y * alpha * T;
        C2 += cd.z * alpha * T;
        if (T > 0.5f && test_T < 0.5f) D = cd.w;
        T = test_T;
        last = (uint32_t)(base + jj + 1);
        if (COUNT) ++nblend;
      }
      if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
    }
  }

  if (inside) {
    const int pix = st.py * a.W + st.px;
    const size_t HW = (size_t)a.W * a.H;
    a.final_T[pix] = T;
    a.n_contrib[pix] = last;
    a.out_color[pix] = C0 + T * __ldg(a.bg + 0);
    a.out_color[HW + pix] = C1 + T * __ldg(a.bg + 1);
    a.out_color[2 * HW + pix] = C2 + T * __ldg(a.bg + 2);
    if (a.out_depth) a.out_depth[pix] = D;
    if (COUNT) a.blended[pix] = nblend;
  }
}

int g_disable_cull = 0;

int launch_render_fwd(const RenderFwdArgs& a, int T, cudaStream_t st) {
  ProfScope prof_(K_RENDER_FWD, st);
  if (T <= 0) return SPLAT_OK;
  if (a.blended) render_fwd_kernel<false, true><<<T, 256, 0, st>>>(a);
  else if (g_disable_cull) render_fwd_kernel<false, false><<<T, 256, 0, st>>>(a);
  else render_fwd_kernel<true, false><<<T, 256, 0, st>>>(a);
  SPLAT_LAUNCH_CHECK("render_fwd_kernel");
  return SPLAT_OK;
}

}  // namespace splat
