// render_fwd.cu -- per-tile front-to-back alpha compositing (colour[3], median depth, final T,
// n_contrib).  Restates renderCUDA<3> forward (HX/cuda_rasterizer/forward.cu:261-393).
//
// Differences in HOW (results identical):
//  * the tile's sorted instance records (48 B each, written by the tile sort in list order) are ONE
//    contiguous byte range, streamed into a RING_NS-stage shared-memory ring by the TMA unit
//    (cp.async.bulk + mbarrier complete_tx).  The warps never meet at a CTA barrier: each waits
//    on the stage's "full" mbarrier and releases the stage through a counter; the warp whose
//    release is the last issues the refill.  Warps drift apart by up to RING_NS stages instead of
//    idling at __syncthreads (the reference re-syncs the whole CTA every 256 Gaussians,
//    forward.cu:311-334);
//  * per-warp 8x4 sub-tiles with an exact conservative footprint test (render.cuh): a warp only
//    evaluates Gaussians whose alpha>=1/255 ellipse can touch its 32 pixels (ballot + ffs walk),
//    two hits at a time so the two exponentials overlap;
//  * colour and depth travel with the record (the reference re-reads colour from global memory per
//    pixel, forward.cu:361);
//  * warp-level early termination, and the ring stops being refilled once all warps of the CTA are done.
// Per-pixel arithmetic keeps the reference's expression order and precise expf so the discrete
// tests (power>0, alpha<1/255, T<1e-4, median crossing) take the same branches.
#include "render.cuh"
#include "kernels.cuh"
#include "pipeline.cuh"

namespace splat {

// DS: fused RGB + depth + silhouette: three more accumulators for the features (z, 1, z*z), z = the
// record's view depth, blended with the SAME alpha/T -- the second rasterizer pass of scripts/splatam.py:257-259
// without its preprocess, sort and alpha evaluation.
template <bool CULL, bool COUNT, bool DS>
__global__ void __launch_bounds__(RING_THREADS) render_fwd_kernel(const RenderFwdArgs a) {
  if (a.hdr->num_rendered > a.Rcap) return;
  __shared__ __align__(128) float4 s_rec[RING_NS][RING_CH * INST_F4];
  __shared__ __align__(8) uint64_t s_full[RING_NS];
  __shared__ int s_rel[RING_NS];
  __shared__ int s_ndone;
  __shared__ int s_stop;

  const int tile = (int)a.order[blockIdx.x / CTA_PARTS], part = blockIdx.x % CTA_PARTS;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const int nch = (n + RING_CH - 1) / RING_CH;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RING_NS; ++s) { mbar_init(&s_full[s], 1); s_rel[s] = 0; }
    s_ndone = 0;
    s_stop = 0x7fffffff;
    mbar_fence_init();
  }
  __syncthreads();

  const char* src = reinterpret_cast<const char*>(a.inst) + (size_t)range.x * INST_BYTES;
  auto fill = [&](int c) {
    const int s = c % RING_NS;
    const uint32_t bytes = (uint32_t)min(RING_CH, n - c * RING_CH) * INST_BYTES;
    mbar_arrive_expect_tx(&s_full[s], bytes);
    tma_load_1d(&s_rec[s][0], src + (size_t)c * RING_CH * INST_BYTES, bytes, &s_full[s]);
  };
  if (tid == 0)
    for (int c = 0; c < min(nch, RING_NS); ++c) fill(c);

  // ------------------------------------------------------------------------------ consumer warps
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  const SubTile st = make_subtile(tile_x, tile_y, part);
  const bool inside = st.px < a.W && st.py < a.H;
  const float pxf = (float)st.px, pyf = (float)st.py;
  const float tile_cx = tile_x * TILE_X + 7.5f, tile_cy = tile_y * TILE_Y + 7.5f;
  const uint32_t rec_base = smem_u32(&s_rec[0][0]);

  // A terminated pixel is encoded as T == 0 (a live T is >= 1e-4): every later candidate then computes
  // test_T == 0 < 1e-4 and changes nothing, so the inner loop needs no separate `done` flag; the
  // transmittance at termination is parked in T_end.
  float T = inside ? 1.0f : 0.f, T_end = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 15.0f;
  float C3 = 0.f, C4 = 0.f, C5 = 0.f;
  uint32_t last = 0, nblend = 0;
  bool warp_done = __all_sync(0xffffffffu, T == 0.f);
  bool counted = false;

  for (int c = 0; c < nch; ++c) {
    const int s = c % RING_NS;
    if (warp_done && !counted) {
      counted = true;
      if (lane == 0) atomicAdd(&s_ndone, 1);
    }
    mbar_wait(&s_full[s], (c / RING_NS) & 1);
    if (c >= *(volatile int*)&s_stop) break;
    if (!warp_done) {
      const int cnt = min(RING_CH, n - c * RING_CH);
      const uint32_t sbase = rec_base + (uint32_t)s * (RING_CH * INST_BYTES);
      const int gbase = c * RING_CH;
      for (int sub = 0; sub < cnt && !warp_done; sub += 32) {
        const int j = sub + lane;
        bool hit = j < cnt;
        if (CULL) {
          if (hit) {
            const float4 q0 = lds128(sbase + j * INST_BYTES);
            const float2 q1 = lds64(sbase + j * INST_BYTES + 16);
            hit = footprint_hits<false>(make_float2(q0.x, q0.y), make_float4(q0.z, q0.w, q1.x, q1.y), tile_cx, tile_cy, st);
          }
        }
        unsigned m = __ballot_sync(0xffffffffu, hit);
        while (m) {
          // pop two hits (front-to-back order); evaluate both alphas, then blend in order
          const int b0 = __ffs(m) - 1;
          m &= m - 1;
          const bool two = m != 0;
          const int b1 = two ? __ffs(m) - 1 : b0;
          m &= m - 1;
          const uint32_t r0 = sbase + (uint32_t)(sub + b0) * INST_BYTES;
          const uint32_t r1 = sbase + (uint32_t)(sub + b1) * INST_BYTES;
          const float4 u0 = lds128(r0), v0 = lds128(r1);
          const float2 u1 = lds64(r0 + 16), v1 = lds64(r1 + 16);
          const float dx0 = u0.x - pxf, dy0 = u0.y - pyf;
          const float dx1 = v0.x - pxf, dy1 = v0.y - pyf;
          const float pw0 = -0.5f * (u0.z * dx0 * dx0 + u1.x * dy0 * dy0) - u0.w * dx0 * dy0;
          const float pw1 = -0.5f * (v0.z * dx1 * dx1 + v1.x * dy1 * dy1) - v0.w * dx1 * dy1;
          const float al0 = fminf(0.99f, u1.y * expf(pw0));
          const float al1 = fminf(0.99f, v1.y * expf(pw1));
          const bool ok0 = !(pw0 > 0.0f) && !(al0 < 1.0f / 255.0f);
          const bool ok1 = two && !(pw1 > 0.0f) && !(al1 < 1.0f / 255.0f);
          if (ok0) {
            const float test_T = T * (1 - al0);
            if (test_T < 0.0001f) {
              if (T != 0.f) { T_end = T; T = 0.f; }
            } else {
              const float4 cq = lds128(r0 + 32);
              C0 += cq.x * al0 * T;
              C1 += cq.y * al0 * T;
              C2 += cq.z * al0 * T;
              if (DS) {
                const float z = lds32(r0 + 24);
                C3 += z * al0 * T;
                C4 += al0 * T;
                C5 += (z * z) * al0 * T;
              }
              if (T > 0.5f && test_T < 0.5f) D = lds32(r0 + 24);
              T = test_T;
              last = (uint32_t)(gbase + sub + b0 + 1);
              if (COUNT) ++nblend;
            }
          }
          if (ok1) {
            const float test_T = T * (1 - al1);
            if (test_T < 0.0001f) {
              if (T != 0.f) { T_end = T; T = 0.f; }
            } else {
              const float4 cq = lds128(r1 + 32);
              C0 += cq.x * al1 * T;
              C1 += cq.y * al1 * T;
              C2 += cq.z * al1 * T;
              if (DS) {
                const float z = lds32(r1 + 24);
                C3 += z * al1 * T;
                C4 += al1 * T;
                C5 += (z * z) * al1 * T;
              }
              if (T > 0.5f && test_T < 0.5f) D = lds32(r1 + 24);
              T = test_T;
              last = (uint32_t)(gbase + sub + b1 + 1);
              if (COUNT) ++nblend;
            }
          }
          if (__all_sync(0xffffffffu, T == 0.f)) { warp_done = true; break; }
        }
      }
    }
    __syncwarp();
    if (lane == 0 && ring_release(&s_rel[s], CTA_WARPS) && c + RING_NS < nch) {
      if (*(volatile int*)&s_ndone == CTA_WARPS) {  // every pixel of the tile has terminated: stop streaming
        atomicMin(&s_stop, c + RING_NS);  // stages already in flight drain; every warp leaves at this chunk
        __threadfence_block();
        mbar_arrive(&s_full[s]);
      } else {
        fill(c + RING_NS);
      }
    }
  }

  if (inside) {
    const int pix = st.py * a.W + st.px;
    const size_t HW = (size_t)a.W * a.H;
    if (T == 0.f) T = T_end;
    a.final_T[pix] = T;
    a.n_contrib[pix] = last;
    a.out_color[pix] = C0 + T * __ldg(a.bg + 0);
    a.out_color[HW + pix] = C1 + T * __ldg(a.bg + 1);
    a.out_color[2 * HW + pix] = C2 + T * __ldg(a.bg + 2);
    if (DS) {
      a.out_color[3 * HW + pix] = C3 + T * __ldg(a.bg + 0);
      a.out_color[4 * HW + pix] = C4 + T * __ldg(a.bg + 1);
      a.out_color[5 * HW + pix] = C5 + T * __ldg(a.bg + 2);
    }
    if (a.out_depth) a.out_depth[pix] = D;
    if (COUNT) a.blended[pix] = nblend;
  }
}

int g_disable_cull = 0;

int launch_render_fwd(const RenderFwdArgs& a, int T, cudaStream_t st) {
  ProfScope prof_(K_RENDER_FWD, st);
  if (T <= 0) return SPLAT_OK;
  if (a.blended) render_fwd_kernel<false, true, false><<<T * CTA_PARTS, RING_THREADS, 0, st>>>(a);
  else if (a.depth_sil) render_fwd_kernel<true, false, true><<<T * CTA_PARTS, RING_THREADS, 0, st>>>(a);
  else if (g_disable_cull) render_fwd_kernel<false, false, false><<<T * CTA_PARTS, RING_THREADS, 0, st>>>(a);
  else render_fwd_kernel<true, false, false><<<T * CTA_PARTS, RING_THREADS, 0, st>>>(a);
  SPLAT_LAUNCH_CHECK("render_fwd_kernel");
  return SPLAT_OK;
}

}  // namespace splat
