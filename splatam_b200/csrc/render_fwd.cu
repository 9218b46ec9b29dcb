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
//  * per-warp 8x8 blocks with an exact conservative footprint test (render.cuh): a warp only
//    evaluates Gaussians whose alpha>=1/255 ellipse can touch its 64 pixels (ballot + ffs walk);
//  * two pixels per lane, blended with packed FP32x2 arithmetic (packed.cuh) including a packed
//    bit-exact expf;
//  * colour and depth travel with the record (the reference re-reads colour from global memory per
//    pixel, forward.cu:361);
//  * warp-level early termination, and the ring stops being refilled once all warps of the CTA are done.
// Per-pixel arithmetic keeps the reference's expression order and precise expf so the discrete
// tests (power>0, alpha<1/255, T<1e-4, median crossing) take the same branches.
#include <type_traits>
#include "render.cuh"
#include "kernels.cuh"
#include "pipeline.cuh"

namespace splat {

// Exact corner test of the footprint cull in the forward too: with 64-pixel hits a rejected hit saves ~40 instructions,
// more than the test costs (r02 A/B at C3: 0.2945 vs 0.3015 ms; with round 1's 32-pixel hits it lost 2%).
#ifndef FWD_EXACT_CULL
#define FWD_EXACT_CULL true
#endif

// DS: fused RGB + depth + silhouette: three more accumulators for the features (z, 1, z*z), z = the
// record's view depth, blended with the SAME alpha/T -- the second rasterizer pass of scripts/splatam.py:257-259
// without its preprocess, sort and alpha evaluation.
//
// Two pixels per lane (render.cuh): the pair's blend is straight-line packed code.  A pixel that fails the alpha test
// runs it with alpha = 0, which leaves its state bit-identical (T * (1 - 0) == T, fma(T, 0 * c, C) == C); a terminated
// pixel carries T == -0.0f (see the kernel body).  r02a capture (profiles/): 3.45 M warp-level hits per launch at C3,
// 91% of them with at least one blending pixel, 87 instructions per hit before this form, ~65 after.
//
// GATHER: the ring stages are not streamed from a re-written copy of the records (emit_records: 100 B per instance of
// extra traffic, 60 us at C3) but GATHERED: the refilling warp reads the chunk's 128 sorted ids (4 coalesced loads) and
// issues three 16-byte cp.async per record straight from the Gaussians' own 48-byte rows -- which preprocess writes in
// exactly the record layout, id included, and which sit in L2 (48 MB at 1M Gaussians) -- completing on the same
// per-stage mbarrier (cp.async.mbarrier.arrive.noinc, expected count 32 lanes instead of the TMA path's 1).
template <bool CULL, bool COUNT, bool DS, bool GATHER>
__global__ void __launch_bounds__(RING_THREADS) render_fwd_kernel(const RenderFwdArgs a) {
  pdl_wait();
  if (a.hdr->num_rendered > a.Rcap) return;
  __shared__ __align__(128) float4 s_rec[RING_NS][RING_CH * INST_F4];
  __shared__ __align__(8) uint64_t s_full[RING_NS];
  __shared__ int s_rel[RING_NS];
  __shared__ int s_ndone;
  __shared__ int s_stop;

  const int tile = (int)a.order[blockIdx.x];
  const int tid = threadIdx.x, lane = tid & 31;
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const int nch = (n + RING_CH - 1) / RING_CH;
  SPLAT_CHECK(range.x <= range.y && range.y <= a.hdr->num_rendered && a.hdr->num_rendered <= a.Rcap);

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RING_NS; ++s) { mbar_init(&s_full[s], GATHER ? 32 : 1); s_rel[s] = 0; }
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
  // whole-warp refill of stage c % RING_NS in gather mode
  const uint32_t* plist = a.plist + range.x;
  auto gather = [&](int c) {
    const int s = c % RING_NS, base = c * RING_CH, cnt = min(RING_CH, n - base);
    const uint32_t dst = smem_u32(&s_rec[s][0]);
    uint32_t id[RING_CH / 32];
#pragma unroll
    for (int k = 0; k < RING_CH / 32; ++k) {
      const int j = lane + 32 * k;
      id[k] = j < cnt ? __ldg(plist + base + j) : 0u;
    }
#pragma unroll
    for (int k = 0; k < RING_CH / 32; ++k) {
      const int j = lane + 32 * k;
      if (j < cnt) {
        SPLAT_CHECK(id[k] < (uint32_t)a.P && base + j < n && j < RING_CH);
        const float4* src = a.grec + (size_t)id[k] * INST_F4;
#pragma unroll
        for (int q = 0; q < INST_F4; ++q) cp_async16(dst + (uint32_t)(j * INST_F4 + q) * 16u, src + q);
      }
    }
    cp_async_mbar_arrive_noinc(&s_full[s]);
  };
  if (GATHER) {
    if (tid < 32)
      for (int c = 0; c < min(nch, RING_NS); ++c) gather(c);
  } else if (tid == 0) {
    for (int c = 0; c < min(nch, RING_NS); ++c) fill(c);
  }

  // ------------------------------------------------------------------------------ consumer warps
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  const SubTile st = make_subtile(tile_x, tile_y);
  const bool in_a = st.px < a.W && st.pya < a.H, in_b = st.px < a.W && st.pyb < a.H;
  const float pxf = (float)st.px;
  const f2 npy2 = pk(-(float)st.pya, -(float)st.pyb);
  const float tile_cx = tile_x * TILE_X + 7.5f, tile_cy = tile_y * TILE_Y + 7.5f;
  const uint32_t rec_base = smem_u32(&s_rec[0][0]);

  // A pixel that has terminated (or lies outside the image) is DEAD and is encoded as T == -0.0f: every product with
  // it is a signed zero, so C += (alpha c) * T and T *= (1 - alpha) leave its state bit-identical with no select, and
  // the termination test below, an UNSIGNED compare of the bits of test_T against those of 1e-4, is false for it.
  // The transmittance it terminated with is parked in Te.
  constexpr uint32_t T_MIN_BITS = 0x38d1b717u;   // 0.0001f (forward.cu:355)
  const float DEAD = __int_as_float(0x80000000);
  constexpr uint32_t LAST_FROZEN = 0x80000000u;
  float Ta = in_a ? 1.0f : DEAD, Tb = in_b ? 1.0f : DEAD, Tea = 1.0f, Teb = 1.0f, Da = 15.0f, Db = 15.0f;
  // loop-carried accumulators live as scalar pairs and are packed at the point of use: ptxas coalesces the pair
  // into the FFMA2 destination that way; carried as 64-bit values it copied each of them with two MOVs per hit
  float c0a = 0.f, c0b = 0.f, c1a = 0.f, c1b = 0.f, c2a = 0.f, c2b = 0.f;
  float c3a = 0.f, c3b = 0.f, c4a = 0.f, c4b = 0.f, c5a = 0.f, c5b = 0.f;
  uint32_t last_a = 0, last_b = 0, nbl_a = 0, nbl_b = 0;
  bool warp_done = __all_sync(0xffffffffu, !in_a && !in_b);
  bool counted = false;
  const ExpConsts ek = exp_consts(a.zero & tid);   // lane-dependent as far as ptxas knows: stays in VECTOR registers

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
            hit = footprint_hits<FWD_EXACT_CULL>(make_float2(q0.x, q0.y), make_float4(q0.z, q0.w, q1.x, q1.y), tile_cx, tile_cy, st);
          }
        }
        // hits in list order = ascending lane = descending bit of the REVERSED ballot: the pop is one FLO (highest
        // set bit t, lane = 31 - t), and record address and list position are affine in t
        unsigned m = __brev(__ballot_sync(0xffffffffu, hit));
        const uint32_t r31 = sbase + (uint32_t)(sub + 31) * INST_BYTES;
        const uint32_t pos31 = (uint32_t)(gbase + sub + 32);
        auto walk = [&](auto median_tag) {
        constexpr bool MEDIAN = decltype(median_tag)::value;
        while (m) {
          uint32_t t;
          asm("bfind.u32 %0, %1;" : "=r"(t) : "r"(m));
          {
            uint32_t below;   // bits under the popped one
            asm("bmsk.clamp.b32 %0, %1, %2;" : "=r"(below) : "r"(0u), "r"(t));
            m &= below;
          }
          const uint32_t r = r31 - t * INST_BYTES;
          SPLAT_CHECK(r >= sbase && r + INST_BYTES <= sbase + RING_CH * INST_BYTES && (int)(sub + 31 - t) < cnt);
          const float4 q0 = lds128(r), q1 = lds128(r + 16);   // (x, y, A, B), (C, opacity, depth, id)
          SPLAT_CHECK(__float_as_uint(q1.w) < (uint32_t)a.P);
          const float4 cq = lds128(r + 32);                     // (r, g, b, -)
          const float dx = q0.x - pxf;
          const f2 dy2 = add2(bc(q0.y), npy2);
          const f2 pw2 = pair_power(dx, dy2, q0.z, q0.w, q1.x);
          float pwa, pwb, ala, alb;
          upk(pw2, pwa, pwb);
          upk(mul2(bc(q1.y), exp2_ref(pw2, ek)), ala, alb);
          // min(0.99, x) < 1/255  <=>  x < 1/255: the test reads the unclamped product
          const bool va = !(pwa > 0.0f) && !(ala < 1.0f / 255.0f);
          const bool vb = !(pwb > 0.0f) && !(alb < 1.0f / 255.0f);
          float aa = va ? fminf(0.99f, ala) : 0.f, ab = vb ? fminf(0.99f, alb) : 0.f;
          const f2 T2 = pk(Ta, Tb);
          f2 om2 = add2(bc(1.0f), pk(-aa, -ab));
          float tta, ttb;
          upk(mul2(T2, om2), tta, ttb);   // test_T = T * (1 - alpha)
          // termination EVENT of a live pixel: 0 < test_T < 1e-4 (a live T is >= 1e-4 and 1 - alpha >= 0.01, so its
          // test_T is > 0; a dead pixel's is -0 = 0x80000000, above the threshold as an unsigned integer)
          const bool ea = __float_as_uint(tta) < T_MIN_BITS, eb = __float_as_uint(ttb) < T_MIN_BITS;
          if (__any_sync(0xffffffffu, ea || eb)) {   // rare: at most once per pixel
            float oma, omb;
            upk(om2, oma, omb);
            // `done = true; continue` (forward.cu:356-359): no contribution, and last_contributor is frozen (top bit)
            if (ea) { Tea = Ta; tta = DEAD; aa = 0.f; oma = DEAD; last_a |= LAST_FROZEN; }
            if (eb) { Teb = Tb; ttb = DEAD; ab = 0.f; omb = DEAD; last_b |= LAST_FROZEN; }
            om2 = pk(oma, omb);
            if (__all_sync(0xffffffffu, __float_as_uint(tta) == 0x80000000u && __float_as_uint(ttb) == 0x80000000u)) {
              warp_done = true;   // this hit's (zero) contribution is still accumulated below; the walk ends with it
              m = 0;
            }
          }
          const f2 al2 = pk(aa, ab);
          upk(fma2(T2, mul2(al2, bc(cq.x)), pk(c0a, c0b)), c0a, c0b);      // C += (c * alpha) * T  (forward.cu:361)
          upk(fma2(T2, mul2(al2, bc(cq.y)), pk(c1a, c1b)), c1a, c1b);
          upk(fma2(T2, mul2(al2, bc(cq.z)), pk(c2a, c2b)), c2a, c2b);
          const float z = q1.z;
          if (DS) {
            upk(fma2(T2, mul2(al2, bc(z)), pk(c3a, c3b)), c3a, c3b);
            upk(fma2(T2, al2, pk(c4a, c4b)), c4a, c4b);
            upk(fma2(T2, mul2(al2, bc(z * z)), pk(c5a, c5b)), c5a, c5b);
          }
          if (MEDIAN) {
            if (Ta > 0.5f && tta < 0.5f) Da = z;          // median depth (forward.cu:365-367); false for a dead pixel
            if (Tb > 0.5f && ttb < 0.5f) Db = z;
          }
          const uint32_t pos1 = pos31 - t;
          // contributor = valid, live and not terminating.  A pixel dies only through a termination event, which froze
          // its last_contributor by setting the top bit: an unsigned max then leaves it alone, so the walk needs no
          // liveness test here (positions are tile-relative and far below 2^31); the stats variant keeps the plain form.
          if (COUNT) {
            if (va && tta >= 0.0001f) { last_a = pos1; ++nbl_a; }
            if (vb && ttb >= 0.0001f) { last_b = pos1; ++nbl_b; }
          } else {
            if (va) last_a = max(last_a, pos1);
            if (vb) last_b = max(last_b, pos1);
          }
          // T <- test_T, formed again IN PLACE (same operands, same bits; an event's factor is -0, which parks the pixel
          // at DEAD): the loop-carried pair keeps its registers, where `T = test_T` cost two to four MOVs per hit.
          // The empty asm hides the factor from common-subexpression elimination.
          asm("" : "+l"(om2));
          upk(mul2(T2, om2), Ta, Tb);
        }
        };
        // The median-depth test can only fire while some pixel of the warp still has T > 0.5 (T never grows): once all
        // 64 are at or below it -- after the first one to three blends on an opaque surface, i.e. for most of the list --
        // the walk runs a body without those six instructions.  Checked once per 32 candidates.
        if (__any_sync(0xffffffffu, Ta > 0.5f || Tb > 0.5f)) walk(std::true_type{});
        else walk(std::false_type{});
      }
    }
    __syncwarp();
    if (GATHER) {
      int act = 0;   // 0: nothing, 1: refill, 2: stop streaming
      if (lane == 0 && ring_release(&s_rel[s], CTA_WARPS) && c + RING_NS < nch) {
        act = 1;
        if (*(volatile int*)&s_ndone == CTA_WARPS) {
          atomicMin(&s_stop, c + RING_NS);
          __threadfence_block();
          act = 2;
        }
      }
      act = __shfl_sync(0xffffffffu, act, 0);
      if (act == 1) gather(c + RING_NS);
      else if (act == 2) mbar_arrive(&s_full[s]);   // all 32 lanes: completes the phase so waiting warps see s_stop
    } else if (lane == 0 && ring_release(&s_rel[s], CTA_WARPS) && c + RING_NS < nch) {
      if (*(volatile int*)&s_ndone == CTA_WARPS) {  // every pixel of the tile has terminated: stop streaming
        atomicMin(&s_stop, c + RING_NS);  // stages already in flight drain; every warp leaves at this chunk
        __threadfence_block();
        mbar_arrive(&s_full[s]);
      } else {
        fill(c + RING_NS);
      }
    }
  }

  pdl_trigger();   // this CTA's list is done: the next kernel's CTAs may start taking the freed slots
  const size_t HW = (size_t)a.W * a.H;
  const float bg0 = __ldg(a.bg + 0), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
  auto store = [&](int py, float T, float Te, float D, uint32_t last, uint32_t nbl, float c0, float c1, float c2,
                   float c3, float c4, float c5) {
    const int pix = py * a.W + st.px;
    if (__float_as_uint(T) == 0x80000000u) T = Te;
    a.final_T[pix] = T;
    a.n_contrib[pix] = last & ~LAST_FROZEN;
    a.out_color[pix] = c0 + T * bg0;
    a.out_color[HW + pix] = c1 + T * bg1;
    a.out_color[2 * HW + pix] = c2 + T * bg2;
    if (DS) {
      a.out_color[3 * HW + pix] = c3 + T * bg0;
      a.out_color[4 * HW + pix] = c4 + T * bg1;
      a.out_color[5 * HW + pix] = c5 + T * bg2;
    }
    if (a.out_depth) a.out_depth[pix] = D;
    if (COUNT) a.blended[pix] = nbl;
  };
  if (in_a) store(st.pya, Ta, Tea, Da, last_a, nbl_a, c0a, c1a, c2a, c3a, c4a, c5a);
  if (in_b) store(st.pyb, Tb, Teb, Db, last_b, nbl_b, c0b, c1b, c2b, c3b, c4b, c5b);
}

int g_disable_cull = 0;

// splat_debug_exp_mismatches: packed exp2_ref vs the device library's expf, bit for bit.
__global__ void __launch_bounds__(256) exp_check_kernel(uint32_t first, unsigned long long pairs, unsigned long long count,
                                                        unsigned long long* out) {
  unsigned long long bad = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < pairs;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t b0 = first + (uint32_t)(2 * i), b1 = b0 + 1u;
    const float x0 = __uint_as_float(b0), x1 = __uint_as_float(b1);
    float e0, e1;
    upk(exp2_ref(pk(x0, x1)), e0, e1);
    const float r0 = expf(x0), r1 = expf(x1);
    const bool ok0 = __float_as_uint(e0) == __float_as_uint(r0) || (e0 != e0 && r0 != r0);
    const bool ok1 = 2 * i + 1 >= count || __float_as_uint(e1) == __float_as_uint(r1) || (e1 != e1 && r1 != r1);
    bad += (ok0 ? 0 : 1) + (ok1 ? 0 : 1);
  }
  bad = __reduce_add_sync(0xffffffffu, (unsigned)bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, bad);
}

int launch_exp_check(uint32_t first, unsigned long long count, unsigned long long* out, cudaStream_t st) {
  if (count == 0) return SPLAT_OK;
  const unsigned long long pairs = (count + 1) / 2;
  exp_check_kernel<<<148 * 8, 256, 0, st>>>(first, pairs, count, out);
  SPLAT_LAUNCH_CHECK("exp_check_kernel");
  return SPLAT_OK;
}

int g_gather = 1;   // measured default (r02 A/B at C3: 1023 vs 982 it/s; mapping iteration 747 vs 704): see splat_set_gather

int launch_render_fwd(const RenderFwdArgs& a, int T, cudaStream_t st) {
  ProfScope prof_(K_RENDER_FWD, st);
  if (T <= 0) return SPLAT_OK;
  const bool g = g_gather != 0 && a.grec && a.plist;
#define FWD_LAUNCH(C, N, D)                                                                  \
  do {                                                                                       \
    if (g) SPLAT_CUDA(launch_k(render_fwd_kernel<C, N, D, true>, dim3(T), dim3(RING_THREADS), 0, st, a));    \
    else SPLAT_CUDA(launch_k(render_fwd_kernel<C, N, D, false>, dim3(T), dim3(RING_THREADS), 0, st, a));  \
  } while (0)
  if (a.blended) FWD_LAUNCH(false, true, false);
  else if (a.depth_sil) FWD_LAUNCH(true, false, true);
  else if (g_disable_cull) FWD_LAUNCH(false, false, false);
  else FWD_LAUNCH(true, false, false);
#undef FWD_LAUNCH
  SPLAT_LAUNCH_CHECK("render_fwd_kernel");
  return SPLAT_OK;
}

}  // namespace splat
