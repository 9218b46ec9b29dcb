// render_bwd.cu -- back-to-front re-blend and per-Gaussian gradient accumulation.
// Restates renderCUDA<3> backward (HX/cuda_rasterizer/backward.cu:586-748; live in DG) and the
// per-pair head of renderCUDAFused<3> (backward.cu:850-1040; live in HX).
//
// Reference: every blended (pixel,Gaussian) pair issues 9 (DG) or 25 (HX) global float atomics and
// the CTA re-syncs every 256 Gaussians.  Here:
//   1. the tile's instance records are streamed BACK TO FRONT through the same TMA + mbarrier ring
//      as the forward pass (no CTA barrier in the loop; the last warp to release a stage refills it);
//   2. each warp walks only the Gaussians whose footprint can touch its 8x4 pixels (render.cuh);
//   3. the per-pair partials of up to GROUP Gaussians are summed over the warp's 32 pixels by a
//      transpose through a padded per-warp shared-memory tile: lane l stores value k at
//      [k][l] (conflict-free), then lane k reads row k with eight 128-bit loads (conflict-free,
//      row stride 36 words) and folds it with packed FADD2;
//   4. lane k issues ONE red.global.add.f32 for its total into a packed per-Gaussian row
//      (12 or 20 floats, so the 9 or 19 atomics of a Gaussian coalesce into 2-3 L2 sectors).
// => one global reduction per (warp, Gaussian) hit instead of one per (pixel, Gaussian) pair.
//
// MODE 0 (power == 1), row = 12 floats of RAW sums over blended pairs, u = G * dL/dalpha:
//     [0..2] alpha*T*dL/dpix[ch]   [3] u*dx  [4] u*dy  [5] u*dx*dx  [6] u*dx*dy  [7] u*dy*dy  [8] u
//   The per-Gaussian factors (opacity, conic, 0.5*W, -0.5) that the reference multiplies in per pair
//   (backward.cu:717-745) are applied once per Gaussian in preprocess_bwd -- the map is linear.
// MODE 2 (power == 1, fused RGB + depth + silhouette): six feature channels (r, g, b, z, 1, z*z) share the
//   pair's alpha and T; row = MODE 0's nine sums + [9] alpha*T*dL/dpix[3] + [10] alpha*T*dL/dpix[5] (the z and z*z
//   "colour" gradients, which preprocess_bwd turns into dL/dmean through z; the constant channel has no parameter).
// MODE 3 = MODE 2 for callers that need no colour / opacity gradient (pose-only tracking: splat_backward with
//   dL_dcolors == dL_dopacity == NULL): the four sums [0..2] and [8] are neither formed nor reduced; same row layout.
// MODE 1 (power == 2), row = 20 floats: dcolor^2[3], dopacity^2, and the 15 second moments of
//   v = (dconic.x, dconic.y, dconic.w, dmean2D.x, dmean2D.y); hessian_finalize turns the moments
//   into sum_pairs (J_k . v)^2 per output k, which is what HX accumulates pair by pair
//   (backward.cu:1093-1137).  The conic triple is taken in the splat's principal-axis frame
//   (d' = (dx + t*dy, dy - t*dx), t from the record) so that the quartic forms stay well
//   conditioned for thin splats; finalize maps J into the same frame.
#include "render.cuh"
#include "kernels.cuh"
#include "pipeline.cuh"

namespace splat {

template <int MODE> struct BwdTraits;
template <> struct BwdTraits<0> { static const int VALS = 9, GROUP = 3, STRIDE = 12; };
template <> struct BwdTraits<1> { static const int VALS = 19, GROUP = 1, STRIDE = 20; };
template <> struct BwdTraits<2> { static const int VALS = 11, GROUP = 2, STRIDE = 12; };
template <> struct BwdTraits<3> { static const int VALS = 7, GROUP = 4, STRIDE = 12; };   // MODE 2 without rows 0-2 and 8

// The backward ring's chunk length is its own (the forward keeps RING_CH): with 64-record stages the CTA's dynamic
// shared memory is 3*64*48 + 4*28*36*4 = 25.3 KB, so BWD_CTAS = 8 CTAs fit an SM (34.5 KB with 128-record stages = 6).
#ifndef BWD_CH
#define BWD_CH 128
#endif
#define TR_STRIDE 36                        // words per transposed row (32 lanes + 4 pad)
#define TR_ROWS 28
#define TR_WARP_BYTES (TR_ROWS * TR_STRIDE * 4)
#define BWD_SMEM (RING_NS * BWD_CH * INST_BYTES + CTA_WARPS * TR_WARP_BYTES)

__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 r;
  asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; "
      "mov.b64 {%0,%1}, rc;}"
      : "=f"(r.x), "=f"(r.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return r;
}

// Sum of the 32 floats of one transposed row (8 x LDS.128, 15 x FADD2 + 1 FADD).
__device__ __forceinline__ float row_sum32(uint32_t addr) {
  float4 q[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) q[t] = lds128(addr + 16 * t);
  float2 s[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) s[t] = fadd2(make_float2(q[t].x, q[t].y), make_float2(q[t].z, q[t].w));
#pragma unroll
  for (int t = 0; t < 4; ++t) s[t] = fadd2(s[t], s[t + 4]);
  s[0] = fadd2(s[0], s[2]);
  s[1] = fadd2(s[1], s[3]);
  s[0] = fadd2(s[0], s[1]);
  return s[0].x + s[0].y;
}

// Resident CTAs per SM the register allocation is held to: r02a (profiles/) had 120 registers = 4 CTAs = 16 warps per SM and
// the kernel waited on latency (issue slots 62% busy); 6 CTAs = 24 warps need <= 80 registers.
#ifndef BWD_CTAS
#define BWD_CTAS 6
#endif
#define BWD_MIN_CTAS(MODE) ((MODE) == 1 ? 4 : BWD_CTAS)

// GATHER: the ring is filled from the Gaussians' own records by id (see render_fwd.cu); no instance-record copy exists.
template <bool CULL, int MODE, bool GATHER>
__global__ void __launch_bounds__(RING_THREADS, BWD_MIN_CTAS(MODE)) render_bwd_kernel(const RenderBwdArgs a) {
  pdl_wait();
  constexpr int VALS = BwdTraits<MODE>::VALS, GROUP = BwdTraits<MODE>::GROUP, STRIDE = BwdTraits<MODE>::STRIDE;
  constexpr int NV = VALS * GROUP;
  // A replay of a captured graph whose forward overflowed its binning capacity left the instance records and the
  // image state untouched (every forward kernel exits on R > Rcap): there is nothing valid to walk.  The accumulator
  // stays zero, so preprocess_bwd writes all-zero gradients for this step (splat_forward_guard reports the overflow).
  if (a.hdr->num_rendered > a.Rcap) return;
  SPLAT_CHECK(a.ranges[a.order[blockIdx.x]].y <= a.hdr->num_rendered);
  extern __shared__ __align__(128) unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t s_full[RING_NS];
  __shared__ int s_rel[RING_NS];
  __shared__ int s_max[CTA_WARPS];

  const int tile = (int)a.order[blockIdx.x];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint2 range = a.ranges[tile];
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  const SubTile st = make_subtile(tile_x, tile_y);
  const bool in_a = st.px < a.W && st.pya < a.H, in_b = st.px < a.W && st.pyb < a.H;
  const int pix_a = st.pya * a.W + st.px, pix_b = st.pyb * a.W + st.px;
  const int last_a = in_a ? (int)a.n_contrib[pix_a] : 0, last_b = in_b ? (int)a.n_contrib[pix_b] : 0;
  const int warp_hi = __reduce_max_sync(0xffffffffu, max(last_a, last_b));

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < RING_NS; ++s) { mbar_init(&s_full[s], GATHER ? 32 : 1); s_rel[s] = 0; }
    mbar_fence_init();
  }
  if (lane == 0) s_max[wid] = warp_hi;
  __syncthreads();
  int hi = 0;  // list positions >= max(n_contrib) over the tile are never touched by anyone
#pragma unroll
  for (int w = 0; w < CTA_WARPS; ++w) hi = max(hi, s_max[w]);
  const int nch = (hi + BWD_CH - 1) / BWD_CH;

  // chunk c = list positions [max(0, hi-(c+1)*BWD_CH), hi - c*BWD_CH), streamed into stage c % RING_NS
  const char* src = reinterpret_cast<const char*>(a.inst) + (size_t)range.x * INST_BYTES;
  auto fill = [&](int c) {
    const int s = c % RING_NS;
    const int top = hi - c * BWD_CH, lo = max(0, top - BWD_CH);
    const uint32_t bytes = (uint32_t)(top - lo) * INST_BYTES;
    mbar_arrive_expect_tx(&s_full[s], bytes);
    tma_load_1d(smem_dyn + s * (BWD_CH * INST_BYTES), src + (size_t)lo * INST_BYTES, bytes, &s_full[s]);
  };
  const uint32_t* plist = a.plist + range.x;
  // (Requesting the next chunk's ids early in EVERY warp, so that the refilling warp finds them in registers, was
  // measured and is slower: 0.520 vs 0.477 ms at C3 -- four redundant loads per warp and chunk, and the extra live
  // registers spill at the 80-register budget.)
  auto load_ids = [&](int c, uint32_t (&id)[BWD_CH / 32]) {
    const int top = hi - c * BWD_CH, lo = max(0, top - BWD_CH), cnt = top - lo;
#pragma unroll
    for (int k = 0; k < BWD_CH / 32; ++k) {
      const int j = lane + 32 * k;
      id[k] = j < cnt ? __ldg(plist + lo + j) : 0u;
    }
  };
  auto gather_ids = [&](int c, const uint32_t (&id)[BWD_CH / 32]) {   // whole warp
    const int s = c % RING_NS;
    const int top = hi - c * BWD_CH, lo = max(0, top - BWD_CH), cnt = top - lo;
    const uint32_t dst = smem_u32(smem_dyn + s * (BWD_CH * INST_BYTES));
#pragma unroll
    for (int k = 0; k < BWD_CH / 32; ++k) {
      const int j = lane + 32 * k;
      if (j < cnt) {
        SPLAT_CHECK(id[k] < (uint32_t)a.P && lo >= 0 && lo + j < (int)(range.y - range.x) && j < BWD_CH);
        const float4* src = a.grec + (size_t)id[k] * INST_F4;
#pragma unroll
        for (int q = 0; q < INST_F4; ++q) cp_async16(dst + (uint32_t)(j * INST_F4 + q) * 16u, src + q);
      }
    }
    cp_async_mbar_arrive_noinc(&s_full[s]);
  };
  auto gather = [&](int c) {
    uint32_t id[BWD_CH / 32];
    load_ids(c, id);
    gather_ids(c, id);
  };
  if (GATHER) {
    if (tid < 32)
      for (int c = 0; c < min(nch, RING_NS); ++c) gather(c);
  } else if (tid == 0) {
    for (int c = 0; c < min(nch, RING_NS); ++c) fill(c);
  }

  // ------------------------------------------------------------------------------ consumer warps
  // Per-pixel state of the lane's two pixels (a, b).  Loop-carried values (T, Rs, MODE 1's running sums) are scalar
  // pairs packed at the point of use: carried as 64-bit values ptxas copies them with two MOVs per hit.
  const float pxf = (float)st.px;
  const f2 npy2 = pk(-(float)st.pya, -(float)st.pyb);
  const float tile_cx = tile_x * TILE_X + 7.5f, tile_cy = tile_y * TILE_Y + 7.5f;
  const size_t HW = (size_t)a.W * a.H;
  constexpr bool DS = (MODE == 2 || MODE == 3);
  constexpr int NCH = DS ? 6 : 3;
  const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
  f2 dp[NCH], nk[NCH];      // dL/dpix per channel; minus (reference level * dL/dpix) (MODE 0/2/3)
  float Ta, Tb, Rsa, Rsb;
  f2 nTb2;
  {
    float d_a[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, d_b[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float k_a[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, k_b[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
      if (in_a) d_a[ch] = a.dL_dpix[ch * HW + pix_a];
      if (in_b) d_b[ch] = a.dL_dpix[ch * HW + pix_b];
    }
    // MODE 0/2/3: features are measured relative to a per-pixel reference level (the rendered value when the caller
    // passes it): k_ch = ref_ch * dL/dpix_ch.  Rs = sum over the pairs behind of alpha*T*((c - ref) . dL/dpix), seeded
    // with the background term.
    if (MODE != 1 && a.out_color != nullptr) {
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        if (in_a) k_a[ch] = a.out_color[ch * HW + pix_a] * d_a[ch];
        if (in_b) k_b[ch] = a.out_color[ch * HW + pix_b] * d_b[ch];
      }
    }
    const float Tf_a = in_a ? a.final_T[pix_a] : 0.f, Tf_b = in_b ? a.final_T[pix_b] : 0.f;
    float bgd_a = bg0 * d_a[0] + bg1 * d_a[1] + bg2 * d_a[2], bgd_b = bg0 * d_b[0] + bg1 * d_b[1] + bg2 * d_b[2];
    if (DS) {
      bgd_a += bg0 * d_a[3] + bg1 * d_a[4] + bg2 * d_a[5];
      bgd_b += bg0 * d_b[3] + bg1 * d_b[4] + bg2 * d_b[5];
    }
    Ta = Tf_a; Tb = Tf_b;
    nTb2 = pk(-Tf_a * bgd_a, -Tf_b * bgd_b);
    Rsa = Tf_a * (bgd_a - (((k_a[0] + k_a[1]) + k_a[2]) + ((k_a[3] + k_a[4]) + k_a[5])));
    Rsb = Tf_b * (bgd_b - (((k_b[0] + k_b[1]) + k_b[2]) + ((k_b[3] + k_b[4]) + k_b[5])));
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) { dp[ch] = pk(d_a[ch], d_b[ch]); nk[ch] = pk(-k_a[ch], -k_b[ch]); }
  }
  const f2 c4_2 = DS ? add2(dp[DS ? 4 : 0], nk[DS ? 4 : 0]) : bc(0.f);   // the constant channel's (1 - ref) * dL/dpix
  // MODE 1: the reference's running accum_rec per channel
  float ar0a = 0.f, ar0b = 0.f, ar1a = 0.f, ar1b = 0.f, ar2a = 0.f, ar2b = 0.f, laa = 0.f, lab = 0.f;
  float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
  const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;
  const uint32_t rec_base = smem_u32(smem_dyn);
  const uint32_t tr_base = rec_base + RING_NS * BWD_CH * INST_BYTES + wid * TR_WARP_BYTES;
  const uint32_t tr_st = tr_base + lane * 4;                 // this lane's column
  const uint32_t tr_ld = tr_base + lane * (TR_STRIDE * 4);   // this lane's row (lane < NV)
  const int my_g = lane / VALS, my_k = lane - my_g * VALS;
  // MODE 3 keeps MODE 2's row layout: values 0-4 -> slots 3-7, values 5, 6 -> slots 9, 10
  const int slot = (MODE == 3) ? (my_k < 5 ? my_k + 3 : my_k + 4) : my_k;
  const ExpConsts ek = exp_consts(a.zero & tid);   // lane-dependent as far as ptxas knows: stays in vector registers

  for (int c = 0; c < nch; ++c) {
    const int s = c % RING_NS;
    mbar_wait(&s_full[s], (c / RING_NS) & 1);
    const int top = hi - c * BWD_CH, lo = max(0, top - BWD_CH), cnt = top - lo;
    if (warp_hi > lo) {
      const uint32_t sbase = rec_base + (uint32_t)s * (BWD_CH * INST_BYTES);
      for (int sub = ((cnt - 1) >> 5) << 5; sub >= 0; sub -= 32) {
        const int j = sub + lane;
        bool hit = j < cnt && (lo + j) < warp_hi;
        if (CULL) {
          if (hit) {
            const float4 q0 = lds128(sbase + j * INST_BYTES);
            const float2 q1 = lds64(sbase + j * INST_BYTES + 16);
            hit = footprint_hits<true>(make_float2(q0.x, q0.y), make_float4(q0.z, q0.w, q1.x, q1.y), tile_cx, tile_cy, st);
          }
        }
        unsigned m = __ballot_sync(0xffffffffu, hit);
        const uint32_t rsub = sbase + (uint32_t)sub * INST_BYTES;
        const int pos_sub = lo + sub;
        while (m) {
          uint32_t ids[GROUP];
          bool anyg[GROUP];
          bool any = false;
#pragma unroll
          for (int gsl = 0; gsl < GROUP; ++gsl) {
            ids[gsl] = 0xffffffffu;
            anyg[gsl] = false;
            if (m) {
              uint32_t b;
              asm("bfind.u32 %0, %1;" : "=r"(b) : "r"(m));   // back to front: highest list position first
              m ^= 1u << b;   // (UBMSK pop, no id preset, group mask instead of flags: each measured at C3, 0 to +1.5 % time)
              const uint32_t r = rsub + b * INST_BYTES;
              const int pos = pos_sub + (int)b;
              SPLAT_CHECK(r >= sbase && r + INST_BYTES <= sbase + BWD_CH * INST_BYTES && sub + (int)b < cnt && pos < hi);
              // (Reading the next hit's record one hit ahead -- q0/q1, or all three quads -- to hide this latency costs 8-12
              // loop-carried register copies per hit: 0.54 ms instead of 0.476 at C3.)
              const float4 q0 = lds128(r), q1 = lds128(r + 16);
              ids[gsl] = __float_as_uint(q1.w);
              SPLAT_CHECK(ids[gsl] < (uint32_t)a.P);
              const float dx = q0.x - pxf;
              const f2 dy2 = add2(bc(q0.y), npy2);
              // power, exp and alpha exactly as the forward pass formed them: the blend decisions below are the same
              const f2 pw2 = pair_power(dx, dy2, q0.z, q0.w, q1.x);
              const f2 G2 = exp2_ref(pw2, ek);
              float pwa, pwb, ala, alb, Ga, Gb;
              upk(pw2, pwa, pwb);
              upk(G2, Ga, Gb);
              upk(mul2(bc(q1.y), G2), ala, alb);
              // min(0.99, x) < 1/255  <=>  x < 1/255: the test reads the unclamped product
              const bool va = pos < last_a && !(pwa > 0.0f) && !(ala < 1.0f / 255.0f);
              const bool vb = pos < last_b && !(pwb > 0.0f) && !(alb < 1.0f / 255.0f);
              // Warp-uniform branch, then straight-line packed code for all 64 pixels: a pixel that does not
              // blend this Gaussian runs the same arithmetic with alpha = G = 0, which leaves its
              // running state bit-identical (1/(1-0) == 1 exactly, T*1 == T, and the deferred
              // accum_rec update 0*c + 1*accum is exact) and makes all of its partials exactly 0.
              anyg[gsl] = __any_sync(0xffffffffu, va || vb);
              if (anyg[gsl]) {
                any = true;
                const f2 al2 = pk(va ? fminf(0.99f, ala) : 0.f, vb ? fminf(0.99f, alb) : 0.f);
                const f2 Gv2 = pk(va ? Ga : 0.f, vb ? Gb : 0.f);
                const f2 rc2 = rcp_newton2(add2(bc(1.f), neg2(al2)));
                const f2 T2 = mul2(pk(Ta, Tb), rc2);
                upk(T2, Ta, Tb);
                const float4 col = lds128(r + 32);
                const f2 w2 = mul2(al2, T2);
                f2 dLda2;
                if (MODE == 1) {
                  // the reference's running form (backward.cu:694-713): accum_rec per channel
                  const f2 la2 = pk(laa, lab);
                  const f2 ola2 = add2(bc(1.f), neg2(la2));
                  const f2 ar0 = fma2(la2, bc(lc0), mul2(ola2, pk(ar0a, ar0b)));
                  const f2 ar1 = fma2(la2, bc(lc1), mul2(ola2, pk(ar1a, ar1b)));
                  const f2 ar2 = fma2(la2, bc(lc2), mul2(ola2, pk(ar2a, ar2b)));
                  upk(ar0, ar0a, ar0b); upk(ar1, ar1a, ar1b); upk(ar2, ar2a, ar2b);
                  lc0 = col.x; lc1 = col.y; lc2 = col.z;
                  dLda2 = mul2(add2(bc(col.x), neg2(ar0)), dp[0]);
                  dLda2 = fma2(add2(bc(col.y), neg2(ar1)), dp[1], dLda2);
                  dLda2 = fma2(add2(bc(col.z), neg2(ar2)), dp[2], dLda2);
                  dLda2 = fma2(dLda2, T2, mul2(nTb2, rc2));
                  upk(al2, laa, lab);
                } else {
                  // Same quantity with ONE scalar suffix sum instead of a running average per channel:
                  //   T_i * accum_rec_i[ch] = (1/(1-alpha_i)) * sum_{j behind i} alpha_j T_j c_j[ch]
                  //   => dL/dalpha_i = T_i (c_i . dp) - (1/(1-alpha_i)) (sum_{j behind i} w_j (c_j . dp) + T_final bg . dp)
                  // each channel's (c - ref) * dp is ONE fma (exact product, single rounding of the difference)
                  f2 sdot = add2(add2(fma2(bc(col.x), dp[0], nk[0]), fma2(bc(col.y), dp[1], nk[1])),
                                 fma2(bc(col.z), dp[2], nk[2]));
                  if (DS) {
                    const float z = q1.z;
                    sdot = add2(sdot, add2(add2(fma2(bc(z), dp[DS ? 3 : 0], nk[DS ? 3 : 0]), c4_2),
                                           fma2(bc(z * z), dp[DS ? 5 : 0], nk[DS ? 5 : 0])));
                  }
                  const f2 Rs2 = pk(Rsa, Rsb);
                  dLda2 = fma2(T2, sdot, neg2(mul2(rc2, Rs2)));
                  upk(fma2(w2, sdot, Rs2), Rsa, Rsb);
                }
                // the lane's partials (summed over its two pixels) go straight into the warp's transpose tile: column =
                // lane, row = value (the previous group's row sums were read before the __syncwarp below)
                const uint32_t tr_row = tr_st + (uint32_t)(gsl * VALS) * (TR_STRIDE * 4);
                auto put = [&](int i, f2 x) {
                  SPLAT_CHECK(gsl * VALS + i < TR_ROWS && tr_row + (uint32_t)i * (TR_STRIDE * 4) + 4 <= tr_base + TR_WARP_BYTES);
                  sts32(tr_row + (uint32_t)i * (TR_STRIDE * 4), hsum(x));
                };
                if (MODE == 3) {
                  const f2 u2 = mul2(Gv2, dLda2), udx = mul2(u2, bc(dx)), udy = mul2(u2, dy2);
                  put(0, udx); put(1, udy);
                  put(2, mul2(udx, bc(dx))); put(3, mul2(udx, dy2)); put(4, mul2(udy, dy2));
                  put(5, mul2(w2, dp[DS ? 3 : 0])); put(6, mul2(w2, dp[DS ? 5 : 0]));
                } else if (MODE == 0 || MODE == 2) {
                  const f2 u2 = mul2(Gv2, dLda2), udx = mul2(u2, bc(dx)), udy = mul2(u2, dy2);
                  put(0, mul2(w2, dp[0])); put(1, mul2(w2, dp[1])); put(2, mul2(w2, dp[2]));
                  put(3, udx); put(4, udy);
                  put(5, mul2(udx, bc(dx))); put(6, mul2(udx, dy2)); put(7, mul2(udy, dy2));
                  put(8, u2);
                  if (MODE == 2) { put(9, mul2(w2, dp[DS ? 3 : 0])); put(10, mul2(w2, dp[DS ? 5 : 0])); }
                } else {
                  const f2 c0 = mul2(w2, dp[0]), c1 = mul2(w2, dp[1]), c2 = mul2(w2, dp[2]);
                  const f2 dL_dG = mul2(bc(q1.y), dLda2);
                  const f2 gdx = mul2(Gv2, bc(dx)), gdy = mul2(Gv2, dy2);
                  // -(gdx*A + gdy*B) * dL_dG * 0.5W ;  -(gdy*C + gdx*B) * dL_dG * 0.5H
                  const f2 mx = mul2(mul2(dL_dG, neg2(fma2(gdx, bc(q0.z), mul2(gdy, bc(q0.w))))), bc(ddelx_dx));
                  const f2 my = mul2(mul2(dL_dG, neg2(fma2(gdy, bc(q1.x), mul2(gdx, bc(q0.w))))), bc(ddely_dy));
                  // conic moments in the splat's own frame (slope col.w, preprocess.cu frame_slope)
                  const f2 ex = fma2(bc(col.w), dy2, bc(dx)), ey = fma2(bc(-col.w), bc(dx), dy2);
                  const f2 hq = mul2(mul2(bc(-0.5f), Gv2), dL_dG), hx = mul2(hq, ex);
                  const f2 qx = mul2(hx, ex), qy = mul2(hx, ey), qw = mul2(mul2(hq, ey), ey);
                  const f2 op = mul2(Gv2, dLda2);
                  put(0, mul2(c0, c0)); put(1, mul2(c1, c1)); put(2, mul2(c2, c2)); put(3, mul2(op, op));
                  const f2 u5[5] = {qx, qy, qw, mx, my};
                  int k = 4;
#pragma unroll
                  for (int p = 0; p < 5; ++p)
#pragma unroll
                    for (int q = p; q < 5; ++q) put(k++, mul2(u5[p], u5[q]));
                }
              }
            }
          }
          if (!any) continue;
          // Gaussians of the group that no pixel of this warp blended are skipped on both sides (warp-uniform).
          __syncwarp();
          uint32_t id = ids[0];
          bool mine = anyg[0];
#pragma unroll
          for (int q = 1; q < GROUP; ++q) if (my_g == q) { id = ids[q]; mine = anyg[q]; }
          mine = mine && lane < NV;
          float tot = 0.f;
          if (mine) tot = row_sum32(tr_ld);
          __syncwarp();
          SPLAT_CHECK(!mine || (id < (uint32_t)a.P && slot < STRIDE));
          if (mine && tot != 0.f) red_add_f32(a.acc + (size_t)id * STRIDE + slot, tot);
        }
      }
    }
    __syncwarp();
    if (GATHER) {
      const bool refill = __shfl_sync(0xffffffffu, (int)(lane == 0 && ring_release(&s_rel[s], CTA_WARPS) && c + RING_NS < nch), 0) != 0;
      if (refill) gather(c + RING_NS);
    } else if (lane == 0 && ring_release(&s_rel[s], CTA_WARPS) && c + RING_NS < nch) {
      fill(c + RING_NS);
    }
  }
  pdl_trigger();
}

// cudaFuncSetAttribute is per DEVICE: remember which devices have had the dynamic shared memory limit raised.
static int raise_smem_limits() {
  static bool done[64] = {};
  int dev = 0;
  SPLAT_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || done[dev]) return SPLAT_OK;
#define BWD_ATTR(C, M)                                                                                                      \
  SPLAT_CUDA(cudaFuncSetAttribute(render_bwd_kernel<C, M, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));  \
  SPLAT_CUDA(cudaFuncSetAttribute(render_bwd_kernel<C, M, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM))
  BWD_ATTR(true, 0); BWD_ATTR(false, 0); BWD_ATTR(true, 1); BWD_ATTR(false, 1); BWD_ATTR(true, 2); BWD_ATTR(true, 3);
#undef BWD_ATTR
  done[dev] = true;
  return SPLAT_OK;
}

int launch_render_bwd(const RenderBwdArgs& a, int T, int mode, cudaStream_t st) {
  ProfScope prof_(K_RENDER_BWD, st);
  if (T <= 0) return SPLAT_OK;
  if (int rc = raise_smem_limits()) return rc;
  const bool g = g_gather != 0 && a.grec && a.plist;
#define BWD_LAUNCH(C, M)                                                                       \
  do {                                                                                         \
    if (g) SPLAT_CUDA(launch_k(render_bwd_kernel<C, M, true>, dim3(T), dim3(RING_THREADS), (size_t)BWD_SMEM, st, a));    \
    else SPLAT_CUDA(launch_k(render_bwd_kernel<C, M, false>, dim3(T), dim3(RING_THREADS), (size_t)BWD_SMEM, st, a));  \
  } while (0)
  if (mode == 3) BWD_LAUNCH(true, 3);
  else if (mode == 2) BWD_LAUNCH(true, 2);
  else if (mode == 0) { if (g_disable_cull) BWD_LAUNCH(false, 0); else BWD_LAUNCH(true, 0); }
  else { if (g_disable_cull) BWD_LAUNCH(false, 1); else BWD_LAUNCH(true, 1); }
#undef BWD_LAUNCH
  SPLAT_LAUNCH_CHECK("render_bwd_kernel");
  return SPLAT_OK;
}

}  // namespace splat
