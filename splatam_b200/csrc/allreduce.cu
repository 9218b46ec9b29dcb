// allreduce.cu -- the mapping step's ONE exchange: sum of the flat gradient bucket over the ranks of a
// single NVSwitch node, written against peer memory (no NCCL on this path).
//
// The reference has no multi-GPU code; this is the exchange step of SURVEY.md 8(e).  Every rank's
// bucket lives in symmetric memory (same size on every rank, every rank holds the peers' device
// pointers).  Two-shot all-reduce in stream order, ONE kernel per rank:
//   1. arrival barrier -- every rank's gradient kernels are complete and visible system-wide;
//   2. reduce          -- rank r owns slice r of the bucket: it reads that slice from every rank over
//                         NVLink (128-bit volatile loads, fixed rank order so all replicas get bit-identical sums),
//                         adds, and stores the sum into every rank's bucket.  Per GPU and direction the
//                         links carry (N-1)/N of the bucket, loads and stores in opposite directions.
//                         From 4 ranks on the sum is formed inside the NVSwitch instead (multimem.ld_reduce /
//                         multimem.st on the bucket's multicast address);
//   3. completion barrier -- every rank's stores have landed before anyone reads its bucket.
// Barriers are epoch counters in a symmetric flag pad: rank r writes `epoch` into word [slot][r] of
// every peer's pad (st.release.sys) and spins on its own pad until all words reach `epoch`
// (ld.acquire.sys).  They run inside the exchange kernel (fused_begin / fused_end below); with
// SPLAT_PEER_FUSED_BARRIER=0 they are two one-warp kernels around it, as in round 1.  A spin that outlives
// SPLAT_PEER_TIMEOUT_S (default 60 s) traps instead of hanging the GPU.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "common.cuh"

namespace splat {

#define PEER_MAX 16
#define PAD_STRIDE 32   // words per barrier slot

struct PeerArgs {
  float4* buf[PEER_MAX];
  uint32_t* pad[PEER_MAX];
  int rank, world;
  uint32_t epoch;
  size_t n4;            // bucket length in float4
  long long timeout_cycles;   // a peer that has not arrived after this many SM cycles: trap (0 = wait forever)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__global__ void __launch_bounds__(32) peer_barrier_kernel(const PeerArgs a, int slot) {
  const int j = threadIdx.x;
  if (j >= a.world) return;
  __threadfence_system();
  st_release_sys(a.pad[j] + slot * PAD_STRIDE + a.rank, a.epoch);
  const uint32_t* mine = a.pad[a.rank] + slot * PAD_STRIDE + j;
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(mine) - a.epoch) < 0) {
    if (a.timeout_cycles && clock64() - t0 > a.timeout_cycles) __trap();   // a peer never arrived: fail loudly
    __nanosleep(64);
  }
}

// The same two barriers INSIDE the exchange kernel (FUSED): no separate launches, and the waits overlap the kernel's own
// ramp-up.  begin: thread j of CTA 0 posts this rank's arrival to peer j; every CTA then spins on the LOCAL pad.  end: every
// thread fences its stores, the last CTA to finish (counter in word 63 of the local pad) posts the completion flags and
// waits for the peers' -- the kernel, hence the stream, does not get past it before every replica is complete.
#define PAD_COUNTER (2 * PAD_STRIDE - 1)
__device__ __forceinline__ void spin_until(const uint32_t* flag, uint32_t epoch, long long timeout_cycles) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {   // a few threads per CTA on a local line: no back-off needed
    if (timeout_cycles && clock64() - t0 > timeout_cycles) __trap();
  }
}
__device__ __forceinline__ void fused_begin(const PeerArgs& a) {
  const int j = threadIdx.x;
  if (j < a.world) {
    if (blockIdx.x == 0) st_release_sys(a.pad[j] + a.rank, a.epoch);   // release.sys: the stream's earlier kernels are visible first
    spin_until(a.pad[a.rank] + j, a.epoch, a.timeout_cycles);
  }
  __syncthreads();
}
__device__ __forceinline__ void fused_end(const PeerArgs& a) {
  __threadfence_system();            // this thread's stores to the replicas are ordered before the flags below
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    uint32_t* cnt = a.pad[a.rank] + PAD_COUNTER;
    __threadfence();
    s_last = atomicAdd(cnt, 1u) == gridDim.x - 1;
    if (s_last) *(volatile uint32_t*)cnt = 0u;    // nobody else touches it until the next call
  }
  __syncthreads();
  if (!s_last) return;
  const int j = threadIdx.x;
  if (j < a.world) {
    st_release_sys(a.pad[j] + PAD_STRIDE + a.rank, a.epoch);
    spin_until(a.pad[a.rank] + PAD_STRIDE + j, a.epoch, a.timeout_cycles);
  }
}

__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <int RED_UNROLL, bool CG, bool FUSED>
__global__ void __launch_bounds__(512) peer_reduce_kernel(const PeerArgs a) {
  if (FUSED) fused_begin(a);
  const size_t per = (a.n4 + a.world - 1) / a.world;
  const size_t lo = (size_t)a.rank * per, hi = min(a.n4, lo + per);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t base = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < hi; base += stride * RED_UNROLL) {
    float4 s[RED_UNROLL];
#pragma unroll
    for (int u = 0; u < RED_UNROLL; ++u) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 0; j < a.world; ++j) {        // fixed order: every owner sums rank 0, 1, ... -> replicas stay identical
      float4 v[RED_UNROLL];
#pragma unroll
      for (int u = 0; u < RED_UNROLL; ++u) {
        const size_t i = base + (size_t)u * stride;
        v[u] = i < hi ? (CG ? ld_cg_f4(a.buf[j] + i) : ld_volatile_f4(a.buf[j] + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < RED_UNROLL; ++u) { s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w; }
    }
    for (int j = 0; j < a.world; ++j) {
#pragma unroll
      for (int u = 0; u < RED_UNROLL; ++u) {
        const size_t i = base + (size_t)u * stride;
        if (i < hi) a.buf[j][i] = s[u];
      }
    }
  }
  if (FUSED) fused_end(a);
}

// NVSwitch in-switch reduction (NVLS): the bucket is also mapped at a MULTICAST address.  A
// multimem.ld_reduce on it returns the sum over all ranks' copies, added inside the switch; a
// multimem.st writes every rank's copy.  Rank r does that for slice r.  Per GPU and direction the links
// carry the bucket once (N-1)/N in + 1/N out and vice versa) instead of 2(N-1)/N of it.  The switch's
// summation order is its own, but every replica receives the SAME sums (one owner per slice).
template <int U, bool FUSED>
__global__ void __launch_bounds__(512) multimem_reduce_kernel(float4* mc, const PeerArgs a) {
  if (FUSED) fused_begin(a);
  const int rank = a.rank, world = a.world;
  const size_t n4 = a.n4;
  const size_t per = (n4 + world - 1) / world;
  const size_t lo = (size_t)rank * per, hi = min(n4, lo + per);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t base = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < hi; base += stride * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(mc + i) : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     ::"l"(mc + i), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
    }
  }
  if (FUSED) fused_end(a);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused exchange + optimizer step of the keyframe-sharded mapping iteration: reduce-scatter -> Adam -> all-gather in ONE
// kernel over peer memory.  The plain path (all-reduce the gradient bucket, then run Adam on every replica) moves the
// bucket twice over the links (sum out, sum back), runs the same 280 MB Adam pass on all N GPUs and needs N full copies
// of the moments.  Here rank r
//     sums slice r of the GRADIENT bucket   (in-switch: multimem.ld_reduce; peer path: loads from every rank, rank order)
//     runs Adam on slice r only             (its 1/N shard of exp_avg / exp_avg_sq lives on this rank alone)
//     stores the updated PARAMETERS of slice r into every replica (multimem.st, or one store per peer)
// so the links carry gradient slices in and parameter slices out (the same bytes as one all-reduce), the optimizer pass
// shrinks to 1/N per GPU, and one launch + two barriers replace two launches + Adam's launch + two barriers.  Every
// replica receives identical parameters (one owner per slice), so replicas cannot drift.
#define ADAM_MAX_SEG 8
struct ReduceAdamArgs {
  PeerArgs p;                  // gradient buckets, flag pads, rank / world / epoch / n4
  float4* param[PEER_MAX];     // parameter buckets (same layout and length as the gradient bucket)
  float4* grad_mc; float4* param_mc;   // NVSwitch multicast addresses of both buckets, or NULL (peer path)
  float4* m; float4* v;        // this rank's shard of the moments: float4 index i - lo
  int nseg;
  long long seg_end[ADAM_MAX_SEG];     // element (float) index where segment k ends: per-tensor learning rates
  float lr1[ADAM_MAX_SEG];             // lr / bias_correction1 per segment
  float w1, w2, beta2, inv_b2, eps;
};

template <bool MC, bool FUSED>
__global__ void __launch_bounds__(512) reduce_adam_kernel(const ReduceAdamArgs a) {
  if (FUSED) fused_begin(a.p);
  const size_t per = (a.p.n4 + a.p.world - 1) / a.p.world;
  const size_t lo = (size_t)a.p.rank * per, hi = min(a.p.n4, lo + per);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const float4* mine = a.param[a.p.rank];
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    float4 g;
    if (MC) {
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "l"(a.grad_mc + i) : "memory");
    } else {
      g = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < a.p.world; ++j) {        // fixed rank order: the sum every replica would have formed
        const float4 t = ld_cg_f4(a.p.buf[j] + i);
        g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
      }
    }
    float4 m = a.m[i - lo], v = a.v[i - lo], p = mine[i];
    const long long e = (long long)i * 4;
    int s0 = 0;
    while (s0 < a.nseg - 1 && e >= a.seg_end[s0]) ++s0;
    float lr[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {                  // a float4 may straddle a tensor boundary when P is not a multiple of 4
      int sc = s0;
      while (sc < a.nseg - 1 && e + c >= a.seg_end[sc]) ++sc;
      lr[c] = a.lr1[sc];
    }
    adam_update(g.x, m.x, v.x, p.x, a.w1, a.w2, a.beta2, lr[0], a.inv_b2, a.eps);
    adam_update(g.y, m.y, v.y, p.y, a.w1, a.w2, a.beta2, lr[1], a.inv_b2, a.eps);
    adam_update(g.z, m.z, v.z, p.z, a.w1, a.w2, a.beta2, lr[2], a.inv_b2, a.eps);
    adam_update(g.w, m.w, v.w, p.w, a.w1, a.w2, a.beta2, lr[3], a.inv_b2, a.eps);
    a.m[i - lo] = m;
    a.v[i - lo] = v;
    if (MC) {
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                   ::"l"(a.param_mc + i), "f"(p.x), "f"(p.y), "f"(p.z), "f"(p.w) : "memory");
    } else {
      for (int j = 0; j < a.p.world; ++j) a.param[j][i] = p;
    }
  }
  if (FUSED) fused_end(a.p);
}

}  // namespace splat

using namespace splat;

// Grid size (0 = default) and loads in flight per thread of the in-switch kernels: measurement knobs.
// SPLAT_PEER_FUSED_BARRIER=0: the two barriers as separate one-warp kernels around the exchange kernel (the round-1 form).
static bool fused_barriers() {
  static const bool on = [] { const char* e = getenv("SPLAT_PEER_FUSED_BARRIER"); return !(e && e[0] == '0'); }();
  return on;
}
static int g_peer_blocks = 0, g_peer_unroll = 0;
extern "C" void splat_peer_set_tuning(int32_t blocks, int32_t unroll) {
  g_peer_blocks = blocks > 0 ? blocks : 0;
  g_peer_unroll = (unroll == 1 || unroll == 2 || unroll == 4 || unroll == 8) ? unroll : 0;
}

extern "C" int splat_peer_allreduce_sum(const uint64_t* buffers, const uint64_t* flag_pads, uint64_t multicast_buffer,
                                        int32_t rank, int32_t world, uint32_t epoch, int64_t n_floats, void* stream) {
  if (!buffers || !flag_pads) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer table");
  if (world < 1 || world > PEER_MAX || rank < 0 || rank >= world) return fail(SPLAT_ERR_BAD_ARG, "bad rank/world %d/%d", rank, world);
  if (n_floats < 0 || (n_floats & 3)) return fail(SPLAT_ERR_BAD_ARG, "bucket length must be a multiple of 4 floats");
  if (world == 1 || n_floats == 0) return SPLAT_OK;
  PeerArgs a;
  for (int j = 0; j < world; ++j) {
    if (!buffers[j] || !flag_pads[j] || (buffers[j] & 15)) return fail(SPLAT_ERR_BAD_ARG, "peer %d: NULL or misaligned buffer", j);
    a.buf[j] = reinterpret_cast<float4*>(buffers[j]);
    a.pad[j] = reinterpret_cast<uint32_t*>(flag_pads[j]);
  }
  a.rank = rank; a.world = world; a.epoch = epoch; a.n4 = (size_t)n_floats / 4;
  {  // SPLAT_PEER_TIMEOUT_S: seconds a rank waits for its peers before the barrier traps (default 60; 0 = forever, as
     // NCCL would).  A rank can legitimately stall for seconds: first-iteration allocation, a checkpoint, a debugger.
    static const double timeout_s = [] { const char* e = getenv("SPLAT_PEER_TIMEOUT_S"); return e ? atof(e) : 60.0; }();
    a.timeout_cycles = timeout_s > 0.0 ? (long long)(timeout_s * 2.0e9) : 0;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool fused = fused_barriers();
  if (!fused) {
    peer_barrier_kernel<<<1, 32, 0, st>>>(a, 0);
    SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  }
  const size_t per = (a.n4 + world - 1) / world;
  if (multicast_buffer) {
    if (multicast_buffer & 15) return fail(SPLAT_ERR_BAD_ARG, "misaligned multicast address");
    // (A hybrid split -- part of each slice through the switch, the rest through peer loads/stores in the same
    // launch -- was swept at N=8: 0.184 ms at 90/10 rising to 0.234 ms at 35/65 vs 0.178 ms all in-switch: both
    // halves contend for the same links.)
    // Few CTAs that LOOP over the slice, not one pass of many: with the whole slice in flight at once every GPU first
    // only loads (its links busy outbound with the data the switch sums) and then only stores (busy inbound with the
    // multicast), one direction idle each time.  Looping interleaves the two.  N=8 sweep (tests/peer_tune_sweep.py,
    // profiles/r02i_peer_tune_n8.log): 0.178 ms at 213 CTAs x 4 loads in flight, 0.142-0.144 ms anywhere in 32-64 CTAs x 1-2.
    const int U = g_peer_unroll > 0 ? g_peer_unroll : 2;
    int mblocks = (int)std::min<size_t>((per + 512 * U - 1) / (512 * U), (size_t)(g_peer_blocks > 0 ? g_peer_blocks : 48));
    if (mblocks < 1) mblocks = 1;
    float4* mc = reinterpret_cast<float4*>(multicast_buffer);
#define MM_LAUNCH(UU)                                                                       \
  do {                                                                                      \
    if (fused) multimem_reduce_kernel<UU, true><<<mblocks, 512, 0, st>>>(mc, a);            \
    else multimem_reduce_kernel<UU, false><<<mblocks, 512, 0, st>>>(mc, a);                 \
  } while (0)
    if (U == 1) MM_LAUNCH(1);
    else if (U == 4) MM_LAUNCH(4);
    else if (U == 8) MM_LAUNCH(8);
    else MM_LAUNCH(2);
#undef MM_LAUNCH
    SPLAT_LAUNCH_CHECK("multimem_reduce_kernel");
    if (!fused) {
      peer_barrier_kernel<<<1, 32, 0, st>>>(a, 1);
      SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
    }
    return SPLAT_OK;
  }
  // Link-bound: at N=2 the 56 MB bucket moves 56 MB per direction per GPU in ~80 us (~700 GB/s, NVLink 5's
  // practical rate); .cg vs volatile loads, unroll 8 and 1 vs 2 CTAs per SM all measure the same +-2%.
  int blocks = (int)std::min<size_t>((per + 512 * 4 - 1) / (512 * 4), (size_t)(g_peer_blocks > 0 ? g_peer_blocks : 148 * 2));
  if (blocks < 1) blocks = 1;
  if (fused) peer_reduce_kernel<4, false, true><<<blocks, 512, 0, st>>>(a);
  else peer_reduce_kernel<4, false, false><<<blocks, 512, 0, st>>>(a);
  SPLAT_LAUNCH_CHECK("peer_reduce_kernel");
  if (!fused) {
    peer_barrier_kernel<<<1, 32, 0, st>>>(a, 1);
    SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  }
  return SPLAT_OK;
}

extern "C" size_t splat_peer_adam_shard_floats(int64_t n_floats, int32_t world) {
  if (n_floats <= 0 || world <= 0) return 0;
  const size_t n4 = ((size_t)n_floats + 3) / 4;
  return ((n4 + world - 1) / world) * 4;
}

extern "C" int splat_peer_reduce_adam(const uint64_t* grad_buffers, const uint64_t* param_buffers, const uint64_t* flag_pads,
                                      uint64_t grad_multicast, uint64_t param_multicast, int32_t rank, int32_t world,
                                      uint32_t epoch, int64_t n_floats, const int64_t* segment_ends, const float* segment_lrs,
                                      int32_t n_segments, int32_t step, double beta1, double beta2, double eps,
                                      float* exp_avg_shard, float* exp_avg_sq_shard, void* stream) {
  if (!grad_buffers || !param_buffers || !flag_pads) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer table");
  if (world < 1 || world > PEER_MAX || rank < 0 || rank >= world) return fail(SPLAT_ERR_BAD_ARG, "bad rank/world %d/%d", rank, world);
  if (n_floats <= 0 || (n_floats & 3)) return fail(SPLAT_ERR_BAD_ARG, "bucket length must be a positive multiple of 4 floats");
  if (!segment_ends || !segment_lrs || n_segments < 1 || n_segments > ADAM_MAX_SEG)
    return fail(SPLAT_ERR_BAD_ARG, "need 1..%d (segment end, learning rate) pairs", ADAM_MAX_SEG);
  if (step < 1) return fail(SPLAT_ERR_BAD_ARG, "step must be >= 1");
  if (!exp_avg_shard || !exp_avg_sq_shard || (((uintptr_t)exp_avg_shard | (uintptr_t)exp_avg_sq_shard) & 15))
    return fail(SPLAT_ERR_BAD_ARG, "moment shards must be 16-byte aligned device buffers of splat_peer_adam_shard_floats() floats");
  if ((grad_multicast != 0) != (param_multicast != 0) || ((grad_multicast | param_multicast) & 15))
    return fail(SPLAT_ERR_BAD_ARG, "multicast addresses: both or neither, 16-byte aligned");
  ReduceAdamArgs a = {};
  for (int j = 0; j < world; ++j) {
    if (!grad_buffers[j] || !param_buffers[j] || !flag_pads[j] || ((grad_buffers[j] | param_buffers[j]) & 15))
      return fail(SPLAT_ERR_BAD_ARG, "peer %d: NULL or misaligned buffer", j);
    a.p.buf[j] = reinterpret_cast<float4*>(grad_buffers[j]);
    a.param[j] = reinterpret_cast<float4*>(param_buffers[j]);
    a.p.pad[j] = reinterpret_cast<uint32_t*>(flag_pads[j]);
  }
  long long prev = 0;
  for (int k = 0; k < n_segments; ++k) {
    if (segment_ends[k] < prev) return fail(SPLAT_ERR_BAD_ARG, "segment ends must be non-decreasing");
    prev = a.seg_end[k] = segment_ends[k];
  }
  a.p.rank = rank; a.p.world = world; a.p.epoch = epoch; a.p.n4 = (size_t)n_floats / 4;
  {
    static const double timeout_s = [] { const char* e = getenv("SPLAT_PEER_TIMEOUT_S"); return e ? atof(e) : 60.0; }();
    a.p.timeout_cycles = timeout_s > 0.0 ? (long long)(timeout_s * 2.0e9) : 0;
  }
  // bias corrections exactly as splat_adam_step forms them (double pow, rounded once)
  const float bc1 = (float)(1.0 - pow(beta1, (double)step));
  const float bc2s = (float)sqrt(1.0 - pow(beta2, (double)step));
  for (int k = 0; k < n_segments; ++k) a.lr1[k] = segment_lrs[k] / bc1;
  a.nseg = n_segments;
  a.w1 = (float)(1.0 - beta1); a.w2 = (float)(1.0 - beta2); a.beta2 = (float)beta2; a.inv_b2 = 1.f / bc2s; a.eps = (float)eps;
  a.grad_mc = reinterpret_cast<float4*>(grad_multicast); a.param_mc = reinterpret_cast<float4*>(param_multicast);
  a.m = reinterpret_cast<float4*>(exp_avg_shard); a.v = reinterpret_cast<float4*>(exp_avg_sq_shard);
  cudaStream_t st = (cudaStream_t)stream;
  const bool fused = fused_barriers();
  if (!fused) {
    peer_barrier_kernel<<<1, 32, 0, st>>>(a.p, 0);          // every rank's gradients are complete
    SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  }
  const size_t per = (a.p.n4 + world - 1) / world;
  // in-switch: 64 looping CTAs (0.121 ms at N=8 against 0.133 ms at 296, same sweep); peer loads / stores: 444 (N=2: 0.094 ms against 0.099 at 296)
  int blocks = (int)std::min<size_t>((per + 511) / 512, (size_t)(g_peer_blocks > 0 ? g_peer_blocks : (grad_multicast ? 64 : 148 * 3)));
  if (blocks < 1) blocks = 1;
  if (grad_multicast) {
    if (fused) reduce_adam_kernel<true, true><<<blocks, 512, 0, st>>>(a);
    else reduce_adam_kernel<true, false><<<blocks, 512, 0, st>>>(a);
  } else {
    if (fused) reduce_adam_kernel<false, true><<<blocks, 512, 0, st>>>(a);
    else reduce_adam_kernel<false, false><<<blocks, 512, 0, st>>>(a);
  }
  SPLAT_LAUNCH_CHECK("reduce_adam_kernel");
  if (!fused) {
    peer_barrier_kernel<<<1, 32, 0, st>>>(a.p, 1);          // every replica holds the updated parameters
    SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  }
  return SPLAT_OK;
}
