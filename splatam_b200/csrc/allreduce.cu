// allreduce.cu -- the mapping step's ONE exchange: sum of the flat gradient bucket over the ranks of a
// single NVSwitch node, written against peer memory (no NCCL on this path).
//
// The reference has no multi-GPU code; this is the exchange step of SURVEY.md 8(e).  Every rank's
// bucket lives in symmetric memory (same size on every rank, every rank holds the peers' device
// pointers).  Two-shot all-reduce in stream order:
//   1. peer_barrier  -- every rank's gradient kernels are complete and visible system-wide;
//   2. peer_reduce   -- rank r owns slice r of the bucket: it reads that slice from every rank over
//                       NVLink (128-bit volatile loads, fixed rank order so all replicas get bit-identical sums),
//                       adds, and stores the sum into every rank's bucket.  Per GPU and direction the
//                       links carry (N-1)/N of the bucket, loads and stores in opposite directions;
//   3. peer_barrier  -- every rank's stores have landed before anyone reads its bucket.
// Barriers are epoch counters in a symmetric flag pad: rank r writes `epoch` into word [slot][r] of
// every peer's pad (st.release.sys) and spins on its own pad until all words reach `epoch`
// (ld.acquire.sys).  A spin that outlives ~4 s traps instead of hanging the GPU.
#include <algorithm>
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

__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <int RED_UNROLL, bool CG>
__global__ void __launch_bounds__(512) peer_reduce_kernel(const PeerArgs a) {
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
}

// NVSwitch in-switch reduction (NVLS): the bucket is also mapped at a MULTICAST address.  A
// multimem.ld_reduce on it returns the sum over all ranks' copies, added inside the switch; a
// multimem.st writes every rank's copy.  Rank r does that for slice r.  Per GPU and direction the links
// carry the bucket once (N-1)/N in + 1/N out and vice versa) instead of 2(N-1)/N of it.  The switch's
// summation order is its own, but every replica receives the SAME sums (one owner per slice).
__global__ void __launch_bounds__(512) multimem_reduce_kernel(float4* mc, int rank, int world, size_t n4) {
  const size_t per = (n4 + world - 1) / world;
  const size_t lo = (size_t)rank * per, hi = min(n4, lo + per);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t base = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < hi; base += stride * 4) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t i = base + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(mc + i) : "memory");
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t i = base + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     ::"l"(mc + i), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
    }
  }
}

}  // namespace splat

using namespace splat;

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
  peer_barrier_kernel<<<1, 32, 0, st>>>(a, 0);
  SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  const size_t per = (a.n4 + world - 1) / world;
  if (multicast_buffer) {
    if (multicast_buffer & 15) return fail(SPLAT_ERR_BAD_ARG, "misaligned multicast address");
    // (A hybrid split -- part of each slice through the switch, the rest through peer loads/stores in the same
    // launch -- was swept at N=8: 0.184 ms at 90/10 rising to 0.234 ms at 35/65 vs 0.178 ms all in-switch: both
    // halves contend for the same links.)
    int mblocks = (int)std::min<size_t>((per + 512 * 4 - 1) / (512 * 4), (size_t)148 * 2);
    if (mblocks < 1) mblocks = 1;
    multimem_reduce_kernel<<<mblocks, 512, 0, st>>>(reinterpret_cast<float4*>(multicast_buffer), rank, world, a.n4);
    SPLAT_LAUNCH_CHECK("multimem_reduce_kernel");
    peer_barrier_kernel<<<1, 32, 0, st>>>(a, 1);
    SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
    return SPLAT_OK;
  }
  // Link-bound: at N=2 the 56 MB bucket moves 56 MB per direction per GPU in ~80 us (~700 GB/s, NVLink 5's
  // practical rate); .cg vs volatile loads, unroll 8 and 1 vs 2 CTAs per SM all measure the same +-2%.
  int blocks = (int)std::min<size_t>((per + 512 * 4 - 1) / (512 * 4), (size_t)148 * 2);
  if (blocks < 1) blocks = 1;
  peer_reduce_kernel<4, false><<<blocks, 512, 0, st>>>(a);
  SPLAT_LAUNCH_CHECK("peer_reduce_kernel");
  peer_barrier_kernel<<<1, 32, 0, st>>>(a, 1);
  SPLAT_LAUNCH_CHECK("peer_barrier_kernel");
  return SPLAT_OK;
}
