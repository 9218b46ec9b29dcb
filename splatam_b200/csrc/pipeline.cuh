// pipeline.cuh -- mbarrier + TMA (cp.async.bulk, 1-D) primitives for the blend kernels' instance
// ring.  sm_100a inline PTX; a CTA launched without a cluster is its own 1-CTA cluster, so the
// shared::cluster destination form addresses this CTA's shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace splat {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// (A suspend-time hint on the probe -- SASS `NANOSLEEP.SYNCS` instead of re-issuing TRYWAIT every ~0.2 us, which is 9 %
// of the forward kernel's executed instructions -- measured no change at C3, 20 us and 1 ms hints: the re-probes only
// take issue slots nobody else wanted.)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// Stage release (one lane per warp): returns true for the LAST of the CTA's `nwarps` releases of this
// stage, which then owns the refill.  The fences order every warp's shared-memory reads of the
// stage before the refill's TMA writes.
__device__ __forceinline__ bool ring_release(int* cnt, int nwarps) {
  __threadfence_block();
  const bool last = atomicAdd(cnt, 1) == nwarps - 1;
  if (last) {
    *(volatile int*)cnt = 0;
    __threadfence_block();
  }
  return last;
}

// Polling wait with nanosleep back-off (kept for single-lane producers).
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
  uint32_t ns = 64;
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(ns);
    if (ns < 1024) ns <<= 1;
  }
}

// 1-D bulk copy global -> shared through the TMA unit; completes `bytes` on `bar`.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 16-byte asynchronous copy global -> shared (LDGSTS, L2 only) and the mbarrier arrival that fires once all of the
// issuing thread's earlier cp.async have landed (.noinc: the barrier's expected count must include this arrival).
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// 1/x for x in [0.01, 1]: MUFU.RCP seed + one Newton step (error < 1 ulp, no slow-path branch).
__device__ __forceinline__ float rcp_newton(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  const float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}
__device__ __forceinline__ void red_add_f32(float* addr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}

}  // namespace splat
