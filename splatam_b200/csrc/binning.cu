// binning.cu -- tile binning without a device-wide sort.
//
// The reference bins by (1) prefix-summing tiles_touched over all P Gaussians, (2) emitting one
// 64-bit key (tile<<32 | depth bits) per instance, (3) a device-wide stable LSD radix sort over
// 32+log2(T) bits (6 passes at Replica size) and (4) a boundary-detect pass for the tile ranges
// (HX/cuda_rasterizer/rasterizer_impl.cu:70-138,277-319).  The sorted order that pipeline
// defines is: by tile, then by depth bits, ties by ascending Gaussian index (emission order +
// stable sort).  That order is a TOTAL order, so any algorithm that produces it is bit-identical.
//
// B200 design: counts per tile are already known after preprocess (RED.ADD in L2), so
//   tile_scan     one CTA: exclusive scan over T tile counts -> ranges (bit-exact, (0,0) for
//                 empty tiles as identifyTileRanges leaves them), cursors, R.
//   scatter       one thread per Gaussian: atomic cursor per tile, writes (depth<<32 | id) into
//                 the tile's bucket (bucket order is arbitrary, fixed up by the sort).
//   tile_sort     one CTA per tile: LSD radix sort of the bucket on the depth word inside shared
//                 memory (only the 8-bit digits that actually vary inside the bucket, typically
//                 3 of 4), then an exact tie fix-up by Gaussian id; writes point_list.
// HBM traffic: 8 B/instance write + 8 B read + 4 B write, instead of ~150 B/instance for the
// 6-pass device-wide sort.  Buckets that do not fit in shared memory take the same code path on
// global ping-pong buffers.
#include "kernels.cuh"
#include <cooperative_groups.h>

namespace splat {

// ------------------------------------------------------------------------------------------ scan
// One thread-block CLUSTER of 8 CTAs x 512 threads.  A single CTA was bound by one SM's memory pipeline:
// the counters and cursors sit one per 256-byte block, so T=3225 tiles are ~13k separate sector requests
// (count load, cursor store, range store, order store) at ~1.5 cycles each = 12.4 us (clock64 timeline,
// scratch/scan_probe.cu).  Eight SMs each take a contiguous slice of T/8 tiles; the slice totals, the
// maximum and the per-class histograms are exchanged through distributed shared memory.
#define SCAN_CTAS 8
#define SCAN_THREADS 512
#define SCAN_KEEP 2048    // counts of a slice kept in shared memory (slices longer than this re-read global memory)
#ifdef SPLAT_SCAN_PROBE   // scratch/scan_probe.cu: phase timeline (thread 0 of cluster rank 0)
__device__ long long g_scan_stamps[16];
#define SCAN_STAMP(i) do { if (threadIdx.x == 0 && blockIdx.x == 0) g_scan_stamps[i] = clock64(); } while (0)
#else
#define SCAN_STAMP(i) do {} while (0)
#endif
__global__ void __cluster_dims__(SCAN_CTAS, 1, 1) __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(int T, const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_cursor,
                 uint2* __restrict__ ranges, ImageHeader* hdr, uint32_t* __restrict__ num_rendered_out,
                 uint32_t sort_cap, uint32_t* __restrict__ tile_order) {
  pdl_wait();
  pdl_trigger();
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  constexpr int NW = SCAN_THREADS / 32;
  __shared__ uint32_t warp_tot[NW], warp_max[NW], warp_big[NW];
  __shared__ uint32_t s_tot[4];           // this slice: sum, max, #big   (read by the other CTAs)
  __shared__ uint32_t s_glob[4];          // base of this slice, R, global max, global #big
  __shared__ uint32_t s_hist[64];         // this slice's length-class histogram (read by the other CTAs)
  __shared__ uint32_t s_off[64];          // placement cursors of this slice
  __shared__ uint32_t s_cnt[SCAN_KEEP];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int rank = (int)cluster.block_rank();
  const int per = (T + SCAN_CTAS - 1) / SCAN_CTAS;
  const int lo = min(T, rank * per), n = min(T, lo + per) - lo;
  SCAN_STAMP(0);
  if (tid < 64) s_hist[tid] = 0;
  // ---- 1a: slice totals
  uint32_t sum = 0, lmax = 0, lbig = 0;
  for (int i = tid; i < n; i += SCAN_THREADS) {
    const uint32_t c = tile_count[(size_t)(lo + i) * CSTRIDE];
    if (i < SCAN_KEEP) s_cnt[i] = c;
    sum += c;
    lmax = max(lmax, c);
    lbig += c > sort_cap;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
    lbig += __shfl_xor_sync(0xffffffffu, lbig, o);
  }
  if (lane == 0) { warp_tot[wid] = sum; warp_max[wid] = lmax; warp_big[wid] = lbig; }
  __syncthreads();
  if (wid == 0) {
    uint32_t a = lane < NW ? warp_tot[lane] : 0u, m = lane < NW ? warp_max[lane] : 0u, g = lane < NW ? warp_big[lane] : 0u;
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
      g += __shfl_xor_sync(0xffffffffu, g, o);
    }
    if (lane == 0) { s_tot[0] = a; s_tot[1] = m; s_tot[2] = g; }
  }
  SCAN_STAMP(1);
  cluster.sync();
  if (wid == 0) {  // lanes 0..7 fetch the eight slice totals
    uint32_t a = 0, m = 0, g = 0;
    if (lane < SCAN_CTAS) {
      const uint32_t* r = cluster.map_shared_rank(s_tot, lane);
      a = r[0]; m = r[1]; g = r[2];
    }
    uint32_t before = lane < rank ? a : 0u;
#pragma unroll
    for (int o = 4; o; o >>= 1) {
      before += __shfl_xor_sync(0xffffffffu, before, o);
      a += __shfl_xor_sync(0xffffffffu, a, o);
      m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
      g += __shfl_xor_sync(0xffffffffu, g, o);
    }
    if (lane == 0) { s_glob[0] = before; s_glob[1] = a; s_glob[2] = m; s_glob[3] = g; carry_s = before; }
  }
  __syncthreads();
  SCAN_STAMP(2);
  if (rank == 0 && tid == 0) {
    hdr->num_rendered = s_glob[1];
    hdr->max_tile_count = s_glob[2];
    hdr->big_tile_count = s_glob[3];
    *num_rendered_out = s_glob[1];   // may be mapped pinned host memory: the host reads it after the stream event
  }
  // ---- 1b: exclusive scan of the slice, starting at its base; cursors, ranges and the class histogram
  const float cls_scale = 64.f / ((float)s_glob[2] + 1.f);   // length class = 63 - min(63, floor(c * 64 / (max + 1)))
  for (int base = 0; base < n; base += SCAN_THREADS) {
    const int i = base + tid;
    const uint32_t c = i < n ? (i < SCAN_KEEP ? s_cnt[i] : tile_count[(size_t)(lo + i) * CSTRIDE]) : 0u;
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += nb;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    if (wid == 0) {
      uint32_t w = lane < NW ? warp_tot[lane] : 0u;
#pragma unroll
      for (int o = 1; o < NW; o <<= 1) {
        const uint32_t nb = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += nb;
      }
      if (lane < NW) warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t incl = carry_s + v + (wid ? warp_tot[wid - 1] : 0u);
    if (i < n) {
      const uint32_t start = incl - c;
      tile_cursor[(size_t)(lo + i) * CSTRIDE] = start;
      ranges[lo + i] = c ? make_uint2(start, incl) : make_uint2(0u, 0u);
      atomicAdd(&s_hist[63u - min(63u, (uint32_t)((float)c * cls_scale))], 1u);
    }
    __syncthreads();
    if (tid == SCAN_THREADS - 1) carry_s = incl;
    __syncthreads();
  }
  SCAN_STAMP(3);
  // ---- 2: launch order for the blend kernels: longest lists first (counting sort into 64 length classes), so
  // the last CTAs of the grid are the cheap ones and the SMs drain together instead of waiting for a heavy tile
  // that happened to sit at the end of the row-major order.  Results do not depend on it.
  cluster.sync();   // all eight slice histograms complete
  if (tid < 64) {
    uint32_t tot = 0, before = 0;
#pragma unroll
    for (int r = 0; r < SCAN_CTAS; ++r) {
      const uint32_t h = cluster.map_shared_rank(s_hist, r)[tid];
      tot += h;
      before += r < rank ? h : 0u;
    }
    // exclusive scan of the 64 class totals over two warps
    uint32_t v = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += nb;
    }
    if (tid == 31) warp_max[0] = v;       // total of classes 0..31 (warp_max is free by now)
    s_off[tid] = v - tot + before;
  }
  __syncthreads();
  if (tid >= 32 && tid < 64) s_off[tid] += warp_max[0];
  __syncthreads();
  SCAN_STAMP(4);
  for (int i = tid; i < n; i += SCAN_THREADS) {
    const uint32_t c = i < SCAN_KEEP ? s_cnt[i] : tile_count[(size_t)(lo + i) * CSTRIDE];
    tile_order[atomicAdd(&s_off[63u - min(63u, (uint32_t)((float)c * cls_scale))], 1u)] = (uint32_t)(lo + i);
  }
  SCAN_STAMP(5);
  cluster.sync();   // no CTA may exit while another still reads its shared memory
  SCAN_STAMP(6);
}

// --------------------------------------------------------------------------------------- scatter
#ifndef SCATTER_ILP
#define SCATTER_ILP 4
#endif
__global__ void __launch_bounds__(256) scatter_kernel(int P, int gx, int gy, const int* __restrict__ radii,
                                                      const float4* __restrict__ grec,
                                                      uint32_t* __restrict__ tile_cursor,
                                                      uint64_t* __restrict__ keys,
                                                      const ImageHeader* __restrict__ hdr, uint32_t Rcap) {
  pdl_wait();
  pdl_trigger();
  if (hdr->num_rendered > Rcap) return;  // capacity exceeded: caller re-runs phase 2
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
  uint64_t kv = 0;
  bool vis = false;
  if (idx < P) {
    const int r = radii[idx];
    if (r > 0) {
      vis = true;
      const float2 p = __ldg(reinterpret_cast<const float2*>(grec + (size_t)idx * INST_F4));
      const float depth = __ldg(reinterpret_cast<const float*>(grec + (size_t)idx * INST_F4 + 1) + 2);
      tile_rect(p, r, gx, gy, x0, y0, x1, y1);
      kv = ((uint64_t)__float_as_uint(depth) << 32) | (uint32_t)idx;
    }
  }
  const int ntile = vis ? (x1 - x0) * (y1 - y0) : 0;
  const bool big = ntile > 32;
  // (Sharing one cursor atomic between the lanes of a warp that sit on the same tile -- match_any +
  // popc -- was measured and is SLOWER here, 63 vs 50 us at C3: the match costs more than the
  // contention it removes.)
  // A slot is one round trip to the tile's cursor in L2 (~0.6 us) and the store needs its answer: the rectangle is
  // walked four tiles at a time, the four atomics in flight together (one at a time: 47 us at C3, SCATTER_ILP 1).
  if (vis && !big) {
    int x = x0, y = y0;
    for (int t0 = 0; t0 < ntile; t0 += SCATTER_ILP) {
      uint32_t pos[SCATTER_ILP];
#pragma unroll
      for (int u = 0; u < SCATTER_ILP; ++u) {
        if (t0 + u < ntile) {
          SPLAT_CHECK(y * gx + x < gx * gy);
          pos[u] = atomicAdd(tile_cursor + (size_t)(y * gx + x) * CSTRIDE, 1u);
          if (++x == x1) { x = x0; ++y; }
        }
      }
#pragma unroll
      for (int u = 0; u < SCATTER_ILP; ++u) {
        if (t0 + u < ntile) {
          SPLAT_CHECK(pos[u] < Rcap);
          keys[pos[u]] = kv;
        }
      }
    }
  }
  unsigned bigmask = __ballot_sync(0xffffffffu, big);
  while (bigmask) {
    const int src = __ffs(bigmask) - 1;
    bigmask &= bigmask - 1;
    const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
    const int bx1 = __shfl_sync(0xffffffffu, x1, src), by1 = __shfl_sync(0xffffffffu, y1, src);
    const uint64_t bkv = __shfl_sync(0xffffffffu, kv, src);
    const int w = bx1 - bx0, n = w * (by1 - by0);
    for (int t0 = lane; t0 < n; t0 += 32 * SCATTER_ILP) {
      uint32_t pos[SCATTER_ILP];
#pragma unroll
      for (int u = 0; u < SCATTER_ILP; ++u) {
        const int t = t0 + 32 * u;
        if (t < n) pos[u] = atomicAdd(tile_cursor + (size_t)((by0 + t / w) * gx + bx0 + t % w) * CSTRIDE, 1u);
      }
#pragma unroll
      for (int u = 0; u < SCATTER_ILP; ++u) {
        if (t0 + 32 * u < n) {
          SPLAT_CHECK(pos[u] < Rcap);
          keys[pos[u]] = bkv;
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------- tile sort
// One CTA per tile.  The bucket's keys ((depth bits << 32) | id) are sorted by a ONE-pass bucket sort:
// depth bits are mapped monotonically onto SORT_NB bins spanning [min, max] of the bucket (a tile's
// depths occupy a narrow range, so bins hold ~1 key on average), a shared-memory histogram + scan
// gives every bin its slice, keys are dropped into their slice in arbitrary order, and each key then
// ranks itself inside its (tiny) bin by full 64-bit comparison.  (depth bits, id) is a total order,
// so the result equals the reference's stable radix sort bit for bit; 7 CTA barriers instead of the
// ~18 of a 3-pass LSD radix sort.  Lists longer than the shared-memory capacity take the same code
// path on global ping-pong buffers.
#define SORT_THREADS 256
#define SORT_NB 2048
// Bins above HEAVY_BIN keys get their own second-level bucket pass.  The pass costs ~10 CTA barriers (~1 us), pairwise
// ranking of an m-key bin costs m compares per key: break-even is a few hundred keys.  With the threshold at 64 a C5
// map whose depths cluster (many bins of 100-200 keys) spent 81 us in this kernel instead of 34 us (r02 A/B).
#define HEAVY_BIN 384
#define MAX_HEAVY 32     // heavy bins handled per tile (beyond that the rest fall back to pairwise ranking)
#define SUB_NB 512       // sub-bins of one heavy bin (2 KB of shared memory: the CTA count per SM stays at 5)

struct BinMap {
  uint32_t dmin, scale;
  bool direct;
  __device__ __forceinline__ uint32_t operator()(uint64_t key) const {
    const uint32_t x = (uint32_t)(key >> 32) - dmin;
    return direct ? x : __umulhi(x, scale);
  }
};

__global__ void __launch_bounds__(SORT_THREADS) tile_sort_kernel(const uint2* __restrict__ ranges,
                                                                 uint64_t* __restrict__ keys,
                                                                 uint64_t* __restrict__ keys_alt,
                                                                 uint32_t* __restrict__ point_list,
                                                                 const ImageHeader* __restrict__ hdr,
                                                                 uint32_t Rcap, int smem_cap) {
  pdl_wait();
  pdl_trigger();
  if (hdr->num_rendered > Rcap) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t bins[SORT_NB];
  __shared__ uint32_t bins2[SUB_NB];                       // sub-bins of one heavy bin
  __shared__ uint32_t red_min[8], red_max[8], wsum[8];
  __shared__ uint64_t red_kmin[8], red_kmax[8];
  __shared__ uint32_t heavy_lo[MAX_HEAVY], heavy_hi[MAX_HEAVY];
  __shared__ int n_heavy;
  const uint2 rg = ranges[blockIdx.x];
  const int n = (int)(rg.y - rg.x);
  if (n <= 0) return;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint64_t* gsrc = keys + rg.x;
  uint32_t* out = point_list + rg.x;
  if (n == 1) {
    if (tid == 0) out[0] = (uint32_t)gsrc[0];
    return;
  }
  const bool in_smem = n <= smem_cap;
  uint64_t* A = in_smem ? reinterpret_cast<uint64_t*>(smem_raw) : gsrc;
  uint64_t* B = in_smem ? A + smem_cap : keys_alt + rg.x;
  // load + depth range of the bucket
  uint32_t dmin = 0xffffffffu, dmax = 0u;
  for (int i = tid; i < n; i += SORT_THREADS) {
    const uint64_t k = gsrc[i];
    if (in_smem) A[i] = k;
    const uint32_t d = (uint32_t)(k >> 32);
    dmin = min(dmin, d); dmax = max(dmax, d);
  }
  for (int i = tid; i < SORT_NB; i += SORT_THREADS) bins[i] = 0;
  dmin = __reduce_min_sync(0xffffffffu, dmin);
  dmax = __reduce_max_sync(0xffffffffu, dmax);
  if (lane == 0) { red_min[wid] = dmin; red_max[wid] = dmax; }
  __syncthreads();
#pragma unroll
  for (int w = 0; w < 8; ++w) { dmin = min(dmin, red_min[w]); dmax = max(dmax, red_max[w]); }
  BinMap bin;
  const uint32_t range = dmax - dmin;
  bin.dmin = dmin;
  bin.direct = range < SORT_NB;
  bin.scale = bin.direct ? 0u : (uint32_t)(((uint64_t)SORT_NB << 32) / ((uint64_t)range + 1ull));
  // histogram
  for (int i = tid; i < n; i += SORT_THREADS) { SPLAT_CHECK(bin(A[i]) < SORT_NB); atomicAdd(&bins[bin(A[i])], 1u); }
  __syncthreads();
  // exclusive scan over SORT_NB bins: thread t owns bins [8t, 8t+8)
  {
    uint32_t c[8], tot = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) { c[q] = bins[tid * 8 + q]; tot += c[q]; }
    uint32_t v = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t nn = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += nn;
    }
    if (lane == 31) wsum[wid] = v;
    __syncthreads();
    uint32_t base = v - tot;
    for (int w = 0; w < wid; ++w) base += wsum[w];
#pragma unroll
    for (int q = 0; q < 8; ++q) { bins[tid * 8 + q] = base; base += c[q]; }
  }
  __syncthreads();
  // placement: bins[] are cursors (start -> end)
  for (int i = tid; i < n; i += SORT_THREADS) {
    const uint64_t k = A[i];
    const uint32_t dstp = atomicAdd(&bins[bin(k)], 1u);
    SPLAT_CHECK(dstp < (uint32_t)n);
    B[dstp] = k;
  }
  __syncthreads();
  // Heavy bins.  A tile whose depths are bit-equal or clustered -- the frame-0 map back-projected from quantised
  // sensor depth and seen fronto-parallel (TUM depth is 16-bit / 5000: hundreds of Gaussians per tile at ONE view.z), or
  // one far outlier stretching [dmin, dmax] -- puts hundreds to thousands of keys in one bin, and ranking them by
  // pairwise comparison is quadratic (r02: 87 us instead of 17 us for a 640x480 wall; a 20k-key bin would take
  // milliseconds).  Bins above HEAVY_BIN keys therefore get a second bucket pass of their own, by the whole CTA, on the
  // FULL 64-bit key (depth bits, id) -- the id is unique within a tile, so the second level has no ties: keys spread over
  // SUB_NB sub-bins between the bin's min and max key, and the final pairwise ranking runs inside sub-bins of ~1 key.
  // bins[b] now holds the END of bin b.
  {
    if (tid == 0) n_heavy = 0;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int b = tid * 8 + q;
      const uint32_t lo = b ? bins[b - 1] : 0u, hi = bins[b];
      if (hi - lo > HEAVY_BIN) {
        const int slot = atomicAdd(&n_heavy, 1);
        if (slot < MAX_HEAVY) { heavy_lo[slot] = lo; heavy_hi[slot] = hi; }
      }
    }
    __syncthreads();
    const int nh = min(n_heavy, MAX_HEAVY);   // (more than n / HEAVY_BIN heavy bins cannot exist; MAX_HEAVY covers the shared-memory path)
    for (int h = 0; h < nh; ++h) {
      const int lo = (int)heavy_lo[h], hi = (int)heavy_hi[h], m = hi - lo;
      // range of the full keys
      uint64_t kmin = ~0ull, kmax = 0ull;
      for (int i = lo + tid; i < hi; i += SORT_THREADS) { const uint64_t k = B[i]; kmin = min(kmin, k); kmax = max(kmax, k); }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
      }
      if (lane == 0) { red_kmin[wid] = kmin; red_kmax[wid] = kmax; }
      for (int i = tid; i < SUB_NB; i += SORT_THREADS) bins2[i] = 0;
      __syncthreads();
#pragma unroll
      for (int w = 0; w < 8; ++w) { kmin = min(kmin, red_kmin[w]); kmax = max(kmax, red_kmax[w]); }
      // monotone map of (key - kmin) onto [0, SUB_NB): both reduced to 21 bits so the product fits 32
      const uint64_t range64 = kmax - kmin;
      const int sh = max(0, 43 - __clzll((long long)(range64 | 1ull)));   // (range >> sh) < 2^21
      const uint32_t rdiv = (uint32_t)(range64 >> sh) + 1u;
      auto sub = [&](uint64_t k) { return (uint32_t)(((uint32_t)((k - kmin) >> sh) * (uint32_t)SUB_NB) / rdiv); };
      for (int i = lo + tid; i < hi; i += SORT_THREADS) { SPLAT_CHECK(sub(B[i]) < SUB_NB); atomicAdd(&bins2[sub(B[i])], 1u); }
      __syncthreads();
      {   // exclusive scan of bins2 (thread t owns sub-bins [2t, 2t+2)), offset by lo
        constexpr int PER = SUB_NB / SORT_THREADS;
        uint32_t c[PER], tot = 0;
#pragma unroll
        for (int q = 0; q < PER; ++q) { c[q] = bins2[tid * PER + q]; tot += c[q]; }
        uint32_t v = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t nn = __shfl_up_sync(0xffffffffu, v, o);
          if (lane >= o) v += nn;
        }
        if (lane == 31) wsum[wid] = v;
        __syncthreads();
        uint32_t base = (uint32_t)lo + v - tot;
        for (int w = 0; w < wid; ++w) base += wsum[w];
#pragma unroll
        for (int q = 0; q < PER; ++q) { bins2[tid * PER + q] = base; base += c[q]; }
      }
      __syncthreads();
      // placement into A (free since the first-level placement), then ranking inside the sub-bins
      for (int i = lo + tid; i < hi; i += SORT_THREADS) {
        const uint64_t k = B[i];
        const uint32_t dstp = atomicAdd(&bins2[sub(k)], 1u);
        SPLAT_CHECK(dstp >= (uint32_t)lo && dstp < (uint32_t)hi);
        A[dstp] = k;
      }
      __syncthreads();
      for (int i = lo + tid; i < hi; i += SORT_THREADS) {
        const uint64_t k = A[i];
        const uint32_t b2 = sub(k);
        const int l2 = b2 ? (int)bins2[b2 - 1] : lo, h2 = (int)bins2[b2];
        int rank = 0;
        for (int q = l2; q < h2; ++q) rank += A[q] < k;
        SPLAT_CHECK(l2 >= lo && h2 <= hi && l2 + rank < hi);
        out[l2 + rank] = (uint32_t)k;
      }
      __syncthreads();
      (void)m;
    }
  }
  // in-bin ranking by the full key for the light bins
  for (int p = tid; p < n; p += SORT_THREADS) {
    const uint64_t k = B[p];
    const uint32_t b = bin(k);
    const int lo = b ? (int)bins[b - 1] : 0, hi = (int)bins[b];
    if (hi - lo > HEAVY_BIN && n_heavy <= MAX_HEAVY) continue;   // already written by its own pass
    int rank = 0;
    for (int q = lo; q < hi; ++q) rank += B[q] < k;
    SPLAT_CHECK(lo >= 0 && hi <= n && lo + rank < hi);
    out[lo + rank] = (uint32_t)k;
  }
}

// Sorted instance records: one thread per 16-byte quad (three per list position) copies its third of the
// Gaussian's 48-byte record; the middle quad carries the id.  Stores are fully coalesced (thread t
// writes quad t) and each record is read by three adjacent lanes; the grec array is L2 resident.
__global__ void __launch_bounds__(256) emit_records_kernel(const uint32_t* __restrict__ point_list,
                                                           const float4* __restrict__ grec,
                                                           float4* __restrict__ inst,
                                                           const ImageHeader* __restrict__ hdr, uint32_t Rcap) {
  pdl_wait();
  pdl_trigger();
  const uint32_t R = hdr->num_rendered;
  if (R > Rcap) return;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)R * INST_F4) return;
  const uint32_t i = (uint32_t)(t / INST_F4), q = (uint32_t)(t - (size_t)i * INST_F4);
  const uint32_t id = __ldg(point_list + i);
  float4 v = __ldg(grec + (size_t)id * INST_F4 + q);
  if (q == 1) v.w = __uint_as_float(id);
  inst[t] = v;
}

int launch_tile_scan(int T, const ImageView& iv, uint32_t* num_rendered_out, uint32_t sort_cap, cudaStream_t st) {
  ProfScope prof_(K_TILE_SCAN, st);
  SPLAT_CUDA(launch_k(tile_scan_kernel, dim3(SCAN_CTAS), dim3(SCAN_THREADS), 0, st, T, iv.tile_count, iv.tile_cursor, iv.ranges, iv.hdr,
                      num_rendered_out, sort_cap, iv.tile_order));
  SPLAT_LAUNCH_CHECK("tile_scan_kernel");
  return SPLAT_OK;
}

int launch_scatter(int P, int gx, int gy, const int* radii, const GeomView& g, const ImageView& iv,
                   const BinView& b, uint32_t Rcap, cudaStream_t st) {
  ProfScope prof_(K_SCATTER, st);
  if (P <= 0) return SPLAT_OK;
  SPLAT_CUDA(launch_k(scatter_kernel, dim3((P + 255) / 256), dim3(256), 0, st, P, gx, gy, radii, g.grec, iv.tile_cursor, b.keys, iv.hdr, Rcap));
  SPLAT_LAUNCH_CHECK("scatter_kernel");
  return SPLAT_OK;
}

static const int kSortSmemCap = 2048;  // keys per shared buffer (2 x 16 KB); longer lists sort in global memory
int sort_smem_cap() { return kSortSmemCap; }

int launch_tile_sort(int T, const ImageView& iv, const GeomView& g, const BinView& b, uint32_t Rcap, cudaStream_t st) {
  {
    ProfScope prof_(K_TILE_SORT, st);
    static bool attr_set[64] = {};   // cudaFuncSetAttribute is per device
    const int smem = kSortSmemCap * 8 * 2;
    int dev = 0;
    SPLAT_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      SPLAT_CUDA(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      attr_set[dev] = true;
    }
    SPLAT_CUDA(launch_k(tile_sort_kernel, dim3(T), dim3(SORT_THREADS), (size_t)smem, st, iv.ranges, b.keys, b.keys_alt, b.point_list, iv.hdr, Rcap,
                        kSortSmemCap));
    SPLAT_LAUNCH_CHECK("tile_sort_kernel");
  }
  if (Rcap > 0 && !g_gather) {   // gather mode: the blend kernels fetch the Gaussians' own records by id
    ProfScope prof_(K_EMIT, st);
    const size_t quads = (size_t)Rcap * INST_F4;
    SPLAT_CUDA(launch_k(emit_records_kernel, dim3((unsigned)((quads + 255) / 256)), dim3(256), 0, st, b.point_list, g.grec, b.inst, iv.hdr, Rcap));
    SPLAT_LAUNCH_CHECK("emit_records_kernel");
  }
  return SPLAT_OK;
}

}  // namespace splat
