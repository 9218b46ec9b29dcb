// densify.cu -- per-frame map surgery: pruning and growing the Gaussian map, Adam moments included.
//
// The reference does this with torch boolean indexing / torch.cat on every tensor separately
// (utils/slam_external.py:123-165 remove_points / cat_params_to_optimizer, scripts/splatam.py:384-426
// add_new_gaussians): per tensor one nonzero() + index kernel + allocation, ~20 tensors per prune, each a host sync.
// Here: ONE flag pass, ONE scan, ONE order-preserving compaction over all row tensors, into the
// other half of a double-buffered arena (splatam_b200/densify.py) -- HBM-bound data movement, bit-exact by
// construction (rows are copied, never recomputed; the order equals tensor[keep]).
//
//   prune_flags      keep[i] = !(sigmoid(logit) < thr  ||  max_j exp(log_scale_j) > big)       (prune_gaussians :172-186)
//   flag_count       1024 flags per block -> block_count
//   block_scan       one CTA: exclusive scan of the block counts -> block_base, total (to pinned host memory)
//   compact_rows     one block per 1024 rows: ranks from the flags once, then every tensor's rows streamed (16-byte
//                    loads) and scattered to rank order
//   err_hist x3      exact lower median of depth_error (= torch.median) by an 11/11/10-bit radix select
//   spawn_mask       non-presence mask of add_new_gaussians (:390-408) + block counts
//   spawn            back-projection (get_pointcloud :73-123), initialize_new_params (:354-381), appended in
//                    raster order behind row P
#include "kernels.cuh"

namespace splat {

#define ROWS_PER_BLOCK 1024   // 256 threads x 4 flags

// exclusive rank of every kept row inside a 1024-row block; s_rank[r] valid where flag set.  Returns block total.
__device__ __forceinline__ uint32_t block_ranks(const uint8_t* __restrict__ flags, int64_t n, int64_t row0,
                                                uint8_t* s_flag, uint16_t* s_rank, uint32_t* s_warp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint32_t f4 = 0;
  const int64_t r = row0 + 4 * tid;
  if (r + 3 < n) f4 = *reinterpret_cast<const uint32_t*>(flags + r);
  else
    for (int j = 0; j < 4; ++j)
      if (r + j < n) f4 |= (uint32_t)flags[r + j] << (8 * j);
  uint32_t b[4], mine = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) { b[j] = ((f4 >> (8 * j)) & 0xffu) ? 1u : 0u; mine += b[j]; }
  uint32_t v = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += nb;
  }
  if (lane == 31) s_warp[wid] = v;
  __syncthreads();
  uint32_t before = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const uint32_t t = s_warp[w];
    before += w < wid ? t : 0u;
    total += t;
  }
  uint32_t run = before + v - mine;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    s_flag[4 * tid + j] = (uint8_t)b[j];
    s_rank[4 * tid + j] = (uint16_t)run;
    run += b[j];
  }
  __syncthreads();
  return total;
}

__global__ void __launch_bounds__(256) prune_flags_kernel(int P, int scale_dim, const float* __restrict__ logit,
                                                          const float* __restrict__ log_scales, float opa_thr,
                                                          float big_thr, uint8_t* __restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float op = 1.f / (1.f + expf(-logit[i]));
  bool remove = op < opa_thr;
  if (big_thr >= 0.f) {
    float m = expf(log_scales[(size_t)i * scale_dim]);
    for (int j = 1; j < scale_dim; ++j) m = fmaxf(m, expf(log_scales[(size_t)i * scale_dim + j]));
    remove = remove || (m > big_thr);
  }
  keep[i] = remove ? 0 : 1;
}

__global__ void __launch_bounds__(256) flag_count_kernel(int64_t n, const uint8_t* __restrict__ flags,
                                                         uint32_t* __restrict__ block_count) {
  __shared__ uint32_t s_warp[8];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t r = (int64_t)blockIdx.x * ROWS_PER_BLOCK + 4 * tid;
  uint32_t c = 0;
  for (int j = 0; j < 4; ++j) c += (r + j < n && flags[r + j]) ? 1u : 0u;
#pragma unroll
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) s_warp[wid] = c;
  __syncthreads();
  if (tid == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += s_warp[w];
    block_count[blockIdx.x] = t;
  }
}

// one CTA: block_base = exclusive scan(block_count); *total_dev = sum; optional copy-outs (device or mapped host)
__global__ void __launch_bounds__(1024) block_scan_kernel(int nblk, const uint32_t* __restrict__ block_count,
                                                          uint32_t* __restrict__ block_base, uint32_t* total_dev,
                                                          uint32_t* out, int out_slot, const uint32_t* extra_src) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    const int i = base + tid;
    const uint32_t c = i < nblk ? block_count[i] : 0u;
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t nb = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += nb;
    }
    if (lane == 31) warp_tot[wid] = v;
    __syncthreads();
    if (wid == 0) {
      uint32_t w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t nb = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += nb;
      }
      warp_tot[lane] = w;
    }
    __syncthreads();
    const uint32_t incl = carry_s + v + (wid ? warp_tot[wid - 1] : 0u);
    if (i < nblk) block_base[i] = incl - c;
    __syncthreads();
    if (tid == 1023) carry_s = incl;
    __syncthreads();
  }
  if (tid == 0) {
    *total_dev = carry_s;
    if (out) {
      out[out_slot] = carry_s;
      if (extra_src) out[0] = *extra_src;
    }
  }
}

#define COMPACT_MAX_TENSORS 24
struct CompactTensorK { const float* src; float* dst; int width; };
struct CompactArgs {
  int64_t n;
  const uint8_t* keep;
  const uint32_t* block_base;
  CompactTensorK t[COMPACT_MAX_TENSORS];
};

// One block per 1024 rows, ALL tensors: the ranks are formed once, then every tensor's 1024*w contiguous source
// floats are streamed with 16-byte loads (four in flight per thread) and scattered to rank order -- consecutive kept
// elements land on consecutive addresses, so the stores coalesce whenever runs of rows survive.
__global__ void __launch_bounds__(256) compact_rows_kernel(const __grid_constant__ CompactArgs a, int count) {
  __shared__ uint8_t s_flag[ROWS_PER_BLOCK];
  __shared__ uint16_t s_rank[ROWS_PER_BLOCK];
  __shared__ uint32_t s_warp[8];
  const int64_t row0 = (int64_t)blockIdx.x * ROWS_PER_BLOCK;
  const uint32_t total = block_ranks(a.keep, a.n, row0, s_flag, s_rank, s_warp);
  if (total == 0) return;
  const int rows = (int)min((int64_t)ROWS_PER_BLOCK, a.n - row0);
  const size_t base = a.block_base[blockIdx.x];
  const bool all = total == (uint32_t)rows;
  for (int ti = 0; ti < count; ++ti) {
    const CompactTensorK t = a.t[ti];
    const int w = t.width;
    const float* __restrict__ src = t.src + (size_t)row0 * w;
    float* __restrict__ dst = t.dst + base * w;
    const int elems = rows * w;
    const bool vec = ((elems & 3) == 0) && ((((uintptr_t)src) & 15) == 0);
    if (all) {                              // nothing dropped in this block: straight copy
      if (vec && ((((uintptr_t)dst) & 15) == 0)) {
        for (int e = threadIdx.x; e < elems / 4; e += 256)
          reinterpret_cast<float4*>(dst)[e] = __ldg(reinterpret_cast<const float4*>(src) + e);
      } else {
        for (int e = threadIdx.x; e < elems; e += 256) dst[e] = src[e];
      }
      continue;
    }
    if (vec) {
      const int quads = elems / 4;
      for (int q0 = threadIdx.x; q0 < quads; q0 += 4 * 256) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int q = q0 + u * 256;
          if (q < quads) v[u] = __ldg(reinterpret_cast<const float4*>(src) + q);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int q = q0 + u * 256;
          if (q >= quads) break;
          const float f[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int e = 4 * q + c;
            const int row = w == 1 ? e : (w == 3 ? e / 3 : (w == 4 ? e >> 2 : e / w));
            if (s_flag[row]) dst[(int)s_rank[row] * w + (e - row * w)] = f[c];
          }
        }
      }
    } else {
#pragma unroll 4
      for (int e = threadIdx.x; e < elems; e += 256) {
        const int row = w == 1 ? e : (w == 3 ? e / 3 : (w == 4 ? e >> 2 : e / w));
        const float v = src[e];
        if (s_flag[row]) dst[(int)s_rank[row] * w + (e - row * w)] = v;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ median select
// scratch words (uint32): [0,2048) hist0  [2048,4096) hist1  [4096,5120) hist2  5120 nan_count  5121 pre_count
// 5122 total; from byte align_up(SEL_BLOCKS*4): block_count[nblk] then block_base[nblk]
#define SEL_H0 0
#define SEL_H1 2048
#define SEL_H2 4096
#define SEL_NAN 5120
#define SEL_PRE 5121
#define SEL_TOTAL 5122
#define SEL_BLOCKS 5124

__device__ __forceinline__ float depth_error_of(float gt, float rd) {
  return fabsf(gt - rd) * (gt > 0.f ? 1.f : 0.f);   // torch.abs(gt - render) * (gt > 0): NaN/Inf * 0 stays NaN
}

// Every block redundantly locates the bucket rank `k` falls into (256 threads, nb = 2048 or 1024 bins): returns the
// bucket and rewrites k to the rank inside it.
__device__ __forceinline__ uint32_t select_bucket(const uint32_t* __restrict__ hist, int nb, uint32_t& k, uint32_t* s_tmp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int per = nb / 256;              // 8 or 4
  uint32_t c[8], mine = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { c[j] = j < per ? hist[tid * per + j] : 0u; mine += c[j]; }
  uint32_t v = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t nbv = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += nbv;
  }
  if (lane == 31) s_tmp[wid] = v;
  if (tid == 0) { s_tmp[8] = 0; s_tmp[9] = 0; }
  __syncthreads();
  uint32_t before = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) before += w < wid ? s_tmp[w] : 0u;
  const uint32_t incl = before + v, excl = incl - mine;
  if (k >= excl && k < incl) {           // exactly one thread (none if k >= the histogram total: NaN case)
    uint32_t run = excl;
    int j = 0;
#pragma unroll
    for (int jj = 0; jj < 7; ++jj) {
      if (j == jj && jj < per - 1 && k >= run + c[jj]) { run += c[jj]; j = jj + 1; }
    }
    s_tmp[8] = (uint32_t)(tid * per + j);
    s_tmp[9] = k - run;
  }
  __syncthreads();
  const uint32_t bucket = s_tmp[8];
  k = s_tmp[9];
  __syncthreads();
  return bucket;
}

template <int PASS>
__global__ void __launch_bounds__(256) err_hist_kernel(int N, const float* __restrict__ render_depth,
                                                       const float* __restrict__ gt_depth, uint32_t* __restrict__ sel) {
  __shared__ uint32_t s_hist[2048];
  __shared__ uint32_t s_tmp[10];
  for (int i = threadIdx.x; i < 2048; i += 256) s_hist[i] = 0;
  uint32_t prefix = 0;
  if (PASS >= 1) {
    uint32_t k = (uint32_t)((N - 1) / 2);
    prefix = select_bucket(sel + SEL_H0, 2048, k, s_tmp);
    if (PASS == 2) prefix = (prefix << 11) | select_bucket(sel + SEL_H1, 2048, k, s_tmp);
  }
  __syncthreads();
  uint32_t nans = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < N; i += gridDim.x * 256) {
    const float e = depth_error_of(gt_depth[i], render_depth[i]);
    if (e != e) { nans += (PASS == 0); continue; }
    const uint32_t b = __float_as_uint(e) & 0x7fffffffu;
    if (PASS == 0) atomicAdd(&s_hist[b >> 21], 1u);
    else if (PASS == 1) { if ((b >> 21) == prefix) atomicAdd(&s_hist[(b >> 10) & 2047u], 1u); }
    else { if ((b >> 10) == prefix) atomicAdd(&s_hist[b & 1023u], 1u); }
  }
  __syncthreads();
  uint32_t* g = sel + (PASS == 0 ? SEL_H0 : PASS == 1 ? SEL_H1 : SEL_H2);
  for (int i = threadIdx.x; i < (PASS == 2 ? 1024 : 2048); i += 256)
    if (s_hist[i]) atomicAdd(&g[i], s_hist[i]);
  if (PASS == 0 && nans) atomicAdd(&sel[SEL_NAN], nans);
}

__global__ void __launch_bounds__(256) spawn_mask_kernel(int N, const float* __restrict__ depth_sil,
                                                         const float* __restrict__ gt_depth, float sil_thres,
                                                         uint8_t* __restrict__ flags, uint32_t* __restrict__ sel,
                                                         uint32_t* __restrict__ block_count) {
  __shared__ uint32_t s_tmp[10];
  __shared__ uint32_t s_cnt[2];
  uint32_t k = (uint32_t)((N - 1) / 2);
  uint32_t bits = select_bucket(sel + SEL_H0, 2048, k, s_tmp);
  bits = (bits << 11) | select_bucket(sel + SEL_H1, 2048, k, s_tmp);
  bits = (bits << 10) | select_bucket(sel + SEL_H2, 1024, k, s_tmp);
  // torch.median propagates NaN; with NaN every `depth_error > 50 * median` is false
  const float median = sel[SEL_NAN] ? __uint_as_float(0x7fc00000u) : __uint_as_float(bits);
  const float far = 50.f * median;
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  uint32_t pre = 0, post = 0;
  const int r = blockIdx.x * ROWS_PER_BLOCK + 4 * threadIdx.x;
  for (int j = 0; j < 4; ++j) {
    const int i = r + j;
    if (i >= N) break;
    const float rd = depth_sil[i], sil = depth_sil[(size_t)N + i], gt = gt_depth[i];
    const float e = depth_error_of(gt, rd);
    const bool p = (sil < sil_thres) || ((rd > gt) && (e > far));
    const bool q = p && (gt > 0.f);
    flags[i] = q ? 1 : 0;
    pre += p; post += q;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    pre += __shfl_xor_sync(0xffffffffu, pre, o);
    post += __shfl_xor_sync(0xffffffffu, post, o);
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&s_cnt[0], pre); atomicAdd(&s_cnt[1], post); }
  __syncthreads();
  if (threadIdx.x == 0) {
    block_count[blockIdx.x] = s_cnt[1];
    if (s_cnt[0]) atomicAdd(&sel[SEL_PRE], s_cnt[0]);
  }
}

struct SpawnArgs {
  int W, H;
  const uint8_t* flags;
  const uint32_t* block_base;
  const float* im;          // [3,H,W]
  const float* gt_depth;    // [H,W]
  float fx, fy, cx, cy;
  const float* cam_unnorm_rots;   // [1,4,n_frames]
  const float* cam_trans;         // [1,3,n_frames]
  int n_frames, time_idx, scale_dim;
  int64_t P_old;
  float *means3D, *rgb, *rot, *logit, *log_scales, *timestep;
};

__global__ void __launch_bounds__(256) spawn_kernel(const __grid_constant__ SpawnArgs a) {
  __shared__ uint8_t s_flag[ROWS_PER_BLOCK];
  __shared__ uint16_t s_rank[ROWS_PER_BLOCK];
  __shared__ uint32_t s_warp[8];
  __shared__ float s_c2w[12];
  const int N = a.W * a.H;
  if (threadIdx.x == 0) {   // c2w = [R^T | -R^T t] of the frame's pose (build_rotation of the normalised quaternion)
    float q[4], t[3];
    for (int c = 0; c < 4; ++c) q[c] = a.cam_unnorm_rots[c * a.n_frames + a.time_idx];
    for (int c = 0; c < 3; ++c) t[c] = a.cam_trans[c * a.n_frames + a.time_idx];
    const float inv = 1.f / fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
    const float r = q[0] * inv, x = q[1] * inv, y = q[2] * inv, z = q[3] * inv;
    const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                        2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                        2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)};
    for (int i = 0; i < 3; ++i) {
      s_c2w[4 * i + 0] = R[0 + i]; s_c2w[4 * i + 1] = R[3 + i]; s_c2w[4 * i + 2] = R[6 + i];
      s_c2w[4 * i + 3] = -(R[0 + i] * t[0] + R[3 + i] * t[1] + R[6 + i] * t[2]);
    }
  }
  const int64_t row0 = (int64_t)blockIdx.x * ROWS_PER_BLOCK;
  const uint32_t total = block_ranks(a.flags, N, row0, s_flag, s_rank, s_warp);
  if (total == 0) return;
  const int64_t out0 = a.P_old + a.block_base[blockIdx.x];
  const float half_f = (a.fx + a.fy) / 2.f;
  for (int j = 0; j < 4; ++j) {
    const int l = 4 * threadIdx.x + j;
    if (!s_flag[l]) continue;
    const int i = (int)row0 + l;
    const int px = i % a.W, py = i / a.W;
    const float z = a.gt_depth[i];
    const float xc = ((float)px - a.cx) / a.fx * z, yc = ((float)py - a.cy) / a.fy * z;
    const int64_t o = out0 + s_rank[l];
    for (int c = 0; c < 3; ++c) {
      a.means3D[o * 3 + c] = fmaf(s_c2w[4 * c], xc, fmaf(s_c2w[4 * c + 1], yc, fmaf(s_c2w[4 * c + 2], z, s_c2w[4 * c + 3])));
      a.rgb[o * 3 + c] = a.im[(size_t)c * N + i];
    }
    *reinterpret_cast<float4*>(a.rot + o * 4) = make_float4(1.f, 0.f, 0.f, 0.f);
    a.logit[o] = 0.f;
    const float sg = z / half_f;
    const float ls = logf(sqrtf(sg * sg));
    for (int c = 0; c < a.scale_dim; ++c) a.log_scales[o * a.scale_dim + c] = ls;
    if (a.timestep) a.timestep[o] = (float)a.time_idx;
  }
}

}  // namespace splat

// ---------------------------------------------------------------------------------------------------- C ABI
using namespace splat;

static inline int64_t nblocks_of(int64_t n) { return (n + ROWS_PER_BLOCK - 1) / ROWS_PER_BLOCK; }
static inline size_t blocks_bytes(int64_t n) { return align_up((size_t)nblocks_of(n) * 4) * 2; }

extern "C" {

int splat_prune_flags(int32_t P, int32_t scale_dim, const float* logit_opacities, const float* log_scales,
                      float opacity_threshold, float big_threshold, uint8_t* keep, void* stream) {
  if (P < 0 || (scale_dim != 1 && scale_dim != 3)) return fail(SPLAT_ERR_BAD_ARG, "P < 0 or scale_dim not 1/3");
  if (P == 0) return SPLAT_OK;
  if (!logit_opacities || !log_scales || !keep) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  ProfScope prof_(K_SURGERY, (cudaStream_t)stream);
  prune_flags_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, scale_dim, logit_opacities, log_scales,
                                                                       opacity_threshold, big_threshold, keep);
  SPLAT_LAUNCH_CHECK("prune_flags_kernel");
  return SPLAT_OK;
}

size_t splat_compact_scratch_bytes(int64_t n) { return 256 + blocks_bytes(n > 0 ? n : 1); }

int splat_compact_rows(int64_t n, const uint8_t* keep, const SplatRowTensor* tensors, int32_t count, void* scratch,
                       size_t scratch_bytes, uint32_t* n_kept_out, void* stream) {
  if (n < 0 || count < 0 || count > COMPACT_MAX_TENSORS)
    return fail(SPLAT_ERR_BAD_ARG, "n < 0 or count not in 0..%d", COMPACT_MAX_TENSORS);
  if (n >= (int64_t)1 << 31) return fail(SPLAT_ERR_UNSUPPORTED, "n must be < 2^31 rows");
  if (!scratch || ((uintptr_t)scratch & 15)) return fail(SPLAT_ERR_BAD_ARG, "scratch is NULL or not 16-byte aligned");
  if (scratch_bytes < splat_compact_scratch_bytes(n)) return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* total = (uint32_t*)scratch;
  if (n == 0) {
    SPLAT_CUDA(cudaMemsetAsync(total, 0, 4, st));
    if (n_kept_out) SPLAT_CUDA(cudaMemcpyAsync(n_kept_out, total, 4, cudaMemcpyDefault, st));
    return SPLAT_OK;
  }
  if (!keep || (count && !tensors)) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  if ((uintptr_t)keep & 3) return fail(SPLAT_ERR_BAD_ARG, "keep must be 4-byte aligned");
  CompactArgs a = {};
  a.n = n; a.keep = keep;
  for (int i = 0; i < count; ++i) {
    if (!tensors[i].src || !tensors[i].dst || tensors[i].width < 1)
      return fail(SPLAT_ERR_BAD_ARG, "tensor %d: NULL pointer or width < 1", i);
    if (tensors[i].src == tensors[i].dst)
      return fail(SPLAT_ERR_BAD_ARG, "tensor %d: compaction is out of place (src == dst)", i);
    a.t[i].src = tensors[i].src; a.t[i].dst = tensors[i].dst; a.t[i].width = tensors[i].width;
  }
  const int nblk = (int)nblocks_of(n);
  uint32_t* block_count = (uint32_t*)((char*)scratch + 256);
  uint32_t* block_base = (uint32_t*)((char*)block_count + align_up((size_t)nblk * 4));
  a.block_base = block_base;
  ProfScope prof_(K_SURGERY, st);
  flag_count_kernel<<<nblk, 256, 0, st>>>(n, keep, block_count);
  SPLAT_LAUNCH_CHECK("flag_count_kernel");
  block_scan_kernel<<<1, 1024, 0, st>>>(nblk, block_count, block_base, total, n_kept_out, 0, nullptr);
  SPLAT_LAUNCH_CHECK("block_scan_kernel");
  if (count) {
    compact_rows_kernel<<<nblk, 256, 0, st>>>(a, count);
    SPLAT_LAUNCH_CHECK("compact_rows_kernel");
  }
  return SPLAT_OK;
}

size_t splat_spawn_scratch_bytes(int32_t W, int32_t H) {
  const int64_t n = (int64_t)(W > 0 ? W : 1) * (H > 0 ? H : 1);
  return align_up((size_t)SEL_BLOCKS * 4) + blocks_bytes(n);
}

int splat_new_gaussian_mask(int32_t W, int32_t H, const float* depth_sil, const float* gt_depth, float sil_thres,
                            uint8_t* flags, void* scratch, size_t scratch_bytes, uint32_t* counts_out, void* stream) {
  if (W <= 0 || H <= 0) return fail(SPLAT_ERR_BAD_ARG, "W, H must be positive");
  if (!depth_sil || !gt_depth || !flags) return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  if (!scratch || ((uintptr_t)scratch & 15) || ((uintptr_t)flags & 3))
    return fail(SPLAT_ERR_BAD_ARG, "scratch NULL / not 16-byte aligned, or flags not 4-byte aligned");
  if (scratch_bytes < splat_spawn_scratch_bytes(W, H)) return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "scratch too small");
  const int N = W * H;
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* sel = (uint32_t*)scratch;
  const int nblk = (int)nblocks_of(N);
  uint32_t* block_count = (uint32_t*)((char*)scratch + align_up((size_t)SEL_BLOCKS * 4));
  uint32_t* block_base = (uint32_t*)((char*)block_count + align_up((size_t)nblk * 4));
  ProfScope prof_(K_SURGERY, st);
  SPLAT_CUDA(cudaMemsetAsync(sel, 0, (size_t)SEL_BLOCKS * 4, st));
  const int hb = nblk * 4 < 148 * 8 ? nblk * 4 : 148 * 8;
  err_hist_kernel<0><<<hb, 256, 0, st>>>(N, depth_sil, gt_depth, sel);
  SPLAT_LAUNCH_CHECK("err_hist_kernel<0>");
  err_hist_kernel<1><<<hb, 256, 0, st>>>(N, depth_sil, gt_depth, sel);
  SPLAT_LAUNCH_CHECK("err_hist_kernel<1>");
  err_hist_kernel<2><<<hb, 256, 0, st>>>(N, depth_sil, gt_depth, sel);
  SPLAT_LAUNCH_CHECK("err_hist_kernel<2>");
  spawn_mask_kernel<<<nblk, 256, 0, st>>>(N, depth_sil, gt_depth, sil_thres, flags, sel, block_count);
  SPLAT_LAUNCH_CHECK("spawn_mask_kernel");
  block_scan_kernel<<<1, 1024, 0, st>>>(nblk, block_count, block_base, sel + SEL_TOTAL, counts_out, 1, sel + SEL_PRE);
  SPLAT_LAUNCH_CHECK("block_scan_kernel");
  return SPLAT_OK;
}

int splat_spawn_gaussians(int32_t W, int32_t H, const uint8_t* flags, const void* scratch, const float* im,
                          const float* gt_depth, float fx, float fy, float cx, float cy, const float* cam_unnorm_rots,
                          const float* cam_trans, int32_t n_frames, int32_t time_idx, int32_t scale_dim, int64_t P_old,
                          float* means3D, float* rgb_colors, float* unnorm_rotations, float* logit_opacities,
                          float* log_scales, float* timestep, void* stream) {
  if (W <= 0 || H <= 0 || P_old < 0) return fail(SPLAT_ERR_BAD_ARG, "W, H must be positive, P_old >= 0");
  if (scale_dim != 1 && scale_dim != 3) return fail(SPLAT_ERR_BAD_ARG, "scale_dim must be 1 or 3");
  if (n_frames <= 0 || time_idx < 0 || time_idx >= n_frames) return fail(SPLAT_ERR_BAD_ARG, "time_idx outside [0, n_frames)");
  if (!flags || !scratch || !im || !gt_depth || !cam_unnorm_rots || !cam_trans || !means3D || !rgb_colors ||
      !unnorm_rotations || !logit_opacities || !log_scales)
    return fail(SPLAT_ERR_BAD_ARG, "NULL pointer");
  if ((uintptr_t)unnorm_rotations & 15) return fail(SPLAT_ERR_BAD_ARG, "unnorm_rotations must be 16-byte aligned");
  const int N = W * H;
  const int nblk = (int)nblocks_of(N);
  const uint32_t* block_count = (const uint32_t*)((const char*)scratch + align_up((size_t)SEL_BLOCKS * 4));
  SpawnArgs a = {};
  a.W = W; a.H = H; a.flags = flags;
  a.block_base = (const uint32_t*)((const char*)block_count + align_up((size_t)nblk * 4));
  a.im = im; a.gt_depth = gt_depth; a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy;
  a.cam_unnorm_rots = cam_unnorm_rots; a.cam_trans = cam_trans; a.n_frames = n_frames; a.time_idx = time_idx;
  a.scale_dim = scale_dim; a.P_old = P_old;
  a.means3D = means3D; a.rgb = rgb_colors; a.rot = unnorm_rotations; a.logit = logit_opacities;
  a.log_scales = log_scales; a.timestep = timestep;
  ProfScope prof_(K_SURGERY, (cudaStream_t)stream);
  spawn_kernel<<<nblk, 256, 0, (cudaStream_t)stream>>>(a);
  SPLAT_LAUNCH_CHECK("spawn_kernel");
  return SPLAT_OK;
}

}  // extern "C"
