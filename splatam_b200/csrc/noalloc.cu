// noalloc.cu -- splat_slam_track_loss_noalloc: the tracking loss with library-owned scratch.
//
// No counterpart in the reference; added for callers that
// cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  The kernels are the ones splat_slam_loss(tracking = 1) launches
// (slam_ops.cu: track_loss_kernel + loss_finalize_kernel); the only difference is where their scratch
// lives: a file-scope __device__ array instead of a caller buffer.  Nothing here allocates or frees.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the device address
// is looked up with cudaGetSymbolAddress once per device (with that device current) and cached.
// Every GPU that loads the library holds the array until exit.  All calls share it: calls that
// overlap on the GPU (different streams, or several host threads) race on it.  Nothing in it has
// to start at zero -- every partial row a call reads is written by that same call's first kernel,
// and the stats block by its finalize kernel -- so there is nothing to clear between calls.
#include <mutex>
#include "kernels.cuh"

namespace splat {

// Scratch layout = splat_slam_loss(tracking = 1)'s: [partials: blocks x 4 floats][stats: 256 bytes].
// slam_loss_blocks caps the tracking grid at 148 * 8 blocks, so 148*8*16 + 256 = 19,200 bytes cover every
// image up to the stated maximum; the entry point re-derives the requirement and refuses anything larger.
#define NOALLOC_MAX_BLOCKS (148 * 8)
#define NOALLOC_SCRATCH_BYTES (NOALLOC_MAX_BLOCKS * 4 * 4 + 256)
#define NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) unsigned char g_track_loss_scratch[NOALLOC_SCRATCH_BYTES];

static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of the array, per device ordinal

// Device address of the scratch array on the CURRENT device (looked up once per device).
static int noalloc_scratch(void** out) {
  int dev = -1;
  SPLAT_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(SPLAT_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    SPLAT_CUDA(cudaGetSymbolAddress(&p, g_track_loss_scratch));
    if (!p) return fail(SPLAT_ERR_UNSUPPORTED, "cudaGetSymbolAddress returned NULL for the scratch array");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return SPLAT_OK;
}

}  // namespace splat

using namespace splat;

extern "C" int splat_slam_track_loss_noalloc(int32_t W, int32_t H, int32_t use_sil_for_loss, float sil_thres, float w_im,
                                             float w_depth, const float* im, const float* depth_sil, const float* gt_im,
                                             const float* gt_depth, float* losses, float* dL_dim, float* dL_ddepth_sil,
                                             void* stream) {
  if (W <= 0 || H <= 0) return fail(SPLAT_ERR_BAD_ARG, "bad image size %dx%d", W, H);
  if ((int64_t)W * (int64_t)H > (int64_t)SPLAT_NOALLOC_MAX_PIXELS)
    return fail(SPLAT_ERR_SCRATCH_TOO_SMALL,
                "splat_slam_track_loss_noalloc: image %dx%d = %lld pixels exceeds the built-in scratch's maximum of %d "
                "pixels; use splat_slam_loss with a caller-owned scratch buffer",
                W, H, (long long)W * (long long)H, (int)SPLAT_NOALLOC_MAX_PIXELS);
  const int nb = slam_loss_blocks(W, H, 1);
  const size_t part_bytes = align_up((size_t)nb * 4 * sizeof(float));
  if (nb > NOALLOC_MAX_BLOCKS || part_bytes + 256 > (size_t)NOALLOC_SCRATCH_BYTES)
    return fail(SPLAT_ERR_SCRATCH_TOO_SMALL, "splat_slam_track_loss_noalloc: %d partial rows exceed the built-in scratch (%d rows)",
                nb, (int)NOALLOC_MAX_BLOCKS);
  if (!im || !depth_sil || !gt_im || !gt_depth || !losses || !dL_dim || !dL_ddepth_sil)
    return fail(SPLAT_ERR_BAD_ARG, "image / loss / gradient pointers must be non-NULL");
  void* base = nullptr;
  if (int rc = noalloc_scratch(&base)) return rc;
  LossArgs a = {};
  a.W = W; a.H = H; a.tracking = 1; a.use_sil = use_sil_for_loss != 0;
  a.sil_thres = sil_thres; a.w_im = w_im; a.w_depth = w_depth;
  a.im = im; a.depth_sil = depth_sil; a.gt_im = gt_im; a.gt_depth = gt_depth;
  a.losses = losses; a.g_im = dL_dim; a.g_depth_sil = dL_ddepth_sil;
  a.dmaps = nullptr;
  float* partial = reinterpret_cast<float*>(base);
  float* stats = reinterpret_cast<float*>(reinterpret_cast<char*>(base) + part_bytes);
  return launch_slam_loss(a, partial, stats, (cudaStream_t)stream);
}
