"""splat_slam_track_loss_noalloc (include/splat_b200.h; no counterpart in the reference):
the tracking loss with its scratch in a static __device__ array inside the library.  Checked through
the C ABI against the existing entry point splat_slam_loss(tracking=1) with a caller-owned scratch, on two calls in a
row with different inputs (the shared array keeps the first call's contents).  Collected last on purpose."""
import ctypes as C
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MAX_PIXELS = 4096 * 4096   # SPLAT_NOALLOC_MAX_PIXELS


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _inputs(W, H, seed):
    g = torch.Generator().manual_seed(seed)
    im = torch.rand(3, H, W, generator=g)
    gt_im = torch.rand(3, H, W, generator=g)
    depth = 1.0 + 2.0 * torch.rand(H, W, generator=g)
    sil = torch.rand(H, W, generator=g)                       # ~1% of pixels fall under the 0.99 threshold's complement
    sil = torch.where(sil > 0.3, torch.ones_like(sil), sil)
    d2 = depth * depth + 0.01 * torch.rand(H, W, generator=g)
    depth_sil = torch.stack([depth, sil, d2])
    gt_depth = depth + 0.1 * torch.randn(H, W, generator=g)
    gt_depth[torch.rand(H, W, generator=g) < 0.05] = 0.0      # invalid sensor depth
    depth_sil[0][torch.rand(H, W, generator=g) < 0.001] = float("nan")
    return [t.contiguous().cuda() for t in (im, depth_sil, gt_im, gt_depth)]


def _run_existing(lib, check, W, H, use_sil, thres, w_im, w_depth, im, depth_sil, gt_im, gt_depth):
    losses = torch.full((4,), -1.0, device="cuda")
    g_im = torch.full((3, H, W), 7.0, device="cuda")
    g_ds = torch.full((3, H, W), 7.0, device="cuda")
    nbytes = lib.splat_slam_loss_scratch_bytes(W, H, 1)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib.splat_slam_loss(W, H, 1, use_sil, thres, w_im, w_depth, _vp(im), _vp(depth_sil), _vp(gt_im), _vp(gt_depth),
                              _vp(losses), _vp(g_im), _vp(g_ds), _vp(scratch), nbytes, st))
    torch.cuda.synchronize()
    return losses, g_im, g_ds


def _run_noalloc(lib, check, W, H, use_sil, thres, w_im, w_depth, im, depth_sil, gt_im, gt_depth):
    losses = torch.full((4,), -1.0, device="cuda")
    g_im = torch.full((3, H, W), 7.0, device="cuda")
    g_ds = torch.full((3, H, W), 7.0, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib.splat_slam_track_loss_noalloc(W, H, use_sil, thres, w_im, w_depth, _vp(im), _vp(depth_sil), _vp(gt_im),
                                            _vp(gt_depth), _vp(losses), _vp(g_im), _vp(g_ds), st))
    torch.cuda.synchronize()
    return losses, g_im, g_ds


@pytest.mark.gpu
def test_noalloc_track_loss_matches_existing_entry_point_twice():
    from splatam_b200._lib import lib, check
    # two calls in a row, different sizes and data: the second runs over the first call's leftovers in the shared array
    for W, H, use_sil, seed in ((640, 480, 1, 11), (1200, 680, 0, 12)):
        ins = _inputs(W, H, seed)
        ref = _run_existing(lib, check, W, H, use_sil, 0.99, 0.5, 1.0, *ins)
        out = _run_noalloc(lib, check, W, H, use_sil, 0.99, 0.5, 1.0, *ins)
        # same kernels, same deterministic two-stage reduction: held to the forward tolerance of the parity tests (1e-5
        # relative), in practice bit-identical
        assert torch.isfinite(ref[0]).all() and float(ref[0][3]) > 0
        assert torch.allclose(out[0], ref[0], rtol=1e-5, atol=0.0), (out[0], ref[0])
        assert torch.equal(out[1], ref[1])
        assert torch.equal(out[2], ref[2])
    # and once more on the first shape: results do not depend on what an earlier, larger call left behind
    ins = _inputs(640, 480, 11)
    ref = _run_existing(lib, check, 640, 480, 1, 0.99, 0.5, 1.0, *ins)
    out = _run_noalloc(lib, check, 640, 480, 1, 0.99, 0.5, 1.0, *ins)
    assert torch.allclose(out[0], ref[0], rtol=1e-5, atol=0.0)


def test_noalloc_track_loss_rejects_oversized_images_before_touching_cuda():
    """Validation precedes every CUDA call, so this runs on a machine without a GPU too."""
    from splatam_b200._lib import lib
    null = C.c_void_p(0)
    rc = lib.splat_slam_track_loss_noalloc(4096, 4097, 1, 0.99, 0.5, 1.0, null, null, null, null, null, null, null, null)
    assert rc != 0 and lib.splat_strerror(rc) == b"scratch buffer too small"
    msg = lib.splat_last_error().decode()
    assert "exceeds the built-in scratch's maximum of %d pixels" % MAX_PIXELS in msg
    rc = lib.splat_slam_track_loss_noalloc(0, 480, 1, 0.99, 0.5, 1.0, null, null, null, null, null, null, null, null)
    assert rc != 0 and lib.splat_strerror(rc) == b"bad argument"
    rc = lib.splat_slam_track_loss_noalloc(640, 480, 1, 0.99, 0.5, 1.0, null, null, null, null, null, null, null, null)
    assert rc != 0 and b"non-NULL" in lib.splat_last_error()
