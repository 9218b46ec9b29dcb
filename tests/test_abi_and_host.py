"""CPU tests (-m "not gpu") of the drop-in boundary: the C-ABI library loads and exports every symbol
include/splat_b200.h declares, host-only entry points (sizes, error strings, argument validation
that happens before any CUDA call) behave, and the Python shims keep the reference's API surface
(HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py:140-204,
DG/diff_gaussian_rasterization/__init__.py).  No compute call is made here."""
import ctypes as C
import inspect
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "splat_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(splat_[A-Za-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from splatam_b200 import _lib
    names = _declared_symbols()
    assert len(names) >= 20
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libsplat_b200.so does not export %s" % n
    # and the ctypes signature table covers the whole header (no entry point left unbound)
    assert set(names) == set(_lib.EXPORTED_SYMBOLS)


def test_abi_version_and_error_strings():
    from splatam_b200._lib import lib
    assert lib.splat_abi_version() == 7
    assert lib.splat_strerror(0) == b"ok"
    assert lib.splat_strerror(1) == b"bad argument"
    assert lib.splat_strerror(2) == b"scratch buffer too small"
    assert lib.splat_strerror(3) == b"unsupported configuration"
    assert lib.splat_strerror(1000 + 2) == b"CUDA runtime error"


def test_state_sizes_are_monotone_and_aligned():
    from splatam_b200._lib import lib
    prev = 0
    for P in (0, 1, 255, 256, 10_000, 1_000_000):
        b = lib.splat_geom_state_bytes(P)
        assert b % 256 == 0 and b >= prev and b >= P * (4 + 8 + 16)
        prev = b
    for (w, h) in ((1, 1), (64, 64), (75, 53), (1200, 680), (1752, 1168)):
        b = lib.splat_image_state_bytes(w, h)
        T = ((w + 15) // 16) * ((h + 15) // 16)
        assert b % 256 == 0 and b >= 8 * w * h + 16 * T
    assert lib.splat_binning_state_bytes(64, 64, 0) > 0
    r1, r2 = lib.splat_binning_state_bytes(1200, 680, 1_000_000), lib.splat_binning_state_bytes(1200, 680, 2_000_000)
    assert r2 > r1 >= 1_000_000 * 20


def test_argument_validation_happens_before_any_cuda_call():
    """NULL camera / bad sizes are rejected with SPLAT_ERR_BAD_ARG on a box without a GPU."""
    from splatam_b200._lib import SplatCamera, lib
    null = C.c_void_p(0)
    rc = lib.splat_forward_geometry(None, 1, null, null, null, null, null, null, 0, null, null, null, null, null, null)
    assert rc == 1 and b"camera" in lib.splat_last_error()
    cam = SplatCamera()
    cam.image_width, cam.image_height = 0, 16
    rc = lib.splat_forward_geometry(C.byref(cam), 1, null, null, null, null, null, null, 0, null, null, null, null,
                                    null, null)
    assert rc == 1 and b"image size" in lib.splat_last_error()
    cam.image_width = 16
    rc = lib.splat_forward_render(C.byref(cam), 1, null, null, null, null, null, 10, null, null, null)
    assert rc == 1
    rc = lib.splat_mark_visible(-1, null, null, null, 0.2, null, null)
    assert rc == 1
    assert lib.splat_mark_visible(0, null, null, null, 0.2, null, null) == 0   # P == 0 short-circuits


def test_slam_glue_entry_points_validate_before_cuda():
    """splat_rendervar_* / splat_slam_loss (ABI v4): sizes and argument checks need no GPU."""
    from splatam_b200._lib import lib
    null = C.c_void_p(0)
    one = C.c_void_p(256)   # any non-NULL, 16-byte aligned address: validation fails before it is dereferenced
    assert lib.splat_rendervar_scratch_bytes(0) >= 64 and lib.splat_rendervar_scratch_bytes(0) % 256 == 0
    assert lib.splat_rendervar_scratch_bytes(10_000_000) <= 148 * 8 * 64 + 256
    rc = lib.splat_rendervar_forward(10, 2, one, one, one, one, one, one, 5, 0, one, one, one, one, null)
    assert rc == 1 and b"1 or 3 columns" in lib.splat_last_error()
    rc = lib.splat_rendervar_forward(10, 1, one, one, one, one, one, one, 5, 5, one, one, one, one, null)
    assert rc == 1 and b"time_idx" in lib.splat_last_error()
    rc = lib.splat_rendervar_forward(10, 1, one, one, one, one, null, one, 5, 0, one, one, one, one, null)
    assert rc == 1 and b"cam_" in lib.splat_last_error()
    rc = lib.splat_rendervar_backward(10, 3, one, one, one, one, one, one, 5, 1, null, null, null, null, 1, 0,
                                      null, null, null, null, null, null, null, 0, null)
    assert rc == 1 and b"gaussians_grad" in lib.splat_last_error()
    rc = lib.splat_rendervar_backward(10, 3, one, one, one, one, one, one, 5, 1, null, null, null, null, 0, 1,
                                      null, null, null, null, one, one, null, 0, null)
    assert rc == 2 and b"scratch" in lib.splat_last_error()
    t, m = lib.splat_slam_loss_scratch_bytes(1200, 680, 1), lib.splat_slam_loss_scratch_bytes(1200, 680, 0)
    assert t % 256 == 0 and m % 256 == 0 and m - t >= 9 * 4 * 1200 * 680
    assert lib.splat_slam_loss_scratch_bytes(0, 680, 1) == 0
    rc = lib.splat_slam_loss(0, 4, 1, 1, 0.99, 0.5, 1.0, one, one, one, one, one, one, one, one, 1 << 20, null)
    assert rc == 1 and b"image size" in lib.splat_last_error()
    rc = lib.splat_slam_loss(64, 64, 1, 1, 0.99, 0.5, 1.0, one, one, one, null, one, one, one, one, 1 << 20, null)
    assert rc == 1
    rc = lib.splat_slam_loss(64, 64, 0, 0, 0.5, 0.5, 1.0, one, one, one, one, one, one, one, one, 1024, null)
    assert rc == 2 and b"scratch" in lib.splat_last_error()


def test_adam_entry_point_validates_before_cuda():
    from splatam_b200._lib import SplatAdamTensor, lib
    d = (SplatAdamTensor * 16)()
    assert lib.splat_adam_step(d, 0, 0.9, 0.999, 1e-8, None, None, None) == 0
    assert lib.splat_adam_step(d, 17, 0.9, 0.999, 1e-8, None, None, None) == 1
    assert lib.splat_adam_step(None, 1, 0.9, 0.999, 1e-8, None, None, None) == 1
    d[0].n = 8
    assert lib.splat_adam_step(d, 1, 0.9, 0.999, 1e-8, None, None, None) == 1 and b"NULL" in lib.splat_last_error()
    d[0].param = d[0].grad = d[0].exp_avg = d[0].exp_avg_sq = 256
    assert lib.splat_adam_step(d, 1, 0.9, 0.999, 1e-8, None, None, None) == 1 and b"step" in lib.splat_last_error()
    assert lib.splat_adam_step(d, 1, 0.9, 0.999, 1e-8, C.c_void_p(256), None, None) == 1 and b"scratch" in lib.splat_last_error()
    assert lib.splat_forward_guard(None, 64, 64, 10, C.c_void_p(256), None) == 1


def test_slam_module_mirrors_reference_signatures():
    """splatam_b200.slam.get_loss keeps the parameter list of scripts/splatam.py:220-222 (plus a trailing Renderer=)."""
    from splatam_b200 import slam
    names = list(inspect.signature(slam.get_loss).parameters)
    assert names[:15] == ["params", "curr_data", "variables", "iter_time_idx", "loss_weights", "use_sil_for_loss",
                          "sil_thres", "use_l1", "ignore_outlier_depth_loss", "tracking", "mapping", "do_ba",
                          "plot_dir", "visualize_tracking_loss", "tracking_iteration"]
    with pytest.raises(NotImplementedError):
        slam.get_loss({}, {}, {}, 0, {}, True, 0.99, True, True, tracking=True)
    with pytest.raises(NotImplementedError):
        slam.get_loss({}, {}, {}, 0, {}, True, 0.99, False, False, tracking=True)
    assert list(inspect.signature(slam.build_rendervar).parameters)[:4] == ["params", "time_idx", "gaussians_grad",
                                                                              "camera_grad"]


@pytest.mark.parametrize("pkg,near,has_power", [("diff_gaussian_rasterization", 0.2, False),
                                                ("hessian_diff_gaussian_rasterization_w_depth", 0.001, True)])
def test_python_api_surface(pkg, near, has_power):
    m = __import__(pkg)
    S = m.GaussianRasterizationSettings
    assert S._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
                         "projmatrix", "sh_degree", "campos", "prefiltered")
    R = m.GaussianRasterizer
    assert abs(R.NEAR_PLANE - near) < 1e-9
    params = list(inspect.signature(R.__init__).parameters)
    assert ("backward_power" in params) == has_power
    fwd = list(inspect.signature(R.forward).parameters)
    assert fwd == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations",
                   "cov3D_precomp"]
    assert hasattr(R, "markVisible") and hasattr(m, "rasterize_gaussians") and hasattr(m, "_RasterizeGaussians")


@pytest.mark.parametrize("pkg", ["diff_gaussian_rasterization", "hessian_diff_gaussian_rasterization_w_depth"])
def test_reference_exception_texts(pkg):
    """Same checks and messages as HX/.../__init__.py:174-178 (raised before touching the device)."""
    m = __import__(pkg)
    from splatam_b200 import synth
    rs = synth.make_settings(**synth.C1)
    r = m.GaussianRasterizer(rs)
    p = torch.zeros(4, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=p, means2D=p, opacities=torch.zeros(4, 1), scales=p, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=p, means2D=p, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=p, scales=p,
          rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(means3D=p, means2D=p, opacities=torch.zeros(4, 1), colors_precomp=p)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(means3D=p, means2D=p, opacities=torch.zeros(4, 1), colors_precomp=p, scales=p, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))


def test_product_path_has_no_cpu_fallback():
    """CPU tensors are refused (the reference would crash in the kernel); nothing routes to the oracle."""
    import hessian_diff_gaussian_rasterization_w_depth as m
    from splatam_b200 import synth
    rs = synth.make_settings(**synth.C1)
    p = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        m.GaussianRasterizer(rs)(means3D=p, means2D=p, opacities=torch.zeros(4, 1), colors_precomp=p, scales=p,
                                 rotations=torch.zeros(4, 4))
    pkg_dir = os.path.join(ROOT, "splatam_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
    for shim in ("diff_gaussian_rasterization", "hessian_diff_gaussian_rasterization_w_depth"):
        src = open(os.path.join(ROOT, shim, "__init__.py")).read()
        assert "oracle" not in src


def test_camera_matrices_follow_setup_camera_layout():
    """synth.camera_matrices restates utils/recon_helpers.py:4-26: viewmatrix = w2c^T,
    projmatrix = (P_gl w2c)^T; with ndc2Pix (auxiliary.h:41-44) a point on the optical axis lands at
    (cx - 0.5, cy - 0.5) -- the reference's own half-pixel convention."""
    from splatam_b200 import synth
    cam = synth.REPLICA
    w2c = synth.arc_pose(3, 24)
    view, full, center = synth.camera_matrices(w2c=w2c, **cam)
    assert torch.allclose(view, w2c.t())
    p_cam = torch.tensor([0.0, 0.0, 2.0, 1.0])
    p_world = torch.inverse(w2c) @ p_cam
    clip = p_world @ full          # row-vector times the transposed matrix
    ndc = clip[:2] / clip[3]
    pix = ((ndc + 1.0) * torch.tensor([cam["w"], cam["h"]]) - 1.0) * 0.5
    assert torch.allclose(pix, torch.tensor([cam["cx"] - 0.5, cam["cy"] - 0.5]), atol=1e-3)
    assert torch.allclose(center, torch.inverse(w2c)[:3, 3])


def test_map_surgery_entry_points_validate_before_cuda():
    """splat_prune_flags / splat_compact_rows / splat_new_gaussian_mask / splat_spawn_gaussians (ABI v5)."""
    from splatam_b200._lib import SplatRowTensor, lib
    null, one = C.c_void_p(0), C.c_void_p(256)
    assert lib.splat_prune_flags(0, 1, null, null, 0.005, -1.0, null, null) == 0           # P == 0 short-circuits
    assert lib.splat_prune_flags(10, 2, one, one, 0.005, -1.0, one, null) == 1
    assert lib.splat_prune_flags(10, 3, one, null, 0.005, -1.0, one, null) == 1 and b"NULL" in lib.splat_last_error()
    a, b = lib.splat_compact_scratch_bytes(1_000_000), lib.splat_compact_scratch_bytes(0)
    assert a % 256 == 0 and b >= 256 and a >= 2 * 4 * 977
    d = (SplatRowTensor * 24)()
    assert lib.splat_compact_rows(10, one, d, 25, one, 1 << 20, null, null) == 1
    assert lib.splat_compact_rows(10, one, d, 1, C.c_void_p(260), 1 << 20, null, null) == 1 and b"aligned" in lib.splat_last_error()
    assert lib.splat_compact_rows(10, one, d, 1, one, 16, null, null) == 2
    assert lib.splat_compact_rows(10, one, d, 1, one, 1 << 20, null, null) == 1 and b"tensor 0" in lib.splat_last_error()
    d[0].src = d[0].dst = 512
    d[0].width = 3
    assert lib.splat_compact_rows(10, one, d, 1, one, 1 << 20, null, null) == 1 and b"out of place" in lib.splat_last_error()
    assert lib.splat_compact_rows(10, C.c_void_p(258), d, 0, one, 1 << 20, null, null) == 1 and b"4-byte" in lib.splat_last_error()
    s = lib.splat_spawn_scratch_bytes(1200, 680)
    assert s % 256 == 0 and s >= 5124 * 4 + 2 * 4 * 797
    assert lib.splat_new_gaussian_mask(0, 10, one, one, 0.5, one, one, s, null, null) == 1
    assert lib.splat_new_gaussian_mask(64, 64, one, null, 0.5, one, one, s, null, null) == 1
    assert lib.splat_new_gaussian_mask(1200, 680, one, one, 0.5, one, one, 1024, null, null) == 2
    args = [64, 64, one, one, one, one, 50.0, 50.0, 32.0, 32.0, one, one, 5, 2, 1, 100, one, one, one, one, one, null, null]
    bad = list(args); bad[13] = 5
    assert lib.splat_spawn_gaussians(*bad) == 1 and b"time_idx" in lib.splat_last_error()
    bad = list(args); bad[14] = 2
    assert lib.splat_spawn_gaussians(*bad) == 1 and b"scale_dim" in lib.splat_last_error()
    bad = list(args); bad[18] = C.c_void_p(260)
    assert lib.splat_spawn_gaussians(*bad) == 1 and b"16-byte" in lib.splat_last_error()
    bad = list(args); bad[16] = null
    assert lib.splat_spawn_gaussians(*bad) == 1 and b"NULL" in lib.splat_last_error()


def test_bookkeeping_entry_points_validate_before_cuda():
    from splatam_b200._lib import lib
    null, one = C.c_void_p(0), C.c_void_p(256)
    assert lib.splat_seen_update(0, null, null, null, null) == 0
    assert lib.splat_scale_inplace(null, 0, null, null) == 0 and lib.splat_scale_inplace(one, -1, one, null) == 1
    assert lib.splat_scale_inplace(one, 8, null, null) == 1 and b"NULL" in lib.splat_last_error()
    assert lib.splat_seen_update(-1, one, one, null, null) == 1
    assert lib.splat_seen_update(10, one, null, null, null) == 1 and b"NULL" in lib.splat_last_error()
    assert lib.splat_best_pose_update(one, one, one, one, 5, 5, one, one, null) == 1 and b"time_idx" in lib.splat_last_error()
    assert lib.splat_best_pose_update(one, null, one, one, 5, 1, one, one, null) == 1


def test_peer_exchange_entry_points_validate_before_cuda():
    """splat_peer_allreduce_sum / splat_peer_reduce_adam / splat_peer_set_tuning: argument checks and the
    single-rank no-op need no GPU and launch nothing."""
    from splatam_b200._lib import lib
    null = C.c_void_p(0)
    one_buf = (C.c_uint64 * 1)(256)
    two_buf = (C.c_uint64 * 2)(256, 512)
    misaligned = (C.c_uint64 * 2)(256, 520)
    n0 = lib.splat_kernel_launch_count()
    lib.splat_peer_set_tuning(0, 0)      # defaults
    lib.splat_peer_set_tuning(64, 3)     # an unroll the kernels are not built for falls back to the default
    lib.splat_peer_set_tuning(0, 0)
    assert lib.splat_peer_allreduce_sum(null, null, 0, 0, 2, 1, 1024, null) == 1 and b"NULL" in lib.splat_last_error()
    assert lib.splat_peer_allreduce_sum(two_buf, two_buf, 0, 2, 2, 1, 1024, null) == 1 and b"rank" in lib.splat_last_error()
    assert lib.splat_peer_allreduce_sum(two_buf, two_buf, 0, 0, 17, 1, 1024, null) == 1
    assert lib.splat_peer_allreduce_sum(two_buf, two_buf, 0, 0, 2, 1, 1022, null) == 1 and b"multiple of 4" in lib.splat_last_error()
    assert lib.splat_peer_allreduce_sum(misaligned, two_buf, 0, 0, 2, 1, 1024, null) == 1 and b"misaligned" in lib.splat_last_error()
    assert lib.splat_peer_allreduce_sum(one_buf, one_buf, 0, 0, 1, 1, 1024, null) == 0      # one rank: nothing to exchange
    assert lib.splat_peer_allreduce_sum(two_buf, two_buf, 0, 0, 2, 1, 0, null) == 0         # empty bucket
    assert lib.splat_peer_adam_shard_floats(14_000_000, 8) == 1_750_000
    assert lib.splat_peer_adam_shard_floats(1_000_003, 8) % 4 == 0 and lib.splat_peer_adam_shard_floats(1_000_003, 8) * 8 >= 1_000_003
    assert lib.splat_peer_adam_shard_floats(0, 8) == 0
    ends = (C.c_int64 * 2)(512, 1024)
    lrs = (C.c_float * 2)(1e-3, 1e-2)
    shard = C.c_void_p(1024)
    args = [two_buf, two_buf, two_buf, 0, 0, 0, 2, 1, 1024, ends, lrs, 2, 1, 0.9, 0.999, 1e-15, shard, shard, null]
    bad = list(args); bad[12] = 0
    assert lib.splat_peer_reduce_adam(*bad) == 1 and b"step" in lib.splat_last_error()
    bad = list(args); bad[11] = 0
    assert lib.splat_peer_reduce_adam(*bad) == 1 and b"segment" in lib.splat_last_error()
    bad = list(args); bad[9] = (C.c_int64 * 2)(1024, 512)
    assert lib.splat_peer_reduce_adam(*bad) == 1 and b"non-decreasing" in lib.splat_last_error()
    bad = list(args); bad[3] = 4096      # one multicast address without the other
    assert lib.splat_peer_reduce_adam(*bad) == 1 and b"multicast" in lib.splat_last_error()
    bad = list(args); bad[16] = null
    assert lib.splat_peer_reduce_adam(*bad) == 1 and b"moment" in lib.splat_last_error()
    assert lib.splat_kernel_launch_count() == n0


def test_densify_module_mirrors_reference_signatures():
    """splatam_b200.densify keeps the parameter lists of utils/slam_external.py:143,172 and scripts/splatam.py:384-385."""
    from splatam_b200 import densify
    assert list(inspect.signature(densify.remove_points).parameters) == ["to_remove", "params", "variables", "optimizer"]
    assert list(inspect.signature(densify.prune_gaussians).parameters) == ["params", "variables", "optimizer", "iter", "prune_dict"]
    assert list(inspect.signature(densify.add_new_gaussians).parameters)[:7] == [
        "params", "variables", "curr_data", "sil_thres", "time_idx", "mean_sq_dist_method", "gaussian_distribution"]
    p = {"means3D": None, "log_scales": type("T", (), {"shape": (4, 1)})()}
    with pytest.raises(ValueError):
        densify.add_new_gaussians(p, {}, {}, 0.5, 0, "nearest", "isotropic")
    with pytest.raises(ValueError):
        densify.add_new_gaussians(p, {}, {}, 0.5, 0, "projective", "anisotropic")
