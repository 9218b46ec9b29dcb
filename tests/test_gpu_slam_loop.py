"""GPU: the hot path COMPOSED the way scripts/splatam.py:690-900 drives it, two frames long --
  tracking iterations (get_loss -> backward -> Adam on the pose)  ->  add_new_gaussians  ->  mapping iterations
  (get_loss -> backward -> Adam on the map) with prune_gaussians in the loop --
once with this repo's fused pieces (slam.get_loss, slam.FusedAdam, densify.*) and once as the reference flow
(oracle glue in torch, two three-channel passes through the drop-in rasterizer, torch.optim.Adam, the oracle's
boolean-index surgery).  Same seeds, same schedule; the map size must agree after every surgery step and poses /
parameters / Adam moments must agree within 2e-3 of each array's scale after the whole sequence (float32 gradients
at 1e-3, a dozen Adam steps)."""
import pytest
import torch

from oracle import densify_oracle as DO
from oracle import slam_oracle as SO

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CAM = dict(w=160, h=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5)
TRACK_LRS = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
                 cam_unnorm_rots=4e-4, cam_trans=2e-3)
MAP_LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3,
               cam_unnorm_rots=0.0, cam_trans=0.0)
PRUNE = dict(start_after=0, remove_big_after=0, stop_after=20, prune_every=3, removal_opacity_threshold=0.05,
             final_removal_opacity_threshold=0.05, reset_opacities=False, reset_opacities_every=500)
WTS = dict(im=0.5, depth=1.0)
N_TRACK, N_MAP, FRAMES = 4, 6, (1, 2)


def _scene():
    from splatam_b200 import synth
    w2c0 = torch.eye(4)
    rs = synth.make_settings(device=DEV, w2c=w2c0, **CAM)
    P = 2500
    sc = synth.surface_scene(P, CAM["w"], CAM["h"], CAM["fx"], CAM["fy"], CAM["cx"], CAM["cy"], seed=21)
    g = torch.Generator().manual_seed(21)
    nf = 3
    rots = torch.tensor([1.0, 0, 0, 0]).view(1, 4, 1).repeat(1, 1, nf) + 0.004 * torch.randn(1, 4, nf, generator=g)
    # bimodal opacities: nothing comes within reach of the 0.05 prune threshold in a dozen Adam steps, so a 1e-6
    # difference between the two flows cannot flip a discrete decision
    u = torch.rand(P, 1, generator=g)
    opac = torch.where(torch.rand(P, 1, generator=g) < 0.15, 0.001 + 0.019 * u, 0.3 + 0.68 * u)
    params = {"means3D": sc["means3D"], "rgb_colors": sc["colors_precomp"], "unnorm_rotations": sc["rotations"],
              "logit_opacities": torch.logit(opac), "log_scales": torch.log(sc["scales"][:, :1]),
              "cam_unnorm_rots": rots, "cam_trans": 0.004 * torch.randn(1, 3, nf, generator=g)}
    params = {k: v.contiguous().to(DEV) for k, v in params.items()}
    frames = {}
    for t in FRAMES:
        im = torch.rand(3, CAM["h"], CAM["w"], generator=g)
        d = 2.5 + 0.5 * torch.rand(1, CAM["h"], CAM["w"], generator=g)
        d[0, :4] = 0.0
        frames[t] = {"cam": rs, "im": im.to(DEV), "depth": d.to(DEV), "w2c": w2c0.to(DEV),
                     "intrinsics": torch.tensor([[CAM["fx"], 0, CAM["cx"]], [0, CAM["fy"], CAM["cy"]], [0, 0, 1.0]]).to(DEV)}
    variables = {"means2D_gradient_accum": torch.zeros(P, device=DEV), "denom": torch.zeros(P, device=DEV),
                 "max_2D_radius": torch.zeros(P, device=DEV), "timestep": torch.zeros(P, device=DEV),
                 "scene_radius": torch.tensor(3.0)}
    return params, frames, variables


def _optimizer(cls, params, lrs, tracking):
    groups = [{"params": [v], "name": k, "lr": lrs[k]} for k, v in params.items()]
    return cls(groups) if tracking else cls(groups, lr=0.0, eps=1e-15)        # scripts/splatam.py:166-172


def _run_ours(record):
    from splatam_b200 import densify, slam
    base, frames, variables = _scene()
    params = {k: torch.nn.Parameter(v.clone()) for k, v in base.items()}
    for t in FRAMES:
        curr = frames[t]
        opt = _optimizer(slam.FusedAdam, params, TRACK_LRS, True)
        for _ in range(N_TRACK):
            loss, variables, _ = slam.get_loss(params, curr, variables, t, WTS, True, 0.99, True, False, tracking=True)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
        params, variables = densify.add_new_gaussians(params, variables, curr, 0.5, t, "projective", "isotropic")
        record.append(("grow", t, params["means3D"].shape[0]))
        opt = _optimizer(slam.FusedAdam, params, MAP_LRS, False)
        for it in range(N_MAP):
            loss, variables, _ = slam.get_loss(params, curr, variables, t, WTS, False, 0.5, True, False, mapping=True)
            loss.backward()
            with torch.no_grad():
                params, variables = densify.prune_gaussians(params, variables, opt, it, PRUNE)
                opt.step()
                opt.zero_grad(set_to_none=True)
            record.append(("map", t, it, params["means3D"].shape[0]))
    return params, opt


def _ref_get_loss(params, curr, variables, t, tracking):
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    gg, cg = (False, True) if tracking else (True, False)
    tg = SO.transform_to_frame(params, t, gg, cg)
    rv = SO.transformed_params2rendervar(params, tg)
    m2d = torch.zeros_like(params["means3D"], requires_grad=True)
    im, radius, _ = Renderer(raster_settings=curr["cam"])(means2D=m2d, **rv)
    rv2 = dict(rv)
    rv2["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], curr["w2c"])
    ds, _, _ = Renderer(raster_settings=curr["cam"])(means2D=torch.zeros_like(m2d, requires_grad=True), **rv2)
    loss, _ = SO.image_loss(im, ds, curr["im"], curr["depth"], WTS, tracking, 0.99 if tracking else 0.5, True, False, tracking)
    seen = radius > 0
    variables["max_2D_radius"][seen] = torch.max(radius[seen].float(), variables["max_2D_radius"][seen])
    return loss, ds.detach()


def _ref_surgery(fn, params, opt, variables):
    """Run a functional oracle surgery step on (params, Adam state, variables) and re-key the optimizer."""
    state = {}
    for k in DO.GAUSS_KEYS:
        st = opt.state.get(params[k]) if opt is not None else None
        if st:
            state[k] = {"exp_avg": st["exp_avg"], "exp_avg_sq": st["exp_avg_sq"]}
    plain = {k: v.detach() for k, v in params.items()}
    new_p, new_s, new_v = fn(plain, state, variables)
    out = {}
    for k, v in new_p.items():
        p = torch.nn.Parameter(v.clone())
        if opt is not None:
            grp = [g for g in opt.param_groups if g["name"] == k][0]
            old = grp["params"][0]
            st = opt.state.pop(old, None)
            if st is not None:
                if k in new_s:
                    st["exp_avg"], st["exp_avg_sq"] = new_s[k]["exp_avg"].clone(), new_s[k]["exp_avg_sq"].clone()
                opt.state[p] = st
            grp["params"][0] = p
            p.grad = None
        out[k] = p
    return out, new_v


def _run_reference_flow(record):
    base, frames, variables = _scene()
    params = {k: torch.nn.Parameter(v.clone()) for k, v in base.items()}
    for t in FRAMES:
        curr = frames[t]
        opt = _optimizer(torch.optim.Adam, params, TRACK_LRS, True)
        for _ in range(N_TRACK):
            loss, _ = _ref_get_loss(params, curr, variables, t, True)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
        with torch.no_grad():
            _, ds = _ref_get_loss(params, curr, dict(variables, max_2D_radius=variables["max_2D_radius"].clone()), t, False)
        params, variables = _ref_surgery(
            lambda p, s, v: DO.add_new_gaussians(p, s, v, ds, curr["im"], curr["depth"], curr["intrinsics"], 0.5, t),
            params, None, variables)
        record.append(("grow", t, params["means3D"].shape[0]))
        opt = _optimizer(torch.optim.Adam, params, MAP_LRS, False)
        for it in range(N_MAP):
            loss, _ = _ref_get_loss(params, curr, variables, t, False)
            loss.backward()
            grads = {k: v.grad for k, v in params.items()}
            with torch.no_grad():
                keep = None
                mask = DO.prune_mask({k: v.detach() for k, v in params.items()}, variables, it, PRUNE)
                if mask is not None:
                    keep = ~mask
                params, variables = _ref_surgery(lambda p, s, v: DO.prune_gaussians(p, s, v, it, PRUNE), params, opt, variables)
                for k, v in params.items():      # the reference drops .grad with the old Parameter objects
                    g = grads[k]
                    if g is not None and keep is not None and k in DO.GAUSS_KEYS:
                        g = None
                    v.grad = g
                opt.step()
                opt.zero_grad(set_to_none=True)
            record.append(("map", t, it, params["means3D"].shape[0]))
    return params, opt


def test_two_frame_slam_sequence_matches_reference_flow():
    rec_a, rec_b = [], []
    pa, opt_a = _run_ours(rec_a)
    pb, opt_b = _run_reference_flow(rec_b)
    assert rec_a == rec_b, (rec_a, rec_b)                     # same map size after every grow / prune step
    assert rec_a[0][2] > 2500 and rec_a[-1][3] < rec_a[0][2] + 10000
    # Isotropic Gaussians: the render does not depend on unnorm_rotations, their gradient is rounding noise (~1e-11)
    # and Adam with eps=1e-15 turns noise into +-lr steps -- a random walk in BOTH flows (and in the reference).
    skip = ("unnorm_rotations",)
    for k in pa:
        if k in skip:
            assert pa[k].shape == pb[k].shape
            continue
        a, b = pa[k].detach().double(), pb[k].detach().double()
        scale = float(b.abs().max()) + 1e-30
        assert float((a - b).abs().max()) <= 2e-3 * scale, (k, float((a - b).abs().max()), scale)
    for k in DO.GAUSS_KEYS:
        if k in skip:
            continue
        sa, sb = opt_a.state[pa[k]], opt_b.state[pb[k]]
        assert float(sa["step"]) == float(sb["step"])
        for m in ("exp_avg", "exp_avg_sq"):
            a, b = sa[m].double(), sb[m].double()
            scale = float(b.abs().max()) + 1e-30
            assert float((a - b).abs().max()) <= 5e-3 * scale, (k, m, float((a - b).abs().max()), scale)


def test_sequence_bench_short_run_tracks_the_trajectory():
    """bench_sequence.py (C5) for a few frames with a short schedule: the fused loop runs end to end (graph-replayed
    tracking, arena growth, pruning) and the estimated trajectory stays on the ground truth."""
    import bench_sequence as BS
    dev = torch.device(DEV)
    rs, frames, intr, gt_poses = BS.make_sequence(200, 400_000, dev)
    n = 6
    arm = BS.OursArm(dev, "dg")
    params, ms = BS.run(arm, rs, frames[:n], intr, n, 40, 10, 5, 20, dev)
    assert params["means3D"].shape[0] >= 0.9 * 640 * 480 * 0.5
    assert BS.ate_rmse(params, gt_poses, n) < 5e-3
    assert ms["tracking"] > 0 and ms["mapping"] > 0
