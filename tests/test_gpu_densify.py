"""GPU parity of the per-frame map surgery through the C ABI:
  * prune_gaussians / remove_points / opacity reset, Adam moments included, vs the golden fixtures produced by the
    reference's own utils/slam_external.py on a real torch.optim.Adam  -- BIT-EXACT (rows are moved, not recomputed);
  * add_new_gaussians vs the golden fixtures produced by the reference's own scripts/splatam.py functions: row count,
    colours, rotations, opacities, variables exact; back-projected means / log-scales within 1e-5 (float32 arithmetic
    in a different association than torch.inverse + matmul);
  * the same at full size (1M Gaussians, 1200x680) vs the pinned oracle, and size-independent properties: pruning twice
    changes nothing, add -> prune -> step keeps the optimizer usable, no reallocation once the arena is warm;
  * the radix-select median vs torch.median on ties, even / odd sizes and NaN."""
import os

import numpy as np
import pytest
import torch

from util import GOLDEN_DIR
from oracle import densify_oracle as DO

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _same(a, b, what=""):
    a, b = torch.as_tensor(a).cpu(), torch.as_tensor(b).cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert torch.equal(a, b), what


def _close(a, b, tol, what=""):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    scale = float(b.abs().max()) + 1e-30
    err = float((a - b).abs().max())
    assert err <= tol * scale, "%s: max abs err %.3g vs scale %.3g" % (what, err, scale)


def _setup(P, iso, seed, opt_kind="torch", state=None):
    from splatam_b200 import slam
    p0, s0, v0 = state or DO.random_state(P, iso, seed)
    params = {k: torch.nn.Parameter(v.to(DEV)) for k, v in p0.items()}
    groups = [{"params": [v], "name": k, "lr": 1e-3} for k, v in params.items()]
    opt = (slam.FusedAdam if opt_kind == "fused" else torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
    for k, v in params.items():
        if k in s0:
            opt.state[v] = {"step": torch.tensor(5.0), "exp_avg": s0[k]["exp_avg"].to(DEV),
                            "exp_avg_sq": s0[k]["exp_avg_sq"].to(DEV)}
    variables = {k: (v.to(DEV) if k != "scene_radius" else v) for k, v in v0.items()}
    return params, opt, variables, (p0, s0, v0)


@pytest.mark.parametrize("opt_kind", ["torch", "fused"])
@pytest.mark.parametrize("case", DO.PRUNE_CASES, ids=lambda c: c["name"])
def test_prune_vs_reference_python_golden(case, opt_kind):
    from splatam_b200 import densify
    gold = np.load(os.path.join(GOLDEN_DIR, "densify_prune_%s.npz" % case["name"]))
    params, opt, variables, _ = _setup(case["P"], case["iso"], case["seed"], opt_kind)
    params, variables = densify.prune_gaussians(params, variables, opt, case["iter"], case["pd"])
    for k in DO.GAUSS_KEYS:
        assert isinstance(params[k], torch.nn.Parameter) and params[k].requires_grad
        assert [g for g in opt.param_groups if g["name"] == k][0]["params"][0] is params[k]
        _same(params[k].detach(), gold["p_" + k], "param " + k)
        st = opt.state[params[k]]
        _same(st["exp_avg"], gold["m_" + k], "exp_avg " + k)
        _same(st["exp_avg_sq"], gold["v_" + k], "exp_avg_sq " + k)
        assert float(st["step"]) == 5.0
    for k in DO.VAR_KEYS:
        _same(variables[k], gold["var_" + k], "variable " + k)


@pytest.mark.parametrize("case", DO.ADD_CASES, ids=lambda c: c["name"])
def test_add_new_gaussians_vs_reference_python_golden(case):
    from splatam_b200 import densify
    gold = np.load(os.path.join(GOLDEN_DIR, "densify_add_%s.npz" % case["name"]))
    params, opt, variables, _ = _setup(case["P"], case["iso"], case["seed"], "fused")
    im, gt_depth, depth_sil, intr = DO.frame(case["h"], case["w"], case["seed"])
    curr = {"cam": None, "im": im.to(DEV), "depth": gt_depth.to(DEV), "intrinsics": intr.to(DEV)}
    params, variables = densify.add_new_gaussians(params, variables, curr, case["sil_thres"], case["time_idx"],
                                                  "projective", "isotropic" if case["iso"] else "anisotropic",
                                                  optimizer=opt, depth_sil=depth_sil.to(DEV))
    P = case["P"]
    for k in DO.GAUSS_KEYS:
        assert params[k].shape == gold["p_" + k].shape, k
        _same(params[k].detach()[:P], gold["p_" + k][:P], "old rows " + k)
        st = opt.state[params[k]]
        assert st["exp_avg"].shape == params[k].shape and float(st["exp_avg"][P:].abs().max()) == 0.0
        assert float(st["exp_avg_sq"][P:].abs().max()) == 0.0
    for k in ("rgb_colors", "unnorm_rotations", "logit_opacities"):
        _same(params[k].detach(), gold["p_" + k], k)
    _close(params["means3D"].detach(), gold["p_means3D"], 1e-5, "means3D")
    _close(params["log_scales"].detach(), gold["p_log_scales"], 1e-5, "log_scales")
    for k in DO.VAR_KEYS:
        _same(variables[k], gold["var_" + k], "variable " + k)


def test_full_size_prune_vs_oracle_and_idempotence():
    from splatam_b200 import densify
    P = 1_000_000
    pd = dict(DO.PRUNE_DICT, removal_opacity_threshold=0.4)
    p0, s0, v0 = DO.random_state(P, True, 77)
    # mask parity needs no value within rounding of a threshold (sigmoid / exp differ by an ulp between CPU and GPU)
    op = torch.sigmoid(p0["logit_opacities"].double()).squeeze()
    sc = torch.exp(p0["log_scales"].double()).max(dim=1).values
    tie = ((op - 0.4).abs() < 1e-6) | ((sc - 0.2).abs() < 1e-6)
    p0["logit_opacities"][tie] = 3.0
    p0["log_scales"][tie] = -3.0
    params, opt, variables, _ = _setup(P, True, 77, "fused", state=(p0, s0, v0))
    ref_p, ref_s, ref_v = DO.prune_gaussians(p0, s0, v0, 0, pd)
    params, variables = densify.prune_gaussians(params, variables, opt, 0, pd)
    assert 0 < params["means3D"].shape[0] < P
    for k in DO.GAUSS_KEYS:
        _same(params[k].detach(), ref_p[k], "param " + k)
        _same(opt.state[params[k]]["exp_avg"], ref_s[k]["exp_avg"], "exp_avg " + k)
        _same(opt.state[params[k]]["exp_avg_sq"], ref_s[k]["exp_avg_sq"], "exp_avg_sq " + k)
    for k in DO.VAR_KEYS:
        _same(variables[k], ref_v[k], "variable " + k)
    # idempotence + the arena is warm: a second prune keeps everything and allocates nothing
    arena = variables[densify.ARENA_KEY]
    n_alloc, P1 = arena.reallocations, params["means3D"].shape[0]
    before = {k: params[k].detach().clone() for k in DO.GAUSS_KEYS}
    params, variables = densify.prune_gaussians(params, variables, opt, 0, pd)
    assert params["means3D"].shape[0] == P1 and arena.reallocations == n_alloc
    for k in DO.GAUSS_KEYS:
        _same(params[k].detach(), before[k], "second prune " + k)
    # the optimizer still steps on the surviving map
    for k in DO.GAUSS_KEYS:
        params[k].grad = torch.ones_like(params[k])
    opt.step()
    assert float((params["means3D"].detach() - before["means3D"]).abs().max()) > 0


def test_full_size_add_vs_oracle_then_prune_chain():
    from splatam_b200 import densify
    P, h, w = 300_000, 680, 1200
    params, opt, variables, (p0, s0, v0) = _setup(P, True, 78, "fused")
    im, gt_depth, depth_sil, intr = DO.frame(h, w, 5)
    ref_p, ref_s, ref_v = DO.add_new_gaussians(p0, s0, v0, depth_sil, im, gt_depth, intr, 0.5, 2)
    curr = {"cam": None, "im": im.to(DEV), "depth": gt_depth.to(DEV), "intrinsics": intr}
    params, variables = densify.add_new_gaussians(params, variables, curr, 0.5, 2, "projective", "isotropic",
                                                  optimizer=opt, depth_sil=depth_sil.to(DEV))
    assert params["means3D"].shape[0] == ref_p["means3D"].shape[0] > P
    for k in ("rgb_colors", "unnorm_rotations", "logit_opacities"):
        _same(params[k].detach(), ref_p[k], k)
    _close(params["means3D"].detach(), ref_p["means3D"], 1e-5, "means3D")
    _close(params["log_scales"].detach(), ref_p["log_scales"], 1e-5, "log_scales")
    for k in DO.VAR_KEYS:
        _same(variables[k], ref_v[k], "variable " + k)
    # chain: a step, a prune, another frame -- shapes stay consistent, the optimizer state follows the map
    for k in DO.GAUSS_KEYS:
        params[k].grad = torch.full_like(params[k], 0.1)
    opt.step()
    pd = dict(DO.PRUNE_DICT, removal_opacity_threshold=0.3)
    params, variables = densify.prune_gaussians(params, variables, opt, 0, pd)
    n1 = params["means3D"].shape[0]
    assert float(torch.sigmoid(params["logit_opacities"].detach()).min()) >= 0.3 - 1e-6
    params, variables = densify.add_new_gaussians(params, variables, curr, 0.5, 3, "projective", "isotropic",
                                                  optimizer=opt, depth_sil=depth_sil.to(DEV))
    n2 = params["means3D"].shape[0]
    assert n2 - n1 == ref_p["means3D"].shape[0] - P          # same frame, same mask -> same number of new Gaussians
    for k in DO.GAUSS_KEYS:
        assert params[k].shape[0] == n2 and opt.state[params[k]]["exp_avg"].shape == params[k].shape
    for k in DO.VAR_KEYS:
        assert variables[k].shape[0] == n2
    assert float(variables["timestep"][n1:].min()) == 3.0 == float(variables["timestep"][n1:].max())


@pytest.mark.parametrize("n,mode", [(4097, "plain"), (4096, "plain"), (70001, "ties"), (5000, "nan"), (3, "plain"), (1, "plain")])
def test_new_gaussian_mask_and_median_vs_torch(n, mode):
    from splatam_b200 import densify
    g = torch.Generator().manual_seed(n)
    h, w = 1, n
    gt = 2.0 + torch.rand(h, w, generator=g)
    rd = gt + 0.05 * torch.randn(h, w, generator=g)
    rd[torch.rand(h, w, generator=g) < 0.1] += 4.0
    if mode == "ties":
        gt[torch.rand(h, w, generator=g) < 0.6] = 0.0        # more than half the errors are exactly 0
        rd = (rd * 64).round() / 64
        gt = (gt * 64).round() / 64
    if mode == "nan":
        rd[0, 17] = float("nan")
    sil = torch.rand(h, w, generator=g)
    depth_sil = torch.stack([rd, sil, rd * rd], 0)
    pre, post = DO.non_presence_mask(depth_sil, gt[None], 0.2)
    flags, n_pre, n_post = densify.new_gaussian_mask(depth_sil.to(DEV), gt[None].to(DEV), 0.2)
    assert n_pre == int(pre.sum()) and n_post == int(post.sum())
    _same(flags.bool(), post, "mask")


def test_remove_points_edge_cases():
    from splatam_b200 import densify
    params, opt, variables, (p0, s0, v0) = _setup(2500, False, 79, "torch")
    none = torch.zeros(2500, dtype=torch.bool, device=DEV)
    params, variables = densify.remove_points(none, params, variables, opt)
    for k in DO.GAUSS_KEYS:
        _same(params[k].detach(), p0[k], "nothing removed " + k)
    some = torch.zeros(2500, dtype=torch.bool, device=DEV)
    some[::3] = True
    some[1024:2048] = True
    ref_p, ref_s, ref_v = DO.remove_points(some.cpu(), p0, s0, v0)
    params, variables = densify.remove_points(some, params, variables, opt)
    for k in DO.GAUSS_KEYS:
        _same(params[k].detach(), ref_p[k], "some removed " + k)
        _same(opt.state[params[k]]["exp_avg_sq"], ref_s[k]["exp_avg_sq"], "exp_avg_sq " + k)
    everything = torch.ones(params["means3D"].shape[0], dtype=torch.bool, device=DEV)
    params, variables = densify.remove_points(everything, params, variables, opt)
    for k in DO.GAUSS_KEYS:
        assert params[k].shape[0] == 0
    assert variables["timestep"].shape[0] == 0


def test_add_new_gaussians_with_own_silhouette_render_vs_reference_flow(tmp_path):
    """The default path (no depth_sil given): builder -> six-plane render -> mask -> spawn, vs the reference flow
    (oracle builder, a three-channel silhouette render through the drop-in rasterizer, oracle add_new_gaussians).
    Then the reference's params.npz format (utils/common_utils.py:25-42) round-trips from the arena-backed tensors."""
    from splatam_b200 import densify, synth
    from oracle import slam_oracle as SO
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    cam = dict(w=160, h=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5)
    w2c0 = torch.eye(4)
    rs = synth.make_settings(device=DEV, w2c=w2c0, **cam)
    P = 3000                                   # a sparse map: part of the frame is not covered yet
    sc = synth.surface_scene(P, cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], seed=9)
    g = torch.Generator().manual_seed(9)
    nf = 3
    rots = torch.tensor([1.0, 0, 0, 0]).view(1, 4, 1).repeat(1, 1, nf) + 0.01 * torch.randn(1, 4, nf, generator=g)
    base = {"means3D": sc["means3D"], "rgb_colors": sc["colors_precomp"], "unnorm_rotations": sc["rotations"] * 1.7,
            "logit_opacities": torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)), "log_scales": torch.log(sc["scales"][:, :1]),
            "cam_unnorm_rots": rots, "cam_trans": 0.01 * torch.randn(1, 3, nf, generator=g)}
    base = {k: v.contiguous().to(DEV) for k, v in base.items()}
    gt_im = torch.rand(3, cam["h"], cam["w"], generator=g).to(DEV)
    gt_d = (2.5 + 0.5 * torch.rand(1, cam["h"], cam["w"], generator=g)).to(DEV)
    gt_d[0, :5] = 0.0
    intr = torch.tensor([[cam["fx"], 0, cam["cx"]], [0, cam["fy"], cam["cy"]], [0, 0, 1.0]])
    curr = {"cam": rs, "im": gt_im, "depth": gt_d, "w2c": w2c0.to(DEV), "intrinsics": intr.to(DEV)}
    variables = {"means2D_gradient_accum": torch.zeros(P, device=DEV), "denom": torch.zeros(P, device=DEV),
                 "max_2D_radius": torch.zeros(P, device=DEV), "timestep": torch.zeros(P, device=DEV)}
    # reference flow
    with torch.no_grad():
        tg = SO.transform_to_frame(base, 1, False, False)
        rv = SO.transformed_params2rendervar(base, tg)
        rv["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], curr["w2c"])
        ds, _, _ = Renderer(raster_settings=rs)(means2D=torch.zeros_like(base["means3D"]), **rv)
    pre, post = DO.non_presence_mask(ds, gt_d, 0.5)
    err = (gt_d[0] - ds[0]).abs() * (gt_d[0] > 0)
    margin = ((ds[1] - 0.5).abs() < 1e-4) | (((err - 50 * err.median()).abs() < 1e-4) & (ds[0] > gt_d[0]))
    ref_p, _, ref_v = DO.add_new_gaussians(base, {}, variables, ds, gt_im, gt_d, intr.to(DEV), 0.5, 1)
    # ours
    params = {k: torch.nn.Parameter(v.clone()) for k, v in base.items()}
    params, variables = densify.add_new_gaussians(params, dict(variables), curr, 0.5, 1, "projective", "isotropic")
    n_ref, n = ref_p["means3D"].shape[0], params["means3D"].shape[0]
    assert n_ref > P + 100
    # the two silhouette renders agree to ~1e-6, so only pixels sitting on a threshold may flip
    assert abs(n - n_ref) <= int(margin.sum())
    if n == n_ref:
        _close(params["means3D"].detach(), ref_p["means3D"], 1e-5, "means3D")
        _same(params["rgb_colors"].detach(), ref_p["rgb_colors"], "rgb")
        _close(params["log_scales"].detach(), ref_p["log_scales"], 1e-5, "log_scales")
        _same(variables["timestep"], ref_v["timestep"], "timestep")
    assert float(variables["denom"].abs().max()) == 0.0 and variables["denom"].shape[0] == n
    # params.npz: the reference's own conversion works on arena views
    to_save = {k: v.detach().cpu().contiguous().numpy() for k, v in params.items()}
    np.savez(os.path.join(tmp_path, "params.npz"), **to_save)
    back = dict(np.load(os.path.join(tmp_path, "params.npz")))
    for k in params:
        assert back[k].shape == tuple(params[k].shape) and np.array_equal(back[k], to_save[k])


def test_torch_adam_steps_on_arena_views_like_on_plain_tensors():
    """torch.optim.Adam keeps working on what prune hands back (Parameters and moments that are VIEWS of the arena):
    two steps after a prune give bit-identical parameters and moments to the same steps on the reference flow's
    freshly allocated tensors."""
    from splatam_b200 import densify
    pd = dict(DO.PRUNE_DICT, removal_opacity_threshold=0.3)
    params, opt, variables, (p0, s0, v0) = _setup(4000, False, 81, "torch")
    params, variables = densify.prune_gaussians(params, variables, opt, 0, pd)
    rp, rs, _ = DO.prune_gaussians({k: v.to(DEV) for k, v in p0.items()}, {k: {n: t.to(DEV) for n, t in d.items()} for k, d in s0.items()},
                                   {k: (v.to(DEV) if k != "scene_radius" else v) for k, v in v0.items()}, 0, pd)
    ref = {k: torch.nn.Parameter(v.clone()) for k, v in rp.items()}
    ropt = torch.optim.Adam([{"params": [v], "name": k, "lr": 1e-3} for k, v in ref.items()], lr=0.0, eps=1e-15)
    for k in DO.GAUSS_KEYS:
        ropt.state[ref[k]] = {"step": torch.tensor(5.0), "exp_avg": rs[k]["exp_avg"].clone(), "exp_avg_sq": rs[k]["exp_avg_sq"].clone()}
    g = torch.Generator().manual_seed(5)
    for _ in range(2):
        for k in DO.GAUSS_KEYS:
            grad = torch.randn(ref[k].shape, generator=g).to(DEV)
            params[k].grad, ref[k].grad = grad.clone(), grad.clone()
        opt.step()
        ropt.step()
    for k in DO.GAUSS_KEYS:
        _same(params[k].detach(), ref[k].detach(), "param " + k)
        _same(opt.state[params[k]]["exp_avg"], ropt.state[ref[k]]["exp_avg"], "exp_avg " + k)
        _same(opt.state[params[k]]["exp_avg_sq"], ropt.state[ref[k]]["exp_avg_sq"], "exp_avg_sq " + k)
        assert float(opt.state[params[k]]["step"]) == 7.0
