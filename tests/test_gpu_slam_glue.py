"""GPU parity of the fused SplaTAM glue through the C ABI:
  * render-variable builder fwd + bwd (incl. the in-kernel pose-gradient reduction) vs the golden fixtures produced by
    the reference's own utils/slam_helpers.py, and vs the torch oracle at full size (1M Gaussians);
  * tracking / mapping image loss + gradient vs the golden fixtures produced by the reference's own get_loss, and vs
    the torch oracle at 1200x680;
  * fused get_loss (builder -> six-plane render -> loss -> backward) vs the reference flow (oracle glue, two
    three-channel renders) on pose and Gaussian gradients.
Tolerances: forward 1e-5 relative to the array's max; gradients 1e-3 (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

from util import GOLDEN_DIR
from oracle import slam_oracle as SO

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _close(a, b, tol, what="", atol=0.0):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    scale = float(b.abs().max()) + 1e-30
    err = float((a - b).abs().max())
    assert err <= tol * scale + atol, "%s: max abs err %.3g vs scale %.3g (tol %.1g)" % (what, err, scale, tol)


def _params(case, n_frames=5, dev=DEV, P=None):
    p = SO.random_params(P or case["P"], n_frames, case["iso"], case["seed"])
    return {k: v.to(dev).requires_grad_(True) for k, v in p.items()}


@pytest.mark.parametrize("case", SO.RV_CASES, ids=lambda c: c["name"])
def test_rendervar_vs_reference_python_golden(case):
    from splatam_b200 import slam
    gold = np.load(os.path.join(GOLDEN_DIR, "slam_glue_rv_%s.npz" % case["name"]))
    params = _params(case)
    rv = slam.build_rendervar(params, 3, case["gg"], case["cg"])
    ds = SO.get_depth_and_silhouette(rv["means3D"], torch.from_numpy(gold["w2c"]).to(DEV))
    g = torch.Generator().manual_seed(100 + case["seed"])
    seeds = {k: torch.randn(rv[k].shape, generator=g).to(DEV) for k in ("means3D", "rotations", "opacities", "scales")}
    seeds["depth_sil"] = torch.randn(ds.shape, generator=g).to(DEV)
    scalar = sum((rv[k] * seeds[k]).sum() for k in seeds if k != "depth_sil") + (ds * seeds["depth_sil"]).sum()
    scalar.backward()
    for k in ("means3D", "rotations", "opacities", "scales"):
        _close(rv[k].detach(), gold["rv_" + k], 1e-5, "rv " + k)
    assert rv["colors_precomp"] is params["rgb_colors"]
    for k, v in params.items():
        if k == "rgb_colors":
            continue
        got = v.grad if v.grad is not None else torch.zeros_like(v)
        _close(got, gold["g_" + k], 1e-4, "grad " + k)


@pytest.mark.parametrize("iso,gg,cg", [(True, False, True), (True, True, False), (False, True, True)])
def test_rendervar_full_size_vs_oracle(iso, gg, cg):
    """1M Gaussians: kernel vs the torch oracle run on the GPU (same ops as the reference, torch autograd)."""
    from splatam_b200 import slam
    case = dict(P=1_000_003, iso=iso, seed=21)
    pa = _params(case, n_frames=7)
    pb = {k: v.detach().clone().requires_grad_(True) for k, v in pa.items()}
    ra = slam.build_rendervar(pa, 4, gg, cg)
    rb = SO.build_rendervar(pb, 4, gg, cg)
    g = torch.Generator().manual_seed(5)
    tot_a = tot_b = 0.0
    for k in ("means3D", "rotations", "opacities", "scales"):
        _close(ra[k].detach(), rb[k].detach(), 1e-5, k)
        w = torch.randn(ra[k].shape, generator=g).to(DEV)
        tot_a = tot_a + (ra[k] * w).sum()
        tot_b = tot_b + (rb[k] * w).sum()
    tot_a.backward()
    tot_b.backward()
    for k in pa:
        if k == "rgb_colors":
            continue
        ga = pa[k].grad if pa[k].grad is not None else torch.zeros_like(pa[k])
        gb = pb[k].grad if pb[k].grad is not None else torch.zeros_like(pb[k])
        # the pose gradient is a sum of 1M terms: torch reduces it in float32 in another order
        _close(ga, gb, 1e-3 if k.startswith("cam_") else 1e-4, "grad " + k)
    # deterministic (two-stage reduction, no atomics)
    pc = {k: v.detach().clone().requires_grad_(True) for k, v in pa.items()}
    rc = slam.build_rendervar(pc, 4, gg, cg)
    g = torch.Generator().manual_seed(5)
    tot_c = 0.0
    for k in ("means3D", "rotations", "opacities", "scales"):
        tot_c = tot_c + (rc[k] * torch.randn(rc[k].shape, generator=g).to(DEV)).sum()
    tot_c.backward()
    for k in pa:
        if pa[k].grad is not None:
            assert torch.equal(pa[k].grad, pc[k].grad), k


def test_rendervar_empty_and_single():
    from splatam_b200 import slam
    for P in (0, 1):
        pa = _params(dict(P=P, iso=True, seed=3), P=P)
        rv = slam.build_rendervar(pa, 0, True, True)
        assert rv["means3D"].shape == (P, 3) and rv["scales"].shape == (P, 3)
        (rv["means3D"].sum() + rv["opacities"].sum()).backward()
        assert pa["cam_trans"].grad is not None and pa["cam_trans"].grad.shape == pa["cam_trans"].shape
        if P == 0:
            assert float(pa["cam_trans"].grad.abs().max()) == 0.0


@pytest.mark.parametrize("case", SO.LOSS_CASES, ids=lambda c: c["name"])
@pytest.mark.parametrize("six", [True, False], ids=["fused6", "two3"])
def test_image_loss_vs_reference_get_loss_golden(case, six):
    from splatam_b200 import slam
    gold = np.load(os.path.join(GOLDEN_DIR, "slam_glue_loss_%s.npz" % case["name"]))
    im, ds, gt_im, gt_d = [t.to(DEV) for t in SO.images(case["h"], case["w_"], case["seed"])]
    if six:
        r6 = torch.cat([im, ds], 0).requires_grad_(True)
        loss, wl = slam.image_loss(r6, None, gt_im, gt_d, case["w"], case["use_sil"], case["sil_thres"], case["tracking"])
        loss.backward()
        g_im, g_ds = r6.grad[:3], r6.grad[3:]
    else:
        im.requires_grad_(True)
        ds.requires_grad_(True)
        loss, wl = slam.image_loss(im, ds, gt_im, gt_d, case["w"], case["use_sil"], case["sil_thres"], case["tracking"])
        (loss * 1.0).backward()
        g_im, g_ds = im.grad, ds.grad
    _close(loss.detach(), gold["loss"], 1e-5, "loss")
    _close(wl["im"], gold["l_im"], 1e-5, "im loss")
    _close(wl["depth"], gold["l_depth"], 1e-5, "depth loss")
    _close(g_im, gold["g_im"], 1e-3, "dL/dim")
    _close(g_ds, gold["g_depth_sil"], 1e-3, "dL/ddepth_sil")


@pytest.mark.parametrize("tracking", [True, False], ids=["tracking", "mapping"])
def test_image_loss_full_size_vs_oracle(tracking):
    from splatam_b200 import slam
    h, w = 680, 1200
    im, ds, gt_im, gt_d = [t.to(DEV) for t in SO.images(h, w, 9)]
    wts = dict(im=0.5, depth=1.0)
    r6 = torch.cat([im, ds], 0).requires_grad_(True)
    loss, wl = slam.image_loss(r6, None, gt_im, gt_d, wts, tracking, 0.99, tracking)
    (loss * 3.0).backward()           # a non-unit incoming gradient must scale the stored one
    im_o, ds_o = im.clone().requires_grad_(True), ds.clone().requires_grad_(True)
    loss_o, wl_o = SO.image_loss(im_o, ds_o, gt_im, gt_d, wts, tracking, 0.99, True, False, tracking)
    (loss_o * 3.0).backward()
    _close(loss.detach(), loss_o.detach(), 2e-5, "loss")
    _close(wl["im"], wl_o["im"].detach(), 2e-5)
    _close(wl["depth"], wl_o["depth"].detach(), 2e-5)
    _close(r6.grad[:3], im_o.grad, 1e-3, "dL/dim")
    _close(r6.grad[3:], ds_o.grad, 1e-3, "dL/ddepth_sil")
    assert float(r6.grad[4:].abs().max()) == 0.0


def _slam_scene(P, tracking, seed=4):
    from splatam_b200 import synth
    cam = dict(w=160, h=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5)
    w2c0 = torch.eye(4)
    rs = synth.make_settings(device=DEV, w2c=w2c0, **cam)
    sc = synth.surface_scene(P, cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], seed=seed)
    g = torch.Generator().manual_seed(seed)
    nf = 3
    rots = torch.tensor([1.0, 0, 0, 0]).view(1, 4, 1).repeat(1, 1, nf) + 0.01 * torch.randn(1, 4, nf, generator=g)
    params = {
        "means3D": sc["means3D"], "rgb_colors": sc["colors_precomp"],
        "unnorm_rotations": sc["rotations"] * 1.7,
        "logit_opacities": torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)),
        "log_scales": torch.log(sc["scales"][:, :1]),
        "cam_unnorm_rots": rots, "cam_trans": 0.01 * torch.randn(1, 3, nf, generator=g),
    }
    params = {k: v.contiguous().to(DEV) for k, v in params.items()}
    gt_im = torch.rand(3, cam["h"], cam["w"], generator=g).to(DEV)
    gt_d = (2.5 + 0.5 * torch.rand(1, cam["h"], cam["w"], generator=g)).to(DEV)
    gt_d[0, :5] = 0.0
    curr = {"cam": rs, "im": gt_im, "depth": gt_d, "w2c": w2c0.to(DEV)}
    return params, curr


@pytest.mark.parametrize("mode", ["tracking", "mapping", "ba"])
def test_fused_get_loss_vs_reference_flow(mode):
    """Whole iteration: fused get_loss vs (oracle builder -> two 3-channel renders -> oracle loss -> torch autograd)."""
    from splatam_b200 import slam
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    tracking = mode == "tracking"
    base, curr = _slam_scene(6000, tracking)
    wts = dict(im=0.5, depth=1.0)
    sil_thres = 0.99 if tracking else 0.5
    # ---- fused
    pa = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    va = {"max_2D_radius": torch.zeros(6000, device=DEV)}
    loss_a, va, wl_a = slam.get_loss(pa, curr, va, 1, wts, tracking, sil_thres, True, False, tracking=tracking,
                                     mapping=not tracking, do_ba=mode == "ba")
    loss_a.backward()
    # ---- reference flow (scripts/splatam.py:226-296 with the oracle's glue)
    pb = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    gg, cg = (False, True) if tracking else ((True, True) if mode == "ba" else (True, False))
    tg = SO.transform_to_frame(pb, 1, gg, cg)
    rv = SO.transformed_params2rendervar(pb, tg)
    m2d = torch.zeros_like(pb["means3D"], requires_grad=True)
    im, radius, _ = Renderer(raster_settings=curr["cam"])(means2D=m2d, **rv)
    rv2 = dict(rv)
    rv2["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], curr["w2c"])
    ds, _, _ = Renderer(raster_settings=curr["cam"])(means2D=torch.zeros_like(m2d, requires_grad=True), **rv2)
    loss_b, wl_b = SO.image_loss(im, ds, curr["im"], curr["depth"], wts, tracking, sil_thres, True, False, tracking)
    loss_b.backward()
    _close(loss_a.detach(), loss_b.detach(), 1e-4, "loss")
    _close(wl_a["im"], wl_b["im"].detach(), 1e-4, "im")
    _close(wl_a["depth"], wl_b["depth"].detach(), 1e-4, "depth")
    assert torch.equal(va["seen"], radius > 0)
    assert torch.equal(va["max_2D_radius"], radius.float())
    for k in pa:
        ga = pa[k].grad if pa[k].grad is not None else torch.zeros_like(pa[k])
        gb = pb[k].grad if pb[k].grad is not None else torch.zeros_like(pb[k])
        # isotropic Gaussians: the rotation gradient is numerically zero in both flows (1e-11 of rounding noise)
        _close(ga, gb, 2e-3, "grad " + k, atol=1e-8 if k == "unnorm_rotations" else 0.0)
        if k.startswith("cam_"):
            assert (float(ga.abs().max()) > 0) == cg, k


def test_pose_only_tracking_matches_full_backward():
    """tracking_pose_only skips the Gaussian-side gradients (blend-backward MODE 3, no colour / opacity sums, no scale /
    rotation chain): the pose gradient and the loss must not move; the Gaussian .grad fields stay None."""
    from splatam_b200 import slam
    base, curr = _slam_scene(6000, True)
    wts = dict(im=0.5, depth=1.0)
    out = {}
    for lite in (False, True):
        p = {k: v.clone().requires_grad_(True) for k, v in base.items()}
        va = {"max_2D_radius": torch.zeros(6000, device=DEV)}
        loss, va, wl = slam.get_loss(p, curr, va, 1, wts, True, 0.99, True, False, tracking=True,
                                     tracking_pose_only=lite)
        loss.backward()
        out[lite] = (loss.detach(), p)
    _close(out[True][0], out[False][0], 1e-7, "loss")
    for k in ("cam_unnorm_rots", "cam_trans"):
        assert float(out[False][1][k].grad.abs().max()) > 0
        _close(out[True][1][k].grad, out[False][1][k].grad, 1e-4, "pose-only grad " + k)
    for k in ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales"):
        assert out[True][1][k].grad is None, k
    assert out[False][1]["rgb_colors"].grad is not None and out[False][1]["logit_opacities"].grad is not None


@pytest.mark.parametrize("six", [False, True], ids=["3ch", "6ch"])
def test_rasterizer_backward_skips_unneeded_gradients(six):
    """Only means3D requires a gradient: splat_backward gets NULL for every other array; dL/dmeans3D is unchanged."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import build_case
    rs, sc, dL = build_case("surface_small", "cuda")
    if six:
        g = torch.Generator().manual_seed(3)
        dL = torch.cat([dL, torch.randn(3, rs.image_height, rs.image_width, generator=g).to(DEV) * 0.3], 0)
    grads = {}
    for only_means in (False, True):
        leaves = {k: v.clone().requires_grad_(k == "means3D" or not only_means) for k, v in sc.items()}
        m2d = torch.zeros_like(leaves["means3D"], requires_grad=not only_means)
        r = GaussianRasterizer(rs)
        img, radii, _ = (r.render_rgbd_sil if six else r.forward)(means2D=m2d, **leaves)
        img.backward(dL)
        grads[only_means] = leaves
    _close(grads[True]["means3D"].grad, grads[False]["means3D"].grad, 1e-4, "dL/dmeans3D")
    for k in ("scales", "rotations", "opacities", "colors_precomp"):
        assert grads[True][k].grad is None and grads[False][k].grad is not None


@pytest.mark.parametrize("eps,lr0", [(1e-15, 0.0), (1e-8, 1e-3)], ids=["mapping", "tracking"])
def test_fused_adam_vs_torch_adam(eps, lr0):
    """splat_adam_step / slam.FusedAdam against torch.optim.Adam built the way initialize_optimizer does
    (scripts/splatam.py:166-172): ragged sizes (vector + scalar tails), a group without a gradient, several steps."""
    from splatam_b200 import slam
    g = torch.Generator().manual_seed(8)
    shapes = {"means3D": (100003, 3), "rgb_colors": (100003, 3), "unnorm_rotations": (100003, 4),
              "logit_opacities": (100003, 1), "log_scales": (100003, 1), "cam_unnorm_rots": (1, 4, 7), "cam_trans": (1, 3, 7)}
    lrs = dict(means3D=0.0001, rgb_colors=0.0025, unnorm_rotations=0.001, logit_opacities=0.05, log_scales=0.001,
               cam_unnorm_rots=0.0004, cam_trans=0.002)
    init = {k: torch.randn(*s, generator=g).to(DEV) for k, s in shapes.items()}
    pa = {k: v.clone().requires_grad_(True) for k, v in init.items()}
    pb = {k: v.clone().requires_grad_(True) for k, v in init.items()}
    mk = lambda cls, p: cls([{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in p.items()], lr=lr0, eps=eps)
    oa, ob = mk(slam.FusedAdam, pa), mk(torch.optim.Adam, pb)
    for step in range(12):
        for k in pa:
            if k == "log_scales" and step % 3 == 0:      # no gradient this step: its state must not advance
                pa[k].grad = pb[k].grad = None
                continue
            gr = torch.randn(*shapes[k], generator=g).to(DEV) * (10.0 ** (step % 4 - 2))
            pa[k].grad, pb[k].grad = gr.clone(), gr.clone()
        oa.step()
        ob.step()
    for k in pa:
        _close(pa[k].detach(), pb[k].detach(), 2e-6, "param " + k)
        sa, sb = oa.state[pa[k]], ob.state[pb[k]]
        assert float(sa["step"]) == float(sb["step"])
        _close(sa["exp_avg"], sb["exp_avg"], 2e-6, "exp_avg " + k)
        _close(sa["exp_avg_sq"], sb["exp_avg_sq"], 2e-6, "exp_avg_sq " + k)
    # the reference's optimizer surgery keeps working: cat_params_to_optimizer pattern (utils/slam_external.py:122-136)
    grp = [x for x in oa.param_groups if x["name"] == "means3D"][0]
    st = oa.state.get(grp["params"][0])
    extra = torch.zeros(5, 3, device=DEV)
    st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(extra)), 0)
    st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(extra)), 0)
    del oa.state[grp["params"][0]]
    grp["params"][0] = torch.nn.Parameter(torch.cat((grp["params"][0], extra), 0).requires_grad_(True))
    oa.state[grp["params"][0]] = st
    grp["params"][0].grad = torch.ones_like(grp["params"][0])
    oa.step()
    assert float(oa.state[grp["params"][0]]["step"]) == 13.0 and bool(torch.isfinite(grp["params"][0]).all())


def test_graphed_tracking_iterations_match_eager():
    """slam.GraphedIteration: 3 eager (warm-up) + 5 replayed, then 1 eager + 2 replayed tracking iterations (get_loss + backward + FusedAdam.step inside one
    CUDA graph) land on the same pose, optimizer state and step count as 11 eager iterations."""
    from splatam_b200 import slam
    base, curr = _slam_scene(6000, True)
    wts = dict(im=0.5, depth=1.0)
    lrs = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
               cam_unnorm_rots=0.0004, cam_trans=0.002)
    res = {}
    for graphed in (False, True):
        p = {k: v.clone().requires_grad_(True) for k, v in base.items()}
        va = {"max_2D_radius": torch.zeros(6000, device=DEV)}
        opt = slam.FusedAdam([{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in p.items()])
        best = slam.BestPose(p, 1)
        seen = []

        def it():
            loss, _, _ = slam.get_loss(p, curr, va, 1, wts, True, 0.99, True, False, tracking=True,
                                       tracking_pose_only=True)
            loss.backward()
            opt.step()
            best.update(loss, p)
            return loss

        if graphed:
            g = slam.GraphedIteration(it, p, opt, warmup=3)
            for _ in range(5):
                loss = g.replay()
            g.check()
            assert float(opt.state[p["cam_trans"]]["step"]) == 8.0
            g2 = slam.GraphedIteration(it, p, opt, warmup=1)     # a second capture with the same optimizer continues
            for _ in range(2):
                loss = g2.replay()
            g2.check()
        else:
            for _ in range(11):
                opt.zero_grad(set_to_none=True)
                loss = it()
        torch.cuda.synchronize()
        res[graphed] = (p, opt, float(loss), best)
    (pe, oe, le, be), (pg, og, lg, bg) = res[False], res[True]
    # device-side candidate-pose bookkeeping (scripts/splatam.py:726-731) agrees between the two runs
    assert float(be.min_loss) < 1e19 and abs(float(be.min_loss) - float(bg.min_loss)) <= 1e-4 * float(be.min_loss)
    _close(bg.rot, be.rot, 1e-5, "best rot")
    _close(bg.tran, be.tran, 1e-4, "best tran")
    assert abs(le - lg) <= 1e-4 * abs(le)
    for k in ("cam_unnorm_rots", "cam_trans"):
        assert float((pe[k] - base[k]).abs().max()) > 1e-4          # the pose did move
        _close(pg[k].detach(), pe[k].detach(), 1e-5, "pose after 11 iterations " + k)
        assert float(og.state[pg[k]]["step"]) == float(oe.state[pe[k]]["step"]) == 11.0
        _close(og.state[pg[k]]["exp_avg"], oe.state[pe[k]]["exp_avg"], 1e-3, "exp_avg " + k)
    assert torch.equal(pg["means3D"].detach(), base["means3D"])


def test_fused_mapping_iteration_writes_the_raw_parameter_bucket():
    """mapping.GradBucket(direct=True, layout=raw_layout()): the builder's backward and the rasterizer write every raw
    parameter's gradient straight into the flat exchange bucket; values equal the ordinary autograd gradients."""
    from splatam_b200 import mapping, slam
    base, curr = _slam_scene(6000, False)
    wts = dict(im=0.5, depth=1.0)
    pa = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    va = {"max_2D_radius": torch.zeros(6000, device=DEV)}
    loss_a, _, _ = slam.get_loss(pa, curr, va, 1, wts, False, 0.5, True, False, mapping=True)
    loss_a.backward()
    bk = mapping.GradBucket({k: base[k] for k, _ in mapping.raw_layout(1)}, DEV, direct=True, layout=mapping.raw_layout(1))
    bk.grads.fill_(float("nan"))            # every row must be overwritten
    pb = dict(bk.leaves(), cam_unnorm_rots=base["cam_unnorm_rots"].clone().requires_grad_(True),
              cam_trans=base["cam_trans"].clone().requires_grad_(True))
    vb = {"max_2D_radius": torch.zeros(6000, device=DEV)}
    loss_b, _, _ = slam.get_loss(pb, curr, vb, 1, wts, False, 0.5, True, False, mapping=True)
    loss_b.backward()
    _close(loss_b.detach(), loss_a.detach(), 1e-6, "loss")
    assert bool(torch.isfinite(bk.grads).all())
    for k, _ in mapping.raw_layout(1):
        assert pb[k].grad.data_ptr() == bk.gviews[k].data_ptr()
        _close(bk.gviews[k], pa[k].grad, 1e-3, "bucket grad " + k, atol=1e-8 if k == "unnorm_rotations" else 0.0)
    opt = slam.FusedAdam([{'params': [v], 'name': k, 'lr': 1e-3} for k, v in pb.items()], lr=0.0, eps=1e-15)
    before = bk.params.clone()
    opt.step()                               # steps the flat parameter buffer through the views
    assert float((bk.params - before).abs().max()) > 0


def test_bookkeeping_kernels_vs_torch():
    """splat_seen_update / splat_best_pose_update vs the torch statements they replace (scripts/splatam.py:348-350,
    :726-731), including never-seen Gaussians, ties and a NaN already stored in max_2D_radius."""
    from splatam_b200 import slam
    g = torch.Generator().manual_seed(3)
    P = 10_007
    radius = torch.randint(0, 40, (P,), generator=g, dtype=torch.int32)
    radius[torch.rand(P, generator=g) < 0.3] = 0
    m0 = torch.rand(P, generator=g) * 40
    m0[5] = float("nan")
    m0[6] = float(radius[6])
    seen_ref = radius > 0
    m_ref = m0.clone()
    m_ref[seen_ref] = torch.max(radius[seen_ref].float(), m_ref[seen_ref])
    m = m0.to(DEV)
    seen = torch.zeros(P, dtype=torch.bool, device=DEV)
    from splatam_b200._lib import lib, check
    import ctypes as C
    vp = lambda t: C.c_void_p(t.data_ptr())                                     # noqa: E731
    check(lib.splat_seen_update(P, vp(radius.to(DEV)), vp(m), vp(seen), None))
    torch.cuda.synchronize()
    assert torch.equal(seen.cpu(), seen_ref)
    assert torch.equal(torch.nan_to_num(m.cpu(), nan=-1.0), torch.nan_to_num(m_ref, nan=-1.0))
    # candidate pose: strictly-less comparison, pose taken from column time_idx
    params = {"cam_unnorm_rots": torch.randn(1, 4, 6, generator=g).to(DEV), "cam_trans": torch.randn(1, 3, 6, generator=g).to(DEV)}
    best = slam.BestPose(params, 4)
    rot0, tran0 = best.rot.clone(), best.tran.clone()
    for loss, moved in ((5.0, True), (5.0, False), (7.0, False), (4.5, True)):
        params["cam_unnorm_rots"].add_(1.0)
        params["cam_trans"].sub_(1.0)
        best.update(torch.tensor(loss, device=DEV), params)
        torch.cuda.synchronize()
        if moved:
            rot0, tran0 = params["cam_unnorm_rots"][..., 4].clone(), params["cam_trans"][..., 4].clone()
        assert torch.equal(best.rot, rot0) and torch.equal(best.tran, tran0)
    assert float(best.min_loss) == 4.5
    params["cam_unnorm_rots"].zero_()
    best.write_back(params)
    assert torch.equal(params["cam_unnorm_rots"][..., 4], rot0)


@pytest.mark.parametrize("h,w,six", [(45, 70, True), (33, 41, True), (37, 53, False)])
def test_loss_gradient_follows_the_incoming_gradient(h, w, six):
    """loss.backward() (incoming gradient exactly 1: the scale kernel exits without touching the image) and
    (2.5 * loss).backward() (every element scaled once) -- splat_scale_inplace, incl. sizes that are not multiples of 4."""
    from splatam_b200 import slam
    im, ds, gt_im, gt_d = SO.images(h, w, 9)
    wts = dict(im=0.5, depth=1.0)
    grads = {}
    for scale in (1.0, 2.5):
        if six:
            r = torch.cat([im, ds], 0).to(DEV).requires_grad_(True)
            loss, _ = slam.image_loss(r, None, gt_im.to(DEV), gt_d.to(DEV), wts, True, 0.99, True)
            leaves = [r]
        else:
            a, b = im.to(DEV).requires_grad_(True), ds.to(DEV).requires_grad_(True)
            loss, _ = slam.image_loss(a, b, gt_im.to(DEV), gt_d.to(DEV), wts, False, 0.5, False)
            leaves = [a, b]
        (loss * scale).backward() if scale != 1.0 else loss.backward()
        grads[scale] = [x.grad.clone() for x in leaves]
    for g1, g2 in zip(grads[1.0], grads[2.5]):
        assert float(g1.abs().max()) > 0
        assert torch.equal(g2, g1 * 2.5)


def test_graph_replay_overflow_is_detected_and_harmless():
    """a replay that renders more instances than the captured binning capacity must not walk stale
    state.  The Gaussians are inflated AFTER capture so every replay overflows: the blend kernels exit untouched, the
    gradients of those steps are exactly zero (so Adam only coasts on its momentum), everything stays finite, and check()
    raises CaptureOverflow so the caller can restore the pose and re-run the frame eagerly."""
    from splatam_b200 import rasterizer as R
    from splatam_b200 import slam
    base, curr = _slam_scene(6000, True)
    wts = dict(im=0.5, depth=1.0)
    lrs = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
               cam_unnorm_rots=0.0004, cam_trans=0.002)
    p = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    va = {"max_2D_radius": torch.zeros(6000, device=DEV)}
    opt = slam.FusedAdam([{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in p.items()])

    def it():
        loss, _, _ = slam.get_loss(p, curr, va, 1, wts, True, 0.99, True, False, tracking=True, tracking_pose_only=True)
        loss.backward()
        opt.step()
        return loss

    saved = (R._CAP_SLACK, R._CAP_STEP, R._CAP_GROWTH)
    R._CAP_SLACK, R._CAP_STEP, R._CAP_GROWTH = 0, 1, 1.0      # capacity == exactly the warm-up's num_rendered
    try:
        R._capacity_hint.clear()
        g = slam.GraphedIteration(it, p, opt, warmup=2)
        g.replay()
        g.check()                                              # same scene: fits
        pose_before = (p["cam_unnorm_rots"].detach().clone(), p["cam_trans"].detach().clone())
        with torch.no_grad():
            p["log_scales"] += 1.2                             # 3.3x larger splats: far more (tile, Gaussian) instances
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        for k in ("cam_unnorm_rots", "cam_trans"):
            assert torch.isfinite(p[k]).all()
            assert float(p[k].grad.abs().max()) == 0.0, "an overflowed replay must produce all-zero pose gradients"
        drift = float((p["cam_trans"].detach() - pose_before[1]).abs().max())
        assert drift <= 3 * lrs["cam_trans"] * 1.01            # at most lr per step: momentum only
        with pytest.raises(slam.CaptureOverflow):
            g.check()
    finally:
        R._CAP_SLACK, R._CAP_STEP, R._CAP_GROWTH = saved
        R._capacity_hint.clear()
