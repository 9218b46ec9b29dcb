"""GPU parity tests proper: the CUDA path (through the Python shim -> C ABI -> sm_100a kernels)
against (1) the CPU oracle, (2) the live, unmodified reference extensions in oracle/_ref and
(3) the committed golden fixtures.  Tolerances are BASELINE.json's: forward 1e-5 relative,
integer/index outputs bit-exact, gradients and Hessian terms 1e-3 relative (atomic-order
nondeterminism), each with an absolute floor relative to the tensor's max (util.rel_err)."""
import os

import numpy as np
import pytest
import torch

from util import CASES, GOLDEN_DIR, build_case, frac_bad, rel_err

pytestmark = pytest.mark.gpu

FWD_TOL = 1e-5
GRAD_TOL = 1e-3
# |a-b| <= 1e-3*max(|a|,|b|) + 1e-6*||ref||_inf  (1e-3 relative with an absolute floor),
# plus -- in the live comparisons -- twice the reference's own run-to-run difference on that entry.
# Calibrated on a B200 (profiles/r01_grad_noise_calibration.txt, scripts/calibrate_grad_noise.py): the
# REFERENCE extension's own run-to-run differences (atomic order, amplified by the cancellation in the
# cov2D/cov3D chain of thin Gaussians) reach 5.6e-3 with a 1e-8*||ref||_inf floor and 0.9e-3..1.1e-3
# with a 1e-3*||ref||_inf one, i.e. the reference does not reproduce ITSELF to 1e-3 on a few entries.
# Live tests therefore run the reference twice and add its self-difference to the allowance; our own
# run-to-run spread is ~4e-5 at this floor (one reduction per warp instead of one atomic per pixel).
GRAD_FLOOR = 1e-3


def _pkg(variant):
    if variant == "hx":
        import hessian_diff_gaussian_rasterization_w_depth as m
    else:
        import diff_gaussian_rasterization as m
    return m


def run_ours(variant, rs, sc, dL, power=1):
    from splatam_b200 import rasterizer as R
    R.keep_last_state = True   # the parity tests read the call's internal state (ranges, point_list, n_contrib ...)
    m = _pkg(variant)
    leaves = {k: v.clone().requires_grad_(True) for k, v in sc.items()}
    means2D = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rast = m.GaussianRasterizer(rs, backward_power=power) if variant == "hx" else m.GaussianRasterizer(rs)
    color, radii, depth = rast(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                               colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                               rotations=leaves["rotations"])
    st = R.last_state
    views = {k: v.clone() for k, v in st.views().items()}
    color.backward(dL)
    torch.cuda.synchronize()
    grads = dict(means3D=leaves["means3D"].grad, means2D=means2D.grad, colors=leaves["colors_precomp"].grad,
                 opacities=leaves["opacities"].grad, scales=leaves["scales"].grad,
                 rotations=leaves["rotations"].grad)
    return dict(color=color.detach(), radii=radii, depth=depth.detach(), grads=grads, views=views,
                num_rendered=st.num_rendered)


def run_ref(variant, rs, sc, dL, power=1):
    from oracle import ref_ext
    if not ref_ext.available(variant):
        pytest.skip("oracle/_ref/ref_%s_C.so not built" % variant)
    f = ref_ext.forward(variant, rs, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                        sc["rotations"])
    g = ref_ext.backward(f, dL, power)
    torch.cuda.synchronize()
    fT, nc, rg = ref_ext.image_state_views(f)
    return dict(color=f["color"], radii=f["radii"], depth=f["depth"], grads=g, num_rendered=f["num_rendered"],
                final_T=fT.clone(), n_contrib=nc.clone(), ranges=rg.clone(),
                point_list=ref_ext.binning_point_list(f).clone())


def _check_grads(ours, ref, keys=("means3D", "means2D", "colors", "opacities", "scales", "rotations"), ref2=None):
    """|ours - ref| <= GRAD_TOL * (max(|ours|,|ref|) + GRAD_FLOOR*||ref||_inf) [+ 2 * spread of the reference].
    `ref2` is one more run (dict) or several more runs (list of dicts) of the SAME reference build on the same inputs:
    where the reference does not reproduce itself (atomic order x cancellation), the spread max-min of its own
    results on that entry, doubled, is added to the allowance."""
    worst = {}
    more = [] if ref2 is None else (list(ref2) if isinstance(ref2, (list, tuple)) else [ref2])
    for k in keys:
        a, b = ours[k].double().flatten(), ref[k].double().flatten()
        allow = GRAD_TOL * (torch.maximum(a.abs(), b.abs()) + GRAD_FLOOR * float(b.abs().max()) + 1e-30)
        if more:
            runs = torch.stack([b] + [r[k].double().flatten() for r in more])
            allow = allow + 2.0 * (runs.max(0)[0] - runs.min(0)[0])
        worst[k] = float(((a - b).abs() / allow).max()) * GRAD_TOL
    bad = {k: v for k, v in worst.items() if not v <= GRAD_TOL}
    assert not bad, "gradient mismatch (rel err): %s (all: %s)" % (bad, worst)


@pytest.mark.parametrize("variant", ["dg", "hx"])
@pytest.mark.parametrize("case", list(CASES))
def test_vs_reference_extension(case, variant):
    rs, sc, dL = build_case(case, "cuda")
    ours = run_ours(variant, rs, sc, dL)
    ref = run_ref(variant, rs, sc, dL)
    H, W = rs.image_height, rs.image_width
    T = ((W + 15) // 16) * ((H + 15) // 16)
    assert torch.equal(ours["radii"], ref["radii"]), "radii differ at %d Gaussians" % int(
        (ours["radii"] != ref["radii"]).sum())
    assert ours["num_rendered"] == ref["num_rendered"]
    assert torch.equal(ours["views"]["ranges"], ref["ranges"][:T]), "tile ranges not bit-exact"
    assert torch.equal(ours["views"]["point_list"], ref["point_list"]), "sorted point list differs"
    assert torch.equal(ours["views"]["n_contrib"].flatten(), ref["n_contrib"]), "n_contrib differs at %d px" % int(
        (ours["views"]["n_contrib"].flatten() != ref["n_contrib"]).sum())
    assert rel_err(ours["views"]["final_T"], ref["final_T"], 1e-6) <= FWD_TOL
    assert rel_err(ours["color"], ref["color"], 1e-6) <= FWD_TOL
    assert rel_err(ours["depth"], ref["depth"], 1e-6) <= FWD_TOL
    # small cases: four more reference runs to see how far the reference is from ITSELF on each entry
    _check_grads(ours["grads"], ref["grads"], ref2=[run_ref(variant, rs, sc, dL)["grads"] for _ in range(4)])


@pytest.mark.parametrize("case", ["cloud_small", "surface_small", "c1_cloud"])
def test_hessian_vs_reference_extension(case):
    """backward_power=2 (compute_Hessian pattern, scripts/ros_handler.py:884-885)."""
    rs, sc, _ = build_case(case, "cuda")
    dL = torch.ones(3, rs.image_height, rs.image_width, device="cuda") * 1e-3
    ours = run_ours("hx", rs, sc, dL, power=2)
    ref = run_ref("hx", rs, sc, dL, power=2)
    _check_grads(ours["grads"], ref["grads"], ref2=run_ref("hx", rs, sc, dL, power=2)["grads"])


@pytest.mark.parametrize("case", ["cloud_small", "cloud_ragged", "surface_small"])
def test_vs_cpu_oracle(case):
    from oracle import raster_oracle as O
    rs_c, sc_c, dL_c = build_case(case, "cpu")
    rs, sc, dL = build_case(case, "cuda")
    ours = run_ours("dg", rs, sc, dL)
    ctx = O.forward(rs_c, 0.2, sc_c["means3D"], sc_c["colors_precomp"], sc_c["opacities"], sc_c["scales"],
                    sc_c["rotations"])
    bl = ctx["pre"]["borderline"]
    same = (ours["radii"].cpu() == ctx["radii"]) | bl
    assert bool(same.all()), "radii differ from the oracle at non-borderline Gaussians"
    if bool((ours["radii"].cpu() == ctx["radii"]).all()):
        assert torch.equal(ours["views"]["ranges"].cpu().long(), ctx["ranges"])
        assert torch.equal(ours["views"]["point_list"].cpu().long(), ctx["point_list"])
        # float32 CPU vs GPU (FMA contraction) can flip a threshold at a handful of pixels
        nc_diff = (ours["views"]["n_contrib"].cpu().long() != ctx["fwd"]["n_contrib"]).float().mean()
        assert float(nc_diff) < 2e-3
        assert frac_bad(ours["color"].cpu(), ctx["color"], 1e-4) < 2e-3
        g = O.backward(ctx, dL_c, 1)
        for k in ("means3D", "means2D", "colors", "opacities", "scales", "rotations"):
            assert frac_bad(ours["grads"][k].cpu(), g[k], 5e-3, 1e-4) < 5e-3, k


def test_cull_is_exact():
    """The per-warp footprint cull must not change any output bit (forward) -- compared by running
    the library in a subprocess with SPLAT_B200_NO_CULL=1."""
    import subprocess
    import sys
    code = r'''
import sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
from util import build_case
from test_gpu_parity import run_ours
rs, sc, dL = build_case("c1_cloud", "cuda")
o = run_ours("hx", rs, sc, dL)
torch.save(dict(color=o["color"].cpu(), nc=o["views"]["n_contrib"].cpu(), T=o["views"]["final_T"].cpu(),
                g={k: v.cpu() for k, v in o["grads"].items()}), sys.argv[1])
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    import tempfile
    outs = []
    for flag in ("0", "1"):
        with tempfile.NamedTemporaryFile(suffix=".pt", delete=False) as f:
            path = f.name
        env = dict(os.environ, SPLAT_B200_NO_CULL=flag)
        subprocess.run([sys.executable, "-c", code, path], check=True, env=env)
        outs.append(torch.load(path))
        os.unlink(path)
    a, b = outs
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["nc"], b["nc"]) and torch.equal(a["T"], b["T"])
    for k in a["g"]:
        assert rel_err(a["g"][k], b["g"][k], GRAD_FLOOR) <= GRAD_TOL, k


def test_against_golden_fixtures():
    files = sorted(f for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith(("slam_glue_", "densify_")))
    if not files:
        pytest.skip("no golden fixtures committed yet")
    for fn in files:
        z = np.load(os.path.join(GOLDEN_DIR, fn))
        case, variant, power = str(z["case"]), str(z["variant"]), int(z["power"])
        rs, sc, dL = build_case(case, "cuda")
        if power == 2:
            dL = torch.ones_like(dL) * 1e-3
        ours = run_ours(variant, rs, sc, dL, power)
        assert np.array_equal(ours["radii"].cpu().numpy(), z["radii"]), fn
        assert np.array_equal(ours["views"]["ranges"].cpu().numpy(), z["ranges"]), fn
        assert np.array_equal(ours["views"]["n_contrib"].cpu().numpy().ravel(), z["n_contrib"].ravel()), fn
        assert rel_err(ours["color"].cpu(), torch.from_numpy(z["color"]), 1e-6) <= FWD_TOL, fn
        # a fixture is ONE stored reference run (its atomic-order noise is frozen in): 99.9% of the entries
        # must meet the 1e-3 bar and none may be off by more than 5e-3
        for k in ("means3D", "means2D", "colors", "opacities", "scales", "rotations"):
            g_ref = torch.from_numpy(z["g_" + k])
            assert frac_bad(ours["grads"][k].cpu(), g_ref, GRAD_TOL, GRAD_FLOOR) <= 1e-3, (fn, k)
            assert rel_err(ours["grads"][k].cpu(), g_ref, GRAD_FLOOR) <= 5 * GRAD_TOL, (fn, k)


def test_edge_cases():
    import hessian_diff_gaussian_rasterization_w_depth as m
    rs, sc, dL = build_case("cloud_small", "cuda")
    # P == 0 (HX/rasterize_points.cu:81)
    e3 = torch.zeros(0, 3, device="cuda")
    color, radii, depth = m.GaussianRasterizer(rs)(means3D=e3, means2D=e3, opacities=torch.zeros(0, 1, device="cuda"),
                                                   colors_precomp=e3, scales=e3, rotations=torch.zeros(0, 4, device="cuda"))
    assert color.shape == (3, 64, 64) and float(color.abs().max()) == 0 and radii.numel() == 0
    # everything behind the camera: R == 0, background only
    sc2 = dict(sc); sc2["means3D"] = sc["means3D"].clone(); sc2["means3D"][:, 2] = -1.0
    rs_bg = rs._replace(bg=torch.tensor([0.25, 0.5, 0.75], device="cuda"))
    leaves = {k: v.clone().requires_grad_(True) for k, v in sc2.items()}
    color, radii, depth = m.GaussianRasterizer(rs_bg)(means3D=leaves["means3D"], means2D=torch.zeros_like(sc["means3D"]),
                                                      opacities=leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                                                      scales=leaves["scales"], rotations=leaves["rotations"])
    assert int(radii.max()) == 0
    assert torch.allclose(color[:, 0, 0], rs_bg.bg) and float(depth.min()) == 15.0
    color.sum().backward()
    assert float(leaves["means3D"].grad.abs().max()) == 0
    # markVisible
    vis = m.GaussianRasterizer(rs).markVisible(sc["means3D"])
    assert torch.equal(vis, sc["means3D"][:, 2] > 0.001)
    # capacity regrowth: a tiny scene first (sets the hint), then a much bigger one
    from splatam_b200 import rasterizer as R
    R._capacity_hint.clear()
    small = {k: v[:10].contiguous() for k, v in sc.items()}
    run = lambda s: m.GaussianRasterizer(rs)(means3D=s["means3D"], means2D=torch.zeros_like(s["means3D"]),
                                             opacities=s["opacities"], colors_precomp=s["colors_precomp"],
                                             scales=s["scales"], rotations=s["rotations"])[0]
    run(small)
    R._capacity_hint[(torch.cuda.current_device(), 64, 64)] = 1   # force the speculative launch to overflow
    slack, R._CAP_SLACK = R._CAP_SLACK, 0
    step, R._CAP_STEP = R._CAP_STEP, 1          # no rounding up: capacity 1 < R, phase 2 must be re-run
    a = run(sc)
    R._CAP_SLACK, R._CAP_STEP = slack, step
    R._capacity_hint.clear()
    b = run(sc)
    assert torch.equal(a, b)


def _cov3d_from(sc):
    """A symmetric PSD cov3D_precomp [P,6] (xx,xy,xz,yy,yz,zz) built in torch from the case's scales/rotations."""
    q = torch.nn.functional.normalize(sc["rotations"], dim=1)
    r, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    L = R * sc["scales"][:, None, :]
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).contiguous()


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_and_cov3d_precomp_vs_reference_extension(deg):
    """The paths SplaTAM itself never takes: colours from spherical harmonics of degree 0..3 with the
    clamp mask, and a precomputed 3D covariance -- forward and backward against the upstream (DG) extension, whose
    SH backward is the defined one (the fork's fused kernel mis-strides dL_dsh, backward.cu:1067)."""
    import diff_gaussian_rasterization as m
    from oracle import ref_ext
    if not ref_ext.available("dg"):
        pytest.skip("oracle/_ref/ref_dg_C.so not built")
    rs, sc, dL = build_case("cloud_small", "cuda")
    rs = rs._replace(sh_degree=deg)
    P = sc["means3D"].shape[0]
    g = torch.Generator().manual_seed(50 + deg)
    sh = (torch.randn(P, 16, 3, generator=g) * 0.5).cuda()      # M = 16 coefficients stored, (deg+1)^2 active
    sh[:, 0] += 0.3                                              # some colours clamp at 0, most do not
    cov = _cov3d_from(sc)
    leaves = dict(means3D=sc["means3D"].clone().requires_grad_(True), opac=sc["opacities"].clone().requires_grad_(True),
                  sh=sh.clone().requires_grad_(True), cov=cov.clone().requires_grad_(True))
    means2D = torch.zeros_like(sc["means3D"], requires_grad=True)
    color, radii, depth = m.GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opac"],
                                                   shs=leaves["sh"], cov3D_precomp=leaves["cov"])
    color.backward(dL)
    f = ref_ext.forward("dg", rs, sc["means3D"], None, sc["opacities"], None, None, sh=sh, cov3D_precomp=cov)
    gr = ref_ext.backward(f, dL)
    gr2 = ref_ext.backward(f, dL)
    assert torch.equal(radii, f["radii"])
    assert rel_err(color.detach(), f["color"], 1e-6) <= FWD_TOL
    assert rel_err(depth.detach(), f["depth"], 1e-6) <= FWD_TOL
    ours = dict(means3D=leaves["means3D"].grad, means2D=means2D.grad, opacities=leaves["opac"].grad,
                sh=leaves["sh"].grad, cov3D=leaves["cov"].grad)
    _check_grads(ours, gr, keys=("means3D", "means2D", "opacities", "sh", "cov3D"), ref2=gr2)
    act = (deg + 1) ** 2
    assert float(ours["sh"][:, act:].abs().max() if act < 16 else 0.0) == 0.0   # inactive coefficients get no gradient


def test_hx_sh_forward_and_cov3d_backward_vs_reference_extension():
    """Fork (HX) variant: SH colours forward (degree 3), and the fused backward through cov3D_precomp with
    precomputed colours, power 1 and 2."""
    import hessian_diff_gaussian_rasterization_w_depth as m
    from oracle import ref_ext
    if not ref_ext.available("hx"):
        pytest.skip("oracle/_ref/ref_hx_C.so not built")
    rs, sc, dL = build_case("cloud_small", "cuda")
    P = sc["means3D"].shape[0]
    g = torch.Generator().manual_seed(60)
    sh = (torch.randn(P, 16, 3, generator=g) * 0.5).cuda()
    sh[:, 0] += 0.3
    cov = _cov3d_from(sc)
    rs3 = rs._replace(sh_degree=3)
    color, radii, depth = m.GaussianRasterizer(rs3)(means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]),
                                                    opacities=sc["opacities"], shs=sh, scales=sc["scales"],
                                                    rotations=sc["rotations"])
    f = ref_ext.forward("hx", rs3, sc["means3D"], None, sc["opacities"], sc["scales"], sc["rotations"], sh=sh)
    assert torch.equal(radii, f["radii"]) and rel_err(color, f["color"], 1e-6) <= FWD_TOL
    for power in (1, 2):
        leaves = dict(means3D=sc["means3D"].clone().requires_grad_(True), opac=sc["opacities"].clone().requires_grad_(True),
                      col=sc["colors_precomp"].clone().requires_grad_(True), cov=cov.clone().requires_grad_(True))
        means2D = torch.zeros_like(sc["means3D"], requires_grad=True)
        color, radii, depth = m.GaussianRasterizer(rs, backward_power=power)(
            means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opac"], colors_precomp=leaves["col"],
            cov3D_precomp=leaves["cov"])
        color.backward(dL)
        f = ref_ext.forward("hx", rs, sc["means3D"], sc["colors_precomp"], sc["opacities"], None, None, cov3D_precomp=cov)
        gr, gr2 = ref_ext.backward(f, dL, power), ref_ext.backward(f, dL, power)
        assert torch.equal(radii, f["radii"]) and rel_err(color.detach(), f["color"], 1e-6) <= FWD_TOL
        ours = dict(means3D=leaves["means3D"].grad, means2D=means2D.grad, opacities=leaves["opac"].grad,
                    colors=leaves["col"].grad, cov3D=leaves["cov"].grad)
        _check_grads(ours, gr, keys=("means3D", "means2D", "opacities", "colors", "cov3D"), ref2=gr2)


def test_gradient_sink_matches_autograd():
    """mapping.GradBucket(direct=True): the backward kernels write the five Gaussian gradient arrays straight into
    the flat bucket (no autograd accumulate); same numbers as the ordinary autograd path, stale contents overwritten."""
    import hessian_diff_gaussian_rasterization_w_depth as m
    from splatam_b200 import mapping
    rs, sc, dL = build_case("cloud_small", "cuda")
    ours = run_ours("hx", rs, sc, dL)
    bk = mapping.GradBucket(sc, "cuda", direct=True)
    bk.grads.fill_(123.0)                       # must be fully overwritten (rows of culled Gaussians become 0)
    lv = bk.leaves()
    means2D = torch.zeros_like(sc["means3D"], requires_grad=True)
    color, radii, depth = m.GaussianRasterizer(rs)(means3D=lv["means3D"], means2D=means2D, opacities=lv["opacities"],
                                                   colors_precomp=lv["colors_precomp"], scales=lv["scales"],
                                                   rotations=lv["rotations"])
    color.backward(dL)
    torch.cuda.synchronize()
    assert torch.equal(color.detach(), ours["color"])
    got = dict(means3D=bk.gviews["means3D"], colors=bk.gviews["colors_precomp"], opacities=bk.gviews["opacities"],
               scales=bk.gviews["scales"], rotations=bk.gviews["rotations"], means2D=means2D.grad)
    for k in got:
        assert lv.get({"colors": "colors_precomp"}.get(k, k)) is None or \
            lv[{"colors": "colors_precomp"}.get(k, k)].grad.data_ptr() == got[k].data_ptr()
    _check_grads(got, ours["grads"])
    assert float(bk.grads.abs().max()) < 123.0


@pytest.mark.parametrize("case", ["cloud_small", "surface_small", "c1_cloud"])
def test_fused_depth_sil_vs_two_reference_passes(case):
    """render_rgbd_sil() returns, from ONE pass, what scripts/splatam.py:255-259 gets from TWO passes of
    the reference -- the RGB render and the render of colors_precomp = (z, 1, z*z) (utils/slam_helpers.py:124-139) --
    and its backward equals the sum of the two reference backwards plus torch's chain from (z, 1, z*z) to means3D."""
    import hessian_diff_gaussian_rasterization_w_depth as m
    from oracle import ref_ext
    if not ref_ext.available("hx"):
        pytest.skip("oracle/_ref/ref_hx_C.so not built")
    rs, sc, _ = build_case(case, "cuda")
    H, W = rs.image_height, rs.image_width
    g = torch.Generator().manual_seed(9)
    dL6 = torch.randn(6, H, W, generator=g).cuda()
    dL6[3:] *= torch.tensor([0.3, 1.0, 0.05], device="cuda").view(3, 1, 1)   # depth ~ metres, depth^2 larger
    leaves = {k: v.clone().requires_grad_(True) for k, v in sc.items()}
    means2D = torch.zeros_like(sc["means3D"], requires_grad=True)
    color6, radii, depth = m.GaussianRasterizer(rs).render_rgbd_sil(
        means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
        colors_precomp=leaves["colors_precomp"], scales=leaves["scales"], rotations=leaves["rotations"])
    assert color6.shape == (6, H, W)
    color6.backward(dL6)
    ours = dict(means3D=leaves["means3D"].grad, means2D=means2D.grad, colors=leaves["colors_precomp"].grad,
                opacities=leaves["opacities"].grad, scales=leaves["scales"].grad, rotations=leaves["rotations"].grad)

    def two_passes():
        means = sc["means3D"].clone().requires_grad_(True)
        pts4 = torch.cat([means, torch.ones_like(means[:, :1])], 1)
        z = (pts4 @ rs.viewmatrix.reshape(4, 4))[:, 2]          # viewmatrix is w2c^T: row-vector convention
        colB = torch.stack([z, torch.ones_like(z), z * z], 1)
        fA = ref_ext.forward("hx", rs, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"])
        gA = ref_ext.backward(fA, dL6[:3].contiguous())
        fB = ref_ext.forward("hx", rs, sc["means3D"], colB.detach().contiguous(), sc["opacities"], sc["scales"],
                             sc["rotations"])
        gB = ref_ext.backward(fB, dL6[3:].contiguous())
        colB.backward(gB["colors"])
        tot = {k: gA[k] + gB[k] for k in ("means3D", "means2D", "opacities", "scales", "rotations")}
        tot["means3D"] = tot["means3D"] + means.grad
        tot["colors"] = gA["colors"]
        return fA, fB, tot

    fA, fB, ref = two_passes()
    _, _, ref2 = two_passes()
    assert torch.equal(radii, fA["radii"])
    assert rel_err(color6[:3].detach(), fA["color"], 1e-6) <= FWD_TOL
    assert rel_err(color6[3:].detach(), fB["color"], 1e-6) <= FWD_TOL
    assert rel_err(depth.detach(), fA["depth"], 1e-6) <= FWD_TOL
    _check_grads(ours, ref, ref2=ref2)


@pytest.mark.parametrize("variant", ["dg", "hx"])
def test_scale_modifier_and_background_vs_reference_extension(variant):
    """Settings fields SplaTAM leaves at their defaults: scale_modifier != 1 (forward.cu:122-126, backward.cu:436-441)
    and a non-black background (forward.cu:388, backward.cu:655-657 -- the bg term of dL/dalpha)."""
    rs, sc, dL = build_case("cloud_small", "cuda")
    rs = rs._replace(scale_modifier=0.7, bg=torch.tensor([0.2, 0.5, 0.9], device="cuda"))
    ours = run_ours(variant, rs, sc, dL)
    ref = run_ref(variant, rs, sc, dL)
    assert torch.equal(ours["radii"], ref["radii"])
    assert torch.equal(ours["views"]["point_list"], ref["point_list"])
    assert torch.equal(ours["views"]["n_contrib"].flatten(), ref["n_contrib"])
    assert rel_err(ours["color"], ref["color"], 1e-6) <= FWD_TOL
    assert rel_err(ours["depth"], ref["depth"], 1e-6) <= FWD_TOL
    _check_grads(ours["grads"], ref["grads"], ref2=[run_ref(variant, rs, sc, dL)["grads"] for _ in range(4)])


def test_mark_visible_vs_reference_extension():
    import hessian_diff_gaussian_rasterization_w_depth as hx
    import diff_gaussian_rasterization as dg
    from oracle import ref_ext
    rs, sc, _ = build_case("c1_cloud", "cuda")
    for variant, m in (("hx", hx), ("dg", dg)):
        if not ref_ext.available(variant):
            pytest.skip("reference build missing")
        ref = ref_ext.load(variant).mark_visible(sc["means3D"], rs.viewmatrix, rs.projmatrix)
        assert torch.equal(m.GaussianRasterizer(rs).markVisible(sc["means3D"]), ref)


# ------------------------------------------------------------------------------------------------
# Full-size cases (BASELINE.json configs C2/C3/C4 shapes).  The CPU oracle would need minutes here, so
# the comparand is the live reference extension on the same box plus size-independent properties.
def _full_case(P, cam_name, layers, aniso=0.0, seed=2, kf=0, nkf=24):
    from splatam_b200 import synth
    cam = getattr(synth, cam_name)
    sc = synth.surface_scene(P, cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], seed=seed,
                             layers=layers, aniso_frac=aniso)
    rs = synth.make_settings(w2c=synth.arc_pose(kf, nkf), device="cuda", **cam)
    sc = {k: v.cuda() for k, v in sc.items()}
    g = torch.Generator().manual_seed(77)
    dL = (torch.randn(3, cam["h"], cam["w"], generator=g) / (3.0 * cam["h"] * cam["w"])).cuda()
    return rs, sc, dL


@pytest.mark.parametrize("variant,P,cam_name,layers,kf", [("dg", 1_000_000, "REPLICA", 3, 5),
                                                         ("hx", 1_000_000, "REPLICA", 3, 17),
                                                         ("dg", 300_000, "REPLICA", 1, 0)])
def test_full_size_vs_reference_extension(variant, P, cam_name, layers, kf):
    rs, sc, dL = _full_case(P, cam_name, layers, kf=kf)
    ours = run_ours(variant, rs, sc, dL)
    ref = run_ref(variant, rs, sc, dL)
    H, W = rs.image_height, rs.image_width
    T = ((W + 15) // 16) * ((H + 15) // 16)
    assert torch.equal(ours["radii"], ref["radii"])
    assert ours["num_rendered"] == ref["num_rendered"]
    assert torch.equal(ours["views"]["ranges"], ref["ranges"][:T])
    assert torch.equal(ours["views"]["point_list"], ref["point_list"])
    assert torch.equal(ours["views"]["n_contrib"].flatten(), ref["n_contrib"])
    assert rel_err(ours["color"], ref["color"], 1e-6) <= FWD_TOL
    assert rel_err(ours["depth"], ref["depth"], 1e-6) <= FWD_TOL
    # The "surface" scene is isotropic (SplaTAM's own initialisation): dL/drotation is EXACTLY zero in exact
    # arithmetic, and both implementations return pure cancellation noise for it (the reference's is atomic-order
    # noise), so that array is compared against its natural scale |dL/dscale| * |scale| instead of against itself.
    _check_grads(ours["grads"], ref["grads"], keys=("means3D", "means2D", "colors", "opacities", "scales"),
                 ref2=run_ref(variant, rs, sc, dL)["grads"])
    nat = float((ref["grads"]["scales"].abs() * sc["scales"]).max())
    assert float(ours["grads"]["rotations"].abs().max()) <= 1e-3 * nat
    assert float(ref["grads"]["rotations"].abs().max()) <= 1e-3 * nat


def test_full_size_hessian_vs_reference_extension():
    """C4 shape: 1752x1168, 2M Gaussians (20% anisotropic), backward_power=2 with dL = 1e-3 (ros_handler.py:885)."""
    rs, sc, _ = _full_case(2_000_000, "SCANNETPP", 3, aniso=0.2, seed=3, kf=2, nkf=8)
    dL = torch.ones(3, rs.image_height, rs.image_width, device="cuda") * 1e-3
    ours = run_ours("hx", rs, sc, dL, power=2)
    ref = run_ref("hx", rs, sc, dL, power=2)
    assert torch.equal(ours["radii"], ref["radii"])
    assert torch.equal(ours["views"]["n_contrib"].flatten(), ref["n_contrib"])
    assert rel_err(ours["color"], ref["color"], 1e-6) <= FWD_TOL
    _check_grads(ours["grads"], ref["grads"], ref2=run_ref("hx", rs, sc, dL, power=2)["grads"])


def test_full_size_properties():
    """Size-independent properties at the C3 shape: sorted lists, R = sum of tile counts, deterministic
    forward, backward linear in dL/dcolour, zero gradient in -> zero gradient out."""
    rs, sc, dL = _full_case(1_000_000, "REPLICA", 3, kf=11)
    a = run_ours("hx", rs, sc, dL)
    v = a["views"]
    rg = v["ranges"].long()
    lens = rg[:, 1] - rg[:, 0]
    assert int(lens.sum()) == a["num_rendered"] and int(lens.min()) >= 0
    nz = lens > 0
    starts = torch.sort(rg[nz, 0])[0]
    ends = torch.sort(rg[nz, 1])[0]
    assert int(starts[0]) == 0 and int(ends[-1]) == a["num_rendered"] and torch.equal(starts[1:], ends[:-1])
    # inside every tile list: ascending (depth bits, id)
    ids = v["point_list"].long()
    depth_bits = v["depths"][ids].contiguous().view(torch.int32).long()
    key = depth_bits * (1 << 32) + ids
    tile_of = torch.repeat_interleave(torch.arange(rg.shape[0], device="cuda"), lens)
    same_tile = tile_of[1:] == tile_of[:-1]
    assert bool((key[1:][same_tile] > key[:-1][same_tile]).all())
    # n_contrib never exceeds the tile's list length
    H, W = rs.image_height, rs.image_width
    gx = (W + 15) // 16
    ty, tx = torch.meshgrid(torch.arange(H, device="cuda") // 16, torch.arange(W, device="cuda") // 16, indexing="ij")
    assert bool((v["n_contrib"].long() <= lens[ty * gx + tx]).all())
    # deterministic forward
    b = run_ours("hx", rs, sc, dL)
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["depth"], b["depth"])
    # linearity / zero
    c = run_ours("hx", rs, sc, 2.0 * dL)
    for k in a["grads"]:
        if k == "rotations":  # isotropic scene: exactly zero in exact arithmetic, atomic-order noise here
            continue
        assert rel_err(c["grads"][k], 2.0 * a["grads"][k], GRAD_FLOOR) <= GRAD_TOL, k
    z = run_ours("hx", rs, sc, torch.zeros_like(dL))
    for k in z["grads"]:
        assert float(z["grads"][k].abs().max()) == 0.0, k


# ------------------------------------------------------------------------------------------------
# Round-2 parity holes (VERDICT r01 "What's weak" #1): the two branches of the tile sort that no seeded scene reached,
# and the packed exponential of the two-pixels-per-lane blend kernels.
def test_packed_exp_matches_expf():
    """csrc/packed.cuh exp2_ref == the device library's expf (what the reference build evaluates, forward.cu:349) on
    EVERY float bit pattern: the alpha >= 1/255 / T < 1e-4 / median-depth decisions rest on it."""
    import ctypes as C
    from splatam_b200._lib import lib, check
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for first in (0x00000000, 0x40000000, 0x80000000, 0xC0000000):
        check(lib.splat_debug_exp_mismatches(first, 1 << 30, C.c_void_p(bad.data_ptr()), st))
    torch.cuda.synchronize()
    assert int(bad.item()) == 0, "%d float inputs where the packed exp differs from expf" % int(bad.item())


def _compare_forward_exact(ours, ref, rs):
    H, W = rs.image_height, rs.image_width
    T = ((W + 15) // 16) * ((H + 15) // 16)
    assert torch.equal(ours["radii"], ref["radii"])
    assert ours["num_rendered"] == ref["num_rendered"]
    assert torch.equal(ours["views"]["ranges"], ref["ranges"][:T]), "tile ranges not bit-exact"
    assert torch.equal(ours["views"]["point_list"], ref["point_list"]), "sorted point list differs at %d positions" % int(
        (ours["views"]["point_list"] != ref["point_list"]).sum())
    assert torch.equal(ours["views"]["n_contrib"].flatten(), ref["n_contrib"])
    assert rel_err(ours["color"], ref["color"], 1e-6) <= FWD_TOL
    assert rel_err(ours["depth"], ref["depth"], 1e-6) <= FWD_TOL


def test_long_tile_lists_take_the_global_memory_sort_path():
    """40k Gaussians on a 64x64 image: every tile list is longer than the 2048 keys the tile sort holds in shared
    memory, so the bucket sort runs on the global ping-pong buffers (csrc/binning.cu, `keys_alt`)."""
    cam = synth_mod().C1
    rs = synth_mod().make_settings(device="cuda", **cam)
    sc = {k: v.cuda() for k, v in synth_mod().cloud_scene(40000, cam["w"], cam["h"], cam["fx"], cam["fy"], seed=21).items()}
    dL = torch.randn(3, cam["h"], cam["w"], generator=torch.Generator().manual_seed(5)).cuda()
    ours = run_ours("dg", rs, sc, dL)
    ref = run_ref("dg", rs, sc, dL)
    lens = (ours["views"]["ranges"][:, 1] - ours["views"]["ranges"][:, 0])
    assert int(lens.max()) > 2048 * 2, "scene does not exceed the shared-memory sort capacity: max list %d" % int(lens.max())
    _compare_forward_exact(ours, ref, rs)
    _check_grads(ours["grads"], ref["grads"], ref2=run_ref("dg", rs, sc, dL)["grads"])


def synth_mod():
    from splatam_b200 import synth
    return synth


def quantised_depth_scene(cam, stride=1, seed=9, wall=True):
    """SplaTAM's frame-0 map (scripts/splatam.py:89-116,131-132): one Gaussian per pixel, back-projected from a depth
    image quantised to 1/5000 m (TUM 16-bit PNG) with w2c = I -- view-space depths are bit-equal for hundreds of
    Gaussians per tile, so the sort order inside a tile is decided by the Gaussian index (stable sort of the
    reference, cub::DeviceRadixSort, rasterizer_impl.cu:304)."""
    W, H, fx, fy, cx, cy = cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"]
    v, u = torch.meshgrid(torch.arange(0, H, stride, dtype=torch.float32), torch.arange(0, W, stride, dtype=torch.float32),
                          indexing="ij")
    u, v = u.reshape(-1), v.reshape(-1)
    z = 2.0 + 0.35 * torch.sin(u / 97.0) + 0.25 * torch.cos(v / 61.0)
    if wall:
        z = torch.where(u < W / 2, torch.full_like(z, 1.8), z)       # a fronto-parallel wall: every depth equal
    z = torch.round(z * 5000.0) / 5000.0
    P = z.numel()
    g = torch.Generator().manual_seed(seed)
    means = torch.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
    scale = (z / ((fx + fy) / 2)).unsqueeze(1).repeat(1, 3) * (1.0 + 0.5 * torch.rand(P, 1, generator=g))
    rot = torch.zeros(P, 4); rot[:, 0] = 1.0
    return dict(means3D=means.contiguous(), scales=scale.contiguous(), rotations=rot,
                opacities=torch.full((P, 1), 0.5), colors_precomp=torch.rand(P, 3, generator=g))


def test_quantised_depth_ties_sort_by_gaussian_index():
    cam = dict(w=160, h=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5)
    rs = synth_mod().make_settings(device="cuda", **cam)
    sc = {k: v.cuda() for k, v in quantised_depth_scene(cam).items()}
    dL = torch.randn(3, cam["h"], cam["w"], generator=torch.Generator().manual_seed(6)).cuda()
    ours = run_ours("dg", rs, sc, dL)
    ref = run_ref("dg", rs, sc, dL)
    # the scene really has ties: fewer distinct depth values than Gaussians on the wall
    z = ours["views"]["depths"][ours["radii"] > 0]
    assert z.unique().numel() < z.numel() // 4
    _compare_forward_exact(ours, ref, rs)


def test_quantised_depth_ties_full_frame_sort_time():
    """TUM-size first frame (640x480, 307,200 Gaussians, half of them on a wall at ONE depth): point_list equal to the
    reference's, and the tile sort's time recorded (gpurun_out/tie_sort_time.json) and bounded -- the in-bin ranking
    must not degrade when a whole tile shares one depth value."""
    import ctypes as C
    import json
    from splatam_b200._lib import lib
    cam = synth_mod().TUM
    rs = synth_mod().make_settings(device="cuda", **cam)
    sc = {k: v.cuda() for k, v in quantised_depth_scene(cam).items()}
    dL = torch.randn(3, cam["h"], cam["w"], generator=torch.Generator().manual_seed(7)).cuda()
    ours = run_ours("dg", rs, sc, dL)
    ref = run_ref("dg", rs, sc, dL)
    _compare_forward_exact(ours, ref, rs)
    times = {}
    for name, wall in (("ties", True), ("smooth", False)):
        s2 = {k: v.cuda() for k, v in quantised_depth_scene(cam, wall=wall).items()}
        run_ours("dg", rs, s2, dL)
        lib.splat_profile_enable(1)
        for _ in range(5):
            run_ours("dg", rs, s2, dL)
        n = 16
        ms = (C.c_float * n)(); cnt = (C.c_int32 * n)()
        lib.splat_profile_read(ms, cnt, n)
        lib.splat_profile_enable(0)
        names = [lib.splat_profile_kernel_name(k).decode() for k in range(n)]
        times[name] = {names[k]: ms[k] / cnt[k] for k in range(n) if cnt[k]}
    lens = ours["views"]["ranges"][:, 1] - ours["views"]["ranges"][:, 0]
    rec = dict(max_list=int(lens.max()), mean_list=float(lens.float().mean()), kernel_ms=times)
    if os.path.isdir(os.path.join(os.path.dirname(GOLDEN_DIR), "..", "gpurun_out")):
        with open(os.path.join(os.path.dirname(GOLDEN_DIR), "..", "gpurun_out", "tie_sort_time.json"), "w") as f:
            json.dump(rec, f, indent=1)
    assert times["ties"]["tile_sort"] < 4.0 * times["smooth"]["tile_sort"] + 0.05, rec


def test_gather_and_streamed_records_give_identical_results():
    """splat_set_gather: blend kernels fed by cp.async gathers of the Gaussians' own records vs by the TMA stream of the
    re-written instance records -- same records, same order, so forward outputs are bit-identical and gradients agree
    to atomic-order noise; in streamed mode the instance records equal the gathered rows (id in slot 7)."""
    from splatam_b200._lib import lib
    rs, sc, dL = build_case("surface_small", "cuda")
    prev = lib.splat_set_gather(1)
    try:
        g = run_ours("hx", rs, sc, dL)
        lib.splat_set_gather(0)
        t = run_ours("hx", rs, sc, dL)
        inst = t["views"]["instances"]
        ids = t["views"]["point_list"].long()
        assert torch.equal(inst[:, 7].contiguous().view(torch.int32).long(), ids)
        assert torch.equal(inst[:, :7], torch.cat([t["views"]["means2D"], t["views"]["conic_opacity"],
                                                    t["views"]["depths"][:, None]], 1)[ids])
    finally:
        lib.splat_set_gather(prev)
    for k in ("color", "depth", "radii"):
        assert torch.equal(g[k], t[k]), k
    for k in ("n_contrib", "final_T", "point_list", "ranges"):
        assert torch.equal(g["views"][k], t["views"][k]), k
    _check_grads(g["grads"], t["grads"])


@pytest.mark.parametrize("P,step", [(40000, 0.5), (90000, 0.15)])
def test_depth_ties_on_long_lists(P, step):
    """Depth ties on lists that take the global-memory sort path (> 2048 keys per tile): bins of hundreds of bit-equal
    depths get the tile sort's second-level pass on the global ping-pong buffers; the second case has more heavy bins
    per tile than the pass handles (MAX_HEAVY), so the rest fall back to pairwise ranking.  point_list must equal the
    reference's stable radix sort either way."""
    cam = synth_mod().C1
    rs = synth_mod().make_settings(device="cuda", **cam)
    sc = synth_mod().cloud_scene(P, cam["w"], cam["h"], cam["fx"], cam["fy"], seed=31)
    z = sc["means3D"][:, 2]
    zq = torch.where(z > 0.2, torch.clamp(torch.round(z / step), min=1.0) * step, z)   # everything in front of the near plane
    sc["means3D"] = torch.stack([sc["means3D"][:, 0] / z * zq, sc["means3D"][:, 1] / z * zq, zq], 1).contiguous()
    sc = {k: v.cuda() for k, v in sc.items()}
    dL = torch.randn(3, cam["h"], cam["w"], generator=torch.Generator().manual_seed(8)).cuda()
    ours = run_ours("dg", rs, sc, dL)
    ref = run_ref("dg", rs, sc, dL)
    lens = ours["views"]["ranges"][:, 1] - ours["views"]["ranges"][:, 0]
    assert int(lens.max()) > 2048
    vis = ours["radii"] > 0
    assert ours["views"]["depths"][vis].unique().numel() <= 64
    _compare_forward_exact(ours, ref, rs)
