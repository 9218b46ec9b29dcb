"""Generates tests/golden/slam_glue_*.npz by running the reference's OWN Python, unmodified, on CPU:
  * utils/slam_helpers.py  (transform_to_frame, transformed_params2rendervar, get_depth_and_silhouette)
  * utils/slam_external.py (build_rotation, calc_ssim)
  * the `get_loss` function of scripts/splatam.py -- the module itself is not importable here (cv2, wandb,
    datasets ...), so the function's source is cut out of the file with `ast` and exec'd in a namespace that
    holds the reference helpers and a stub `Renderer` returning two leaf images (so the image-space
    gradients dL/dim and dL/ddepth_sil of the reference code can be read off).
The reference code hard-codes `.cuda()` / device="cuda"; those are redirected to the CPU for the duration of
the run (device placement only, no arithmetic changes).

Run HERE (needs /root/reference):  python tests/golden/make_slam_golden.py
Inputs are regenerated from seeds (oracle/slam_oracle.py::random_params / images / *_CASES); outputs are stored.
"""
import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from oracle import slam_oracle as SO  # noqa: E402  (only for the seeded input generators)


def _redirect_cuda():
    torch.Tensor.cuda = lambda self, *a, **k: self
    for name in ("zeros", "zeros_like", "ones", "eye"):
        orig = getattr(torch, name)

        def wrap(*a, _orig=orig, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k.pop("device")
            return _orig(*a, **k)
        setattr(torch, name, wrap)


def main():
    _redirect_cuda()
    sys.path.insert(0, REF)
    from utils import slam_helpers as RH          # the reference's own module
    from utils import slam_external as RE

    # ---- render-variable builder -----------------------------------------------------------------
    for c in SO.RV_CASES:
        params = SO.random_params(c["P"], 5, c["iso"], c["seed"])
        params = {k: v.clone().requires_grad_(True) for k, v in params.items()}
        t = 3
        tg = RH.transform_to_frame(params, t, gaussians_grad=c["gg"], camera_grad=c["cg"])
        rv = RH.transformed_params2rendervar(params, tg)
        w2c = torch.eye(4)
        w2c[:3, :3] = RE.build_rotation(torch.tensor([[0.9, 0.1, -0.2, 0.05]]))[0]
        w2c[:3, 3] = torch.tensor([0.1, -0.2, 0.3])
        ds = RH.get_depth_and_silhouette(tg["means3D"], w2c)
        g = torch.Generator().manual_seed(100 + c["seed"])
        seeds = {k: torch.randn(rv[k].shape, generator=g) for k in ("means3D", "rotations", "opacities", "scales")}
        seeds["depth_sil"] = torch.randn(ds.shape, generator=g)
        scalar = sum((rv[k] * seeds[k]).sum() for k in ("means3D", "rotations", "opacities", "scales")) \
            + (ds * seeds["depth_sil"]).sum()
        scalar.backward()
        out = {"rv_" + k: rv[k].detach().numpy() for k in ("means3D", "rotations", "opacities", "scales")}
        out["depth_sil"] = ds.detach().numpy()
        out["w2c"] = w2c.numpy()
        for k, v in params.items():
            out["g_" + k] = (v.grad if v.grad is not None else torch.zeros_like(v)).numpy()
        np.savez_compressed(os.path.join(HERE, "slam_glue_rv_%s.npz" % c["name"]), **out)
        print("wrote rv", c["name"])

    # ---- get_loss, cut out of scripts/splatam.py ---------------------------------------------------
    src = open(os.path.join(REF, "scripts", "splatam.py")).read()
    fn = [n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "get_loss"][0]
    code = ast.get_source_segment(src, fn)

    class Renderer:                                 # stub: hands back the two leaf images in call order
        queue = []

        def __init__(self, raster_settings=None):
            pass

        def __call__(self, **kw):
            return Renderer.queue.pop(0), torch.ones(kw["means3D"].shape[0], dtype=torch.int32), None

    ns = dict(torch=torch, transform_to_frame=RH.transform_to_frame,
              transformed_params2rendervar=RH.transformed_params2rendervar,
              transformed_params2depthplussilhouette=RH.transformed_params2depthplussilhouette,
              l1_loss_v1=RH.l1_loss_v1, calc_ssim=RE.calc_ssim, Renderer=Renderer)
    exec(compile(code, "reference:get_loss", "exec"), ns)
    get_loss = ns["get_loss"]
    for c in SO.LOSS_CASES:
        im, ds, gt_im, gt_d = SO.images(c["h"], c["w_"], c["seed"])
        im.requires_grad_(True)
        ds.requires_grad_(True)
        Renderer.queue = [im, ds]
        params = SO.random_params(64, 2, True, 5)
        variables = {"max_2D_radius": torch.zeros(64)}
        curr = {"cam": None, "im": gt_im, "depth": gt_d, "w2c": torch.eye(4)}
        loss, _, wl = get_loss(params, curr, variables, 0, c["w"], c["use_sil"], c["sil_thres"], True, False,
                               tracking=c["tracking"], mapping=not c["tracking"])
        loss.backward()
        np.savez_compressed(os.path.join(HERE, "slam_glue_loss_%s.npz" % c["name"]),
                            loss=loss.detach().numpy(), l_im=wl["im"].detach().numpy(),
                            l_depth=wl["depth"].detach().numpy(), g_im=im.grad.numpy(), g_depth_sil=ds.grad.numpy())
        print("wrote loss", c["name"], float(loss))


if __name__ == "__main__":
    main()
