"""CPU: pins oracle/slam_oracle.py (render-variable builder + image loss restatement) against
tests/golden/slam_glue_*.npz -- outputs of the reference's own Python run unmodified on CPU by
tests/golden/make_slam_golden.py."""
import os

import numpy as np
import pytest
import torch

from util import GOLDEN_DIR
from oracle import slam_oracle as SO


def _close(a, b, tol):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    scale = float(b.abs().max()) + 1e-30
    assert float((a - b).abs().max()) <= tol * scale, (float((a - b).abs().max()), scale)


@pytest.mark.parametrize("case", SO.RV_CASES, ids=lambda c: c["name"])
def test_rendervar_oracle_vs_reference_python(case):
    gold = np.load(os.path.join(GOLDEN_DIR, "slam_glue_rv_%s.npz" % case["name"]))
    params = {k: v.clone().requires_grad_(True) for k, v in SO.random_params(case["P"], 5, case["iso"], case["seed"]).items()}
    tg = SO.transform_to_frame(params, 3, case["gg"], case["cg"])
    rv = SO.transformed_params2rendervar(params, tg)
    ds = SO.get_depth_and_silhouette(tg["means3D"], torch.from_numpy(gold["w2c"]))
    g = torch.Generator().manual_seed(100 + case["seed"])
    seeds = {k: torch.randn(rv[k].shape, generator=g) for k in ("means3D", "rotations", "opacities", "scales")}
    seeds["depth_sil"] = torch.randn(ds.shape, generator=g)
    scalar = sum((rv[k] * seeds[k]).sum() for k in seeds if k != "depth_sil") + (ds * seeds["depth_sil"]).sum()
    scalar.backward()
    for k in ("means3D", "rotations", "opacities", "scales"):
        _close(rv[k].detach(), gold["rv_" + k], 2e-6)
    _close(ds.detach(), gold["depth_sil"], 2e-6)
    for k, v in params.items():
        got = v.grad if v.grad is not None else torch.zeros_like(v)
        _close(got, gold["g_" + k], 1e-5)
        if not case["cg"] and k.startswith("cam_"):
            assert float(got.abs().max()) == 0.0
        if not case["gg"] and k in ("means3D", "unnorm_rotations"):
            assert float(got.abs().max()) == 0.0


@pytest.mark.parametrize("case", SO.LOSS_CASES, ids=lambda c: c["name"])
def test_image_loss_oracle_vs_reference_get_loss(case):
    gold = np.load(os.path.join(GOLDEN_DIR, "slam_glue_loss_%s.npz" % case["name"]))
    im, ds, gt_im, gt_d = SO.images(case["h"], case["w_"], case["seed"])
    im.requires_grad_(True)
    ds.requires_grad_(True)
    loss, wl = SO.image_loss(im, ds, gt_im, gt_d, case["w"], case["use_sil"], case["sil_thres"], True, False,
                             case["tracking"])
    loss.backward()
    _close(loss.detach(), gold["loss"], 1e-6)
    _close(wl["im"].detach(), gold["l_im"], 1e-6)
    _close(wl["depth"].detach(), gold["l_depth"], 1e-6)
    _close(im.grad, gold["g_im"], 1e-5)
    _close(ds.grad, gold["g_depth_sil"], 1e-5)
