"""Generates tests/golden/densify_*.npz by running the reference's OWN Python, unmodified, on CPU:
  * utils/slam_external.py: prune_gaussians (-> remove_points, update_params_and_optimizer) on a real
    torch.optim.Adam whose state was seeded, so the Adam-moment surgery is part of the golden output;
  * scripts/splatam.py: add_new_gaussians / get_pointcloud / initialize_new_params, cut out of the file with `ast`
    (the module itself is not importable here: cv2, wandb, datasets ...) and exec'd in a namespace that holds the
    reference helpers and a stub `Renderer` returning the seeded silhouette render.
`.cuda()` / device="cuda" are redirected to the CPU for the run (device placement only, no arithmetic changes).

Run HERE (needs /root/reference):  python tests/golden/make_densify_golden.py
Inputs are regenerated from seeds (oracle/densify_oracle.py::random_state / frame / *_CASES); outputs are stored.
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

from oracle import densify_oracle as DO  # noqa: E402  (only for the seeded input generators)


def _redirect_cuda():
    torch.Tensor.cuda = lambda self, *a, **k: self
    for name in ("zeros", "zeros_like", "ones", "eye"):
        orig = getattr(torch, name)

        def wrap(*a, _orig=orig, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k.pop("device")
            return _orig(*a, **k)
        setattr(torch, name, wrap)


def _optimizer(params, state):
    """initialize_optimizer (scripts/splatam.py:166-172) + seeded Adam moments."""
    lrs = {k: 1e-3 for k in params}
    groups = [{"params": [v], "name": k, "lr": lrs[k]} for k, v in params.items()]
    opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    for k, v in params.items():
        if k in state:
            opt.state[v] = {"step": torch.tensor(5.0), "exp_avg": state[k]["exp_avg"].clone(),
                            "exp_avg_sq": state[k]["exp_avg_sq"].clone()}
    return opt


def main():
    _redirect_cuda()
    sys.path.insert(0, REF)
    from utils import slam_helpers as RH
    from utils import slam_external as RE

    # ---- prune_gaussians on a live optimizer -------------------------------------------------------------
    for c in DO.PRUNE_CASES:
        p0, s0, v0 = DO.random_state(c["P"], c["iso"], c["seed"])
        params = {k: torch.nn.Parameter(v.clone()) for k, v in p0.items()}
        opt = _optimizer(params, s0)
        variables = {k: v.clone() for k, v in v0.items()}
        params, variables = RE.prune_gaussians(params, variables, opt, c["iter"], c["pd"])
        out = {}
        for k in DO.GAUSS_KEYS:
            out["p_" + k] = params[k].detach().numpy()
            st = opt.state[params[k]]
            out["m_" + k] = st["exp_avg"].numpy()
            out["v_" + k] = st["exp_avg_sq"].numpy()
        for k in DO.VAR_KEYS:
            out["var_" + k] = variables[k].numpy()
        np.savez_compressed(os.path.join(HERE, "densify_prune_%s.npz" % c["name"]), **out)
        print("wrote prune", c["name"], c["P"], "->", out["p_means3D"].shape[0])

    # ---- add_new_gaussians, cut out of scripts/splatam.py ------------------------------------------------
    src = open(os.path.join(REF, "scripts", "splatam.py")).read()
    body = ast.parse(src).body
    code = "\n\n".join(ast.get_source_segment(src, n) for n in body if isinstance(n, ast.FunctionDef)
                       and n.name in ("get_pointcloud", "initialize_new_params", "add_new_gaussians"))

    class Renderer:                                 # stub: hands back the seeded silhouette render
        result = None

        def __init__(self, raster_settings=None):
            pass

        def __call__(self, **kw):
            return Renderer.result, None, None

    ns = {"torch": torch, "np": np, "Renderer": Renderer, "transform_to_frame": RH.transform_to_frame,
          "transformed_params2depthplussilhouette": RH.transformed_params2depthplussilhouette,
          "build_rotation": RE.build_rotation}
    exec(code, ns)
    for c in DO.ADD_CASES:
        p0, _, v0 = DO.random_state(c["P"], c["iso"], c["seed"])
        im, gt_depth, depth_sil, intr = DO.frame(c["h"], c["w"], c["seed"])
        params = {k: torch.nn.Parameter(v.clone()) for k, v in p0.items()}
        variables = {k: v.clone() for k, v in v0.items()}
        Renderer.result = depth_sil
        curr = {"cam": None, "im": im, "depth": gt_depth, "intrinsics": intr, "w2c": torch.eye(4)}
        params, variables = ns["add_new_gaussians"](params, variables, curr, c["sil_thres"], c["time_idx"], "projective",
                                                    "isotropic" if c["iso"] else "anisotropic")
        out = {"p_" + k: params[k].detach().numpy() for k in DO.GAUSS_KEYS}
        for k in DO.VAR_KEYS:
            out["var_" + k] = variables[k].numpy()
        np.savez_compressed(os.path.join(HERE, "densify_add_%s.npz" % c["name"]), **out)
        print("wrote add", c["name"], c["P"], "->", out["p_means3D"].shape[0])


if __name__ == "__main__":
    main()
