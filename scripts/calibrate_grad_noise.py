"""Run-to-run and cross-implementation gradient noise table (GPU box).  Used to choose the
absolute floor of the 1e-3 gradient tolerance: reference-vs-reference (atomic order) sets the
noise level any other implementation is allowed."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import CASES, build_case, rel_err  # noqa: E402
from test_gpu_parity import run_ours, run_ref  # noqa: E402

FLOORS = tuple(float(x) for x in os.environ.get("FLOORS", "1e-6,1e-5,1e-4,1e-3").split(","))
KEYS = ("means3D", "means2D", "colors", "opacities", "scales", "rotations")
for case in CASES:
    for variant, power in (("dg", 1), ("hx", 1), ("hx", 2)):
        rs, sc, dL = build_case(case, "cuda")
        if power == 2:
            dL = torch.ones_like(dL) * 1e-3
        r1 = run_ref(variant, rs, sc, dL, power)["grads"]
        r2 = run_ref(variant, rs, sc, dL, power)["grads"]
        o1 = run_ours(variant, rs, sc, dL, power)["grads"]
        o2 = run_ours(variant, rs, sc, dL, power)["grads"]
        for floor in FLOORS:
            row = "%-14s %s p%d floor=%g |" % (case, variant, power, floor)
            for k in KEYS:
                row += " %s rr=%.1e oo=%.1e or=%.1e |" % (k[:6], rel_err(r1[k], r2[k], floor), rel_err(o1[k], o2[k], floor),
                                                         rel_err(o1[k], r1[k], floor))
            print(row)
