"""Shared helpers for the parity tests (scene construction, tolerance metrics)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from splatam_b200 import synth  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# name -> (camera dict, scene builder kwargs).  Small enough for the CPU oracle to finish in seconds.
CASES = {
    "c1_cloud": dict(cam=synth.C1, kind="cloud", P=10000, seed=0),
    "cloud_small": dict(cam=synth.C1, kind="cloud", P=1500, seed=5),
    "cloud_ragged": dict(cam=dict(w=75, h=53, fx=70.0, fy=66.0, cx=36.2, cy=27.1), kind="cloud", P=3000, seed=7),
    "surface_small": dict(cam=dict(w=160, h=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5), kind="surface", P=6000, seed=3,
                          aniso_frac=0.2),
}


def build_case(name, device="cpu"):
    c = CASES[name]
    cam = c["cam"]
    rs = synth.make_settings(device=device, **cam)
    if c["kind"] == "cloud":
        sc = synth.cloud_scene(c["P"], cam["w"], cam["h"], cam["fx"], cam["fy"], seed=c["seed"])
    else:
        sc = synth.surface_scene(c["P"], cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"],
                                 seed=c["seed"], aniso_frac=c.get("aniso_frac", 0.0))
    sc = {k: v.to(device) for k, v in sc.items()}
    g = torch.Generator().manual_seed(1000 + c["seed"])
    dL = torch.randn(3, cam["h"], cam["w"], generator=g).to(device)
    return rs, sc, dL


def rel_err(a, b, floor_frac=1e-6):
    """max |a-b| / (max(|a|,|b|) + floor), floor = floor_frac * ||b||_inf  (SURVEY.md 7.3)."""
    a, b = a.double().flatten(), b.double().flatten()
    floor = floor_frac * float(b.abs().max()) + 1e-30
    return float(((a - b).abs() / (torch.maximum(a.abs(), b.abs()) + floor)).max())


def frac_bad(a, b, tol, floor_frac=1e-6):
    a, b = a.double().flatten(), b.double().flatten()
    floor = floor_frac * float(b.abs().max()) + 1e-30
    bad = (a - b).abs() > tol * (torch.maximum(a.abs(), b.abs()) + floor)
    return float(bad.double().mean())
