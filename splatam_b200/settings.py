"""GaussianRasterizationSettings -- the 11-field NamedTuple of the reference API
(HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py:140-151), in a module of its own so that scene / camera
helpers (synth.py, mapping.py) can be imported without mapping libsplat_b200.so (bench.py's reference arm does)."""
from typing import NamedTuple

import torch


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
