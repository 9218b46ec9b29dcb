"""Import-name shim: `from diff_gaussian_rasterization import GaussianRasterizer,
GaussianRasterizationSettings` (scripts/splatam.py:37, utils/recon_helpers.py:2 of the reference).

Selects the semantics of the upstream submodule DG/diff_gaussian_rasterization/__init__.py:
near plane 0.2 (DG/cuda_rasterizer/auxiliary.h:154) and the plain two-stage backward; there is
no backward_power argument in this variant.
"""
from splatam_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer as _Base,
                                     rasterize_gaussians as _rasterize, _RasterizeGaussians, NEAR_DG)


class GaussianRasterizer(_Base):
    NEAR_PLANE = NEAR_DG
    FUSED_BACKWARD = False

    def __init__(self, raster_settings):
        super().__init__(raster_settings, backward_power=1)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _rasterize(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                      raster_settings, 1, NEAR_DG, False)


__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]
