"""splatam_b200 -- B200-native (sm_100a) differentiable Gaussian-splat rasterizer.

Drop-in for the rasterizer extensions of fgarciacardenas/SplaTAM.  Import it under the
reference's names (`diff_gaussian_rasterization`, `hessian_diff_gaussian_rasterization_w_depth`;
both packages live at the repo root) or directly from here.
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,
                         _RasterizeGaussians, NEAR_DG, NEAR_HX)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
           "_RasterizeGaussians", "NEAR_DG", "NEAR_HX"]
