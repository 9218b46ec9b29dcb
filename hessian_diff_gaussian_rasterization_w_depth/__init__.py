"""Import-name shim: `from hessian_diff_gaussian_rasterization_w_depth import GaussianRasterizer`
(scripts/ros_handler.py:21, scripts/splatam_realtime.py:37 of the reference).

Selects the fork's semantics (HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py): near
plane 0.001 (HX/cuda_rasterizer/auxiliary.h:154) and `backward_power` (1 = gradient, 2 = the
diagonal Gauss-Newton / "Hessian" accumulation of renderCUDAFused).
"""
from splatam_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer as _Base,
                                     rasterize_gaussians as _rasterize, _RasterizeGaussians, NEAR_HX)


class GaussianRasterizer(_Base):
    NEAR_PLANE = NEAR_HX
    FUSED_BACKWARD = True


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, backward_power):
    return _rasterize(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                      raster_settings, backward_power, NEAR_HX, True)


__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]
