"""splatam_b200 -- B200-native (sm_100a) differentiable Gaussian-splat rasterizer.

Drop-in for the rasterizer extensions of fgarciacardenas/SplaTAM.  Import it under the
reference's names (`diff_gaussian_rasterization`, `hessian_diff_gaussian_rasterization_w_depth`;
both packages live at the repo root) or directly from here.

The rasterizer names resolve lazily (PEP 562): touching any of them imports `.rasterizer`, which maps
libsplat_b200.so and raises if it is missing -- there is no CPU fallback.  Plain helpers (`synth`, `mapping`,
`settings`) import without the library, so bench.py's reference arm never maps it.
"""
from .settings import GaussianRasterizationSettings

_LAZY = ("GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians", "NEAR_DG", "NEAR_HX")
__all__ = ["GaussianRasterizationSettings"] + list(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        from . import rasterizer
        return getattr(rasterizer, name)
    raise AttributeError("module %r has no attribute %r" % (__name__, name))
