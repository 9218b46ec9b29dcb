"""Loader + thin call wrappers for the UNMODIFIED reference CUDA extensions built by
oracle/build_ref.py into oracle/_ref/ (TEST INFRASTRUCTURE ONLY; never imported by the product).

The wrappers below are our own harness around the reference's pybind entry points
(`rasterize_gaussians`, `rasterize_gaussians_backward`, `mark_visible`: HX/ext.cpp:14-18); the
argument order follows HX/rasterize_points.h:18-65.  The reference's Python package
(`__init__.py`) is not used: the GPU box has no /root/reference, only the prebuilt .so files.
"""
import importlib.util
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
_cache = {}


def so_path(variant):
    return os.path.join(HERE, "_ref", "ref_%s_C.so" % variant)


def available(variant):
    return os.path.exists(so_path(variant))


def load(variant):
    """variant: 'hx' (fork, near 0.001, fused backward with power) or 'dg' (upstream, near 0.2)."""
    if variant not in _cache:
        name = "ref_%s_C" % variant
        spec = importlib.util.spec_from_file_location(name, so_path(variant))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cache[variant] = mod
    return _cache[variant]


def _e():
    return torch.Tensor([])


def forward(variant, rs, means3D, colors_precomp, opacities, scales, rotations, sh=None, cov3D_precomp=None):
    """Returns dict(num_rendered, color, radii, geom, binning, img, depth)."""
    mod = load(variant)
    args = (rs.bg, means3D, colors_precomp if colors_precomp is not None else _e(), opacities,
            scales if scales is not None else _e(), rotations if rotations is not None else _e(),
            rs.scale_modifier, cov3D_precomp if cov3D_precomp is not None else _e(), rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
            sh if sh is not None else _e(), rs.sh_degree, rs.campos, rs.prefiltered)
    n, color, radii, geom, binning, img, depth = mod.rasterize_gaussians(*args)
    return dict(num_rendered=n, color=color, radii=radii, geom=geom, binning=binning, img=img, depth=depth,
                inputs=(means3D, colors_precomp, opacities, scales, rotations, sh, cov3D_precomp), rs=rs,
                variant=variant)


def backward(fwd, grad_color, power=1):
    """Returns dict(means2D, colors, opacities, means3D, cov3D, sh, scales, rotations)."""
    mod = load(fwd["variant"])
    rs = fwd["rs"]
    means3D, colors, opac, scales, rots, sh, cov3D = fwd["inputs"]
    args = [rs.bg, means3D, fwd["radii"], colors if colors is not None else _e(),
            scales if scales is not None else _e(), rots if rots is not None else _e(), rs.scale_modifier,
            cov3D if cov3D is not None else _e(), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
            grad_color, sh if sh is not None else _e(), rs.sh_degree, rs.campos, fwd["geom"],
            fwd["num_rendered"], fwd["binning"], fwd["img"]]
    if fwd["variant"] == "hx":
        args.append(int(power))
    elif power != 1:
        raise ValueError("the DG variant has no backward_power")
    m2d, col, op, m3d, cov, dsh, sc, rot = mod.rasterize_gaussians_backward(*args)
    return dict(means2D=m2d, colors=col, opacities=op, means3D=m3d, cov3D=cov, sh=dsh, scales=sc, rotations=rot)


def image_state_views(fwd):
    """Carve final_T / n_contrib / ranges out of the reference's imgBuffer
    (ImageState::fromChunk, HX/cuda_rasterizer/rasterizer_impl.cu:171-179: three arrays of N
    elements, each 128-byte aligned: float accum_alpha, uint32 n_contrib, uint2 ranges)."""
    rs = fwd["rs"]
    N = rs.image_height * rs.image_width
    buf = fwd["img"]
    base = buf.data_ptr()

    def al(x):
        return (x + 127) // 128 * 128
    o0 = al(base) - base
    o1 = al(base + o0 + 4 * N) - base
    o2 = al(base + o1 + 4 * N) - base
    final_T = buf[o0:o0 + 4 * N].view(torch.float32)
    n_contrib = buf[o1:o1 + 4 * N].view(torch.int32)
    ranges = buf[o2:o2 + 8 * N].view(torch.int32).view(N, 2)
    return final_T, n_contrib, ranges


def binning_point_list(fwd):
    """point_list = first array of BinningState (rasterizer_impl.cu:181-194), uint32[R]."""
    R = fwd["num_rendered"]
    buf = fwd["binning"]
    base = buf.data_ptr()
    o0 = (base + 127) // 128 * 128 - base
    return buf[o0:o0 + 4 * R].view(torch.int32)


class _RefRasterize(torch.autograd.Function):
    """autograd edge around the reference extension's two pybind calls -- what the reference's own
    `_RasterizeGaussians` does (HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py:44-138), restated on
    the wrappers above so that the reference's SplaTAM iteration (torch glue + two renders) can be run and timed
    on the GPU box, where /root/reference does not exist."""
    @staticmethod
    def forward(ctx, variant, rs, means3D, means2D, colors_precomp, opacities, scales, rotations):
        f = forward(variant, rs, means3D.contiguous(), colors_precomp.contiguous(), opacities.contiguous(),
                    scales.contiguous(), rotations.contiguous())
        ctx.f = f
        ctx.mark_non_differentiable(f["radii"], f["depth"])
        return f["color"], f["radii"], f["depth"]

    @staticmethod
    def backward(ctx, g_color, _r, _d):
        g = backward(ctx.f, g_color.contiguous(), 1)
        ctx.f = None
        return None, None, g["means3D"], g["means2D"], g["colors"], g["opacities"], g["scales"], g["rotations"]


class RefRenderer:
    """`Renderer(raster_settings=cam)(**rendervar)` of scripts/splatam.py:255 on the unmodified reference extension."""
    variant = "dg"   # scripts/splatam.py:37 imports diff_gaussian_rasterization

    def __init__(self, raster_settings):
        self.rs = raster_settings

    def __call__(self, means3D, means2D, opacities, colors_precomp, scales, rotations):
        return _RefRasterize.apply(self.variant, self.rs, means3D, means2D, colors_precomp, opacities, scales, rotations)


def renderer_class(variant):
    return type("RefRenderer_%s" % variant, (RefRenderer,), {"variant": variant})
