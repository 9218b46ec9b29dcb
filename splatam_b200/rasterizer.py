"""Host-side mirror of the reference rasterizer's Python API on top of libsplat_b200.

Same names, argument meaning, exception texts and gradient order as
HX/hessian_diff_gaussian_rasterization_w_depth/__init__.py:17-204 (and the DG twin,
DG/diff_gaussian_rasterization/__init__.py), so `scripts/splatam.py` and `scripts/ros_handler.py`
run unchanged.  PyTorch is only plumbing here: it owns device memory and the stream; every kernel
is in libsplat_b200.so and is reached through the C ABI (splatam_b200/_lib.py).

What is different from the reference host code, by design:
  * no blocking `cudaMemcpy` of num_rendered in the middle of the forward
    (HX/cuda_rasterizer/rasterizer_impl.cu:281-282): phase 2 is launched speculatively into a
    binning buffer sized from the previous call's R; the host only waits for the 4-byte R that
    phase 1 produced (the GPU is already busy with phase 2 by then) and relaunches phase 2 in the
    rare case the capacity was exceeded;
  * one flat gradient allocation per backward instead of nine `torch::zeros`
    (HX/rasterize_points.cu:151-159).
"""
import ctypes as C
import time
import threading
import torch
import torch.nn as nn

from . import _lib
from ._lib import SplatCamera, check, lib
from .settings import GaussianRasterizationSettings  # noqa: F401  (re-exported: the reference API's name)

NEAR_DG = 0.2     # DG/cuda_rasterizer/auxiliary.h:154
NEAR_HX = 0.001   # HX/cuda_rasterizer/auxiliary.h:154



_tls = threading.local()
# (device index, H, W) -> slowly decaying maximum of the recent num_rendered values.  The speculative binning capacity is
# derived from it and rounded up to _CAP_STEP instances, so consecutive calls (the 24 keyframes of a mapping window have
# 24 different R) ask the caching allocator for the SAME block size instead of a new one every step, and a caller that
# alternates a small and a large map at one image size keeps the large capacity instead of re-running phase 2.
_capacity_hint = {}
_CAP_GROWTH, _CAP_SLACK, _CAP_STEP, _CAP_DECAY = 1.25, 65536, 1 << 20, 0.98


def _spec_capacity(hint):
    cap = int(hint * _CAP_GROWTH) + _CAP_SLACK
    return (cap + _CAP_STEP - 1) // _CAP_STEP * _CAP_STEP


def _sticky_capacity(key, want):
    """The capacity actually allocated: the previous one for this key as long as it still covers `want` and is not more
    than twice too large.  Without the hysteresis the slowly decaying hint crosses a _CAP_STEP boundary every few dozen
    calls, the allocator is asked for a new block size, and a cudaMalloc (0.5 - 35 ms measured) lands in the caller's
    loop -- bench.py's 20-step mapping leg caught exactly one of those at its second timed step."""
    bkey = ("block",) + key
    cur = _capacity_hint.get(bkey)
    if cur is None or want > cur or want * 2 < cur:
        cur = want
        _capacity_hint[bkey] = cur
    return cur


def _pinned_scalar(device):
    cache = getattr(_tls, "pinned", None)
    if cache is None:
        cache = _tls.pinned = {}
    key = device.index
    if key not in cache:
        cache[key] = torch.zeros(1, dtype=torch.int32).pin_memory()
    return cache[key]


_guard = {}


def capture_guard(device):
    """Pinned uint32[2] per device: [0] = some captured forward exceeded its binning capacity, [1] = largest
    num_rendered seen by a captured forward (written by splat_forward_guard; see slam.GraphedIteration)."""
    key = torch.device(device).index
    if key not in _guard:
        _guard[key] = torch.zeros(2, dtype=torch.int32).pin_memory()
    return _guard[key]


def _f32(t, name):
    if t.dtype != torch.float32:
        raise RuntimeError("%s must be float32, got %s" % (name, t.dtype))
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor" % name)
    return t.contiguous()


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NULL = _NullCtx()


def _on_device(dev):
    """torch.cuda.device(dev) only when dev is not already current (the context manager costs ~15 us per use)."""
    return _NULL if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def _camera(rs, near, keep, depth_sil=False):
    cam = SplatCamera()
    cam.depth_sil = 1 if depth_sil else 0
    cam.image_height = int(rs.image_height)
    cam.image_width = int(rs.image_width)
    cam.tanfovx = float(rs.tanfovx)
    cam.tanfovy = float(rs.tanfovy)
    bg = _f32(rs.bg, "bg")
    vm = _f32(rs.viewmatrix, "viewmatrix")
    pm = _f32(rs.projmatrix, "projmatrix")
    keep += [bg, vm, pm]
    cam.bg = bg.data_ptr()
    cam.scale_modifier = float(rs.scale_modifier)
    cam.viewmatrix = vm.data_ptr()
    cam.projmatrix = pm.data_ptr()
    cam.sh_degree = int(rs.sh_degree)
    if rs.campos is not None and rs.campos.numel() >= 3 and rs.campos.is_cuda:
        cp = _f32(rs.campos, "campos")
        keep.append(cp)
        cam.campos = cp.data_ptr()
    else:
        cam.campos = None
    cam.prefiltered = int(bool(rs.prefiltered))
    cam.near_plane = float(near)
    return cam


class RasterState:
    """Opaque per-call state saved for the backward pass (the reference's geomBuffer /
    binningBuffer / imgBuffer byte tensors, HX/rasterize_points.cu:73-78)."""
    __slots__ = ("geom", "image", "binning", "capacity", "num_rendered", "P", "H", "W", "cam", "keep", "plan")

    def _view(self, buf, ptr, nbytes, dtype):
        off = ptr - buf.data_ptr()
        return buf[off:off + nbytes].view(dtype)

    def views(self):
        """Typed views into the state for the parity tests (splat_state_* of the C ABI):
        ranges int32[T,2], n_contrib int32[H,W], final_T float32[H,W], point_list int32[R],
        instances float32[R,12], and (strided, rows with radii > 0 only) depths [P], means2D [P,2],
        conic_opacity [P,4], colors [P,3]."""
        W, H, P, R = self.W, self.H, self.P, self.num_rendered
        T = ((W + 15) // 16) * ((H + 15) // 16)
        img, geo, bn = self.image, self.geom, self.binning
        ip, gp, bp = C.c_void_p(img.data_ptr()), C.c_void_p(geo.data_ptr()), C.c_void_p(bn.data_ptr())
        grec = self._view(geo, lib.splat_state_gaussian_records(gp, P), 48 * P, torch.float32).view(P, 12)
        return dict(
            ranges=self._view(img, lib.splat_state_ranges(ip, W, H), 8 * T, torch.int32).view(T, 2),
            n_contrib=self._view(img, lib.splat_state_n_contrib(ip, W, H), 4 * W * H, torch.int32).view(H, W),
            final_T=self._view(img, lib.splat_state_final_T(ip, W, H), 4 * W * H, torch.float32).view(H, W),
            point_list=self._view(bn, lib.splat_state_point_list(bp, W, H, self.capacity), 4 * R, torch.int32),
            instances=self._view(bn, lib.splat_state_instance_records(bp, W, H, self.capacity), 48 * R,
                                 torch.float32).view(R, 12),
            depths=grec[:, 6], means2D=grec[:, 0:2], conic_opacity=grec[:, 2:6], colors=grec[:, 8:11])


# Debug / parity tests only: with keep_last_state = True the most recent RasterState is kept in last_state (it pins that
# call's geometry / image / binning buffers, ~200 MB at Replica size, until the next call).  Off by default.
keep_last_state = False
last_state = None


def rasterize_forward(rs, near, means3D, colors_precomp, opacities, scales, rotations, cov3D_precomp, sh, bwd=None,
                      depth_sil=False):
    """Runs both forward phases.  Returns (color, radii, depth, RasterState).  bwd = (power, fused, sinks) when a
    backward pass will follow: its buffers are then prepared here, ahead of the blocking read-back."""
    global last_state
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    P = means3D.shape[0]
    H, W = int(rs.image_height), int(rs.image_width)
    st = RasterState()
    st.P, st.H, st.W = P, H, W
    st.plan = None
    if P == 0:  # HX/rasterize_points.cu:81 -- outputs stay at their fill value 0
        st.geom = st.image = st.binning = torch.empty(0, dtype=torch.uint8, device=dev)
        st.capacity = st.num_rendered = 0
        return (torch.zeros(6 if depth_sil else 3, H, W, device=dev), torch.zeros(0, dtype=torch.int32, device=dev),
                torch.zeros(1, H, W, device=dev), st)
    keep = []
    cam = _camera(rs, near, keep, depth_sil)
    st.cam, st.keep = cam, keep   # the backward pass reuses the struct (and keeps its tensors alive)
    with _on_device(dev):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        M = sh.shape[1] if sh.numel() else 0
        out_color = torch.empty(6 if depth_sil else 3, H, W, dtype=torch.float32, device=dev)
        out_depth = torch.empty(1, H, W, dtype=torch.float32, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        st.geom = torch.empty(lib.splat_geom_state_bytes(P), dtype=torch.uint8, device=dev)
        st.image = torch.empty(lib.splat_image_state_bytes(W, H), dtype=torch.uint8, device=dev)
        # R is written by the scan kernel STRAIGHT into pinned host memory (UVA: the pinned page is device-addressable).
        # A cudaMemcpyAsync of the word would queue behind whatever bulk device->host copy is in flight on the copy
        # engine (e.g. the previous step's 60 MB gradient read-back) and stall the host for its whole duration.
        pinned = _pinned_scalar(dev)
        check(lib.splat_forward_geometry(C.byref(cam), P, _ptr(means3D), _ptr(scales), _ptr(rotations),
                                         _ptr(opacities), _ptr(cov3D_precomp), _ptr(sh), M,
                                         _ptr(colors_precomp), _ptr(st.geom), _ptr(st.image), _ptr(radii),
                                         C.c_void_p(pinned.data_ptr()), stream))
        capturing = torch.cuda.is_current_stream_capturing()
        if not capturing:
            ev = torch.cuda.Event()
            ev.record()

        def phase2(cap):
            st.capacity = cap
            st.binning = torch.empty(lib.splat_binning_state_bytes(W, H, cap), dtype=torch.uint8, device=dev)
            check(lib.splat_forward_render(C.byref(cam), P, _ptr(colors_precomp), _ptr(radii), _ptr(st.geom),
                                           _ptr(st.image), _ptr(st.binning), cap, _ptr(out_color),
                                           _ptr(out_depth), stream))

        key = (dev.index, H, W)
        hint = _capacity_hint.get(key)
        if capturing:
            # CUDA-graph capture: no host wait for num_rendered.  Phase 2 runs with the capacity the warm-up iterations
            # established; the guard kernel records (device side) whether a replay ever exceeds it.
            if hint is None:
                raise RuntimeError("run one un-captured forward at this image size before capturing a CUDA graph")
            phase2(_sticky_capacity(key, _spec_capacity(hint)))
            check(lib.splat_forward_guard(_ptr(st.image), W, H, st.capacity, C.c_void_p(capture_guard(dev).data_ptr()),
                                          stream))
            if bwd is not None:
                st.plan = _plan_backward(st, dev, bwd[0], bwd[1], means3D, colors_precomp, scales, rotations,
                                         cov3D_precomp, sh, radii, bwd[2], bwd[3] if len(bwd) > 3 else frozenset())
            st.num_rendered = hint
            if keep_last_state:
                last_state = st
            return out_color, radii, out_depth, st
        launched = False
        if hint is not None:
            phase2(_sticky_capacity(key, _spec_capacity(hint)))
            launched = True
        if bwd is not None:
            st.plan = _plan_backward(st, dev, bwd[0], bwd[1], means3D, colors_precomp, scales, rotations,
                                     cov3D_precomp, sh, radii, bwd[2], bwd[3] if len(bwd) > 3 else frozenset())
        t_w = time.perf_counter()
        ev.synchronize()
        host_stats["wait_s"] += time.perf_counter() - t_w
        host_stats["forwards"] += 1
        R = int(pinned.item())
        _capacity_hint[key] = max(R, int(hint * _CAP_DECAY)) if hint is not None else R
        if not launched or R > st.capacity:
            phase2(_sticky_capacity(key, _spec_capacity(R)))
        st.num_rendered = R
    if keep_last_state:
        last_state = st
    return out_color, radii, out_depth, st


# Host-side accounting: seconds the host spent blocked on the num_rendered read-back.  A step whose wait is ~0
# means the GPU was waiting for the host, not the other way round (bench.py reports the per-step mean).
host_stats = {"wait_s": 0.0, "forwards": 0}

SINK_ATTR = "_splat_grad_sink"
_SINK_SHAPES = {"means3D": 3, "colors_precomp": 3, "opacities": 1, "scales": 3, "rotations": 4}


def _collect_sinks(**named):
    """Gradient sinks (opt-in, mapping.GradBucket): an input tensor carrying a `_splat_grad_sink` attribute gets
    its gradient WRITTEN (assigned, not accumulated -- every row is written by the backward kernels) straight
    into that buffer, and autograd receives None for it: no AccumulateGrad add, no pre-zeroing."""
    sinks = {}
    for k, t in named.items():
        sk = getattr(t, SINK_ATTR, None)
        if sk is None:
            continue
        w = _SINK_SHAPES[k]
        if (sk.dtype != torch.float32 or not sk.is_contiguous() or tuple(sk.shape) != (t.shape[0], w)
                or sk.device != t.device or sk.data_ptr() % (16 if w == 4 else 4)):
            raise RuntimeError("gradient sink of %s must be a contiguous float32 [P,%d] tensor on the same device" % (k, w))
        sinks[k] = sk
    return sinks or None


def _plan_backward(st, dev, power, fused, means3D, colors_precomp, scales, rotations, cov3D_precomp, sh, radii, sinks,
                   skip=frozenset()):
    """Everything of the backward call that does not depend on dL/dcolour: the flat gradient/scratch allocation, the
    returned views and the packed C arguments.  rasterize_forward runs this BEFORE it blocks on the num_rendered
    read-back, so that after the wake-up the host only has to hand the backward launch to the (then busy) GPU."""
    P = st.P
    M = sh.shape[1] if sh.numel() else 0
    hess = bool(fused) and int(power) != 1
    n_scratch = lib.splat_backward_scratch_bytes(P, 2 if hess else 1) // 4
    sinks = sinks if (sinks and P) else {}
    # `skip`: gradient arrays nobody asked for (autograd's needs_input_grad) are neither allocated nor written
    # (splat_backward takes NULL for them); only the plain power-1, non-SH backward supports that
    if hess or M:
        skip = frozenset()
    if skip:
        skip = skip | {"conic"}
    # one flat allocation: [scratch | rotations 4 | conic 4 | means3D 3 | means2D 3 | colors 3 | opacity 1 |
    # scales 3 | cov3D 6 | sh 3M]; float4-accessed arrays first so they stay 16-byte aligned for any P (torch
    # allocations are 512-byte aligned).  Arrays that have a sink are not carved at all.
    layout = (("rotations", 4, (P, 4)), ("conic", 4, None), ("means3D", 3, (P, 3)), ("means2D", 3, (P, 3)),
              ("colors_precomp", 3, (P, 3)), ("opacities", 1, (P, 1)), ("scales", 3, (P, 3)), ("cov3D", 6, (P, 6)),
              ("sh", 3 * M, (P, M, 3)))
    total = n_scratch + sum(P * k for name, k, _ in layout if name not in sinks and name not in skip)
    flat = torch.empty(total, dtype=torch.float32, device=dev)
    base = flat.data_ptr()
    ptrs, outs = {}, {}
    off = n_scratch
    for name, k, shape in layout:
        if name in sinks:
            ptrs[name] = sinks[name].data_ptr()
            continue
        if name in skip:
            ptrs[name] = 0
            continue
        ptrs[name] = base + 4 * off if P * k else 0
        if shape is not None:
            outs[name] = flat[off:off + P * k].view(shape)
        off += P * k
    result = (outs.get("means3D"), outs.get("means2D"), outs.get("sh"), outs.get("colors_precomp"), outs.get("opacities"),
              outs.get("scales"), outs.get("rotations"), outs.get("cov3D"))
    vp = C.c_void_p
    head = (_ptr(means3D), _ptr(scales), _ptr(rotations), _ptr(cov3D_precomp), _ptr(sh), M, _ptr(colors_precomp),
            _ptr(radii))
    tail = (int(power), int(bool(fused)), vp(ptrs["means3D"]), vp(ptrs["means2D"]), vp(ptrs["colors_precomp"]),
            vp(ptrs["opacities"]), vp(ptrs["scales"]), vp(ptrs["rotations"]), vp(ptrs["cov3D"]), vp(ptrs["conic"]),
            vp(ptrs["sh"]), vp(base), n_scratch * 4)
    return dict(flat=flat, result=result, head=head, tail=tail, key=(int(power), bool(fused), skip))


def rasterize_backward(rs, near, st, power, fused, means3D, colors_precomp, scales, rotations,
                       cov3D_precomp, sh, radii, grad_out_color, sinks=None, out_color=None, skip=frozenset()):
    dev = means3D.device
    P = st.P
    with _on_device(dev):
        plan = st.plan
        st.plan = None   # a plan's buffers are handed to autograd once
        if plan is None or plan["key"][:2] != (int(power), bool(fused)):
            plan = _plan_backward(st, dev, power, fused, means3D, colors_precomp, scales, rotations, cov3D_precomp, sh,
                                  radii, sinks, skip)
        if P == 0:
            return plan["result"]
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        grad = _f32(grad_out_color, "grad_out_color")
        check(lib.splat_backward(C.byref(st.cam), P, st.capacity, *plan["head"], _ptr(st.geom), _ptr(st.image),
                                 _ptr(st.binning), _ptr(grad), _ptr(out_color), *plan["tail"], stream))
    return plan["result"]


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, backward_power, near, fused, sinks=None, depth_sil=False):
        # the reference passes tensors through .contiguous() (HX/rasterize_points.cu:94-112)
        means3D = _f32(means3D, "means3D")
        opacities = _f32(opacities, "opacities")
        sh = sh.contiguous()
        colors_precomp = colors_precomp.contiguous()
        scales = scales.contiguous()
        rotations = rotations.contiguous()
        if rotations.numel() and rotations.data_ptr() % 16:
            rotations = rotations.clone()   # the kernels read quaternions as float4
        cov3Ds_precomp = cov3Ds_precomp.contiguous()
        for name, t in (("sh", sh), ("colors_precomp", colors_precomp), ("scales", scales),
                        ("rotations", rotations), ("cov3D_precomp", cov3Ds_precomp)):
            if t.numel() and (t.dtype != torch.float32 or not t.is_cuda):
                raise RuntimeError("%s must be a float32 CUDA tensor" % name)
        # inputs that do not require a gradient (e.g. everything but the means in pose-only tracking) are skipped by
        # the backward kernels; means3D's gradient is always formed
        need = ctx.needs_input_grad
        skip = frozenset(n for n, i in (("means2D", 1), ("colors_precomp", 3), ("opacities", 4), ("scales", 5),
                                        ("rotations", 6), ("cov3D", 7)) if not need[i])
        if not {"colors_precomp", "opacities"} <= skip:
            skip = skip - {"colors_precomp", "opacities"}   # the blend backward drops these sums only together
        ctx.skip = skip
        bwd = (backward_power, fused, sinks, skip) if any(need) else None
        color, radii, depth, st = rasterize_forward(raster_settings, near, means3D, colors_precomp, opacities,
                                                    scales, rotations, cov3Ds_precomp, sh, bwd, depth_sil)
        ctx.depth_sil = depth_sil
        ctx.raster_settings = raster_settings
        ctx.num_rendered = st.num_rendered
        ctx.backward_power = backward_power
        ctx.state = st
        ctx.near = near
        ctx.fused = fused
        ctx.sinks = sinks
        # the rendered image conditions the blend backward's sums (splat_backward's optional out_color).  A detached
        # alias, not save_for_backward: a later in-place edit of the image by the caller is harmless (any values
        # are valid reference levels) and must not trip autograd's version check.
        ctx.out_color = color.detach() if bwd is not None else None
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh)
        # radii and depth are not differentiable in the reference (only grad_out_color is read,
        # HX/.../__init__.py:91-123): do not let autograd materialise zero gradients for them every backward
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_depth):
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh = ctx.saved_tensors
        nch = 6 if ctx.depth_sil else 3
        if grad_out_color is None:
            grad_out_color = torch.zeros(nch, ctx.state.H, ctx.state.W, device=means3D.device)
        if tuple(grad_out_color.shape) != (nch, ctx.state.H, ctx.state.W):
            raise RuntimeError("grad of the rendered image must be [%d,%d,%d]" % (nch, ctx.state.H, ctx.state.W))
        (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D) = rasterize_backward(
            ctx.raster_settings, ctx.near, ctx.state, ctx.backward_power, ctx.fused, means3D, colors_precomp,
            scales, rotations, cov3Ds_precomp, sh, radii, grad_out_color, ctx.sinks, ctx.out_color, ctx.skip)
        ctx.out_color = None
        sunk = ctx.sinks or {}
        return (g_means3D if "means3D" not in sunk else None, g_means2D,
                g_sh if sh.numel() else None,
                g_colors if colors_precomp.numel() and "colors_precomp" not in sunk else None,
                g_opac if "opacities" not in sunk else None,
                g_scales if scales.numel() and "scales" not in sunk else None,
                g_rots if rotations.numel() and "rotations" not in sunk else None,
                g_cov3D if cov3Ds_precomp.numel() else None,
                None, None, None, None, None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, backward_power=1, near=NEAR_HX, fused=True, depth_sil=False):
    if depth_sil and int(backward_power) != 1:
        raise RuntimeError("render_depth_sil supports backward_power == 1 only")
    sinks = _collect_sinks(means3D=means3D, colors_precomp=colors_precomp, opacities=opacities, scales=scales,
                           rotations=rotations)
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, backward_power, near, fused, sinks, bool(depth_sil))


class GaussianRasterizer(nn.Module):
    """Drop-in for both reference `GaussianRasterizer` classes.  `NEAR_PLANE` / `FUSED_BACKWARD` are
    set by the import-name packages (diff_gaussian_rasterization: 0.2 / DG two-stage backward;
    hessian_diff_gaussian_rasterization_w_depth: 0.001 / HX fused semantics with backward_power)."""
    NEAR_PLANE = NEAR_HX
    FUSED_BACKWARD = True

    def __init__(self, raster_settings, backward_power: int = 1):
        super().__init__()
        self.raster_settings = raster_settings
        self.backward_power = backward_power

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            positions = _f32(positions, "positions")
            P = positions.shape[0]
            present = torch.zeros(P, dtype=torch.bool, device=positions.device)
            if P:
                vm = _f32(rs.viewmatrix, "viewmatrix")
                pm = _f32(rs.projmatrix, "projmatrix")
                with torch.cuda.device(positions.device):
                    check(lib.splat_mark_visible(P, _ptr(positions), _ptr(vm), _ptr(pm), float(self.NEAR_PLANE),
                                                 _ptr(present), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        """Same call as the reference (HX/.../__init__.py:170-204)."""
        return self._render(means3D, means2D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, False)

    def render_rgbd_sil(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                        cov3D_precomp=None):
        """Extension, same arguments as forward(): the returned image has SIX planes -- planes 0..2 the
        colours, planes 3..5 exactly what a second pass with colors_precomp = get_depth_and_silhouette(means3D, w2c) =
        (z, 1, z*z) renders (utils/slam_helpers.py:124-139, scripts/splatam.py:257-259) -- from ONE preprocess / sort /
        alpha evaluation; the gradient of planes 3..5 reaches means3D through z.  backward_power must be 1."""
        return self._render(means3D, means2D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, True)

    def _render(self, means3D, means2D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, depth_sil):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        if shs is None:
            shs = empty
        if colors_precomp is None:
            colors_precomp = empty
        if scales is None:
            scales = empty
        if rotations is None:
            rotations = empty
        if cov3D_precomp is None:
            cov3D_precomp = empty
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings, self.backward_power, self.NEAR_PLANE,
                                   self.FUSED_BACKWARD, depth_sil)
