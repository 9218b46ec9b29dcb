"""Host-side mirror of the SplaTAM glue on either side of the rasterizer, on top of
libsplat_b200's `splat_rendervar_*` and `splat_slam_loss` entry points.  Same function names, argument meaning
and return values as the reference:

  transform_to_frame + transformed_params2rendervar  (utils/slam_helpers.py:253-304, :123-139)  -> build_rendervar
  get_loss                                           (scripts/splatam.py:220-353)              -> get_loss

`get_loss` here = fused builder -> ONE six-plane render (GaussianRasterizer.render_rgbd_sil) -> fused loss that hands
dL/d(render) straight to the rasterizer's backward; the reference runs ~45 small torch kernels, two rasterizer passes and
a second autograd sweep for the same quantities.  Patch point: `from splatam_b200.slam import get_loss` in
scripts/splatam.py.  PyTorch is plumbing only (memory, streams, autograd edges); no CPU fallback.

Restrictions (raise, never silently differ): use_l1=True and ignore_outlier_depth_loss=False (every shipped config,
configs/*/splatam.py) and no matplotlib visualisation.  curr_data['w2c'] must be the camera's view matrix (it is:
both are first_frame_w2c, scripts/splatam.py:189-200) because the fused render takes depth from the view matrix.
"""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import check, lib
from .rasterizer import SINK_ATTR, _f32, _on_device


# A/B switch for measurements only: SPLAT_TORCH_BOOKKEEPING=1 issues the original torch statements instead of the
# two fused bookkeeping kernels (same results; still on the GPU)
_FUSED_BOOKKEEPING = os.environ.get("SPLAT_TORCH_BOOKKEEPING", "0") != "1"


def _vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ------------------------------------------------------------------------------------------------ 8f-2
class _BuildRendervar(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, unnorm_rotations, logit_opacities, log_scales, cam_unnorm_rots, cam_trans, time_idx,
                gaussians_grad, camera_grad):
        # gradient sinks (mapping.GradBucket(direct=True, layout=mapping.raw_layout())): the backward kernel writes a
        # parameter's gradient straight into the flat exchange bucket and autograd receives None for it
        ctx.sinks = tuple(getattr(t, SINK_ATTR, None) for t in (means3D, unnorm_rotations, logit_opacities, log_scales))
        for sk, t in zip(ctx.sinks, (means3D, unnorm_rotations, logit_opacities, log_scales)):
            if sk is not None and (sk.dtype != torch.float32 or not sk.is_contiguous() or sk.shape != t.shape
                                   or sk.device != t.device or sk.data_ptr() % 16 and t.shape[1] == 4):
                raise RuntimeError("gradient sink must be a contiguous float32 tensor of the parameter's shape")
        means3D = _f32(means3D, "means3D")
        unnorm_rotations = _f32(unnorm_rotations, "unnorm_rotations")
        logit_opacities = _f32(logit_opacities, "logit_opacities")
        log_scales = _f32(log_scales, "log_scales")
        cam_unnorm_rots = _f32(cam_unnorm_rots, "cam_unnorm_rots")
        cam_trans = _f32(cam_trans, "cam_trans")
        P = means3D.shape[0]
        sd = log_scales.shape[1]
        F = cam_unnorm_rots.shape[-1]
        if cam_unnorm_rots.numel() != 4 * F or cam_trans.numel() != 3 * F:
            raise RuntimeError("cam_unnorm_rots / cam_trans must be [1,4,F] / [1,3,F]")
        if unnorm_rotations.data_ptr() % 16:
            unnorm_rotations = unnorm_rotations.clone()
        dev = means3D.device
        with _on_device(dev):
            flat = torch.empty(P * 11, dtype=torch.float32, device=dev)     # [rotations 4 | means 3 | scales 3 | opacity 1]
            rots = flat[:4 * P].view(P, 4)
            means = flat[4 * P:7 * P].view(P, 3)
            scales = flat[7 * P:10 * P].view(P, 3)
            opac = flat[10 * P:].view(P, 1)
            check(lib.splat_rendervar_forward(P, sd, _vp(means3D), _vp(unnorm_rotations), _vp(logit_opacities),
                                              _vp(log_scales), _vp(cam_unnorm_rots), _vp(cam_trans), F, int(time_idx),
                                              _vp(means), _vp(rots), _vp(opac), _vp(scales), _stream()))
        ctx.save_for_backward(means3D, unnorm_rotations, logit_opacities, log_scales, cam_unnorm_rots, cam_trans)
        ctx.cfg = (P, sd, F, int(time_idx), bool(gaussians_grad), bool(camera_grad))
        ctx.set_materialize_grads(False)
        # an output is differentiable only if something it depends on wants a gradient: the rasterizer's backward
        # skips the gradient arrays of inputs that do not (pose-only tracking: everything but the means)
        need = ctx.needs_input_grad
        cam = bool(camera_grad) and (need[4] or need[5])
        dead = []
        if not need[2]:
            dead.append(opac)
        if not need[3]:
            dead.append(scales)
        if not ((gaussians_grad and need[1]) or (sd == 3 and cam)):
            dead.append(rots)
        if not ((gaussians_grad and need[0]) or cam):
            dead.append(means)
        if dead:
            ctx.mark_non_differentiable(*dead)
        return means, rots, opac, scales

    @staticmethod
    def backward(ctx, g_means, g_rots, g_op, g_scales):
        means3D, unnorm_rotations, logit_opacities, log_scales, cam_unnorm_rots, cam_trans = ctx.saved_tensors
        P, sd, F, t, gg, cg = ctx.cfg
        need = ctx.needs_input_grad
        gg = gg and (need[0] or need[1])
        cg = cg and (need[4] or need[5])
        dev = means3D.device
        g_means, g_rots, g_op, g_scales = [None if g is None else _f32(g, "grad") for g in (g_means, g_rots, g_op, g_scales)]
        if g_rots is not None and g_rots.data_ptr() % 16:
            g_rots = g_rots.clone()
        sk_m, sk_r, sk_o, sk_s = ctx.sinks
        with _on_device(dev):
            own = lambda need_it, sk: need_it and sk is None     # noqa: E731
            n = P * ((4 if own(gg, sk_r) else 0) + (3 if own(gg, sk_m) else 0) + (1 if own(need[2], sk_o) else 0)
                     + (sd if own(need[3], sk_s) else 0))
            flat = torch.empty(max(n, 1), dtype=torch.float32, device=dev)
            off = 0
            d_rots = d_means = d_logit = d_ls = None
            if gg:
                d_rots = sk_r
                if d_rots is None:
                    d_rots = flat[:4 * P].view(P, 4)
                    off = 4 * P
                d_means = sk_m
                if d_means is None:
                    d_means = flat[off:off + 3 * P].view(P, 3)
                    off += 3 * P
            if need[2]:
                d_logit = sk_o
                if d_logit is None:
                    d_logit = flat[off:off + P].view(P, 1)
                    off += P
            if need[3]:
                d_ls = sk_s
                if d_ls is None:
                    d_ls = flat[off:off + P * sd].view(P, sd)
            d_cr = d_ct = scratch = None
            nscr = 0
            if cg:
                pose = torch.zeros(7 * F, dtype=torch.float32, device=dev)
                d_cr, d_ct = pose[:4 * F].view(1, 4, F), pose[4 * F:].view(1, 3, F)
                nscr = lib.splat_rendervar_scratch_bytes(P)
                scratch = torch.empty(nscr, dtype=torch.uint8, device=dev)
            check(lib.splat_rendervar_backward(P, sd, _vp(means3D), _vp(unnorm_rotations), _vp(logit_opacities),
                                               _vp(log_scales), _vp(cam_unnorm_rots), _vp(cam_trans), F, t,
                                               _vp(g_means), _vp(g_rots), _vp(g_op), _vp(g_scales), int(gg), int(cg),
                                               _vp(d_means), _vp(d_rots), _vp(d_logit), _vp(d_ls), _vp(d_cr), _vp(d_ct),
                                               _vp(scratch), nscr, _stream()))
        return (d_means if need[0] and sk_m is None else None, d_rots if need[1] and sk_r is None else None,
                d_logit if sk_o is None else None, d_ls if sk_s is None else None,
                d_cr if need[4] else None, d_ct if need[5] else None, None, None, None)


def build_rendervar(params, time_idx, gaussians_grad, camera_grad, pose_only=False):
    """transform_to_frame + transformed_params2rendervar in one kernel (utils/slam_helpers.py:253-304, :123-139).
    Returns the reference's rendervar dict (means3D, colors_precomp, rotations, opacities, scales, means2D).

    pose_only (opt-in, tracking): the reference lets colour / opacity / scale gradients flow during tracking although its
    tracking optimizer has lr = 0 for them and is discarded after the frame (scripts/splatam.py:700,
    configs/*/splatam.py `tracking.lrs`).  With pose_only those inputs are detached, so the backward kernels form the
    pose gradient only; cam_unnorm_rots.grad / cam_trans.grad are unchanged, the Gaussian .grad fields stay None."""
    det = (lambda t: t.detach()) if pose_only else (lambda t: t)
    means, rots, opac, scales = _BuildRendervar.apply(
        params['means3D'], params['unnorm_rotations'], det(params['logit_opacities']), det(params['log_scales']),
        params['cam_unnorm_rots'], params['cam_trans'], time_idx, gaussians_grad, camera_grad)
    return {'means3D': means, 'colors_precomp': det(params['rgb_colors']), 'rotations': rots, 'opacities': opac,
            'scales': scales,
            # a leaf that only ever receives a gradient (the reference adds `+ 0` and calls retain_grad(), :135)
            'means2D': torch.zeros_like(params['means3D'], requires_grad=not pose_only)}


# ------------------------------------------------------------------------------------------------ 8f-3
def _loss_call(dev, H, W, tracking, use_sil, sil_thres, w_im, w_depth, im, depth_sil, gt_im, gt_depth, g_im, g_ds):
    losses = torch.empty(4, dtype=torch.float32, device=dev)
    nscr = lib.splat_slam_loss_scratch_bytes(W, H, int(tracking))
    scratch = torch.empty(nscr, dtype=torch.uint8, device=dev)
    check(lib.splat_slam_loss(W, H, int(tracking), int(use_sil), float(sil_thres), float(w_im), float(w_depth),
                              _vp(im), _vp(depth_sil), _vp(gt_im), _vp(gt_depth), _vp(losses), _vp(g_im), _vp(g_ds),
                              _vp(scratch), nscr, _stream()))
    return losses


def _check_images(render, nplanes, gt_im, gt_depth):
    render = _f32(render, "rendered image")
    gt_im = _f32(gt_im, "curr_data['im']")
    gt_depth = _f32(gt_depth, "curr_data['depth']")
    H, W = render.shape[-2], render.shape[-1]
    if render.shape[0] != nplanes or gt_im.numel() != 3 * H * W or gt_depth.numel() != H * W:
        raise RuntimeError("image shapes do not match: render %s, im %s, depth %s"
                           % (tuple(render.shape), tuple(gt_im.shape), tuple(gt_depth.shape)))
    return render, gt_im, gt_depth, H, W


def _scale_grad(grad, g_loss):
    """grad *= g_loss without a pass over the image in the usual case g_loss == 1 (decided on the device)."""
    if _FUSED_BOOKKEEPING and g_loss.dtype == torch.float32 and g_loss.numel() == 1 and g_loss.is_cuda:
        with _on_device(grad.device):
            check(lib.splat_scale_inplace(_vp(grad), grad.numel(), _vp(g_loss), _stream()))
        return grad
    return grad.mul_(g_loss)


class _ImageLoss6(torch.autograd.Function):
    """Loss on the six-plane fused render: planes 0-2 = im, 3-5 = (depth, silhouette, depth^2)."""
    @staticmethod
    def forward(ctx, render, gt_im, gt_depth, tracking, use_sil, sil_thres, w_im, w_depth):
        render, gt_im, gt_depth, H, W = _check_images(render, 6, gt_im, gt_depth)
        dev = render.device
        with _on_device(dev):
            grad = torch.empty_like(render)
            losses = _loss_call(dev, H, W, tracking, use_sil, sil_thres, w_im, w_depth, render[:3], render[3:], gt_im,
                                gt_depth, grad[:3], grad[3:])
        ctx.grad = grad
        ctx.mark_non_differentiable(losses)
        return losses[2], losses

    @staticmethod
    def backward(ctx, g_loss, _g_losses):
        grad, ctx.grad = ctx.grad, None
        return _scale_grad(grad, g_loss), None, None, None, None, None, None, None


class _ImageLoss33(torch.autograd.Function):
    """Loss on the two images of the reference flow: im [3,H,W] and depth_sil [3,H,W] (two Renderer passes)."""
    @staticmethod
    def forward(ctx, im, depth_sil, gt_im, gt_depth, tracking, use_sil, sil_thres, w_im, w_depth):
        im, gt_im, gt_depth, H, W = _check_images(im, 3, gt_im, gt_depth)
        depth_sil = _f32(depth_sil, "depth_sil")
        if tuple(depth_sil.shape) != (3, H, W):
            raise RuntimeError("depth_sil must be [3,%d,%d]" % (H, W))
        dev = im.device
        with _on_device(dev):
            grad = torch.empty(6, H, W, dtype=torch.float32, device=dev)
            losses = _loss_call(dev, H, W, tracking, use_sil, sil_thres, w_im, w_depth, im, depth_sil, gt_im, gt_depth,
                                grad[:3], grad[3:])
        ctx.grad = grad
        ctx.mark_non_differentiable(losses)
        return losses[2], losses

    @staticmethod
    def backward(ctx, g_loss, _g_losses):
        grad, ctx.grad = ctx.grad, None
        _scale_grad(grad, g_loss)
        return grad[:3], grad[3:], None, None, None, None, None, None, None


def image_loss(im, depth_sil, gt_im, gt_depth, loss_weights, use_sil_for_loss, sil_thres, tracking):
    """Image-space part of get_loss (scripts/splatam.py:262-296,342-346).  `im` may be the six-plane fused render with
    depth_sil=None.  Returns (loss, {'im':, 'depth':, 'loss':}) as device scalars; loss.backward() delivers the
    gradient the kernel already formed (scaled by the incoming gradient)."""
    args = (bool(tracking), bool(use_sil_for_loss), float(sil_thres), float(loss_weights['im']),
            float(loss_weights['depth']))
    if depth_sil is None:
        loss, losses = _ImageLoss6.apply(im, gt_im, gt_depth, *args)
    else:
        loss, losses = _ImageLoss33.apply(im, depth_sil, gt_im, gt_depth, *args)
    return loss, {'im': losses[0], 'depth': losses[1], 'loss': loss}


def _default_renderer():
    from diff_gaussian_rasterization import GaussianRasterizer   # what scripts/splatam.py:37 imports
    return GaussianRasterizer


def get_loss(params, curr_data, variables, iter_time_idx, loss_weights, use_sil_for_loss,
             sil_thres, use_l1, ignore_outlier_depth_loss, tracking=False,
             mapping=False, do_ba=False, plot_dir=None, visualize_tracking_loss=False, tracking_iteration=None,
             Renderer=None, tracking_pose_only=False):
    """Same signature and return value as the reference's get_loss (scripts/splatam.py:220-353).  Two trailing
    extras: Renderer (rasterizer class, default the diff_gaussian_rasterization shim) and tracking_pose_only (see
    build_rendervar: skip the Gaussian-side gradients the reference computes and discards while tracking)."""
    if not use_l1 or ignore_outlier_depth_loss:
        raise NotImplementedError("fused get_loss covers use_l1=True, ignore_outlier_depth_loss=False (all shipped configs)")
    if visualize_tracking_loss:
        raise NotImplementedError("visualize_tracking_loss needs matplotlib/cv2 windows: unavailable on this box")
    if tracking:
        gg, cg = False, True
    elif mapping and do_ba:
        gg, cg = True, True
    else:
        gg, cg = True, False
    rendervar = build_rendervar(params, iter_time_idx, gaussians_grad=gg, camera_grad=cg,
                                pose_only=bool(tracking and tracking_pose_only))
    Renderer = Renderer or _default_renderer()
    out, radius, _ = Renderer(raster_settings=curr_data['cam']).render_rgbd_sil(**rendervar)
    variables['means2D'] = rendervar['means2D']
    # NOTE (deviation): the reference gives its depth/silhouette pass a separate means2D, so
    # only the COLOUR pass feeds means2D.grad (utils/slam_helpers.py:196-249).  The fused six-plane render has one
    # means2D: its gradient also carries the depth and silhouette planes' screen-space terms.  Nothing in SplaTAM's
    # shipped configs reads it (use_gaussian_splatting_densification=False); a caller that enables GS-style
    # densification (utils/slam_external.py:101-102 accumulates means2D.grad) should use the two-pass drop-in path.
    variables['means2D_grad_source'] = 'rgb+depth+silhouette planes (fused render)'
    loss, weighted_losses = image_loss(out, None, curr_data['im'], curr_data['depth'], loss_weights,
                                       use_sil_for_loss, sil_thres, tracking)
    # == `seen = radius > 0; max_2D_radius[seen] = max(radius[seen], max_2D_radius[seen])` (:348-350) without the
    # boolean-index host synchronisation: radius is 0 where not seen and max_2D_radius >= 0
    m2r = variables['max_2D_radius']
    if _FUSED_BOOKKEEPING and m2r.dtype == torch.float32 and m2r.is_contiguous() and radius.dtype == torch.int32 \
            and radius.is_contiguous():
        seen = torch.empty(radius.shape, dtype=torch.bool, device=radius.device)      # one launch for all three ops
        with _on_device(radius.device):
            check(lib.splat_seen_update(radius.numel(), _vp(radius), _vp(m2r), _vp(seen), _stream()))
        variables['seen'] = seen
    else:
        torch.maximum(m2r, radius.to(m2r.dtype), out=m2r)
        variables['seen'] = radius > 0
    return loss, variables, weighted_losses


# ------------------------------------------------------------------------------------------------ optimizer step
class FusedAdam(torch.optim.Optimizer):
    """`torch.optim.Adam(param_groups[, lr=0.0, eps=1e-15])` as SplaTAM builds it (initialize_optimizer,
    scripts/splatam.py:166-172: one param group per tensor with its own lr) with the whole step in ONE kernel launch
    (`splat_adam_step`) and ~40 us of host code; torch's foreach path costs 0.3-0.8 ms per iteration here, more than
    the fused tracking iteration itself.  Same constructor arguments, same `param_groups` / `state[p]` layout
    (`step`, `exp_avg`, `exp_avg_sq`), so the reference's optimizer surgery (cat_params_to_optimizer, remove_points,
    update_params_and_optimizer: utils/slam_external.py:107-161) works on it unchanged.  weight_decay / amsgrad /
    maximize are not supported (SplaTAM never sets them).

    CUDA-graph capture: a step() issued while the stream is capturing uses a device-side replay counter for the
    bias corrections, so each replay is one more optimizer step; GraphedIteration brackets the capture with
    begin_capture() / end_capture() and brings the host-side `state[p]['step']` up to date in check()."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("FusedAdam: weight_decay / amsgrad are not supported")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))
        self._desc = (_lib.SplatAdamTensor * 16)()
        self._replay = {}        # device -> (int32 counter, float32 scratch), allocated outside any capture
        self._captured = []      # state dicts whose step advanced inside the captured step

    def _replay_buffers(self, dev):
        if dev not in self._replay:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("FusedAdam: run one un-captured step() before capturing (allocates the replay counter)")
            self._replay[dev] = (torch.zeros(1, dtype=torch.int32, device=dev), torch.zeros(32, dtype=torch.float32, device=dev))
        return self._replay[dev]

    def begin_capture(self):
        """Called (eagerly) right before a step() is captured: replays count from zero for the new graph."""
        for counter, _ in self._replay.values():
            counter.zero_()

    def end_capture(self):
        """Called right after the capture: the captured step() advanced the host-side step counters although nothing
        executed -- undo that, and hand the caller the state dicts the captured step touches."""
        states, self._captured = self._captured, []
        for st in states:
            st['step'] -= 1
        return states

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        capturing = torch.cuda.is_current_stream_capturing()
        if capturing:
            self._captured = []
        batches = {}
        for group in self.param_groups:
            key = (float(group['betas'][0]), float(group['betas'][1]), float(group['eps']))
            lr = float(group['lr'])
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous():
                    raise RuntimeError("FusedAdam needs contiguous float32 CUDA parameters")
                state = self.state[p]
                if len(state) == 0:
                    state['step'] = torch.tensor(0.0)
                    state['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state['step'] += 1
                if capturing:
                    self._captured.append(state)
                if not g.is_contiguous():
                    g = g.contiguous()
                batches.setdefault((p.device, key), []).append((p, g, state['exp_avg'], state['exp_avg_sq'], lr,
                                                                int(state['step'])))
        for (dev, (b1, b2, eps)), items in batches.items():
            with _on_device(dev):
                counter, scratch = self._replay_buffers(dev)
                stream = _stream()
                for i0 in range(0, len(items), 16):
                    chunk = items[i0:i0 + 16]
                    if capturing and len(items) > 16:
                        raise RuntimeError("FusedAdam: at most 16 tensors per (device, betas, eps) inside a CUDA graph")
                    d = self._desc
                    for i, (p, g, m, v, lr, t) in enumerate(chunk):
                        e = d[i]
                        e.param, e.grad, e.exp_avg, e.exp_avg_sq = p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr()
                        e.n, e.lr, e.step = p.numel(), lr, t
                    check(lib.splat_adam_step(d, len(chunk), b1, b2, eps, _vp(counter) if capturing else None,
                                              _vp(scratch) if capturing else None, stream))
        return loss


# ------------------------------------------------------------------------------------------------ CUDA-graph replay
_GRAPH_POOL = {}
_GRAPH_KEEP = {}


class CaptureOverflow(RuntimeError):
    """A replayed iteration rendered more instances than the binning capacity its graph was captured with.  Every kernel
    of such a replay that depends on the binning state exits untouched (forward AND backward blend: the gradients of that
    step are all zero, never read out of bounds), so nothing is corrupted -- but the steps taken since are not the ones an
    eager loop would have taken: restore the state saved before the frame and re-run it un-captured."""


class GraphedIteration:
    """Captures ONE optimisation iteration -- `fn()` = get_loss + loss.backward() + FusedAdam.step(), no zero_grad --
    into a CUDA graph and replays it: the 40-200 tracking iterations SplaTAM runs on one frame (scripts/splatam.py:
    710-758) are launch- and host-bound once the kernels take 0.65 ms, and differ only in the values of the pose
    parameters, which the captured optimizer step updates in place.

    What makes the iteration capturable: the rasterizer's forward does not wait for num_rendered while capturing (the
    binning capacity is fixed from the warm-up iterations; `splat_forward_guard` records on the device whether any
    replay exceeded it -> `check()`), and FusedAdam keeps its step count on the device.  Everything `fn` allocates lives in
    the graph's private pool; gradients are written (not accumulated) because capture starts with `.grad = None`.
    `fn` must return a tensor or a tuple of tensors (e.g. the loss): the same objects are refreshed by every replay."""

    def __init__(self, fn, params, optimizer=None, warmup=1):
        from . import rasterizer
        self.optimizer, self.replays, self._adam_states = optimizer, 0, []
        tensors = list(params.values()) if isinstance(params, dict) else list(params)
        dev = tensors[0].device
        self.device = dev
        # Warm-up = ordinary eager iterations on the current stream (they count as optimisation iterations): they size
        # the binning capacity for this frame and allocate optimizer state / replay counters.  (On a side stream the
        # caching allocator would start from an empty per-stream pool: measured 4 ms of cudaMalloc per frame.)
        for _ in range(max(1, warmup)):
            for t in tensors:
                t.grad = None
            fn()
        for t in tensors:
            t.grad = None
        rasterizer.capture_guard(dev).zero_()
        if optimizer is not None and hasattr(optimizer, "begin_capture"):
            optimizer.begin_capture()
        # capture_begin / capture_end directly: torch.cuda.graph() would run gc.collect() + empty_cache() on entry
        # (tens of milliseconds -- as much as the 40 iterations of a Replica frame) and give every graph a fresh private
        # pool (cudaMalloc per capture); all GraphedIterations of a device share one pool and one capture stream.
        if dev not in _GRAPH_POOL:
            _GRAPH_POOL[dev] = (torch.cuda.graph_pool_handle(), torch.cuda.Stream(dev))
        pool, side = _GRAPH_POOL[dev]
        cur = torch.cuda.current_stream(dev)
        side.wait_stream(cur)
        self.graph = torch.cuda.CUDAGraph()
        warn = torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch
        warn(False)   # the parameters' AccumulateGrad nodes were created by eager iterations on another stream: harmless here
        try:
            with torch.cuda.stream(side):
                self.graph.capture_begin(pool=pool)
                try:
                    self.out = fn()
                finally:
                    self.graph.capture_end()
        finally:
            warn(True)
        cur.wait_stream(side)
        # the shared pool lives only while some graph uses it: keep the most recent graph of the device alive until
        # the next one has been captured (a caller that drops its GraphedIteration between frames would otherwise
        # free the pool and the next capture_begin on the same handle asserts inside the caching allocator)
        _GRAPH_KEEP[dev] = self.graph
        if optimizer is not None and hasattr(optimizer, "end_capture"):
            self._adam_states = optimizer.end_capture()

    def replay(self):
        self.graph.replay()
        self.replays += 1
        return self.out

    __call__ = replay

    def check(self):
        """Synchronises; raises if any replay needed more binning capacity than the capture had (its results are then
        undefined -- re-run those iterations un-captured); brings the optimizer's host-side step counts up to date."""
        from . import rasterizer
        torch.cuda.synchronize(self.device)
        for st in self._adam_states:      # every replay was one optimizer step
            st['step'] += self.replays
        self.replays = 0
        flags = rasterizer.capture_guard(self.device)
        if int(flags[0]) != 0:
            flags[0] = 0
            raise CaptureOverflow("a replayed iteration rendered %d instances, more than the captured binning capacity"
                                  % int(flags[1]))


class BestPose:
    """The tracking loop's candidate-pose bookkeeping (scripts/splatam.py:701-704, :726-731, :761-764) without its
    per-iteration host synchronisation: the reference evaluates `loss < current_min_loss` on the host (a blocking
    read of a CUDA scalar every iteration); here the comparison and both conditional copies stay on the device
    (three tiny kernels, capturable inside a GraphedIteration).  Call `update(loss, params)` right after
    optimizer.step() -- like the reference it stores the pose AFTER the step whose pre-step loss was the minimum."""

    def __init__(self, params, time_idx):
        self.t = int(time_idx)
        self.rot = params['cam_unnorm_rots'][..., self.t].detach().clone().contiguous()
        self.tran = params['cam_trans'][..., self.t].detach().clone().contiguous()
        self.min_loss = torch.full((), 1e20, dtype=torch.float32, device=self.rot.device)

    @torch.no_grad()
    def update(self, loss, params):
        loss = loss.detach()
        cr, ct = params['cam_unnorm_rots'], params['cam_trans']
        if _FUSED_BOOKKEEPING and loss.dtype == torch.float32 and cr.is_contiguous() and ct.is_contiguous() \
                and cr.dtype == torch.float32:
            with _on_device(loss.device):   # one one-warp kernel instead of a compare, three where() and three copies
                check(lib.splat_best_pose_update(_vp(loss), _vp(self.min_loss), _vp(cr), _vp(ct), cr.shape[-1], self.t,
                                                 _vp(self.rot), _vp(self.tran), _stream()))
            return
        better = loss < self.min_loss
        self.min_loss.copy_(torch.where(better, loss, self.min_loss))
        self.rot.copy_(torch.where(better, cr[..., self.t], self.rot))
        self.tran.copy_(torch.where(better, ct[..., self.t], self.tran))

    @torch.no_grad()
    def write_back(self, params):
        params['cam_unnorm_rots'][..., self.t] = self.rot
        params['cam_trans'][..., self.t] = self.tran
