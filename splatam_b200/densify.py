"""Host-side mirror of SplaTAM's per-frame map surgery on top of libsplat_b200's
`splat_prune_flags` / `splat_compact_rows` / `splat_new_gaussian_mask` / `splat_spawn_gaussians`.  Same function
names, argument meaning and return values as the reference:

  remove_points, prune_gaussians, cat_params_to_optimizer-style growth   (utils/slam_external.py:123-195)
  add_new_gaussians (+ get_pointcloud, initialize_new_params)           (scripts/splatam.py:73-123, 354-426)

The reference boolean-indexes / torch.cat's every parameter, both Adam moments and four per-Gaussian variables one
tensor at a time: ~20 nonzero()+gather launches, as many host syncs and fresh allocations per prune, a full
reallocation of the map per added frame.  Here every per-Gaussian tensor lives in a `GaussianArena` -- two
capacity-doubling buffers per tensor -- and

  prune  = 1 flag kernel + 1 count + 1 scan + ONE compaction launch over all tensors into the other buffer,
           one 4-byte read-back of the new P; the tensors handed back are views of the arena (no allocation);
  grow   = exact median (radix select) + mask + scan, one read-back of the count, ONE kernel that back-projects
           the flagged pixels and writes the new rows in place behind row P.

Rows are copied, never recomputed: the surviving parameters, moments and variables are bit-identical to the
reference's.  Works with torch.optim.Adam and slam.FusedAdam alike (same param_groups / state layout).  PyTorch
is plumbing only; there is no CPU fallback.  GS-style `densify` (clone / split, utils/slam_external.py:198-257) is
switched off in every SplaTAM config (use_gaussian_splatting_densification=False) and is not provided.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, lib
from .rasterizer import _on_device

GAUSS_KEYS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
VAR_KEYS = ("means2D_gradient_accum", "denom", "max_2D_radius", "timestep")
ARENA_KEY = "_splat_arena"        # where the arena rides along: variables[ARENA_KEY]


def _vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class GaussianArena:
    """Two float32 buffers [capacity, width] per per-Gaussian tensor.  `adopt` moves a tensor in (a copy the first
    time, free afterwards: the views handed back by prune / grow are arena memory already); `flip` makes the other
    buffer current after a compaction."""

    def __init__(self, device, growth=2.0, min_capacity=16384):
        self.device = torch.device(device)
        self.growth, self.min_capacity = float(growth), int(min_capacity)
        self.capacity = 0
        self.bufs, self.cur = {}, {}
        self._scratch = None
        self._flags = None
        self._counts = torch.zeros(2, dtype=torch.int32).pin_memory()
        self._desc = (_lib.SplatRowTensor * 24)()
        self.reallocations = 0

    # -- storage ----------------------------------------------------------------------------------------------
    def reserve(self, rows):
        """Capacity for at least `rows` rows in every registered tensor (contents of the current buffers kept)."""
        if rows <= self.capacity:
            return
        cap = max(self.min_capacity, int(rows * self.growth))
        for name, pair in self.bufs.items():
            c = self.cur[name]
            new = torch.empty(cap, pair[c].shape[1], dtype=torch.float32, device=self.device)
            new[:self.capacity].copy_(pair[c])
            pair[c] = new
            pair[1 - c] = None
        self.capacity = cap
        self.reallocations += 1

    def _pair(self, name, width):
        if name not in self.bufs:
            self.bufs[name] = [torch.empty(self.capacity, width, dtype=torch.float32, device=self.device), None]
            self.cur[name] = 0
        pair = self.bufs[name]
        if pair[self.cur[name]].shape[1] != width:
            raise RuntimeError("arena tensor %s changed width" % name)
        return pair

    def adopt(self, name, t):
        """Make the arena's current buffer of `name` hold `t` ([P] or [P,w], float32 CUDA)."""
        P = t.shape[0]
        width = 1 if t.dim() == 1 else t.shape[1]
        self.reserve(P)
        buf = self._pair(name, width)[self.cur[name]]
        if t.data_ptr() != buf.data_ptr() or not t.is_contiguous():
            if t.dtype != torch.float32 or not t.is_cuda:
                raise RuntimeError("%s must be a float32 CUDA tensor" % name)
            buf[:P].copy_(t.detach().reshape(P, width))
        return buf

    def current(self, name):
        return self.bufs[name][self.cur[name]]

    def other(self, name):
        pair, c = self.bufs[name], self.cur[name]
        if pair[1 - c] is None or pair[1 - c].shape[0] != self.capacity:
            pair[1 - c] = torch.empty(self.capacity, pair[c].shape[1], dtype=torch.float32, device=self.device)
        return pair[1 - c]

    def flip(self, name):
        self.cur[name] ^= 1

    def view(self, name, P, one_d):
        buf = self.current(name)
        return buf[:P].view(P) if one_d else buf[:P]

    def scratch(self, nbytes):
        if self._scratch is None or self._scratch.numel() < nbytes:
            self._scratch = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self._scratch

    def flags(self, n):
        if self._flags is None or self._flags.numel() < n:
            self._flags = torch.empty(int(n * 1.25) + 1024, dtype=torch.uint8, device=self.device)
        return self._flags


def _arena(variables, device):
    a = variables.get(ARENA_KEY)
    if a is None or a.device != torch.device(device):
        a = variables[ARENA_KEY] = GaussianArena(device)
    return a


def _group(optimizer, name):
    return [g for g in optimizer.param_groups if g.get("name") == name][0]


def _state_of(optimizer, k):
    if optimizer is None:
        return None
    st = optimizer.state.get(_group(optimizer, k)["params"][0], None)
    return st if st is not None and "exp_avg" in st else None


def _swap_param(params, optimizer, k, new_data, moments):
    """What remove_points / cat_params_to_optimizer do to one tensor (utils/slam_external.py:123-160): a new leaf
    Parameter over `new_data`, the optimizer group re-pointed, the Adam state re-keyed with the new moments."""
    new_p = torch.nn.Parameter(new_data.requires_grad_(True))
    if optimizer is not None:
        group = _group(optimizer, k)
        old = group["params"][0]
        stored = optimizer.state.get(old, None)
        if stored is not None:
            if moments is not None:
                stored["exp_avg"], stored["exp_avg_sq"] = moments
            del optimizer.state[old]
            optimizer.state[new_p] = stored
        group["params"][0] = new_p
    params[k] = new_p


def _compact(arena, keep, n, names):
    """One compaction launch over the arena tensors `names`; returns the new row count (host int)."""
    d = arena._desc
    for i, name in enumerate(names):
        src, dst = arena.current(name), arena.other(name)
        d[i].src, d[i].dst, d[i].width = src.data_ptr(), dst.data_ptr(), src.shape[1]
    nbytes = lib.splat_compact_scratch_bytes(n)
    scratch = arena.scratch(nbytes)
    counts = arena._counts
    check(lib.splat_compact_rows(n, _vp(keep), d, len(names), _vp(scratch), nbytes, _vp(counts), _stream()))
    torch.cuda.current_stream().synchronize()      # the ONE read-back: the new number of Gaussians
    for name in names:
        arena.flip(name)
    return int(counts[0])


def _remove(keep, params, variables, optimizer):
    keys = [k for k in params.keys() if k not in ("cam_unnorm_rots", "cam_trans")]
    P = params[keys[0]].shape[0]
    dev = params[keys[0]].device
    with _on_device(dev):
        arena = _arena(variables, dev)
        names, states = [], {}
        for k in keys:
            arena.adopt("p:" + k, params[k])
            names.append("p:" + k)
            states[k] = _state_of(optimizer, k)
            if states[k] is not None:
                arena.adopt("m:" + k, states[k]["exp_avg"])
                arena.adopt("v:" + k, states[k]["exp_avg_sq"])
                names += ["m:" + k, "v:" + k]
        vkeys = [k for k in VAR_KEYS if k in variables]
        for k in vkeys:
            arena.adopt("var:" + k, variables[k])
            names.append("var:" + k)
        if len(names) > 24:
            raise RuntimeError("more than 24 per-Gaussian tensors")
        P_new = _compact(arena, keep, P, names)
        for k in keys:
            one_d = params[k].dim() == 1
            moments = None
            if states[k] is not None:
                moments = (arena.view("m:" + k, P_new, one_d), arena.view("v:" + k, P_new, one_d))
            _swap_param(params, optimizer, k, arena.view("p:" + k, P_new, one_d), moments)
        for k in vkeys:
            variables[k] = arena.view("var:" + k, P_new, variables[k].dim() == 1)
    return params, variables


def remove_points(to_remove, params, variables, optimizer):
    """utils/slam_external.py:143-165.  `to_remove`: bool [P] CUDA tensor."""
    keep = (~to_remove).to(torch.uint8).contiguous()
    return _remove(keep, params, variables, optimizer)


def prune_gaussians(params, variables, optimizer, iter, prune_dict):
    """utils/slam_external.py:172-195, same arguments and return value."""
    if iter <= prune_dict['stop_after']:
        if (iter >= prune_dict['start_after']) and (iter % prune_dict['prune_every'] == 0):
            if iter == prune_dict['stop_after']:
                remove_threshold = prune_dict['final_removal_opacity_threshold']
            else:
                remove_threshold = prune_dict['removal_opacity_threshold']
            big = -1.0
            if iter >= prune_dict['remove_big_after']:
                big = 0.1 * float(variables['scene_radius'])
            lo, ls = params['logit_opacities'], params['log_scales']
            P = lo.shape[0]
            with _on_device(lo.device):
                keep = _arena(variables, lo.device).flags(P)
                check(lib.splat_prune_flags(P, ls.shape[1], _vp(lo.detach().contiguous()), _vp(ls.detach().contiguous()),
                                            float(remove_threshold), big, _vp(keep), _stream()))
            params, variables = _remove(keep, params, variables, optimizer)
        if iter > 0 and iter % prune_dict['reset_opacities_every'] == 0 and prune_dict['reset_opacities']:
            # update_params_and_optimizer (:107-120): logit(0.01) everywhere, zeroed moments -- in place
            lo = params['logit_opacities']
            x = torch.ones(1, 1, dtype=lo.dtype, device=lo.device) * 0.01
            value = torch.log(x / (1 - x))
            st = _state_of(optimizer, 'logit_opacities')
            moments = (st["exp_avg"].zero_(), st["exp_avg_sq"].zero_()) if st is not None else None
            data = lo.detach()
            data.copy_(value.expand_as(data))
            _swap_param(params, optimizer, 'logit_opacities', data, moments)
    return params, variables


def new_gaussian_mask(depth_sil, gt_depth, sil_thres, arena=None):
    """The flat uint8 mask of pixels add_new_gaussians spawns from and (sum before the valid-depth AND, sum after) --
    scripts/splatam.py:390-408.  depth_sil [3,H,W], gt_depth [1,H,W] or [H,W]."""
    H, W = depth_sil.shape[-2:]
    dev = depth_sil.device
    arena = arena or GaussianArena(dev)
    with _on_device(dev):
        ds = depth_sil.detach().contiguous()
        gt = gt_depth.detach().contiguous()
        if ds.dtype != torch.float32 or gt.dtype != torch.float32 or ds.shape[0] != 3 or gt.numel() != H * W:
            raise RuntimeError("depth_sil must be float32 [3,H,W] and gt_depth float32 [1,H,W]")
        flags = arena.flags(H * W)
        nbytes = lib.splat_spawn_scratch_bytes(W, H)
        scratch = arena.scratch(nbytes)
        counts = arena._counts
        check(lib.splat_new_gaussian_mask(W, H, _vp(ds), _vp(gt), float(sil_thres), _vp(flags), _vp(scratch), nbytes,
                                          _vp(counts), _stream()))
        torch.cuda.current_stream().synchronize()    # the ONE read-back: how many Gaussians to add
    return flags[:H * W], int(counts[0]), int(counts[1])


def add_new_gaussians(params, variables, curr_data, sil_thres, time_idx, mean_sq_dist_method, gaussian_distribution,
                      optimizer=None, depth_sil=None, Renderer=None):
    """scripts/splatam.py:384-426, same leading arguments and return value.  Extras: `optimizer` (a live optimizer
    whose moments grow with the map, as cat_params_to_optimizer would; the reference rebuilds its optimizer after
    this call, scripts/splatam.py:826), `depth_sil` (a [3,H,W] silhouette render to reuse instead of rendering one)
    and `Renderer` (rasterizer class, default the diff_gaussian_rasterization shim)."""
    if mean_sq_dist_method != "projective":
        raise ValueError(f"Unknown mean_sq_dist_method {mean_sq_dist_method}")
    if gaussian_distribution not in ("isotropic", "anisotropic"):
        raise ValueError(f"Unknown gaussian_distribution {gaussian_distribution}")
    scale_dim = 1 if gaussian_distribution == "isotropic" else 3
    if params['log_scales'].shape[1] != scale_dim:
        raise ValueError("gaussian_distribution does not match params['log_scales']")
    dev = params['means3D'].device
    with _on_device(dev):
        arena = _arena(variables, dev)
        if depth_sil is None:
            from . import slam
            with torch.no_grad():
                rendervar = slam.build_rendervar(params, time_idx, gaussians_grad=False, camera_grad=False)
                Renderer = Renderer or slam._default_renderer()
                out, _, _ = Renderer(raster_settings=curr_data['cam']).render_rgbd_sil(**rendervar)
            depth_sil = out[3:6]
        flags, n_pre, M = new_gaussian_mask(depth_sil, curr_data['depth'], sil_thres, arena)
        if n_pre == 0:
            return params, variables
        P = params['means3D'].shape[0]
        H, W = depth_sil.shape[-2:]
        intr = curr_data['intrinsics']
        intr = intr.tolist() if isinstance(intr, torch.Tensor) else intr
        fx, fy, cx, cy = float(intr[0][0]), float(intr[1][1]), float(intr[0][2]), float(intr[1][2])
        states = {k: _state_of(optimizer, k) for k in GAUSS_KEYS}
        arena.reserve(P + M)          # before adopting: one growth step, not one per tensor
        for k in GAUSS_KEYS:
            arena.adopt("p:" + k, params[k])
            if states[k] is not None:
                arena.adopt("m:" + k, states[k]["exp_avg"])
                arena.adopt("v:" + k, states[k]["exp_avg_sq"])
        has_t = 'timestep' in variables
        if has_t:
            arena.adopt("var:timestep", variables['timestep'])
        for k in ("means2D_gradient_accum", "denom", "max_2D_radius"):
            arena._pair("var:" + k, 1)
        im = curr_data['im'].detach().contiguous()
        gt = curr_data['depth'].detach().contiguous()
        cr, ct = params['cam_unnorm_rots'].detach().contiguous(), params['cam_trans'].detach().contiguous()
        cur = arena.current
        check(lib.splat_spawn_gaussians(W, H, _vp(flags), _vp(arena._scratch), _vp(im), _vp(gt), fx, fy, cx, cy,
                                        _vp(cr), _vp(ct), cr.shape[-1], int(time_idx), scale_dim, P,
                                        _vp(cur("p:means3D")), _vp(cur("p:rgb_colors")), _vp(cur("p:unnorm_rotations")),
                                        _vp(cur("p:logit_opacities")), _vp(cur("p:log_scales")),
                                        _vp(cur("var:timestep")) if has_t else None, _stream()))
        num_pts = P + M
        for k in GAUSS_KEYS:
            moments = None
            if states[k] is not None:
                moments = (arena.view("m:" + k, num_pts, False), arena.view("v:" + k, num_pts, False))
                moments[0][P:].zero_()
                moments[1][P:].zero_()
            _swap_param(params, optimizer, k, arena.view("p:" + k, num_pts, False), moments)
        for k in ("means2D_gradient_accum", "denom", "max_2D_radius"):
            variables[k] = arena.view("var:" + k, num_pts, True).zero_()
        if has_t:
            variables['timestep'] = arena.view("var:timestep", num_pts, True)
    return params, variables
