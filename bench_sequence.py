#!/usr/bin/env python
"""bench_sequence.py -- config C5 (BASELINE.json configs[4]): TUM-shape 640x480 synthetic RGB-D sequence,
end-to-end SplaTAM frames/sec: per frame  pose init -> tracking iterations -> add_new_gaussians -> mapping iterations
(random keyframe of the window per iteration, prune_gaussians in the loop) -> keyframe bookkeeping, with the TUM
schedule (configs/tum/splatam.py: 200 tracking / 30 mapping iterations, keyframe every 5, window 20).

The sequence is rendered once from a fixed "room" of Gaussians (wider than the first view) along a smooth trajectory;
SLAM starts from the back-projection of frame 0 and grows the map.  Two arms, same data, same schedule, same seeds:
  --impl ours        slam.get_loss (fused builder / six-plane render / loss), FusedAdam, tracking iterations replayed from
                     a per-frame CUDA graph (pose-only), densify.add_new_gaussians / prune_gaussians over the arena
  --impl reference   the reference flow: torch glue (oracle/slam_oracle.py restates it) around the UNMODIFIED reference
                     extension (oracle/_ref, --variant dg), two passes per iteration, torch.optim.Adam, torch boolean-index
                     surgery (oracle/densify_oracle.py) -- the oracle is the measured thing ONLY in this leg
Control plane, shared by both arms: keyframe selection is the reference's own `keyframe_selection_overlap`
(utils/keyframe_selection.py:41-95), loaded AS-IS from the copy oracle/build_ref.py stages in oracle/_ref/refpy, with the
reference's keyframe-list bookkeeping (scripts/splatam.py:826-843, 931-945); if that copy is missing the run falls back to
"the most recent window-2 keyframes" and says so in its JSON line.  No evaluation / logging.
Prints one JSON line: frames/sec, ms per frame split into tracking / growing / mapping, trajectory error (ATE RMSE),
final map size.
"""
import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import bench  # noqa: E402

TUM = dict(w=640, h=480, fx=517.3, fy=516.5, cx=318.6, cy=255.3)          # configs/data/TUM/freiburg1_desk.yaml:3-8
TRACK_LRS = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
                 cam_unnorm_rots=0.002, cam_trans=0.002)                  # configs/tum/splatam.py:70-78
MAP_LRS = dict(means3D=0.0001, rgb_colors=0.0025, unnorm_rotations=0.001, logit_opacities=0.05, log_scales=0.001,
               cam_unnorm_rots=0.0, cam_trans=0.0)                        # :95-103
PRUNE = dict(start_after=0, remove_big_after=0, stop_after=20, prune_every=20, removal_opacity_threshold=0.005,
             final_removal_opacity_threshold=0.005, reset_opacities=False, reset_opacities_every=500)   # :106-116
WTS = dict(im=0.5, depth=1.0)


def gt_pose(t, n):
    """w2c of frame t relative to frame 0: a slow sway (a few cm, a few degrees over the sequence)."""
    s = 2 * math.pi * t / max(n, 2)
    ang = math.radians(3.0) * math.sin(s)
    c, sn = math.cos(ang), math.sin(ang)
    w2c = torch.eye(4)
    w2c[:3, :3] = torch.tensor([[c, 0, sn], [0, 1, 0], [-sn, 0, c]])
    w2c[:3, 3] = torch.tensor([0.12 * math.sin(s), 0.03 * math.sin(2 * s), 0.04 * (1 - math.cos(s))])
    return w2c


def rot_to_quat(R):
    w = math.sqrt(max(1e-12, 1.0 + R[0, 0] + R[1, 1] + R[2, 2])) / 2.0
    return torch.tensor([w, (R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w)])


def make_sequence(n_frames, room, device):
    """GT frames (colour, depth) rendered from the room with this repo's forward (data generation, untimed)."""
    from splatam_b200 import slam, synth
    from diff_gaussian_rasterization import GaussianRasterizer
    c = TUM
    mx, my = 128, 96                                   # the room extends beyond the first view
    sc = synth.surface_scene(room, c["w"] + 2 * mx, c["h"] + 2 * my, c["fx"], c["fy"], c["cx"] + mx, c["cy"] + my,
                             seed=4, layers=3)
    rs = synth.make_settings(device=device, **c)       # the camera stays at the first frame's w2c = I
    poses = [gt_pose(t, n_frames) for t in range(n_frames)]
    gt = {"means3D": sc["means3D"], "rgb_colors": sc["colors_precomp"], "unnorm_rotations": sc["rotations"],
          "logit_opacities": torch.logit(sc["opacities"].clamp(0.5, 0.999)), "log_scales": torch.log(sc["scales"][:, :1]),
          "cam_unnorm_rots": torch.stack([rot_to_quat(p[:3, :3]) for p in poses], 1).unsqueeze(0),
          "cam_trans": torch.stack([p[:3, 3] for p in poses], 1).unsqueeze(0)}
    gt = {k: v.float().contiguous().to(device) for k, v in gt.items()}
    frames = []
    with torch.no_grad():
        for t in range(n_frames):
            rv = slam.build_rendervar(gt, t, False, False)
            img, _, _ = GaussianRasterizer(raster_settings=rs).render_rgbd_sil(**rv)
            depth = img[3:4] / img[4:5].clamp_min(1e-6)
            depth[img[4:5] < 0.9] = 0.0                # no surface -> invalid depth
            frames.append({"im": img[:3].clamp(0, 1).contiguous(), "depth": depth.contiguous()})
    intr = torch.tensor([[c["fx"], 0, c["cx"]], [0, c["fy"], c["cy"]], [0, 0, 1.0]])
    return rs, frames, intr, torch.stack(poses)


def ate_rmse(params, gt_poses, n):
    """Camera-centre error after the sequence (both trajectories start at the identity)."""
    with torch.no_grad():
        q = F.normalize(params["cam_unnorm_rots"][0, :, :n].t().cpu())
        t = params["cam_trans"][0, :, :n].t().cpu()
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                         2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                         2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).view(-1, 3, 3)
        centre = -(R.transpose(1, 2) @ t.unsqueeze(-1)).squeeze(-1)
        Rg, tg = gt_poses[:n, :3, :3], gt_poses[:n, :3, 3]
        centre_g = -(Rg.transpose(1, 2) @ tg.unsqueeze(-1)).squeeze(-1)
        return float(((centre - centre_g) ** 2).sum(1).mean().sqrt())


def initialize_camera_pose(params, t):
    """scripts/splatam.py:429-448 with forward_prop=True."""
    with torch.no_grad():
        if t > 1:
            r1 = F.normalize(params['cam_unnorm_rots'][..., t - 1].detach())
            r2 = F.normalize(params['cam_unnorm_rots'][..., t - 2].detach())
            params['cam_unnorm_rots'][..., t] = F.normalize(r1 + (r1 - r2))
            t1, t2 = params['cam_trans'][..., t - 1].detach(), params['cam_trans'][..., t - 2].detach()
            params['cam_trans'][..., t] = t1 + (t1 - t2)
        else:
            params['cam_unnorm_rots'][..., t] = params['cam_unnorm_rots'][..., t - 1].detach()
            params['cam_trans'][..., t] = params['cam_trans'][..., t - 1].detach()


def optimizer_of(cls, params, lrs, tracking):
    groups = [{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in params.items()]
    return cls(groups) if tracking else cls(groups, lr=0.0, eps=1e-15)     # scripts/splatam.py:166-172


def empty_map(n_frames, device):
    z = lambda w: torch.zeros(0, w, device=device)                        # noqa: E731
    params = {"means3D": z(3), "rgb_colors": z(3), "unnorm_rotations": z(4), "logit_opacities": z(1), "log_scales": z(1),
              "cam_unnorm_rots": torch.tensor([1.0, 0, 0, 0], device=device).view(1, 4, 1).repeat(1, 1, n_frames).contiguous(),
              "cam_trans": torch.zeros(1, 3, n_frames, device=device)}    # initialize_params, scripts/splatam.py:134-139
    variables = {k: torch.zeros(0, device=device) for k in ("means2D_gradient_accum", "denom", "max_2D_radius", "timestep")}
    return params, variables


class OursArm:
    name = "ours"

    def __init__(self, device, variant):
        from splatam_b200 import densify, slam
        self.slam, self.densify, self.device = slam, densify, device
        self.Adam = slam.FusedAdam

    def init_map(self, curr, n_frames):
        params, variables = empty_map(n_frames, self.device)
        params = {k: torch.nn.Parameter(v) for k, v in params.items()}
        nothing = torch.zeros(3, *curr["depth"].shape[-2:], device=self.device)     # an empty map renders silhouette 0
        return self.densify.add_new_gaussians(params, variables, curr, 0.5, 0, "projective", "isotropic", depth_sil=nothing)

    def track(self, params, variables, curr, t, iters):
        slam = self.slam
        opt = optimizer_of(self.Adam, params, TRACK_LRS, True)
        best = slam.BestPose(params, t)

        def iteration():
            loss, _, _ = slam.get_loss(params, curr, variables, t, WTS, True, 0.99, True, False, tracking=True,
                                       tracking_pose_only=True)
            loss.backward()
            opt.step()
            best.update(loss, params)
            return loss
        # pose of this frame before tracking: the fall-back point if a replay overflows its captured binning capacity
        pose0 = (params["cam_unnorm_rots"].detach()[..., t].clone(), params["cam_trans"].detach()[..., t].clone())
        try:
            it = slam.GraphedIteration(iteration, params, opt, warmup=1)  # 1 eager iteration + capture, then replays
            for _ in range(iters - 1):
                it.replay()
            it.check()
        except slam.CaptureOverflow:
            self.overflow_frames = getattr(self, "overflow_frames", 0) + 1
            with torch.no_grad():
                params["cam_unnorm_rots"][..., t] = pose0[0]
                params["cam_trans"][..., t] = pose0[1]
            opt = optimizer_of(self.Adam, params, TRACK_LRS, True)
            best = slam.BestPose(params, t)
            for _ in range(iters):                                        # the reference's loop, un-captured
                for v in params.values():
                    v.grad = None
                iteration()
        best.write_back(params)
        for v in params.values():
            v.grad = None

    def grow(self, params, variables, curr, t):
        return self.densify.add_new_gaussians(params, variables, curr, 0.5, t, "projective", "isotropic")

    def map_iteration(self, params, variables, data, t_iter, opt, it):
        loss, variables, _ = self.slam.get_loss(params, data, variables, t_iter, WTS, False, 0.5, True, False, mapping=True)
        loss.backward()
        with torch.no_grad():
            params, variables = self.densify.prune_gaussians(params, variables, opt, it, PRUNE)
            opt.step()
            opt.zero_grad(set_to_none=True)
        return params, variables


class ReferenceArm:
    name = "reference"

    def __init__(self, device, variant):
        from oracle import densify_oracle as DO
        from oracle import ref_ext
        from oracle import slam_oracle as SO
        ref_ext.load(variant)
        self.Renderer = ref_ext.renderer_class(variant)
        self.SO, self.DO, self.device = SO, DO, device
        self.Adam = torch.optim.Adam

    def _surgery(self, fn, params, opt, variables):
        """The oracle's functional surgery on (params, Adam moments, variables) + the reference's optimizer re-keying
        (utils/slam_external.py:123-165)."""
        DO = self.DO
        state = {}
        for k in DO.GAUSS_KEYS:
            st = opt.state.get(params[k]) if opt is not None else None
            if st:
                state[k] = {"exp_avg": st["exp_avg"], "exp_avg_sq": st["exp_avg_sq"]}
        new_p, new_s, new_v = fn({k: v.detach() for k, v in params.items()}, state, variables)
        out = {}
        for k, v in new_p.items():
            if k not in DO.GAUSS_KEYS:
                out[k] = params[k]
                continue
            p = torch.nn.Parameter(v.requires_grad_(True))
            if opt is not None:
                grp = [g for g in opt.param_groups if g["name"] == k][0]
                st = opt.state.pop(grp["params"][0], None)
                if st is not None:
                    if k in new_s:
                        st["exp_avg"], st["exp_avg_sq"] = new_s[k]["exp_avg"], new_s[k]["exp_avg_sq"]
                    opt.state[p] = st
                grp["params"][0] = p
            out[k] = p
        return out, new_v

    def init_map(self, curr, n_frames):
        params, variables = empty_map(n_frames, self.device)
        nothing = torch.zeros(3, *curr["depth"].shape[-2:], device=self.device)
        p, _, v = self.DO.add_new_gaussians(params, {}, variables, nothing, curr["im"], curr["depth"], curr["intrinsics"], 0.5, 0)
        return {k: torch.nn.Parameter(t) for k, t in p.items()}, v

    def _get_loss(self, params, curr, variables, t, tracking):      # scripts/splatam.py:226-353 with the restated glue
        SO = self.SO
        gg, cg = (False, True) if tracking else (True, False)
        tg = SO.transform_to_frame(params, t, gg, cg)
        rv = SO.transformed_params2rendervar(params, tg)
        rv["means2D"] = torch.zeros_like(params["means3D"], requires_grad=True) + 0
        if rv["means2D"].requires_grad:
            rv["means2D"].retain_grad()
        im, radius, _ = self.Renderer(curr["cam"])(**rv)
        variables["means2D"] = rv["means2D"]
        rv2 = dict(rv)
        rv2["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], curr["w2c"])
        rv2["means2D"] = torch.zeros_like(params["means3D"], requires_grad=True) + 0
        ds, _, _ = self.Renderer(curr["cam"])(**rv2)
        loss, _ = SO.image_loss(im, ds, curr["im"], curr["depth"], WTS, tracking, 0.99 if tracking else 0.5, True, False,
                                tracking)
        seen = radius > 0
        variables["max_2D_radius"][seen] = torch.max(radius[seen], variables["max_2D_radius"][seen])
        variables["seen"] = seen
        return loss, ds

    def track(self, params, variables, curr, t, iters):            # scripts/splatam.py:699-764
        opt = optimizer_of(self.Adam, params, TRACK_LRS, True)
        cand_rot = params['cam_unnorm_rots'][..., t].detach().clone()
        cand_tran = params['cam_trans'][..., t].detach().clone()
        current_min_loss = float(1e20)
        for _ in range(iters):
            loss, _ = self._get_loss(params, curr, variables, t, True)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
            with torch.no_grad():
                if loss < current_min_loss:
                    current_min_loss = loss
                    cand_rot = params['cam_unnorm_rots'][..., t].detach().clone()
                    cand_tran = params['cam_trans'][..., t].detach().clone()
        with torch.no_grad():
            params['cam_unnorm_rots'][..., t] = cand_rot
            params['cam_trans'][..., t] = cand_tran

    def grow(self, params, variables, curr, t):
        with torch.no_grad():
            scratch = dict(variables, max_2D_radius=variables["max_2D_radius"].clone())
            _, ds = self._get_loss(params, curr, scratch, t, False)
        return self._surgery(lambda p, s, v: self.DO.add_new_gaussians(p, s, v, ds, curr["im"], curr["depth"],
                                                                       curr["intrinsics"], 0.5, t), params, None, variables)

    def map_iteration(self, params, variables, data, t_iter, opt, it):
        loss, _ = self._get_loss(params, data, variables, t_iter, False)
        loss.backward()
        with torch.no_grad():
            if self.DO.prune_mask({k: v.detach() for k, v in params.items()}, variables, it, PRUNE) is not None:
                params, variables = self._surgery(lambda p, s, v: self.DO.prune_gaussians(p, s, v, it, PRUNE), params, opt,
                                                  variables)
            opt.step()
            opt.zero_grad(set_to_none=True)
        return params, variables


def load_keyframe_selection():
    """The reference's keyframe_selection_overlap, unchanged (control plane, SURVEY section 2 row 14: "re-used as-is"),
    from the staged copy; None when it is not there."""
    import importlib.util
    path = os.path.join(ROOT, "oracle", "_ref", "refpy", "utils", "keyframe_selection.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_keyframe_selection", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.keyframe_selection_overlap


def est_w2c(params, t, device):
    """scripts/splatam.py:826-831: the estimated world-to-camera of frame t."""
    with torch.no_grad():
        q = F.normalize(params['cam_unnorm_rots'][..., t].detach())[0]
        r, x, y, z = [float(v) for v in q]
        R = torch.tensor([[1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)],
                          [2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)],
                          [2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)]], device=device)   # build_rotation
        w2c = torch.eye(4, device=device)
        w2c[:3, :3] = R
        w2c[:3, 3] = params['cam_trans'][..., t].detach()[0]
    return w2c


KEYFRAME_POLICY = {"name": None}
TRACK_MS_PER_FRAME = []


def run(arm, rs, frames, intr, n_frames, track_iters, map_iters, kf_every, window, device, seed=0):
    """scripts/splatam.py:660-900 (rgbd_slam main loop), hot path only."""
    rng = np.random.RandomState(seed)
    torch.manual_seed(seed)      # keyframe_selection_overlap draws from the global torch / numpy generators
    np.random.seed(seed)
    select = load_keyframe_selection()
    KEYFRAME_POLICY["name"] = ("reference keyframe_selection_overlap (utils/keyframe_selection.py, as-is; both arms)"
                               if select else "most recent window-2 keyframes + last + current (staged reference copy missing)")
    keyframe_list = []
    first_w2c = torch.eye(4, device=device)
    intr_d = intr.to(device)
    data = lambda t: {"cam": rs, "im": frames[t]["im"], "depth": frames[t]["depth"], "id": t, "intrinsics": intr_d,  # noqa: E731
                      "w2c": first_w2c}
    params, variables = arm.init_map(data(0), n_frames)
    variables["scene_radius"] = torch.max(frames[0]["depth"]).cpu() / 2.0       # scene_radius_depth_ratio = 2
    keyframes = []
    ms = {"tracking": 0.0, "growing": 0.0, "mapping": 0.0}
    TRACK_MS_PER_FRAME.clear()

    def clock(key, t0):
        torch.cuda.synchronize(device)
        ms[key] += (time.perf_counter() - t0) * 1e3

    for t in range(n_frames):
        curr = data(t)
        if t > 0:
            initialize_camera_pose(params, t)
            t0 = time.perf_counter()
            arm.track(params, variables, curr, t, track_iters)
            before = ms["tracking"]
            clock("tracking", t0)
            TRACK_MS_PER_FRAME.append(ms["tracking"] - before)
            t0 = time.perf_counter()
            params, variables = arm.grow(params, variables, curr, t)
            clock("growing", t0)
        t0 = time.perf_counter()
        if select is not None:                                                    # scripts/splatam.py:826-843
            picked = select(curr["depth"], est_w2c(params, t, device), intr_d, keyframe_list[:-1], window - 2)
            selected = [keyframe_list[i]['id'] for i in picked]
            if keyframe_list:
                selected.append(keyframe_list[-1]['id'])
            selected.append(t)
        else:
            selected = [k for k in keyframes[-(window - 2) - 1:-1]] + keyframes[-1:] + [t]
        opt = optimizer_of(arm.Adam, params, MAP_LRS, False)
        for it in range(map_iters):
            t_iter = selected[rng.randint(0, len(selected))]
            params, variables = arm.map_iteration(params, variables, data(t_iter), t_iter, opt, it)
        clock("mapping", t0)
        if t == 0 or (t + 1) % kf_every == 0 or t == n_frames - 2:                # scripts/splatam.py:931-945
            keyframes.append(t)
            keyframe_list.append({'id': t, 'est_w2c': est_w2c(params, t, device), 'color': curr["im"], 'depth': curr["depth"]})
    return params, ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", default="dg", choices=["dg", "hx"])
    ap.add_argument("--frames", type=int, default=None, help="frames to run (default 500 ours, 100 reference)")
    ap.add_argument("--sequence-frames", type=int, default=500, help="length of the trajectory the frames are cut from")
    ap.add_argument("--room", type=int, default=1_500_000, help="Gaussians of the ground-truth room")
    ap.add_argument("--tracking-iters", type=int, default=200)
    ap.add_argument("--mapping-iters", type=int, default=30)
    ap.add_argument("--profile-end", action="store_true",
                    help="ours: after the run, profile 20 eager pose-only tracking iterations on the final state (per-kernel ms)")
    args = ap.parse_args()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    assert torch.cuda.is_available(), "bench_sequence.py needs a CUDA device"
    device = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    n = args.frames or (500 if args.impl == "ours" else 100)
    n = min(n, args.sequence_frames)
    rs, frames, intr, gt_poses = make_sequence(args.sequence_frames, args.room, device)
    frames = frames[:n]
    arm = (OursArm if args.impl == "ours" else ReferenceArm)(device, args.variant)
    from splatam_b200._lib import lib
    # untimed warm-up: two short frames through the same code (library / allocator / cuDNN-free lazy initialisation)
    run(arm, rs, frames, intr, min(2, n), 6, 6, 5, 20, device)
    sampler = bench.ClockSampler(0)
    sampler.start()
    torch.cuda.synchronize()
    l0 = lib.splat_kernel_launch_count()
    sampler.mark_start()
    t0 = time.perf_counter()
    params, ms = run(arm, rs, frames, intr, n, args.tracking_iters, args.mapping_iters, 5, 20, device)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    sampler.mark_end()
    out = {"metric": "end-to-end SplaTAM frames/sec (track + grow + map + prune), TUM shape 640x480", "unit": "frames/s",
           "value": n / wall, "ms_per_frame": wall * 1e3 / n, "higher_is_better": True, "impl": args.impl,
           "frames": n, "n_gpus": 1, "dtype": "f32", "data": "synthetic",
           "ms_per_frame_split": {k: round(v / n, 3) for k, v in ms.items()},
           "ate_rmse_m": ate_rmse(params, gt_poses, n), "final_gaussians": int(params["means3D"].shape[0]),
           "config": {"workload": "tum_640x480_sequence", "image": [TUM["w"], TUM["h"]], "room_gaussians": args.room,
                      "tracking_iters": args.tracking_iters, "mapping_iters": args.mapping_iters, "keyframe_every": 5,
                      "mapping_window": 20, "variant": args.variant if args.impl == "reference" else "fused",
                      "keyframe_selection": KEYFRAME_POLICY["name"]},
           "tracking_ms_per_frame_by_50": [round(sum(TRACK_MS_PER_FRAME[i:i + 50]) / len(TRACK_MS_PER_FRAME[i:i + 50]), 1)
                                           for i in range(0, len(TRACK_MS_PER_FRAME), 50)],
           "gpu_launches": lib.splat_kernel_launch_count() - l0, "clocks": sampler.stop()}
    if args.profile_end and args.impl == "ours":
        import ctypes as C
        from splatam_b200 import slam
        t = n - 1
        curr = {"cam": rs, "im": frames[t]["im"], "depth": frames[t]["depth"], "id": t, "intrinsics": intr.to(device),
                "w2c": torch.eye(4, device=device)}
        variables = {"max_2D_radius": torch.zeros(params["means3D"].shape[0], device=device)}

        def it():
            for v in params.values():
                v.grad = None
            loss, _, _ = slam.get_loss(params, curr, variables, t, WTS, True, 0.99, True, False, tracking=True,
                                       tracking_pose_only=True)
            loss.backward()
        for _ in range(5):
            it()
        lib.splat_profile_enable(1)
        for _ in range(20):
            it()
        k = 16
        msb = (C.c_float * k)(); cnt = (C.c_int32 * k)()
        lib.splat_profile_read(msb, cnt, k)
        lib.splat_profile_enable(0)
        names = [lib.splat_profile_kernel_name(j).decode() for j in range(k)]
        out["final_state_tracking_kernel_ms"] = {names[j]: round(msb[j] / 20, 4) for j in range(k) if cnt[j]}
        with torch.no_grad():
            sc = torch.exp(params["log_scales"])
            out["final_state"] = {"gaussians": int(params["means3D"].shape[0]), "scale_mean": float(sc.mean()),
                                  "scale_median": float(sc.median()), "opacity_mean": float(torch.sigmoid(params["logit_opacities"]).mean())}
    real_stdout.write(json.dumps(out) + "\n")
    real_stdout.flush()


if __name__ == "__main__":
    main()
