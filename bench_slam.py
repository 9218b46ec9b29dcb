#!/usr/bin/env python
"""bench_slam.py -- one full SplaTAM optimisation iteration (get_loss + backward), SURVEY.md 8f-1..3.

What scripts/splatam.py runs 40-200 times per tracked frame and 30-60 times per mapped frame:
    transform_to_frame -> rendervars -> RGB render -> depth+silhouette render -> masked L1 (+SSIM) loss -> backward
Arms (same seeded scene, same step schedule, CUDA events around K steps after W warm-ups, inputs resident in HBM):
  --impl ours        splatam_b200.slam.get_loss: fused builder kernel -> ONE six-plane render -> fused loss kernel ->
                     rasterizer backward -> builder backward (pose gradient reduced in-kernel)
  --impl dropin      the reference's torch glue (oracle/slam_oracle.py restates it) around OUR 3-channel drop-in
                     rasterizer, two passes: what an unmodified scripts/splatam.py gets from this repo
  --impl reference   the same torch glue around the UNMODIFIED reference extension (oracle/_ref, --variant dg|hx)
--mode tracking: C2 shape, 1200x680, 300k isotropic Gaussians, pose-only gradients (sil_thres 0.99, summed L1)
--mode mapping : C3 shape, 1200x680, 1M Gaussians, Gaussian gradients (0.8 L1 + 0.2 (1-SSIM) colour, mean L1 depth)
The optimizer step (torch Adam) is outside the timed iteration in every arm.  One JSON line on stdout.
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

import torch  # noqa: E402

import bench  # noqa: E402  (ClockSampler)


def rot_to_quat(R):
    """(w,x,y,z) of a rotation matrix with positive trace (small camera rotations)."""
    w = math.sqrt(max(1e-12, 1.0 + R[0, 0] + R[1, 1] + R[2, 2])) / 2.0
    return torch.tensor([w, (R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w)])


def build(mode, device, n_frames=24):
    from splatam_b200 import synth
    cam = synth.REPLICA
    P, layers = (300_000, 1) if mode == "tracking" else (1_000_000, 3)
    sc = synth.surface_scene(P, cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], seed=2, layers=layers)
    rs = synth.make_settings(device=device, **cam)            # camera stays at the first frame's w2c = I
    rots, trans = [], []
    for k in range(n_frames):                                 # relative poses of the keyframe window live in params
        w2c = synth.arc_pose(k, n_frames)
        rots.append(rot_to_quat(w2c[:3, :3]) * 1.1)
        trans.append(w2c[:3, 3])
    params = {
        "means3D": sc["means3D"], "rgb_colors": sc["colors_precomp"],
        "unnorm_rotations": sc["rotations"],
        "logit_opacities": torch.logit(sc["opacities"].clamp(1e-4, 1 - 1e-4)),
        "log_scales": torch.log(sc["scales"][:, :1]),
        "cam_unnorm_rots": torch.stack(rots, 1).unsqueeze(0), "cam_trans": torch.stack(trans, 1).unsqueeze(0),
    }
    params = {k: v.float().contiguous().to(device).requires_grad_(True) for k, v in params.items()}
    return cam, rs, params


def make_targets(params, rs, n_frames, device):
    """Ground-truth colour/depth per keyframe: our own fused render at a pose perturbed by (0.5 deg, 5 mm), SURVEY 8d C2."""
    from splatam_b200 import slam
    from diff_gaussian_rasterization import GaussianRasterizer
    g = torch.Generator().manual_seed(5)
    out = []
    with torch.no_grad():
        pert = {k: v.detach().clone() for k, v in params.items()}
        pert["cam_unnorm_rots"] += 0.004 * torch.randn(pert["cam_unnorm_rots"].shape, generator=g).to(device)
        pert["cam_trans"] += 0.005 * torch.randn(pert["cam_trans"].shape, generator=g).to(device)
        for k in range(n_frames):
            rv = slam.build_rendervar(pert, k, False, False)
            img, _, _ = GaussianRasterizer(raster_settings=rs).render_rgbd_sil(**rv)
            depth = img[3:4].clone()
            depth[img[4:5] < 0.5] = 0.0                       # no surface -> invalid depth
            out.append({"cam": rs, "im": img[:3].clamp(0, 1).contiguous(), "depth": depth.contiguous(),
                        "w2c": torch.eye(4, device=device)})
    return out


def make_targets_ref(params, rs, n_frames, device, variant):
    """make_targets for the reference arm: the same perturbed-pose ground truth rendered by the UNMODIFIED reference
    extension (two passes, scripts/splatam.py:255-259) through the restated torch glue -- libsplat_b200.so is not
    touched."""
    from oracle import ref_ext, slam_oracle as SO
    ref_ext.load(variant)
    Renderer = ref_ext.renderer_class(variant)
    g = torch.Generator().manual_seed(5)
    out = []
    with torch.no_grad():
        pert = {k: v.detach().clone() for k, v in params.items()}
        pert["cam_unnorm_rots"] += 0.004 * torch.randn(pert["cam_unnorm_rots"].shape, generator=g).to(device)
        pert["cam_trans"] += 0.005 * torch.randn(pert["cam_trans"].shape, generator=g).to(device)
        w2c = torch.eye(4, device=device)
        for k in range(n_frames):
            tg = SO.transform_to_frame(pert, k, False, False)
            rv = SO.transformed_params2rendervar(pert, tg)
            rv["means2D"] = torch.zeros_like(pert["means3D"])
            im, _, _ = Renderer(rs)(**rv)
            rv2 = dict(rv)
            rv2["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], w2c)
            ds, _, _ = Renderer(rs)(**rv2)
            depth = ds[0:1].clone()
            depth[ds[1:2] < 0.5] = 0.0
            out.append({"cam": rs, "im": im.clamp(0, 1).contiguous(), "depth": depth.contiguous(), "w2c": w2c})
    return out


MAPPING_LRS = dict(means3D=0.0001, rgb_colors=0.0025, unnorm_rotations=0.001, logit_opacities=0.05, log_scales=0.001,
                   cam_unnorm_rots=0.0, cam_trans=0.0)     # configs/replica/splatam.py:92-100
TRACKING_LRS = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0,
                    cam_unnorm_rots=0.0004, cam_trans=0.002)   # configs/replica/splatam.py:71-79


def make_step(impl, variant, params, targets, variables, wts, tracking, sil_thres, pose_only=False):
    """One get_loss + backward (scripts/splatam.py:226-353) as a callable step(kf) -> loss, for impl in
    ours (fused kernels) | dropin (reference torch glue around OUR 3-channel rasterizer) | reference (same glue around
    the unmodified reference extension)."""
    if impl == "ours":
        from splatam_b200 import slam

        def step(kf):
            loss, _, wl = slam.get_loss(params, targets[kf], variables, kf, wts, tracking, sil_thres, True, False,
                                        tracking=tracking, mapping=not tracking, tracking_pose_only=pose_only)
            loss.backward()
            return loss
        return step
    from oracle import slam_oracle as SO
    if impl == "dropin":
        from diff_gaussian_rasterization import GaussianRasterizer as R
        Renderer = lambda raster_settings: R(raster_settings)   # noqa: E731
    else:
        from oracle import ref_ext
        ref_ext.load(variant)
        Renderer = ref_ext.renderer_class(variant)

    def step(kf):   # scripts/splatam.py:226-353 (get_loss) with the restated glue
        cur = targets[kf]
        gg, cg = (False, True) if tracking else (True, False)
        tg = SO.transform_to_frame(params, kf, gg, cg)
        rv = SO.transformed_params2rendervar(params, tg)
        rv["means2D"] = torch.zeros_like(params["means3D"], requires_grad=True) + 0
        rv["means2D"].retain_grad()
        im, radius, _ = Renderer(cur["cam"])(**rv)
        variables["means2D"] = rv["means2D"]
        rv2 = dict(rv)
        rv2["colors_precomp"] = SO.get_depth_and_silhouette(tg["means3D"], cur["w2c"])
        rv2["means2D"] = torch.zeros_like(params["means3D"], requires_grad=True) + 0
        ds, _, _ = Renderer(cur["cam"])(**rv2)
        loss, wl = SO.image_loss(im, ds, cur["im"], cur["depth"], wts, tracking, sil_thres, True, False, tracking)
        seen = radius > 0
        variables["max_2D_radius"][seen] = torch.max(radius[seen], variables["max_2D_radius"][seen])
        variables["seen"] = seen
        loss.backward()
        return loss
    return step


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="tracking", choices=["tracking", "mapping"])
    ap.add_argument("--impl", default="ours", choices=["ours", "dropin", "reference"])
    ap.add_argument("--variant", default="dg", choices=["dg", "hx"])
    ap.add_argument("--pose-only", action="store_true",
                    help="ours, tracking: skip the Gaussian-side gradients the reference computes and discards")
    ap.add_argument("--optimizer", default="none", choices=["none", "torch", "fused"],
                    help="include the per-iteration optimizer step + zero_grad in the timed iteration: torch = "
                         "torch.optim.Adam as the reference builds it, fused = splatam_b200.slam.FusedAdam (ours only)")
    ap.add_argument("--graph", type=int, default=0, metavar="ITERS",
                    help="ours + --optimizer fused: per frame, run ITERS iterations as 1 eager one + one CUDA-graph capture "
                         "+ (ITERS-1) replays (slam.GraphedIteration); 40 = Replica tracking, 200 = TUM; capture cost is "
                         "inside the timed region")
    ap.add_argument("--gpus", type=int, default=1,
                    help="mapping, ours: keyframe-sharded over N ranks (launch under torchrun): every rank runs the fused "
                         "iteration on its own keyframe, the raw-parameter gradient bucket is all-reduced (own NVLink kernel), "
                         "FusedAdam steps on every replica; value = N * steps / time")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--steps", type=int, default=96)
    ap.add_argument("--warmup", type=int, default=8)
    args = ap.parse_args()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    assert torch.cuda.is_available(), "bench_slam.py needs a CUDA device"
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local)
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        assert args.mode == "mapping" and args.impl == "ours" and args.optimizer == "fused" and args.gpus == world, \
            "N > 1: --mode mapping --impl ours --optimizer fused --gpus WORLD_SIZE"
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    nf = 24
    cam, rs, params = build(args.mode, device, nf)
    targets = make_targets(params, rs, nf, device)
    tracking = args.mode == "tracking"
    P = params["means3D"].shape[0]
    variables = {"max_2D_radius": torch.zeros(P, device=device)}
    wts = dict(im=0.5, depth=1.0)                             # configs/replica/splatam.py:67-70, :88-91
    sil_thres = 0.99 if tracking else 0.5

    from splatam_b200 import slam
    from splatam_b200._lib import lib
    bucket, sched, parallelism = None, None, "single GPU"
    if world > 1:
        # replicated raw parameters in one flat bucket; their gradients are written by the backward kernels straight
        # into the (symmetric-memory) gradient bucket, summed over the ranks by ONE exchange, then FusedAdam steps
        from splatam_b200 import mapping
        raw = {k: params[k].detach() for k, _ in mapping.raw_layout(1)}
        why = ""
        if args.collective == "peer":
            try:
                bucket = mapping.GradBucket(raw, device, direct=True, peer_group=dist.group.WORLD, layout=mapping.raw_layout(1))
            except Exception as e:  # noqa: BLE001
                why, bucket = repr(e)[:160], None
            ok = torch.tensor([0.0 if bucket is None else 1.0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok.item()) != 1.0:
                sys.stderr.write("peer-memory all-reduce unavailable (%s); using NCCL\n" % why)
                bucket = None
        if bucket is None:
            bucket = mapping.GradBucket(raw, device, direct=True, layout=mapping.raw_layout(1))
            parallelism = "keyframe-sharded x%d + NCCL all-reduce of the raw-parameter gradient bucket" % world
        else:
            parallelism = "keyframe-sharded x%d + own %s all-reduce of the raw-parameter gradient bucket" % (
                world, "in-switch (multimem)" if bucket.peer.multicast else "NVLink peer-memory")
        leaves = bucket.leaves()
        params = dict(leaves, cam_unnorm_rots=params["cam_unnorm_rots"], cam_trans=params["cam_trans"])
        sched = mapping.keyframe_schedule(nf, max(args.steps, 8) + args.warmup, world, seed=11)
    step = make_step(args.impl, args.variant, params, targets, variables, wts, tracking, sil_thres, args.pose_only)

    opt = None
    if args.optimizer != "none":   # initialize_optimizer (scripts/splatam.py:166-172) with the config's learning rates
        lrs = TRACKING_LRS if tracking else MAPPING_LRS
        Adam = slam.FusedAdam if args.optimizer == "fused" else torch.optim.Adam
        groups = [{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in params.items()]
        opt = Adam(groups) if tracking else Adam(groups, lr=0.0, eps=1e-15)
        inner = step

        def step(kf):   # scripts/splatam.py:724-725 / :889-890
            loss = inner(kf)
            if bucket is not None:
                bucket.all_reduce()
            with torch.no_grad():
                opt.step()
            return loss

    def zero():
        if bucket is not None:     # the bucket views stay attached as .grad; every row is overwritten by the next backward
            return
        if opt is not None:
            opt.zero_grad(set_to_none=True)
            return
        for v in params.values():
            v.grad = None

    def kf_of(i):
        return int(sched[i % sched.shape[0], rank]) if sched is not None else (7 * i) % nf

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = bench.ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(args.warmup):
        zero()
        loss = step(kf_of(i))   # held across the next call, as in the timed loop (same caching-allocator pattern)
    if args.graph:   # one untimed captured frame: first-capture costs (graph pool, instantiate) are per process, not per frame
        assert args.impl == "ours" and args.optimizer == "fused", "--graph needs --impl ours --optimizer fused"
        w = slam.GraphedIteration(lambda: step(0), params, opt, warmup=1)
        for _ in range(3):
            w.replay()
        w.check()
    barrier()
    l0 = lib.splat_kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.mark_start()
    t0 = time.perf_counter()
    e0.record()
    if args.graph:
        assert args.impl == "ours" and args.optimizer == "fused", "--graph needs --impl ours --optimizer fused"
        frames = max(1, args.steps // args.graph)
        args.steps = frames * args.graph
        for f in range(frames):   # one tracked frame: same keyframe for all of its iterations (scripts/splatam.py:710-758)
            kf = (7 * f) % nf
            it = slam.GraphedIteration(lambda: step(kf), params, opt, warmup=1)
            for _ in range(args.graph - 1):
                loss = it.replay()
            it.check()
    else:
        for i in range(args.steps):
            zero()
            loss = step(kf_of(args.warmup + i))
    e1.record()
    barrier()
    wall = time.perf_counter() - t0
    sampler.mark_end()
    ms = e0.elapsed_time(e1)
    if dist is not None:    # max over ranks
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = lib.splat_kernel_launch_count() - l0
    out = {"metric": "SplaTAM %s iterations/sec (get_loss + backward: builder, RGB + depth + silhouette render, loss)"
                     % args.mode,
           "unit": "iters/s", "value": world * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
           "wall_ms_per_step": wall * 1e3 / args.steps, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "higher_is_better": True, "scaling": "weak", "dtype": "f32", "data": "synthetic", "impl": args.impl,
           "pose_only": bool(args.pose_only), "optimizer_in_timed_step": args.optimizer, "graph_iters_per_frame": args.graph,
           "config": {"workload": "replica_1200x680_%s_%s" % ("300k" if tracking else "1M", args.mode),
                      "image": [cam["w"], cam["h"]], "gaussians": P, "keyframes": nf, "parallelism": parallelism,
                      "variant": args.variant if args.impl == "reference" else "dg",
                      "loss": "tracking: masked summed L1, sil_thres 0.99" if tracking else
                              "mapping: 0.8 L1 + 0.2 (1-SSIM 11x11) colour, masked mean L1 depth"},
           "gpu_launches": launches, "last_loss": float(loss.detach()), "clocks": sampler.stop() if rank == 0 else None}
    if world > 1:
        if rank == 0:
            real_stdout.write(json.dumps(out) + "\n")
            real_stdout.flush()
        dist.destroy_process_group()
        return
    if args.impl == "ours":   # per-kernel event times from a second pass (kept out of the timed region)
        import ctypes as C
        lib.splat_profile_enable(1)
        for i in range(args.steps):
            zero()
            step((7 * i) % nf)
        n = 16
        ms_sum = (C.c_float * n)(); cnt = (C.c_int32 * n)()
        lib.splat_profile_read(ms_sum, cnt, n)
        lib.splat_profile_enable(0)
        names = [lib.splat_profile_kernel_name(k).decode() for k in range(n)]
        out["kernel_ms_per_step"] = {names[k]: round(ms_sum[k] / args.steps, 5) for k in range(n) if cnt[k]}
        N = cam["w"] * cam["h"]
        sd = 1
        # algorithmic HBM bytes of the glue kernels
        rv_f = (12 + 16 + 4 + 4 * sd) + (12 + 16 + 4 + 12)
        rv_b = (24 if tracking else 12 + 12 + 16 + 16 + 16) + (4 + 4 + 4) + (12 + 4 + 4 * sd)
        loss_b = (40 + 24) if tracking else (6 * 4 + 9 * 4) + (9 * 4 + 6 * 4 + 6 * 4) + 16 + 12
        peaks, src = bench.measured_peaks()
        roof = {}
        for name, byt in (("rendervar_fwd", rv_f * P), ("rendervar_bwd", rv_b * P), ("slam_loss", loss_b * N)):
            t = out["kernel_ms_per_step"].get(name)
            if t:
                roof[name] = {"bound": "hbm", "algorithmic_bytes": byt, "achieved": byt / (t * 1e-3) / 1e9,
                              "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": byt / (t * 1e-3) / 1e9 / peaks["hbm_gbs"],
                              "kernel_ms": t}
        out["roofline_glue"] = roof
    real_stdout.write(json.dumps(out) + "\n")
    real_stdout.flush()


if __name__ == "__main__":
    main()
