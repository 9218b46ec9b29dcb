"""Run under torchrun on the GPUs of one NVSwitch node (not collected by pytest): sweeps the grid size and the loads in
flight per thread of the two in-switch kernels (splat_peer_set_tuning) on the mapping-size bucket and prints one line
per setting -- plain all-reduce of 14M floats and fused reduce + Adam + all-gather of 1M Gaussians."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from splatam_b200 import mapping  # noqa: E402
from splatam_b200._lib import lib  # noqa: E402

LRS = dict(means3D=1e-4, rgb_colors=2.5e-3, unnorm_rotations=1e-3, logit_opacities=5e-2, log_scales=1e-3)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)

    def timeit(fn, n=40):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    mc = os.environ.get("SWEEP_MULTICAST", "1") != "0"   # 0: the peer load / store kernels (what N = 2 and 4 run)
    pa = mapping.PeerAllReduce(14_000_000, dev, dist.group.WORLD, multicast=mc)
    g = torch.Generator().manual_seed(9)
    scene = {k: torch.randn(1_000_000, w, generator=g).to(dev) for k, w in mapping.raw_layout(1)}
    bucket = mapping.GradBucket(scene, dev, direct=False, peer_group=dist.group.WORLD, layout=mapping.raw_layout(1),
                                shard_optimizer=True, multicast=mc)
    opt = mapping.ShardedAdam(bucket, LRS)
    bucket.grads.normal_()
    if rank == 0:
        print("world %d multicast allreduce=%s reduce_adam=%s" % (world, bool(pa.multicast), bool(bucket.peer.multicast)), flush=True)
    small = mapping.PeerAllReduce(4096, dev, dist.group.WORLD, multicast=bool(pa.multicast))
    for blocks in (0, 16, 24, 32, 48, 64, 96, 148, 296, 444, 592):
        for unroll in ((1, 2, 4, 8) if mc else (4,)):
            lib.splat_peer_set_tuning(blocks, unroll)
            t_ar = timeit(pa)
            t_ra = timeit(opt.step) if unroll == 4 else float("nan")
            if rank == 0:
                print("SWEEP blocks %3d unroll %d   allreduce_ms %.4f   reduce_adam_ms %.4f" % (blocks, unroll, t_ar, t_ra), flush=True)
    lib.splat_peer_set_tuning(0, 0)
    t_fix = timeit(small)
    if rank == 0:
        print("fixed cost (4096 floats) ms %.4f" % t_fix, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
