"""Run under torchrun on >= 2 GPUs of one node (not collected by pytest: needs N GPUs; scripts/gpu_multi_adam.sh runs it).
mapping.ShardedAdam -- reduce-scatter + Adam + all-gather in one kernel per rank (splat_peer_reduce_adam) -- against the
plain path: explicit gather + rank-ordered sum of the gradients, then slam.FusedAdam (full-size moments) on every rank.
Peer path: bit-exact.  In-switch path: the switch's own summation order, so rounding-level tolerance, but every replica
must hold the SAME bits.  P is not a multiple of 4 on purpose (float4 groups straddle tensor boundaries)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from splatam_b200 import mapping, slam  # noqa: E402

LRS = dict(means3D=0.0001, rgb_colors=0.0025, unnorm_rotations=0.001, logit_opacities=0.05, log_scales=0.001)


def scene(P, seed, dev):
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn(P, w, generator=g).to(dev) for k, w in mapping.raw_layout(1)}


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for P, mc in ((100_003, False), (100_003, True)):
        bucket = mapping.GradBucket(scene(P, 7, dev), dev, direct=False, peer_group=dist.group.WORLD,
                                    layout=mapping.raw_layout(1), shard_optimizer=True, multicast=mc)
        if mc and not bucket.peer.multicast:
            if rank == 0:
                print("no multicast mapping on this fabric; in-switch path not tested", flush=True)
            continue
        opt = mapping.ShardedAdam(bucket, LRS)
        ref = {k: v.clone() for k, v in bucket.views.items()}
        ref_leaf = {k: v.requires_grad_(True) for k, v in ref.items()}
        ref_opt = slam.FusedAdam([{'params': [v], 'name': k, 'lr': LRS[k]} for k, v in ref_leaf.items()], lr=0.0, eps=1e-15)
        n = bucket.params.numel()
        for it in range(3):
            g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
            x = torch.randn(n, generator=g, device=dev) * (0.5 ** it)
            bucket.grads.copy_(x)
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            want = parts[0].clone()
            for j in range(1, world):
                want += parts[j]
            off = 0
            for k, w in bucket.layout:
                ref_leaf[k].grad = want[off:off + P * w].view(P, w).clone()
                off += P * w
            with torch.no_grad():
                ref_opt.step()
            torch.cuda.synchronize()
            opt.step()
            torch.cuda.synchronize()
            for k in ref_leaf:
                got, exp = bucket.views[k], ref_leaf[k].detach()
                same = torch.equal(got, exp) if not mc else bool(((got - exp).abs() <= 2e-6 * (exp.abs() + 1.0)).all())
                if not same:
                    print("rank", rank, "mc", mc, "step", it, k, "max diff", float((got - exp).abs().max()), flush=True)
                ok = ok and same
            cs = bucket.params.double().sum().reshape(1)
            lo_, hi_ = cs.clone(), cs.clone()
            dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
            if not bool(lo_ == hi_):
                print("rank", rank, "replicas differ after step", it, flush=True)
                ok = False
        del bucket, opt

    # timing at mapping size (1M Gaussians, 12M floats): fused step vs all-reduce + FusedAdam on every replica
    def timeit(fn, n=50):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    P = 1_000_000
    b1 = mapping.GradBucket(scene(P, 9, dev), dev, direct=False, peer_group=dist.group.WORLD, layout=mapping.raw_layout(1),
                            shard_optimizer=True)
    o1 = mapping.ShardedAdam(b1, LRS)
    b1.grads.normal_()
    t_fused = timeit(o1.step)
    b2 = mapping.GradBucket(scene(P, 9, dev), dev, direct=False, peer_group=dist.group.WORLD, layout=mapping.raw_layout(1))
    lv = b2.leaves()
    o2 = slam.FusedAdam([{'params': [v], 'name': k, 'lr': LRS[k]} for k, v in lv.items()], lr=0.0, eps=1e-15)
    b2.grads.normal_()

    def plain():
        b2.all_reduce()
        with torch.no_grad():
            o2.step()
    t_plain = timeit(plain)
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PEER_REDUCE_ADAM", "OK" if float(flag) == 1.0 else "MISMATCH", "world", world,
              "fused_ms %.4f allreduce+adam_ms %.4f (%s)" % (t_fused, t_plain, "in-switch" if b1.peer.multicast else "peer"),
              flush=True)
    dist.destroy_process_group()
    sys.exit(0 if float(flag) == 1.0 else 1)


if __name__ == "__main__":
    main()
