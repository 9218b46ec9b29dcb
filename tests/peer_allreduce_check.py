"""Run under torchrun on >= 2 GPUs of one node (not collected by pytest: needs N GPUs; scripts/gpu_n2.sh runs it):
the library's peer-memory all-reduce must equal NCCL's all_reduce bit for bit in rank order... (sum order differs from
NCCL's, so the check is against an explicit gather + ordered sum), for an odd bucket size and repeated calls."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from splatam_b200 import mapping  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for numel, mc in ((1_000_003, False), (14_000_000, False), (1_000_003, True), (14_000_000, True)):
        pa = mapping.PeerAllReduce(numel, dev, dist.group.WORLD, multicast=mc)
        if mc and not pa.multicast:
            if rank == 0:
                print("no multicast mapping on this fabric; NVLS path not tested", flush=True)
            continue
        for it in range(4):
            g = torch.Generator(device=dev).manual_seed(100 * it + rank)
            x = torch.randn(numel, generator=g, device=dev)
            pa.buffer.zero_()
            pa.buffer[:numel].copy_(x)
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            want = parts[0].clone()
            for j in range(1, world):
                want += parts[j]
            # (0 + x0) + x1 ... is what the kernel computes; adding to 0 first is exact
            torch.cuda.synchronize()
            pa()
            torch.cuda.synchronize()
            got = pa.buffer[:numel]
            # peer loads: ordered sum, bit-exact.  in-switch: the switch's order, so a rounding-level tolerance -- but
            # every replica must hold the SAME bits (checked below through a MAX/MIN all-reduce of a checksum)
            same = torch.equal(got, want) if not mc else bool(((got - want).abs() <= 1e-5 * (want.abs() + 1.0)).all())
            if mc:
                cs = got.double().sum().reshape(1)
                lo_, hi_ = cs.clone(), cs.clone()
                dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
                same = same and bool(lo_ == hi_)
            ok = ok and same
            if not same:
                print("rank", rank, "numel", numel, "iter", it, "max diff", float((got - want).abs().max()), flush=True)
    # timing of the 56 MB bucket against NCCL
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
    y = torch.randn(14_000_000, device=dev)
    t_nccl = timeit(lambda: dist.all_reduce(y))
    t_peer = timeit(pa)
    pa_p2p = mapping.PeerAllReduce(14_000_000, dev, dist.group.WORLD, multicast=False)
    t_p2p = timeit(pa_p2p)
    if rank == 0:
        print("56 MB bucket: in-switch (multimem) ms %.4f   peer loads/stores ms %.4f" % (t_peer if pa.multicast else float("nan"), t_p2p), flush=True)
    small = mapping.PeerAllReduce(4096, dev, dist.group.WORLD)
    t_small = timeit(small)
    if rank == 0:
        print("fixed cost (4096 floats) ms %.4f" % t_small, flush=True)
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PEER_ALLREDUCE", "OK" if float(flag) == 1.0 else "MISMATCH", "world", world,
              "peer_ms %.4f nccl_ms %.4f" % (t_peer, t_nccl), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if float(flag) == 1.0 else 1)


if __name__ == "__main__":
    main()
