"""Probe (run under torchrun): does torch symmetric memory come up on this box, is there NVSwitch multicast, and how
fast are NCCL / symm_mem all-reduces of the 56 MB gradient bucket?"""
import os, time, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 14 * 1_000_000
x = torch.randn(n, device=dev)
def timeit(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
t_nccl = timeit(lambda: dist.all_reduce(x))
if rank == 0: print("nccl all_reduce 56MB ms", t_nccl, flush=True)
try:
    t = symm_mem.empty(n, dtype=torch.float32, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    if rank == 0:
        try:
            print("symm ok; mc_ptr", hex(hdl.multicast_ptr or 0), "bufs", [hex(p) for p in hdl.buffer_ptrs],
                  "pad size", hdl.signal_pad_size, flush=True)
        except Exception as e:
            print("print failed", repr(e)[:200], flush=True)
    t.copy_(x)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            if name == "one_shot_all_reduce":
                f = lambda: op(t, "sum", dist.group.WORLD.group_name)
            else:
                f = lambda: op(t, "sum", dist.group.WORLD.group_name)
            ms = timeit(f)
            if rank == 0: print(name, "ms", ms, flush=True)
        except Exception as e:
            if rank == 0: print(name, "failed:", repr(e)[:300], flush=True)
except Exception as e:
    print("rank", rank, "symm_mem failed:", repr(e)[:500], flush=True)
dist.destroy_process_group()
