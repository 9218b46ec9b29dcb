"""Where does a fused tracking iteration spend its time?  Per-phase CUDA-event and host timings, gather on / off."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench_slam as BS  # noqa: E402
from splatam_b200 import slam  # noqa: E402
from splatam_b200._lib import lib  # noqa: E402

dev = torch.device("cuda", 0)
cam, rs, params = BS.build("tracking", dev, 24)
targets = BS.make_targets(params, rs, 24, dev)
P = params["means3D"].shape[0]
variables = {"max_2D_radius": torch.zeros(P, device=dev)}
wts = dict(im=0.5, depth=1.0)
for gather in (0, 1, 0, 1):
    lib.splat_set_gather(gather)
    step = BS.make_step("ours", "dg", params, targets, variables, wts, True, 0.99, pose_only=True)
    for i in range(10):
        for v in params.values():
            v.grad = None
        step((7 * i) % 24)
    torch.cuda.synchronize()
    lib.splat_profile_enable(1)
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 48
    for i in range(n):
        for v in params.values():
            v.grad = None
        step((7 * i) % 24)
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3 / n
    k = 16
    ms = (C.c_float * k)(); cnt = (C.c_int32 * k)(); gaps = (C.c_float * k)()
    lib.splat_profile_read(ms, cnt, k)
    lib.splat_profile_read_gaps(gaps, k)
    lib.splat_profile_enable(0)
    names = [lib.splat_profile_kernel_name(j).decode() for j in range(k)]
    print("gather", gather, "gpu ms/step %.3f wall %.3f" % (e0.elapsed_time(e1) / n, wall))
    print("  kernel", {names[j]: round(ms[j] / n, 4) for j in range(k) if cnt[j]})
    print("  gap   ", {names[j]: round(gaps[j] / n, 4) for j in range(k) if cnt[j]})
