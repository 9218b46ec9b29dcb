"""Where does a GraphedIteration capture spend its time?  (host wall clock, per phase)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_slam
from splatam_b200 import slam, rasterizer

dev = torch.device("cuda", 0)
cam, rs, params = bench_slam.build("tracking", dev, 24)
targets = bench_slam.make_targets(params, rs, 24, dev)
variables = {"max_2D_radius": torch.zeros(params["means3D"].shape[0], device=dev)}
wts = dict(im=0.5, depth=1.0)
lrs = dict(means3D=0.0, rgb_colors=0.0, unnorm_rotations=0.0, logit_opacities=0.0, log_scales=0.0, cam_unnorm_rots=0.0004, cam_trans=0.002)
opt = slam.FusedAdam([{'params': [v], 'name': k, 'lr': lrs[k]} for k, v in params.items()])

def it(kf):
    loss, _, _ = slam.get_loss(params, targets[kf], variables, kf, wts, True, 0.99, True, False, tracking=True, tracking_pose_only=True)
    loss.backward()
    opt.step()
    return loss

def sync():
    torch.cuda.synchronize()
    return time.perf_counter()

for kf in range(3):
    opt.zero_grad(set_to_none=True); it(kf)
pool = torch.cuda.graph_pool_handle()
keep = []
for rep in range(6):
    kf = rep
    t0 = sync()
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            opt.zero_grad(set_to_none=True); it(kf)
    torch.cuda.current_stream().wait_stream(side)
    t1 = sync()
    opt.zero_grad(set_to_none=True)
    rasterizer.capture_guard(dev).zero_()
    g = torch.cuda.CUDAGraph()
    t2 = time.perf_counter()
    with torch.cuda.stream(side):
        g.capture_begin(pool=pool)
        out = it(kf)
        t3 = time.perf_counter()
        g.capture_end()
    t4 = time.perf_counter()
    g.replay(); t5 = sync()
    for _ in range(20): g.replay()
    t6 = sync()
    keep = [g, out]   # previous graph dies here
    t7 = time.perf_counter()
    print("rep %d: warmup(3 it) %.2f ms | setup %.2f | capture body %.2f | capture_end(instantiate) %.2f | first replay %.2f | 20 replays %.2f (%.3f each) | drop old graph %.2f"
          % (rep, (t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3, (t5-t4)*1e3, (t6-t5)*1e3, (t6-t5)*1e3/20, (t7-t6)*1e3))
