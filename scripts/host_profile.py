"""cProfile of the host side of the C3 step loop (where do the ~0.45 ms of Python per step go?)."""
import cProfile, pstats, os, sys, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
cam, sc, cams, dL, power = bench.build_workload("replica_1200x680_1M", "cuda")
r = bench.OursRunner(sc, cams, dL, power, "hx", torch.device("cuda", 0))
for i in range(20):
    r.step(i % 24)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for i in range(200):
    r.step(i % 24)
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
print(s.getvalue()[:6000])
