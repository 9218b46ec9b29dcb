"""PCIe probe: H2D alone, D2H alone, both at once (two streams), 64 MB pinned buffers."""
import torch, time
n = 16 * 1024 * 1024
h_in = torch.empty(n, dtype=torch.float32).pin_memory(); h_out = torch.empty(n, dtype=torch.float32).pin_memory()
d_in = torch.empty(n, device="cuda"); d_out = torch.empty(n, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, it=20):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(it):
        if h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / it
    return dt
for _ in range(2): run(True, True, 3)
a, b, c = run(True, False), run(False, True), run(True, True)
gb = n * 4 / 1e9
print("H2D alone %.1f GB/s  D2H alone %.1f GB/s  both: %.3f ms per pair -> %.1f GB/s each direction" % (gb / a, gb / b, c * 1e3, gb / c))
