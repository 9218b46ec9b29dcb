import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import build_case, rel_err
from test_gpu_parity import run_ours, run_ref
case = sys.argv[1] if len(sys.argv) > 1 else "surface_small"
rs, sc, dL = build_case(case, "cuda")
o = run_ours("dg", rs, sc, dL); r = run_ref("dg", rs, sc, dL); r2 = run_ref("dg", rs, sc, dL)
print("cull disabled:", os.environ.get("SPLAT_B200_NO_CULL"))
for k in ("means2D", "opacities", "colors", "means3D", "scales"):
    a, b, b2 = o["grads"][k].double(), r["grads"][k].double(), r2["grads"][k].double()
    mx = b.abs().max()
    err = (a - b).abs() / (torch.maximum(a.abs(), b.abs()) + 1e-5 * mx)
    flat = err.flatten().argsort(descending=True)[:5]
    print(k, "max|ref|=%.3g" % mx)
    for i in flat:
        row = int(i) // a.shape[1]
        print("   idx %d row %d err %.2e ours %.6e ref %.6e ref2 %.6e radius %d opac %.4f scale %.4g xy %s" % (
            int(i), row, float(err.flatten()[i]), float(a.flatten()[i]), float(b.flatten()[i]), float(b2.flatten()[i]),
            int(o["radii"][row]), float(sc["opacities"][row]), float(sc["scales"][row, 0]),
            o["views"]["means2D"][row].tolist()))
