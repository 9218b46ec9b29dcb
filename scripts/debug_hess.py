"""C4 Hessian (backward_power=2): where do our per-Gaussian sums differ from the reference build's, and by how much
relative to the Gaussian's own row?  (Run on the GPU box; needs oracle/_ref.)"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_parity import _full_case, run_ours, run_ref
rs, sc, _ = _full_case(2_000_000, "SCANNETPP", 3, aniso=0.2, seed=3, kf=2, nkf=8)
dL = torch.ones(3, rs.image_height, rs.image_width, device="cuda") * 1e-3
o = run_ours("hx", rs, sc, dL, power=2); r = run_ref("hx", rs, sc, dL, power=2); r2 = run_ref("hx", rs, sc, dL, power=2)
print("colour eq", torch.equal(o["color"], r["color"]), float((o["color"] - r["color"]).abs().max()))
aniso = (sc["scales"].max(1)[0] != sc["scales"].min(1)[0])
for k in ("means3D", "means2D", "colors", "opacities", "scales", "rotations"):
    a, b, b2 = o["grads"][k].double(), r["grads"][k].double(), r2["grads"][k].double()
    if a.dim() == 1: a, b, b2 = a[:, None], b[:, None], b2[:, None]
    gmax = float(b.abs().max())
    err = (a - b).abs()
    self_err = (b - b2).abs()
    rel_el = err / (torch.maximum(a.abs(), b.abs()) + 1e-30)
    rowmax = b.abs().max(1, keepdim=True)[0]
    rel_row = err / (torch.maximum(a.abs(), b.abs()) + 1e-3 * rowmax + 1e-30)
    rel_glob = err / (torch.maximum(a.abs(), b.abs()) + 1e-3 * gmax)
    bad = rel_glob > 1e-3
    print(k, "gmax %.3e" % gmax, "ref self max rel %.2e" % float((self_err / (b.abs() + 1e-3 * gmax)).max()),
          "| elementwise max %.2e" % float(rel_el.max()), "row-floor max %.2e" % float(rel_row.max()),
          "global-floor max %.2e" % float(rel_glob.max()), "n bad(global) %d" % int(bad.sum()),
          "of which aniso rows %d" % int((bad.any(1) & aniso).sum()))
    if int(bad.sum()):
        rows = bad.any(1).nonzero().flatten()[:6]
        for i in rows.tolist():
            print("   row", i, "ours", a[i].tolist(), "ref", b[i].tolist(), "scales", sc["scales"][i].tolist(),
                  "radius", int(o["radii"][i]))
    for q in (0.5, 0.9, 0.99, 0.999, 0.9999):
        nzr = rel_glob.flatten()
        print("     q%.4f rel(global floor) %.2e   rel(row floor) %.2e" % (q, float(torch.quantile(nzr[:4_000_000], q)),
                                                                         float(torch.quantile(rel_row.flatten()[:4_000_000], q))))
