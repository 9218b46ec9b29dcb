import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_parity import _full_case, run_ours, run_ref
rs, sc, _ = _full_case(2_000_000, "SCANNETPP", 3, aniso=0.2, seed=3, kf=2, nkf=8)
dL = torch.ones(3, rs.image_height, rs.image_width, device="cuda") * 1e-3
o = run_ours("hx", rs, sc, dL, power=2); r = run_ref("hx", rs, sc, dL, power=2)
H, W = rs.image_height, rs.image_width
T = ((W + 15) // 16) * ((H + 15) // 16)
print("R", o["num_rendered"], r["num_rendered"], "ranges eq", torch.equal(o["views"]["ranges"], r["ranges"][:T]),
      "pl eq", torch.equal(o["views"]["point_list"], r["point_list"]))
d = (o["color"] - r["color"]).abs()
print("max abs diff", float(d.max()), "n diff px", int((d > 0).sum()), "of", d.numel())
idx = int(d.flatten().argmax()); ch, rem = divmod(idx, H * W); y, x = divmod(rem, W)
print("at ch,y,x", ch, y, x, float(o["color"][ch, y, x]), float(r["color"][ch, y, x]), "n_contrib", int(o["views"]["n_contrib"][y, x]),
      "T", float(o["views"]["final_T"][y, x]), float(r["final_T"][y * W + x]))
lens = (o["views"]["ranges"][:, 1] - o["views"]["ranges"][:, 0])
print("max tile len", int(lens.max()), "tile of px", (y // 16) * ((W + 15) // 16) + x // 16, "len", int(lens[(y // 16) * ((W + 15) // 16) + x // 16]))
dT = (o["views"]["final_T"].flatten() - r["final_T"]).abs()
print("final_T max diff", float(dT.max()), "depth diff", float((o["depth"] - r["depth"]).abs().max()))
rel = d / (torch.maximum(o["color"].abs(), r["color"].abs()) + 1e-6 * float(r["color"].abs().max()))
i2 = int(rel.flatten().argmax()); ch, rem = divmod(i2, H * W); y, x = divmod(rem, W)
print("max rel", float(rel.max()), "at", ch, y, x, float(o["color"][ch, y, x]), float(r["color"][ch, y, x]))

# carve the reference geometry buffer (GeometryState::fromChunk, rasterizer_impl.cu:155-169)
from oracle import ref_ext
f = ref_ext.forward("hx", rs, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"])
P = sc["means3D"].shape[0]
buf = f["geom"]; base = buf.data_ptr()
def al(x): return (x + 127) // 128 * 128
off = al(base) - base
def take(nbytes):
    global off
    o0 = al(base + off) - base
    off = o0 + nbytes
    return buf[o0:o0 + nbytes]
depths = take(4 * P).view(torch.float32)
clamped = take(3 * P)
radii_i = take(4 * P).view(torch.int32)
means2D = take(8 * P).view(torch.float32).view(P, 2)
cov3D = take(24 * P).view(torch.float32).view(P, 6)
conic_o = take(16 * P).view(torch.float32).view(P, 4)
vis = o["radii"] > 0
v = o["views"]
print("depth eq", torch.equal(v["depths"][vis], depths[vis]), "xy eq", torch.equal(v["means2D"][vis], means2D[vis]),
      "conic eq", torch.equal(v["conic_opacity"][vis], conic_o[vis]))
dc = (v["conic_opacity"][vis] - conic_o[vis]).abs() / (conic_o[vis].abs() + 1e-30)
print("conic max rel diff", dc.max(0)[0].tolist(), "n differing rows", int((dc > 0).any(1).sum()), "of", int(vis.sum()))
dxy = (v["means2D"][vis] - means2D[vis]).abs()
print("xy max abs diff", float(dxy.max()), "n", int((dxy > 0).any(1).sum()))
aniso = (sc["scales"][:, 0] != sc["scales"][:, 1])
print("differing rows that are anisotropic:", int(((dc > 0).any(1) & aniso[vis]).sum()), "iso:", int(((dc > 0).any(1) & ~aniso[vis]).sum()))
