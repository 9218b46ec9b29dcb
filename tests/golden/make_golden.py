"""Generates the golden fixtures in tests/golden/ from the UNMODIFIED reference CUDA extensions
(oracle/_ref/ref_{hx,dg}_C.so, built by oracle/build_ref.py from /root/reference).

Run on a GPU box:   python tests/golden/make_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/ and commit them.  Inputs are NOT stored: they
are regenerated from the seeds in tests/util.py::CASES.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from util import build_case  # noqa: E402
from oracle import ref_ext  # noqa: E402

JOBS = [("cloud_small", "dg", 1), ("cloud_small", "hx", 1), ("cloud_small", "hx", 2),
        ("surface_small", "dg", 1), ("surface_small", "hx", 2), ("cloud_ragged", "hx", 1),
        ("c1_cloud", "dg", 1)]


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    for case, variant, power in JOBS:
        rs, sc, dL = build_case(case, "cuda")
        if power == 2:
            dL = torch.ones_like(dL) * 1e-3
        f = ref_ext.forward(variant, rs, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                            sc["rotations"])
        g = ref_ext.backward(f, dL, power)
        torch.cuda.synchronize()
        fT, nc, rg = ref_ext.image_state_views(f)
        H, W = rs.image_height, rs.image_width
        T = ((W + 15) // 16) * ((H + 15) // 16)
        out = dict(case=case, variant=variant, power=power, num_rendered=f["num_rendered"],
                   color=f["color"].cpu().numpy(), depth=f["depth"].cpu().numpy(), radii=f["radii"].cpu().numpy(),
                   ranges=rg[:T].cpu().numpy(), n_contrib=nc.cpu().numpy().reshape(H, W),
                   final_T=fT.cpu().numpy().reshape(H, W), point_list=ref_ext.binning_point_list(f).cpu().numpy())
        for k in ("means3D", "means2D", "colors", "opacities", "scales", "rotations", "cov3D"):
            out["g_" + k] = g[k].cpu().numpy()
        fn = os.path.join(out_dir, "%s_%s_p%d.npz" % (case, variant, power))
        np.savez_compressed(fn, **out)
        print("wrote", fn, "R =", f["num_rendered"])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "_generated"))
