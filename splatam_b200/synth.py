"""Deterministic synthetic cameras and Gaussian scenes.

Everything is generated on the CPU from a seeded torch.Generator so that the CPU oracle and the
GPU see bit-identical inputs; callers move the tensors to the device they need.  There are no
datasets on the GPU box, so these scenes stand in for Replica / ScanNet++ / TUM shapes.
"""
import math

import torch

from .settings import GaussianRasterizationSettings


def camera_matrices(w, h, fx, fy, cx, cy, w2c=None, near=0.01, far=100.0):
    """viewmatrix/projmatrix/campos in the layout of utils/recon_helpers.py:4-26 (setup_camera):
    viewmatrix = w2c^T, projmatrix = (P_gl . w2c)^T, both float32 [4,4]."""
    if w2c is None:
        w2c = torch.eye(4)
    w2c = w2c.float()
    cam_center = torch.inverse(w2c)[:3, 3].contiguous()
    view = w2c.t().contiguous()
    gl = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                       [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                       [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                       [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32).t()
    full = (view @ gl).contiguous()
    return view, full, cam_center


def make_settings(w, h, fx, fy, cx, cy, w2c=None, device="cpu", bg=(0.0, 0.0, 0.0), sh_degree=0):
    view, full, center = camera_matrices(w, h, fx, fy, cx, cy, w2c)
    return GaussianRasterizationSettings(
        image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy),
        bg=torch.tensor(bg, dtype=torch.float32, device=device), scale_modifier=1.0,
        viewmatrix=view.unsqueeze(0).to(device), projmatrix=full.unsqueeze(0).to(device),
        sh_degree=sh_degree, campos=center.to(device), prefiltered=False)


def settings_to(rs, device):
    return rs._replace(bg=rs.bg.to(device), viewmatrix=rs.viewmatrix.to(device),
                       projmatrix=rs.projmatrix.to(device), campos=rs.campos.to(device))


def cloud_scene(P, w, h, fx, fy, seed=0, behind_frac=0.05):
    """C1-style random anisotropic cloud filling the frustum."""
    g = torch.Generator().manual_seed(seed)
    tanx, tany = w / (2 * fx), h / (2 * fy)
    z = torch.rand(P, generator=g) * 5.5 + 0.5
    nb = int(P * behind_frac)
    if nb:
        z[:nb] = torch.rand(nb, generator=g) * 1.3 - 1.0
    x = (torch.rand(P, generator=g) * 2 - 1) * 1.2 * tanx * z
    y = (torch.rand(P, generator=g) * 2 - 1) * 1.2 * tany * z
    means = torch.stack([x, y, z], 1).contiguous()
    lo, hi = math.log(0.005), math.log(0.15)
    scales = torch.exp(torch.rand(P, 3, generator=g) * (hi - lo) + lo)
    q = torch.randn(P, 4, generator=g)
    rots = q / q.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    colors = torch.rand(P, 3, generator=g)
    return dict(means3D=means, scales=scales.contiguous(), rotations=rots.contiguous(),
                opacities=opac.contiguous(), colors_precomp=colors.contiguous())


def surface_depth(u, v):
    return 2.5 + 0.8 * torch.sin(u / 97.0) + 0.6 * torch.cos(v / 61.0) + 0.3 * torch.sin((u + v) / 23.0)


def surface_scene(P, w, h, fx, fy, cx, cy, seed=1, layers=1, aniso_frac=0.0, scale_lo=0.7, scale_hi=3.0):
    """C2-C4 style "surface" scene: P Gaussians back-projected from a smooth synthetic depth map,
    isotropic scale = s*z/f with s ~ LogU(scale_lo, scale_hi) (SplaTAM initialises s=1,
    scripts/splatam.py:103-106), identity rotation, logit-opacity ~ N(0,1), procedural colour.
    `layers` stacks the surface at depth offsets to reach large P with realistic occlusion."""
    g = torch.Generator().manual_seed(seed)
    per = (P + layers - 1) // layers
    chunks = []
    for li in range(layers):
        n = min(per, P - li * per)
        if n <= 0:
            break
        # stratified jitter over the image plane
        side = int(math.ceil(math.sqrt(n * w / h)))
        rows = int(math.ceil(n / side))
        idx = torch.arange(n)
        u = ((idx % side).float() + torch.rand(n, generator=g)) * (w / side)
        v = ((idx // side).float() + torch.rand(n, generator=g)) * (h / rows)
        z = surface_depth(u, v) + 0.35 * li
        chunks.append((u, v, z))
    u = torch.cat([c[0] for c in chunks]); v = torch.cat([c[1] for c in chunks]); z = torch.cat([c[2] for c in chunks])
    n = u.numel()
    perm = torch.randperm(n, generator=g)   # SLAM maps are not depth-ordered in memory
    u, v, z = u[perm], v[perm], z[perm]
    means = torch.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1).contiguous()
    s = torch.exp(torch.rand(n, generator=g) * (math.log(scale_hi) - math.log(scale_lo)) + math.log(scale_lo))
    iso = (s * z / ((fx + fy) / 2)).unsqueeze(1).repeat(1, 3)
    rots = torch.zeros(n, 4); rots[:, 0] = 1.0
    na = int(n * aniso_frac)
    if na:
        iso[:na] = iso[:na] * torch.exp(torch.randn(na, 3, generator=g) * 0.5)
        q = torch.randn(na, 4, generator=g)
        rots[:na] = q / q.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(torch.randn(n, 1, generator=g))
    colors = torch.stack([0.5 + 0.5 * torch.sin(u / 37.0), 0.5 + 0.5 * torch.cos(v / 29.0),
                          0.5 + 0.5 * torch.sin((u - v) / 53.0)], 1)
    return dict(means3D=means, scales=iso.contiguous(), rotations=rots.contiguous(),
                opacities=opac.contiguous(), colors_precomp=colors.contiguous())


def arc_pose(k, n, radius=0.5, target_z=2.5):
    """k-th of n world-to-camera poses on an arc looking at the surface (C3 keyframes)."""
    ang = (k / max(n - 1, 1) - 0.5) * 0.5
    c, s = math.cos(ang), math.sin(ang)
    # camera centre on an arc around (0,0,target_z), looking at it
    cx, cz = radius * math.sin(ang) * 2.0, target_z - target_z * math.cos(ang) * 1.0
    R = torch.tensor([[c, 0.0, -s], [0.0, 1.0, 0.0], [s, 0.0, c]], dtype=torch.float32)
    t = -R @ torch.tensor([cx, 0.0, cz], dtype=torch.float32)
    w2c = torch.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = t
    return w2c


REPLICA = dict(w=1200, h=680, fx=600.0, fy=600.0, cx=599.5, cy=339.5)
SCANNETPP = dict(w=1752, h=1168, fx=1200.0, fy=1200.0, cx=875.5, cy=583.5)
TUM = dict(w=640, h=480, fx=517.3, fy=516.5, cx=318.6, cy=255.3)
C1 = dict(w=64, h=64, fx=64.0, fy=64.0, cx=31.5, cy=31.5)
