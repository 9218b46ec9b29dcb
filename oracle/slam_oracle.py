"""CPU ORACLE (TEST INFRASTRUCTURE ONLY) for SURVEY.md 8f-2 / 8f-3: the PyTorch glue SplaTAM runs on
either side of the rasterizer -- the render-variable builder (`transform_to_frame`,
`transformed_params2rendervar`) and the image-space loss of `get_loss` (masked L1, L1 + SSIM).
Plain torch ops in the reference's order; gradients come from torch autograd, exactly as in the
reference flow.  Nothing under splatam_b200/ may import this file.

Parity pin: tests/golden/slam_glue_*.npz were produced by the reference's OWN Python
(/root/reference/utils/slam_helpers.py, utils/slam_external.py and the `get_loss` function body of
scripts/splatam.py, executed unmodified on CPU by tests/golden/make_slam_golden.py);
tests/test_slam_oracle_golden.py checks this file against them, so this oracle is PINNED.
"""
import math

import torch
import torch.nn.functional as F


# ---------------------------------------------------------------------------------------------
# utils/slam_external.py:25-42
def build_rotation(q):
    norm = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    q = q / norm[:, None]
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rows = [torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], -1),
            torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], -1),
            torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1)]
    return torch.stack(rows, 1)


# utils/slam_helpers.py:21-28
def quat_mult(q1, q2):
    w1, x1, y1, z1 = q1.T
    w2, x2, y2, z2 = q2.T
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack([w, x, y, z]).T


# utils/slam_helpers.py:253-304
def transform_to_frame(params, time_idx, gaussians_grad, camera_grad):
    if camera_grad:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx])
        cam_tran = params['cam_trans'][..., time_idx]
    else:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx].detach())
        cam_tran = params['cam_trans'][..., time_idx].detach()
    R = build_rotation(cam_rot)[0]
    transform_rots = params['log_scales'].shape[1] != 1
    if gaussians_grad:
        pts, unnorm_rots = params['means3D'], params['unnorm_rotations']
    else:
        pts, unnorm_rots = params['means3D'].detach(), params['unnorm_rotations'].detach()
    out = {'means3D': pts @ R.T + cam_tran[0]}       # == (rel_w2c @ [pts,1]^T)^T[:, :3]
    if transform_rots:
        out['unnorm_rotations'] = quat_mult(cam_rot, F.normalize(unnorm_rots))
    else:
        out['unnorm_rotations'] = unnorm_rots
    return out


# utils/slam_helpers.py:123-139 (means2D left to the caller)
def transformed_params2rendervar(params, transformed_gaussians):
    if params['log_scales'].shape[1] == 1:
        log_scales = torch.tile(params['log_scales'], (1, 3))
    else:
        log_scales = params['log_scales']
    return {'means3D': transformed_gaussians['means3D'],
            'colors_precomp': params['rgb_colors'],
            'rotations': F.normalize(transformed_gaussians['unnorm_rotations']),
            'opacities': torch.sigmoid(params['logit_opacities']),
            'scales': torch.exp(log_scales)}


def build_rendervar(params, time_idx, gaussians_grad, camera_grad):
    return transformed_params2rendervar(params, transform_to_frame(params, time_idx, gaussians_grad, camera_grad))


# utils/slam_helpers.py:196-212
def get_depth_and_silhouette(pts_3D, w2c):
    pts4 = torch.cat((pts_3D, torch.ones_like(pts_3D[:, :1])), dim=-1)
    z = (w2c @ pts4.transpose(0, 1)).transpose(0, 1)[:, 2]
    return torch.stack([z, torch.ones_like(z), z * z], 1)


# ---------------------------------------------------------------------------------------------
# utils/slam_external.py:54-97
def _window(window_size, channel):
    g = torch.tensor([math.exp(-(x - window_size // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(window_size)])
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2.expand(channel, 1, window_size, window_size).contiguous()


def calc_ssim(img1, img2, window_size=11):
    channel = img1.size(-3)
    window = _window(window_size, channel).type_as(img1)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + c1) * (2 * sigma12 + c2)) / ((mu1_sq + mu2_sq + c1) * (sigma1_sq + sigma2_sq + c2))
    return ssim_map.mean()


# scripts/splatam.py:262-296, 342-346 -- the image-space part of get_loss: `im` [3,H,W] and `depth_sil` [3,H,W]
# are what the two Renderer calls return (scripts/splatam.py:255-259).
def image_loss(im, depth_sil, gt_im, gt_depth, loss_weights, use_sil_for_loss, sil_thres, use_l1,
               ignore_outlier_depth_loss, tracking):
    losses = {}
    depth = depth_sil[0, :, :].unsqueeze(0)
    silhouette = depth_sil[1, :, :]
    presence_sil_mask = (silhouette > sil_thres)
    depth_sq = depth_sil[2, :, :].unsqueeze(0)
    uncertainty = (depth_sq - depth ** 2).detach()
    nan_mask = (~torch.isnan(depth)) & (~torch.isnan(uncertainty))
    if ignore_outlier_depth_loss:
        depth_error = torch.abs(gt_depth - depth) * (gt_depth > 0)
        mask = (depth_error < 10 * depth_error.median())
        mask = mask & (gt_depth > 0)
    else:
        mask = (gt_depth > 0)
    mask = mask & nan_mask
    if tracking and use_sil_for_loss:
        mask = mask & presence_sil_mask
    if use_l1:
        mask = mask.detach()
        if tracking:
            losses['depth'] = torch.abs(gt_depth - depth)[mask].sum()
        else:
            losses['depth'] = torch.abs(gt_depth - depth)[mask].mean()
    if tracking and (use_sil_for_loss or ignore_outlier_depth_loss):
        color_mask = torch.tile(mask, (3, 1, 1)).detach()
        losses['im'] = torch.abs(gt_im - im)[color_mask].sum()
    elif tracking:
        losses['im'] = torch.abs(gt_im - im).sum()
    else:
        losses['im'] = 0.8 * torch.abs(im - gt_im).mean() + 0.2 * (1.0 - calc_ssim(im, gt_im))
    weighted = {k: v * loss_weights[k] for k, v in losses.items()}
    weighted['loss'] = sum(weighted.values())
    return weighted['loss'], weighted


# ---------------------------------------------------------------------------------------------
def random_params(P, n_frames, iso, seed, dtype=torch.float32):
    """Seeded SplaTAM-style parameter dict (scripts/splatam.py:118-140 shapes)."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g, dtype=dtype)
    params = {
        'means3D': r(P, 3) * torch.tensor([1.5, 1.0, 0.8], dtype=dtype) + torch.tensor([0.0, 0.0, 3.0], dtype=dtype),
        'rgb_colors': torch.rand(P, 3, generator=g, dtype=dtype),
        'unnorm_rotations': r(P, 4) * 0.7 + torch.tensor([1.0, 0, 0, 0], dtype=dtype),
        'logit_opacities': r(P, 1) * 1.5,
        'log_scales': r(P, 1 if iso else 3) * 0.5 - 3.5,
        'cam_unnorm_rots': (r(1, 4, n_frames) * 0.05 + torch.tensor([1.0, 0, 0, 0], dtype=dtype).view(1, 4, 1)) * 1.3,
        'cam_trans': r(1, 3, n_frames) * 0.1,
    }
    return params


def images(h, w, seed):
    """Seeded render-like images: im/gt_im in [0,1], depth ~ 2..4 m with holes, silhouette mostly ~1."""
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    base = 0.5 + 0.3 * torch.sin(xx / 7.0)[None] * torch.cos(yy / 5.0)[None] * torch.tensor([1.0, 0.7, -0.8]).view(3, 1, 1)
    gt_im = (base + 0.05 * torch.randn(3, h, w, generator=g)).clamp(0, 1)
    im = (base + 0.08 * torch.randn(3, h, w, generator=g)).clamp(0, 1.2)
    gt_depth = 3.0 + torch.sin(xx / 11.0) * 0.5 + 0.02 * torch.randn(h, w, generator=g)
    gt_depth[torch.rand(h, w, generator=g) < 0.1] = 0.0            # invalid depth
    depth = 3.0 + torch.sin(xx / 11.0) * 0.5 + 0.05 * torch.randn(h, w, generator=g)
    sil = (1.0 - 0.02 * torch.rand(h, w, generator=g))
    sil[torch.rand(h, w, generator=g) < 0.15] *= 0.6                # empty space
    depth_sq = depth * depth + 0.01 * torch.rand(h, w, generator=g)
    depth = depth.clone()
    depth[3, 5] = float("nan")                                       # exercised by nan_mask
    depth_sil = torch.stack([depth, sil, depth_sq], 0)
    return im, depth_sil, gt_im, gt_depth[None]


LOSS_CASES = [
    dict(name="track_sil", tracking=True, use_sil=True, w=dict(im=0.5, depth=1.0), sil_thres=0.99, h=45, w_=70, seed=1),
    dict(name="track_nosil", tracking=True, use_sil=False, w=dict(im=0.5, depth=1.0), sil_thres=0.99, h=33, w_=40, seed=2),
    dict(name="map", tracking=False, use_sil=False, w=dict(im=0.5, depth=1.0), sil_thres=0.5, h=45, w_=70, seed=3),
    dict(name="map_ragged", tracking=False, use_sil=False, w=dict(im=0.7, depth=0.3), sil_thres=0.5, h=37, w_=53, seed=4),
]
RV_CASES = [
    dict(name="iso_track", iso=True, gg=False, cg=True, P=257, seed=11),
    dict(name="iso_map", iso=True, gg=True, cg=False, P=257, seed=12),
    dict(name="aniso_ba", iso=False, gg=True, cg=True, P=300, seed=13),
    dict(name="aniso_track", iso=False, gg=False, cg=True, P=300, seed=14),
]
