"""CPU ORACLE (TEST INFRASTRUCTURE ONLY) for SURVEY.md 8f-4: the per-frame map surgery SplaTAM does around the
rasterizer -- pruning (`prune_gaussians` / `remove_points`), opacity reset, and growing the map from a frame
(`add_new_gaussians` / `get_pointcloud` / `initialize_new_params`), INCLUDING the Adam-state surgery.  Plain torch
ops in the reference's order.  Nothing under splatam_b200/ may import this file.

Parity pin: tests/golden/densify_*.npz were produced by the reference's OWN Python (utils/slam_external.py and the
`add_new_gaussians` / `get_pointcloud` / `initialize_new_params` bodies of scripts/splatam.py, executed unmodified
on CPU by tests/golden/make_densify_golden.py); tests/test_densify_oracle_golden.py checks this file against
them, so this oracle is PINNED.

State convention here: `params` {name: [P,w] tensor}, `state` {name: {"exp_avg","exp_avg_sq"}} for the Gaussian
tensors that have Adam state, `variables` {means2D_gradient_accum, denom, max_2D_radius, timestep, scene_radius}.
"""
import torch
import torch.nn.functional as F

from oracle.slam_oracle import build_rotation

GAUSS_KEYS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
VAR_KEYS = ("means2D_gradient_accum", "denom", "max_2D_radius", "timestep")


# utils/slam_external.py:143-165
def remove_points(to_remove, params, state, variables):
    keep = ~to_remove
    params = dict(params)
    state = {k: dict(v) for k, v in state.items()}
    variables = dict(variables)
    for k in GAUSS_KEYS:
        if k in state:
            state[k]["exp_avg"] = state[k]["exp_avg"][keep]
            state[k]["exp_avg_sq"] = state[k]["exp_avg_sq"][keep]
        params[k] = params[k][keep]
    for k in VAR_KEYS:
        if k in variables:
            variables[k] = variables[k][keep]
    return params, state, variables


# utils/slam_external.py:172-195
def prune_mask(params, variables, iter, prune_dict):
    """None when this iteration does not prune."""
    if iter > prune_dict["stop_after"]:
        return None
    if not ((iter >= prune_dict["start_after"]) and (iter % prune_dict["prune_every"] == 0)):
        return None
    thr = prune_dict["final_removal_opacity_threshold"] if iter == prune_dict["stop_after"] \
        else prune_dict["removal_opacity_threshold"]
    to_remove = (torch.sigmoid(params["logit_opacities"]) < thr).squeeze()
    if iter >= prune_dict["remove_big_after"]:
        big = torch.exp(params["log_scales"]).max(dim=1).values > 0.1 * variables["scene_radius"]
        to_remove = torch.logical_or(to_remove, big)
    return to_remove


def prune_gaussians(params, state, variables, iter, prune_dict):
    to_remove = prune_mask(params, variables, iter, prune_dict)
    if to_remove is not None:
        params, state, variables = remove_points(to_remove, params, state, variables)
    if iter <= prune_dict["stop_after"] and iter > 0 and iter % prune_dict["reset_opacities_every"] == 0 \
            and prune_dict["reset_opacities"]:
        # update_params_and_optimizer, utils/slam_external.py:107-120: new value, ZEROED Adam moments
        one = torch.ones_like(params["logit_opacities"]) * 0.01
        params = dict(params)
        params["logit_opacities"] = torch.log(one / (1 - one))
        if "logit_opacities" in state:
            state = {k: dict(v) for k, v in state.items()}
            state["logit_opacities"] = {"exp_avg": torch.zeros_like(one), "exp_avg_sq": torch.zeros_like(one)}
    return params, state, variables


# scripts/splatam.py:73-123 (transform_pts=True, mask given, projective mean-square distance)
def get_pointcloud(color, depth, intrinsics, w2c, mask):
    width, height = color.shape[2], color.shape[1]
    CX, CY, FX, FY = intrinsics[0][2], intrinsics[1][2], intrinsics[0][0], intrinsics[1][1]
    dev = depth.device
    x_grid, y_grid = torch.meshgrid(torch.arange(width, device=dev).float(), torch.arange(height, device=dev).float(),
                                    indexing="xy")
    xx = ((x_grid - CX) / FX).reshape(-1)
    yy = ((y_grid - CY) / FY).reshape(-1)
    depth_z = depth[0].reshape(-1)
    pts_cam = torch.stack((xx * depth_z, yy * depth_z, depth_z), dim=-1)
    pts4 = torch.cat((pts_cam, torch.ones(height * width, 1, device=dev)), dim=1)
    c2w = torch.inverse(w2c)
    pts = (c2w @ pts4.T).T[:, :3]
    scale_gaussian = depth_z / ((FX + FY) / 2)
    mean3_sq_dist = scale_gaussian ** 2
    cols = torch.permute(color, (1, 2, 0)).reshape(-1, 3)
    point_cld = torch.cat((pts, cols), -1)
    return point_cld[mask], mean3_sq_dist[mask]


# scripts/splatam.py:384-399: the mask of pixels the map does not explain yet.  depth_sil [3,H,W] is what the
# silhouette render returns.  Returns (mask before the valid-depth AND, mask after it).
def non_presence_mask(depth_sil, gt_depth, sil_thres):
    silhouette = depth_sil[1, :, :]
    non_presence_sil_mask = (silhouette < sil_thres)
    gt = gt_depth[0, :, :]
    render_depth = depth_sil[0, :, :]
    depth_error = torch.abs(gt - render_depth) * (gt > 0)
    non_presence_depth_mask = (render_depth > gt) * (depth_error > 50 * depth_error.median())
    pre = (non_presence_sil_mask | non_presence_depth_mask).reshape(-1)
    return pre, pre & (gt > 0).reshape(-1)


# scripts/splatam.py:384-426 with the render result passed in (the Renderer call itself is the rasterizer, 8a)
def add_new_gaussians(params, state, variables, depth_sil, im, gt_depth, intrinsics, sil_thres, time_idx):
    pre, mask = non_presence_mask(depth_sil, gt_depth, sil_thres)
    if int(pre.sum()) == 0:
        return params, state, variables
    cam_rot = F.normalize(params["cam_unnorm_rots"][..., time_idx].detach())
    cam_tran = params["cam_trans"][..., time_idx].detach()
    dev = gt_depth.device
    w2c = torch.eye(4, device=dev)
    w2c[:3, :3] = build_rotation(cam_rot)
    w2c[:3, 3] = cam_tran
    new_pt_cld, mean3_sq_dist = get_pointcloud(im, gt_depth, intrinsics, w2c, mask)
    n = new_pt_cld.shape[0]
    iso = params["log_scales"].shape[1] == 1
    new = {  # initialize_new_params, scripts/splatam.py:354-381
        "means3D": new_pt_cld[:, :3],
        "rgb_colors": new_pt_cld[:, 3:6],
        "unnorm_rotations": torch.tensor([1.0, 0, 0, 0], device=dev).repeat(n, 1),
        "logit_opacities": torch.zeros(n, 1, device=dev),
        "log_scales": torch.tile(torch.log(torch.sqrt(mean3_sq_dist))[..., None], (1, 1 if iso else 3)),
    }
    params = dict(params)
    state = {k: dict(v) for k, v in state.items()}
    variables = dict(variables)
    for k, v in new.items():
        params[k] = torch.cat((params[k], v), dim=0)
        if k in state:  # the mapping optimizer is rebuilt per frame in the reference (scripts/splatam.py:826); when a
            # live optimizer is carried instead, cat_params_to_optimizer (utils/slam_external.py:123-140) applies:
            state[k]["exp_avg"] = torch.cat((state[k]["exp_avg"], torch.zeros_like(v)), dim=0)
            state[k]["exp_avg_sq"] = torch.cat((state[k]["exp_avg_sq"], torch.zeros_like(v)), dim=0)
    num_pts = params["means3D"].shape[0]
    for k in ("means2D_gradient_accum", "denom", "max_2D_radius"):
        variables[k] = torch.zeros(num_pts, device=dev)
    variables["timestep"] = torch.cat((variables["timestep"], time_idx * torch.ones(n, device=dev)), dim=0)
    return params, state, variables


# ---------------------------------------------------------------------------------------------------------------
def random_state(P, iso, seed, n_frames=5):
    """Seeded map + Adam moments + per-Gaussian variables."""
    from oracle.slam_oracle import random_params
    params = random_params(P, n_frames, iso, seed)
    g = torch.Generator().manual_seed(1000 + seed)
    params["logit_opacities"] = torch.randn(P, 1, generator=g) * 3.0          # a good share below the thresholds
    params["log_scales"] = torch.randn(P, 1 if iso else 3, generator=g) * 0.8 - 2.5
    state = {k: {"exp_avg": torch.randn(params[k].shape, generator=g) * 1e-3,
                 "exp_avg_sq": torch.rand(params[k].shape, generator=g) * 1e-5} for k in GAUSS_KEYS}
    variables = {"means2D_gradient_accum": torch.rand(P, generator=g), "denom": torch.rand(P, generator=g).round() * 3,
                 "max_2D_radius": torch.rand(P, generator=g) * 40, "timestep": torch.randint(0, 7, (P,), generator=g).float(),
                 "scene_radius": torch.tensor(2.0)}
    return params, state, variables


def frame(h, w, seed):
    """Seeded RGB-D frame + a silhouette render of a map that misses part of it."""
    g = torch.Generator().manual_seed(2000 + seed)
    yy, xx = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    im = torch.rand(3, h, w, generator=g)
    gt_depth = 2.5 + 0.8 * torch.sin(xx / 9.0) * torch.cos(yy / 13.0) + 0.01 * torch.randn(h, w, generator=g)
    gt_depth[torch.rand(h, w, generator=g) < 0.08] = 0.0
    render_depth = gt_depth + 0.02 * torch.randn(h, w, generator=g)
    far = torch.rand(h, w, generator=g) < 0.05                               # a new foreground object in front
    render_depth[far] += 3.0 + torch.rand(int(far.sum()), generator=g)
    sil = 1.0 - 0.02 * torch.rand(h, w, generator=g)
    hole = (xx > 0.7 * w) | (torch.rand(h, w, generator=g) < 0.05)           # unobserved strip + speckle
    sil[hole] *= 0.3
    depth_sil = torch.stack([render_depth, sil, render_depth * render_depth], 0)
    fx = 0.9 * w
    intrinsics = torch.tensor([[fx, 0.0, w / 2 - 0.5], [0.0, fx * 1.02, h / 2 - 0.5], [0.0, 0.0, 1.0]])
    return im, gt_depth[None], depth_sil, intrinsics


PRUNE_DICT = dict(start_after=0, remove_big_after=0, stop_after=20, prune_every=20, removal_opacity_threshold=0.005,
                  final_removal_opacity_threshold=0.005, reset_opacities=False, reset_opacities_every=500)
PRUNE_CASES = [
    dict(name="iso_mid", iso=True, P=1500, seed=21, iter=0, pd=dict(PRUNE_DICT, removal_opacity_threshold=0.3)),
    dict(name="aniso_final", iso=False, P=1100, seed=22, iter=20, pd=dict(PRUNE_DICT, final_removal_opacity_threshold=0.5)),
    dict(name="iso_nobig", iso=True, P=700, seed=23, iter=20, pd=dict(PRUNE_DICT, prune_every=10, stop_after=40, remove_big_after=30,
                                                                     removal_opacity_threshold=0.2)),
    dict(name="aniso_reset", iso=False, P=600, seed=24, iter=40, pd=dict(PRUNE_DICT, prune_every=20, stop_after=100,
                                                                        removal_opacity_threshold=0.1, reset_opacities=True,
                                                                        reset_opacities_every=40)),
    dict(name="iso_off_iter", iso=True, P=300, seed=25, iter=7, pd=PRUNE_DICT),
]
ADD_CASES = [
    dict(name="iso", iso=True, P=900, seed=31, h=45, w=70, sil_thres=0.5, time_idx=3),
    dict(name="aniso", iso=False, P=500, seed=32, h=37, w=53, sil_thres=0.5, time_idx=2),
]
