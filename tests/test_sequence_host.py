"""CPU: host-side pieces of bench_sequence.py (C5): trajectory, pose propagation (scripts/splatam.py:429-448), ATE."""
import math

import torch
import torch.nn.functional as F

import bench_sequence as BS
from oracle.slam_oracle import build_rotation


def test_gt_trajectory_and_quaternions_round_trip():
    n = 50
    assert torch.equal(BS.gt_pose(0, n), torch.eye(4))
    for t in (1, 13, 37):
        w2c = BS.gt_pose(t, n)
        R = w2c[:3, :3]
        assert torch.allclose(R @ R.t(), torch.eye(3), atol=1e-6)
        q = BS.rot_to_quat(R)
        assert torch.allclose(build_rotation(q[None])[0], R, atol=1e-6)
        step = (BS.gt_pose(t + 1, n)[:3, 3] - w2c[:3, 3]).norm()
        assert float(step) < 0.03                      # a slow sway: centimetres per frame at n = 50


def test_constant_velocity_pose_initialisation_and_ate():
    n = 6
    poses = torch.stack([BS.gt_pose(t, 500) for t in range(n)])
    params = {"cam_unnorm_rots": torch.stack([BS.rot_to_quat(p[:3, :3]) for p in poses], 1).unsqueeze(0).contiguous(),
              "cam_trans": torch.stack([p[:3, 3] for p in poses], 1).unsqueeze(0).contiguous()}
    assert BS.ate_rmse(params, poses, n) < 1e-6
    est = {k: v.clone() for k, v in params.items()}
    est["cam_unnorm_rots"][..., 1] = 0
    est["cam_trans"][..., 1] = 0
    BS.initialize_camera_pose(est, 1)                  # t == 1: copy of frame 0
    assert torch.equal(est["cam_trans"][..., 1], est["cam_trans"][..., 0])
    est["cam_trans"][..., 3] = 0
    est["cam_unnorm_rots"][..., 3] = 0
    BS.initialize_camera_pose(est, 3)                  # t > 1: constant velocity from frames 1, 2
    t1, t2 = est["cam_trans"][..., 2], est["cam_trans"][..., 1]
    assert torch.allclose(est["cam_trans"][..., 3], t1 + (t1 - t2))
    r1, r2 = F.normalize(est["cam_unnorm_rots"][..., 2]), F.normalize(est["cam_unnorm_rots"][..., 1])
    assert torch.allclose(est["cam_unnorm_rots"][..., 3], F.normalize(r1 + (r1 - r2)))
    assert math.isfinite(BS.ate_rmse(est, poses, n)) and BS.ate_rmse(est, poses, n) < 0.01


def test_reference_schedule_constants_match_tum_config():
    assert BS.TRACK_LRS["cam_unnorm_rots"] == 0.002 and BS.TRACK_LRS["cam_trans"] == 0.002
    assert BS.MAP_LRS["logit_opacities"] == 0.05 and BS.MAP_LRS["cam_trans"] == 0.0
    assert BS.PRUNE["prune_every"] == 20 and BS.PRUNE["stop_after"] == 20
    assert (BS.TUM["w"], BS.TUM["h"]) == (640, 480)


def test_estimated_w2c_matches_build_rotation():
    q = torch.tensor([[0.9, 0.1, -0.2, 0.05]])
    params = {"cam_unnorm_rots": q.view(1, 4, 1).repeat(1, 1, 3).contiguous(),
              "cam_trans": torch.tensor([0.1, -0.2, 0.3]).view(1, 3, 1).repeat(1, 1, 3).contiguous()}
    w2c = BS.est_w2c(params, 1, "cpu")
    assert torch.allclose(w2c[:3, :3], build_rotation(F.normalize(q))[0], atol=1e-6)
    assert torch.allclose(w2c[:3, 3], torch.tensor([0.1, -0.2, 0.3]))
    assert torch.equal(w2c[3], torch.tensor([0.0, 0.0, 0.0, 1.0]))
