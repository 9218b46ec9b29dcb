"""CPU tests (-m "not gpu") of the keyframe-sharded mapping plumbing (splatam_b200/mapping.py):
the schedule, the flat gradient bucket, and a world_size-2 `gloo` run of mapping_step whose
all-reduced bucket must equal the single-process sum over the same keyframes.
The "render" here is a deterministic differentiable stand-in on CPU tensors -- the rasterizer itself
needs a GPU and is covered by the -m gpu tests; this file covers the N>1 host logic only."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from splatam_b200 import mapping, synth


def _scene(P=257, seed=3):
    return synth.cloud_scene(P, 64, 64, 64.0, 64.0, seed=seed)


def _fake_render(cam, leaves):
    """Differentiable stand-in: a loss that depends on every parameter and on the keyframe pose."""
    w = cam  # a [4,4] pose
    pts = leaves["means3D"] @ w[:3, :3].t() + w[:3, 3]
    loss = (pts * leaves["colors_precomp"]).sum() + (leaves["opacities"].squeeze(1) * leaves["scales"].sum(1)).sum() \
        + (leaves["rotations"] ** 2).sum() * float(w[0, 3] + 1.0)
    loss.backward()


def test_schedule_is_deterministic_and_distinct_within_a_step():
    a = mapping.keyframe_schedule(24, 60, 8, seed=5)
    b = mapping.keyframe_schedule(24, 60, 8, seed=5)
    assert a.shape == (60, 8) and torch.equal(a, b)
    assert int(a.min()) >= 0 and int(a.max()) < 24
    for s in range(60):
        assert len(set(a[s].tolist())) == 8
    # every keyframe of the window gets rendered
    assert len(set(a.flatten().tolist())) == 24
    # more ranks than keyframes: still valid ids
    c = mapping.keyframe_schedule(3, 5, 8, seed=1)
    assert c.shape == (5, 8) and int(c.max()) < 3
    assert mapping.keyframe_schedule(24, 0, 2).shape == (0, 2)
    with pytest.raises(ValueError):
        mapping.keyframe_schedule(0, 1, 1)


def test_bucket_views_alias_the_flat_buffers():
    sc = _scene()
    bk = mapping.GradBucket(sc)
    assert bk.params.numel() == bk.P * mapping.FLOATS_PER_GAUSSIAN == bk.grads.numel()
    for k, w in mapping.LAYOUT:
        assert torch.equal(bk.views[k], sc[k])
        assert bk.views[k].data_ptr() % 16 == bk.params.data_ptr() % 16 or k != "rotations"
    assert bk.views["rotations"].data_ptr() == bk.params.data_ptr()   # float4 array first => aligned for any P
    lv = bk.leaves()
    _fake_render(synth.arc_pose(1, 4), lv)
    # autograd accumulated straight into the bucket (no pack step)
    assert float(bk.grads.abs().sum()) > 0
    assert lv["means3D"].grad.data_ptr() == bk.gviews["means3D"].data_ptr()
    bk.zero_grad()
    assert float(bk.grads.abs().sum()) == 0
    with pytest.raises(ValueError):
        mapping.GradBucket(dict(sc, scales=sc["scales"][:, :2]))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, steps, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        sc = _scene()
        cams = [synth.arc_pose(k, 6) for k in range(6)]
        sched = mapping.keyframe_schedule(6, steps, world, seed=9)
        bk = mapping.GradBucket(sc)
        outs = []
        for s in range(steps):
            kf = mapping.mapping_step(bk, _fake_render, cams, sched, s, rank)
            assert kf == int(sched[s, rank])
            outs.append(bk.grads.clone())
        torch.save(torch.stack(outs), os.path.join(out_dir, "rank%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def test_world_size_2_allreduce_equals_single_process_sum(tmp_path):
    world, steps = 2, 3
    mp.spawn(_worker, args=(world, _free_port(), steps, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(str(tmp_path / "rank0.pt"))
    r1 = torch.load(str(tmp_path / "rank1.pt"))
    assert torch.equal(r0, r1), "ranks disagree after the all-reduce"
    # single-process reference: sum of the two keyframes' gradients per step
    sc = _scene()
    cams = [synth.arc_pose(k, 6) for k in range(6)]
    sched = mapping.keyframe_schedule(6, steps, world, seed=9)
    for s in range(steps):
        tot = None
        for r in range(world):
            bk = mapping.GradBucket(sc)
            _fake_render(cams[int(sched[s, r])], bk.leaves())
            tot = bk.grads.clone() if tot is None else tot + bk.grads
        assert torch.allclose(r0[s], tot, rtol=1e-6, atol=1e-6)


def test_single_process_all_reduce_is_a_noop():
    bk = mapping.GradBucket(_scene())
    _fake_render(synth.arc_pose(0, 4), bk.leaves())
    before = bk.grads.clone()
    bk.all_reduce()
    assert torch.equal(before, bk.grads)


def test_direct_bucket_marks_leaves_and_skips_the_fill():
    """direct=True: leaves carry the rasterizer's gradient-sink attribute aliasing the bucket views, and zero_grad
    leaves the buffer alone (the next render overwrites every row)."""
    from splatam_b200 import rasterizer  # noqa: F401  (SINK_ATTR lives with its consumer)
    bk = mapping.GradBucket(_scene(), direct=True)
    lv = bk.leaves()
    for k, _ in mapping.LAYOUT:
        sk = getattr(lv[k], rasterizer.SINK_ATTR)
        assert sk.data_ptr() == bk.gviews[k].data_ptr() and lv[k].grad.data_ptr() == sk.data_ptr()
    bk.grads.fill_(7.0)
    bk.zero_grad()
    assert float(bk.grads.min()) == 7.0
    plain = mapping.GradBucket(_scene())
    assert not hasattr(plain.leaves()["means3D"], rasterizer.SINK_ATTR)
    plain.grads.fill_(7.0)
    plain.zero_grad()
    assert float(plain.grads.abs().max()) == 0.0


def test_raw_parameter_layout_bucket():
    """layout=raw_layout(): SplaTAM's raw parameter names, float4 array first, log_scales [P,1] or [P,3]."""
    from oracle import slam_oracle as SO
    for iso in (True, False):
        p = SO.random_params(101, 3, iso, 1)
        sd = 1 if iso else 3
        bk = mapping.GradBucket({k: p[k] for k, _ in mapping.raw_layout(sd)}, direct=True, layout=mapping.raw_layout(sd))
        assert bk.params.numel() == 101 * (4 + 3 + 3 + 1 + sd)
        lv = bk.leaves()
        assert set(lv) == {"unnorm_rotations", "means3D", "rgb_colors", "logit_opacities", "log_scales"}
        assert lv["unnorm_rotations"].data_ptr() == bk.params.data_ptr() and lv["log_scales"].shape == (101, sd)
        assert torch.equal(bk.views["means3D"], p["means3D"])
        (lv["means3D"].sum() + lv["log_scales"].sum() * 2).backward()   # autograd accumulates into the bucket views
        assert float(bk.gviews["log_scales"].min()) == 2.0 and lv["means3D"].grad.data_ptr() == bk.gviews["means3D"].data_ptr()


def test_cost_balanced_schedule_groups_similar_keyframes():
    """keyframe_schedule(cost=...): every pass still visits each keyframe once, ranks of a step get distinct keyframes,
    and the spread of cost inside a step is the smallest possible (consecutive groups of the sorted costs)."""
    from splatam_b200 import mapping
    n, world = 24, 8
    cost = torch.linspace(1.0, 2.0, n)[torch.randperm(n, generator=torch.Generator().manual_seed(3))]
    s = mapping.keyframe_schedule(n, 12, world, seed=5, cost=cost)
    assert s.shape == (12, world)
    for p in range(4):                                   # 3 steps per pass
        block = s[3 * p:3 * p + 3]
        assert sorted(block.flatten().tolist()) == list(range(n))
    spread = (cost[s].max(1)[0] - cost[s].min(1)[0])
    assert float(spread.max()) <= float((cost.sort()[0][world - 1::world] - cost.sort()[0][0::world]).max()) + 1e-12
    plain = mapping.keyframe_schedule(n, 12, world, seed=5)
    assert float((cost[plain].max(1)[0] - cost[plain].min(1)[0]).mean()) > 2.0 * float(spread.mean())
    # same seed + same cost -> same schedule on every rank; world = 1 ignores the cost
    assert torch.equal(s, mapping.keyframe_schedule(n, 12, world, seed=5, cost=cost.clone()))
    assert torch.equal(mapping.keyframe_schedule(n, 12, 1, seed=5, cost=cost), mapping.keyframe_schedule(n, 12, 1, seed=5))
    with pytest.raises(ValueError):
        mapping.keyframe_schedule(n, 4, world, cost=cost[:5])
