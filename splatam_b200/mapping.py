"""Keyframe-sharded mapping: host-side plumbing for the multi-GPU leg.

The reference is single-GPU: each mapping iteration renders ONE randomly chosen keyframe of the
window against all Gaussians and takes an optimiser step (scripts/splatam.py:848-889).  Renders of
different keyframes are independent given the parameters and their Gaussian gradients add, so the
path shards over keyframes: one process per GPU (torchrun), the Gaussian set replicated, rank r
renders keyframe `schedule[step, r]`, then ONE all-reduce (NCCL over NVLink on the GPU box, gloo in
the CPU tests) of a single flat gradient bucket.  Nothing else is exchanged.

PyTorch is plumbing here (device memory, torch.distributed); the rendering itself is the C-ABI
library.  This module is device-agnostic on purpose so the world_size-2 gloo tests exercise exactly
the code bench.py runs.
"""
import torch

# (name, floats per Gaussian).  float4-accessed arrays first so every view stays 16-byte aligned
# for any P (the kernels read quaternions as float4).
LAYOUT = (("rotations", 4), ("means3D", 3), ("colors_precomp", 3), ("opacities", 1), ("scales", 3))
FLOATS_PER_GAUSSIAN = sum(w for _, w in LAYOUT)


def raw_layout(scale_dim=1):
    """Bucket layout over SplaTAM's RAW Gaussian parameters (scripts/splatam.py:118-140) for the fused mapping iteration
    (slam.get_loss): the builder's backward and the rasterizer write their gradients straight into these views."""
    return (("unnorm_rotations", 4), ("means3D", 3), ("rgb_colors", 3), ("logit_opacities", 1), ("log_scales", int(scale_dim)))


def keyframe_schedule(n_keyframes, steps, world, seed=0, cost=None):
    """int64 [steps, world]: keyframe rendered by each rank at each step.  Identical on every rank
    (seeded, CPU generator) -- it replaces the reference's per-iteration `np.random.randint`
    (scripts/splatam.py:851).  Within a step the ranks get DISTINCT keyframes whenever
    world <= n_keyframes (consecutive slices of concatenated random permutations).

    cost (optional, one positive number per keyframe, the SAME values on every rank -- e.g. the measured time or the
    num_rendered of one render of each keyframe): a step ends when its slowest rank does, so with costs the keyframes of
    a pass are grouped by similar cost (sorted, cut into groups of `world`, groups visited in random order, ranks
    assigned in random order inside a group).  Every pass still renders each (usable) keyframe exactly once; only
    which keyframes share a step changes.  Measured at N = 8 / C3 (r01): keyframes differ by ~6% in cost and the step
    waited for the slowest of eight random ones."""
    if n_keyframes <= 0 or steps < 0 or world <= 0:
        raise ValueError("keyframe_schedule: n_keyframes, world must be positive and steps non-negative")
    if cost is not None:
        cost = torch.as_tensor(cost, dtype=torch.float64).flatten()
        if cost.numel() != n_keyframes:
            raise ValueError("keyframe_schedule: cost must have one entry per keyframe")
    g = torch.Generator().manual_seed(int(seed))
    need = steps * world
    if world <= n_keyframes:
        # pad each permutation to a multiple of `world` from the next one so a step never straddles
        # a permutation boundary with a repeated keyframe
        out = []
        while sum(len(o) for o in out) < need:
            perm = torch.randperm(n_keyframes, generator=g)
            usable = (n_keyframes // world) * world
            perm = perm[:usable]
            if cost is not None and world > 1:
                by_cost = perm[torch.argsort(cost[perm], descending=True, stable=True)].reshape(-1, world)
                by_cost = by_cost[torch.randperm(by_cost.shape[0], generator=g)]            # group order
                perm = torch.stack([row[torch.randperm(world, generator=g)] for row in by_cost]).reshape(-1)
            out.append(perm)
        flat = torch.cat(out)[:need] if need else torch.zeros(0, dtype=torch.int64)
    else:
        flat = torch.randint(0, n_keyframes, (need,), generator=g)
    return flat.reshape(steps, world)


class GradBucket:
    """The replicated Gaussian parameters in ONE flat buffer and their gradients in another, with
    named [P, w] views (LAYOUT), so the exchange is a single all-reduce with no pack kernel.

    direct=True marks every leaf with the rasterizer's gradient-sink attribute: the backward kernels
    then WRITE each gradient array straight into its bucket view (splatam_b200.rasterizer
    `_collect_sinks`) -- no autograd accumulate kernels and no zero fill, because one render assigns
    every row.  Use it when a step is exactly one render per bucket (the mapping step); leave it off
    when several losses accumulate into the same leaves."""

    def __init__(self, scene, device="cpu", direct=False, peer_group=None, layout=LAYOUT, shard_optimizer=False,
                 multicast="auto"):
        """peer_group: a torch.distributed process group whose ranks share one NVSwitch node.  The gradient buffer
        is then allocated in symmetric memory and `all_reduce` runs the library's own peer-memory two-shot kernel
        (csrc/allreduce.cu) on the current stream instead of NCCL.
        shard_optimizer (needs peer_group): the PARAMETER buffer lives in symmetric memory too, so that ShardedAdam can
        run exchange + optimizer step as one fused kernel (reduce-scatter -> Adam on this rank's slice -> all-gather of
        the updated parameters)."""
        self.direct = bool(direct)
        self.peer = None
        self.layout = tuple(layout)
        LAYOUT = self.layout   # noqa: N806  (the rest of the class reads the instance's layout)
        self.P = int(scene["means3D"].shape[0])
        for k, w in LAYOUT:
            if tuple(scene[k].shape) != (self.P, w):
                raise ValueError("GradBucket: %s must be [%d, %d], got %s" % (k, self.P, w, tuple(scene[k].shape)))
        self.params = torch.cat([scene[k].reshape(-1).float() for k, _ in LAYOUT]).to(device)
        if peer_group is not None:
            self.peer = PeerAllReduce(self.params.numel(), self.params.device, peer_group, multicast=multicast,
                                      with_params=bool(shard_optimizer))
            self.grads = self.peer.buffer[:self.params.numel()]
            self.grads.zero_()
            if shard_optimizer:
                flat = self.params
                self.params = self.peer.params[:flat.numel()]
                self.params.copy_(flat)
                torch.cuda.synchronize(self.params.device)
                import torch.distributed as dist
                dist.barrier(peer_group)     # every replica holds the initial parameters before anyone's first step
        else:
            self.grads = torch.zeros_like(self.params)
        self.views, self.gviews = {}, {}
        off = 0
        for k, w in self.layout:
            n = self.P * w
            self.views[k] = self.params[off:off + n].view(self.P, w)
            self.gviews[k] = self.grads[off:off + n].view(self.P, w)
            off += n

    def leaves(self):
        """Autograd leaves aliasing the parameter views, with .grad aliasing the bucket."""
        lv = {k: v.detach().requires_grad_(True) for k, v in self.views.items()}
        for k, leaf in lv.items():
            leaf.grad = self.gviews[k]
            if self.direct:
                setattr(leaf, "_splat_grad_sink", self.gviews[k])
        return lv

    def zero_grad(self):
        if not self.direct:   # a direct bucket is fully overwritten by the next render
            self.grads.zero_()

    def all_reduce(self, group=None, average=False):
        """Sum (or mean) of the gradient bucket over the ranks; no-op without a process group."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return self.grads
        if self.peer is not None:
            self.peer()
        else:
            dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=group)
        if average:
            self.grads.div_(dist.get_world_size(group))
        return self.grads

    def nbytes(self):
        return self.params.numel() * 4


class PeerAllReduce:
    """In-place sum of one flat float32 buffer over the ranks of a single NVSwitch node through peer memory:
    torch symmetric memory provides the allocation and the exchange of device pointers (plumbing); the reduction
    itself is `splat_peer_allreduce_sum` (two-shot over NVLink loads/stores, csrc/allreduce.cu), launched on the
    current stream -- no NCCL stream hop, and every replica receives bit-identical sums."""

    def __init__(self, numel, device, group, multicast="auto", with_params=False):
        import ctypes as C
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.numel = (int(numel) + 3) // 4 * 4
        self.params = None
        if with_params:      # a second symmetric bucket of the same length: the replicated parameters (ShardedAdam)
            self.params = symm_mem.empty(self.numel, dtype=torch.float32, device=device)
            self.params.zero_()
            torch.cuda.synchronize(device)
            hp = symm_mem.rendezvous(self.params, group)
            self._pbufs = (C.c_uint64 * int(hp.world_size))(*[int(p) for p in hp.buffer_ptrs])
            self._param_multicast_ptr = int(getattr(hp, "multicast_ptr", 0) or 0)
            self._param_handle = hp
        self.buffer = symm_mem.empty(self.numel, dtype=torch.float32, device=device)
        self.flags = symm_mem.empty(256, dtype=torch.int32, device=device)
        self.buffer.zero_()
        self.flags.zero_()
        torch.cuda.synchronize(device)
        hb = symm_mem.rendezvous(self.buffer, group)
        hf = symm_mem.rendezvous(self.flags, group)
        self.rank, self.world = int(hb.rank), int(hb.world_size)
        self._bufs = (C.c_uint64 * self.world)(*[int(p) for p in hb.buffer_ptrs])
        self._pads = (C.c_uint64 * self.world)(*[int(p) for p in hf.buffer_ptrs])
        self._handles = (hb, hf)
        # NVSwitch multicast mapping of the bucket, if the fabric has one: sums are then formed inside the switch
        # (multimem.ld_reduce / multimem.st).  Measured on 56 MB (B200, NVLink 5): peer loads/stores 0.100 ms at N=2,
        # 0.165 at N=4, 0.264 at N=8; in-switch 0.171 at N=2, 0.178 at N=8; NCCL 0.129 / 0.160 / 0.241.
        mc = int(getattr(hb, "multicast_ptr", 0) or 0)
        if multicast == "auto":
            # 56 MB bucket, tuned grids (profiles/r02i_peer_tune_*.log): N=2 peer loads / stores 0.100 ms vs in-switch 0.157;
            # N=4 0.149-0.165 vs 0.143; N=8 0.26 vs 0.142
            multicast = self.world >= 4
        self.multicast = mc if multicast else 0
        self.param_multicast = (self._param_multicast_ptr if (with_params and self.multicast) else 0)
        if with_params and self.multicast and not self.param_multicast:
            self.multicast = 0          # both buckets through the switch, or neither
        self.epoch = 0
        dist.barrier(group)   # every rank's flag pad is zeroed before anyone signals into it
        torch.cuda.synchronize(device)

    def __call__(self):
        import ctypes as C
        from ._lib import check, lib
        self.epoch = (self.epoch + 1) & 0xffffffff   # full 32-bit epochs: the kernel's (int32)(flag - epoch) < 0 is wrap-safe
        check(lib.splat_peer_allreduce_sum(self._bufs, self._pads, self.multicast, self.rank, self.world, self.epoch,
                                           self.numel,
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return self.buffer


def mapping_step(bucket, render_fn, cams, schedule, step, rank, group=None, average=False):
    """One keyframe-sharded mapping step: zero the bucket, render + backward this rank's keyframe
    (render_fn(cam, leaves) must leave d(loss)/d(param) in the leaves' .grad, i.e. in the bucket),
    all-reduce.  Returns the keyframe id rendered."""
    kf = int(schedule[step, rank])
    bucket.zero_grad()
    render_fn(cams[kf], bucket.leaves())
    bucket.all_reduce(group, average)
    return kf


class ShardedAdam:
    """Exchange + optimizer step of the keyframe-sharded mapping iteration as ONE kernel per rank
    (`splat_peer_reduce_adam`, csrc/allreduce.cu): rank r sums slice r of the gradient bucket over the ranks (inside the
    NVSwitch when the fabric has a multicast mapping, else over peer loads in rank order), applies Adam -- the arithmetic
    of FusedAdam / torch.optim.Adam as SplaTAM builds it (initialize_optimizer, scripts/splatam.py:166-172: one learning
    rate per tensor, eps 1e-15 while mapping) -- to slice r of the parameter bucket with ITS shard of the moments, and
    stores the updated parameters into every replica.  Replaces `bucket.all_reduce(); optimizer.step()`: the links carry
    the same bytes as one all-reduce, the optimizer pass and the moment memory shrink to 1/N per GPU, and every replica
    receives bit-identical parameters.  `bucket` must have been built with peer_group and shard_optimizer=True."""

    def __init__(self, bucket, lrs, betas=(0.9, 0.999), eps=1e-15):
        import ctypes as C
        from ._lib import lib
        peer = bucket.peer
        if peer is None or peer.params is None:
            raise ValueError("ShardedAdam needs a GradBucket(peer_group=..., shard_optimizer=True)")
        self.bucket, self.peer, self.betas, self.eps = bucket, peer, (float(betas[0]), float(betas[1])), float(eps)
        self.steps = 0
        n = int(lib.splat_peer_adam_shard_floats(peer.numel, peer.world))
        dev = bucket.params.device
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        ends, off = [], 0
        for k, w in bucket.layout:
            off += bucket.P * w
            ends.append(off)
        ends[-1] = peer.numel                       # the padding up to a multiple of 4 floats rides with the last tensor
        self._ends = (C.c_int64 * len(ends))(*ends)
        self._lrs = (C.c_float * len(ends))(*[float(lrs[k]) for k, _ in bucket.layout])
        self._n = len(ends)

    def step(self):
        """Sums the gradient bucket over the ranks and takes one Adam step on every replica (current stream)."""
        import ctypes as C
        from ._lib import check, lib
        peer = self.peer
        self.steps += 1
        peer.epoch = (peer.epoch + 1) & 0xffffffff
        check(lib.splat_peer_reduce_adam(peer._bufs, peer._pbufs, peer._pads, peer.multicast, peer.param_multicast,
                                         peer.rank, peer.world, peer.epoch, peer.numel, self._ends, self._lrs, self._n,
                                         self.steps, self.betas[0], self.betas[1], self.eps,
                                         C.c_void_p(self.exp_avg.data_ptr()), C.c_void_p(self.exp_avg_sq.data_ptr()),
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream)))
