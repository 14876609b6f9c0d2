"""View-parallel training layer: one process per GPU, a full Gaussian replica per rank, views sharded by rank.

The reference trains on one GPU, one view per iteration (trainer.py:323-392) and its torch.distributed helpers
(tools/distributed.py:21-152) are never called, so this layer is new.  Per optimizer step every
rank renders a different view; the only exchange is

  * one all-reduce(SUM) over a single flat gradient bucket holding every per-Gaussian gradient
    (xyz 3, SH 48, opacity 1, scaling 3, rotation 4 = 59 floats per Gaussian at SH degree 3), and
  * the densification statistics of scene/gaussian_model.py:669-671 and trainer.py:345, accumulated locally per view
    and combined only when the model densifies: all-reduce(SUM) of [ sum |grad means2D_densify| * visible, sum visible ]
    and all-reduce(MAX) of max_radii2D, which reproduces exactly what one process would have accumulated over the
    same views.

Backends: "nccl" on GPUs (NVLink 5 / NVSwitch), "gloo" on CPU for the tests.  No collective is issued at world size 1.
"""
import os
from typing import Dict, Optional

import torch
import torch.distributed as dist

# per-Gaussian parameter groups, in bucket order (name, trailing shape)
PARAM_LAYOUT = (("means3D", (3,)), ("shs", (16, 3)), ("opacities", (1,)), ("scales", (3,)), ("rotations", (4,)))


def init_from_env(backend: Optional[str] = None):
    """torch.distributed init from RANK / WORLD_SIZE / MASTER_* (torchrun).  Returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


class GaussianBucket:
    """All per-Gaussian parameters in ONE flat tensor and all their gradients in ONE flat tensor.

    `params[name]` are views into `flat`; their `.grad` are views into `grad_flat`, so autograd accumulates straight
    into the bucket and the gradient exchange is a single collective with no packing copy."""

    def __init__(self, tensors: Dict[str, torch.Tensor], device=None, direct: bool = False):
        """direct=True: the rasterizer's backward writes each parameter gradient straight into `grad_flat` (through the
        `_vcr_grad_out` hook of diff_gaussian_rasterization) and autograd adopts views of it as `.grad` — no per-step
        memset of the bucket and no accumulate pass.  The price: exactly one backward per `zero_grad()`."""
        self.direct = direct
        P = tensors["means3D"].shape[0]
        self.P = P
        self.layout = []
        off = 0
        for name, _ in PARAM_LAYOUT:
            t = tensors[name]
            n = t.numel()
            self.layout.append((name, off, n, tuple(t.shape)))
            off += n
        self.numel = off
        device = device if device is not None else tensors["means3D"].device
        self.flat = torch.empty(off, dtype=torch.float32, device=device)
        self.grad_flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.params = {}
        for name, o, n, shape in self.layout:
            self.flat[o:o + n].copy_(tensors[name].reshape(-1))
            p = self.flat[o:o + n].view(shape).requires_grad_(True)
            if direct:
                p._vcr_grad_out = self.grad_flat[o:o + n].view(shape)
            else:
                p.grad = self.grad_flat[o:o + n].view(shape)
            self.params[name] = p

    def zero_grad(self):
        if self.direct:
            for p in self.params.values():
                p.grad = None     # the next backward overwrites every element of grad_flat
        else:
            self.grad_flat.zero_()

    @property
    def floats_per_gaussian(self):
        return self.numel // max(self.P, 1)


class ViewParallel:
    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def view_for_step(self, step: int, n_views: int, seed: int = 0) -> int:
        """Rank r takes views r, r+world, ... of a seeded permutation of the camera list, reshuffled every epoch
        (replaces viewpoint_stack.pop(randint(...)), trainer.py:326-328).  Identical on every rank."""
        per_epoch = max(n_views // self.world, 1)
        epoch, k = divmod(step, per_epoch)
        g = torch.Generator().manual_seed(seed * 1_000_003 + epoch)
        perm = torch.randperm(n_views, generator=g)
        return int(perm[(k * self.world + self.rank) % n_views])

    def allreduce_gradients(self, bucket: GaussianBucket, average: bool = True, async_op: bool = False):
        """SUM (or mean) of the flat gradient bucket over ranks.  Returns the work handle when async_op."""
        if self.world == 1:
            return None
        work = dist.all_reduce(bucket.grad_flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        if average:
            if async_op:
                work.wait()
                work = None
            bucket.grad_flat.div_(self.world)
        return work

    def accumulate_densification_stats(self, xyz_gradient_accum, denom, max_radii2D, grad_means2D_densify, radii):
        """Per view, LOCAL: add_densification_stats (scene/gaussian_model.py:669-671) + the max_radii2D update
        (trainer.py:345).  No communication: sums and maxima commute, and the statistics are only read when the model
        densifies (every 100 iterations, trainer.py:357-370), so the ranks are combined there, once, by
        sync_densification_stats().  Tensors are updated in place; shapes [P,1], [P,1], [P]."""
        visible = radii > 0
        xyz_gradient_accum.add_((torch.norm(grad_means2D_densify[:, :2], dim=-1) * visible).unsqueeze(1))
        denom.add_(visible.unsqueeze(1).to(denom.dtype))
        max_radii2D.copy_(torch.max(max_radii2D, torch.where(visible, radii, torch.zeros_like(radii)).to(max_radii2D.dtype)))

    def sync_densification_stats(self, xyz_gradient_accum, denom, max_radii2D):
        """Combine the per-rank statistics right before densify/prune: all-reduce(SUM) of the two accumulators and
        all-reduce(MAX) of max_radii2D.  Afterwards every rank holds exactly what ONE process would have accumulated over
        the same views, and the accumulators must then be reset on every rank (the reference zeroes them in
        densification_postfix, scene/gaussian_model.py).  Call it with freshly reset local accumulators between
        densification rounds, i.e. each rank's tensors must contain only its own views since the last sync."""
        if self.world == 1:
            return
        packed = torch.cat([xyz_gradient_accum.reshape(-1), denom.reshape(-1)])
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=self.group)
        n = xyz_gradient_accum.numel()
        xyz_gradient_accum.copy_(packed[:n].view_as(xyz_gradient_accum))
        denom.copy_(packed[n:].view_as(denom))
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=self.group)

    def reduce_densification_stats(self, xyz_gradient_accum, denom, max_radii2D, grad_means2D_densify, radii):
        """Per-step variant (one SUM and one MAX all-reduce every view): accumulate this view on every rank and combine
        immediately.  Same result as accumulate_... + sync_..., at the cost of two small collectives per step; kept for
        callers that read the statistics every iteration."""
        visible = radii > 0
        packed = torch.zeros((radii.shape[0], 2), dtype=torch.float32, device=radii.device)
        packed[:, 0] = torch.norm(grad_means2D_densify[:, :2], dim=-1) * visible
        packed[:, 1] = visible.float()
        rad = torch.where(visible, radii, torch.zeros_like(radii)).to(torch.int32)
        if self.world > 1:
            dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(rad, op=dist.ReduceOp.MAX, group=self.group)
        xyz_gradient_accum.add_(packed[:, 0:1])
        denom.add_(packed[:, 1:2])
        max_radii2D.copy_(torch.max(max_radii2D, rad.to(max_radii2D.dtype)))

    def broadcast_parameters(self, bucket: GaussianBucket, src: int = 0):
        """Make every replica bit-identical to rank `src` (start of training / after a restore)."""
        if self.world > 1:
            dist.broadcast(bucket.flat, src=src, group=self.group)

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)
