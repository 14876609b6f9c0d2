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
import ctypes
import os
from typing import Dict, Optional

import torch
import torch.distributed as dist

# per-Gaussian parameter groups, in bucket order (name, trailing shape)
PARAM_LAYOUT = (("means3D", (3,)), ("shs", (16, 3)), ("opacities", (1,)), ("scales", (3,)), ("rotations", (4,)))
# Every group starts on a 16-byte boundary inside the flat tensors: the kernels read `shs` / `rotations` and write
# `dL_dsh` / `dL_drot` with float4 accesses (include/vcr_gaus_b200.h, "Alignment").  Without the padding the group
# offsets 3P, 51P, ... would be 16-byte aligned only when P is a multiple of 4 — which it never is after a densification.
GROUP_ALIGN_FLOATS = 4


def _pad4(n: int) -> int:
    return (int(n) + GROUP_ALIGN_FLOATS - 1) // GROUP_ALIGN_FLOATS * GROUP_ALIGN_FLOATS


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


def bind_to_gpu_numa_node(device_index: int):
    """Restrict this process to the CPU cores next to its GPU (NVML's ideal CPU affinity for the device, intersected with the
    cores the process may use) so that the pinned host buffers of its data loader are allocated on that NUMA node and the
    per-view host->device copies do not cross the socket interconnect.  One process per GPU makes this a per-process
    setting.  Returns the number of cores kept, or 0 if nothing was changed (no NVML, single-node host, restricted
    cpuset)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        near = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        keep = near & allowed
        if not keep or keep == allowed:
            return 0
        os.sched_setaffinity(0, keep)
        return len(keep)
    except Exception:
        return 0


class GaussianBucket:
    """All per-Gaussian parameters in ONE flat tensor and all their gradients in ONE flat tensor.

    `params[name]` are views into `flat`; their `.grad` are views into `grad_flat`, so autograd accumulates straight
    into the bucket and the gradient exchange is a single collective with no packing copy."""

    def __init__(self, tensors: Dict[str, torch.Tensor], device=None, direct: bool = False, grad_storage=None, names=None):
        """direct=True: the rasterizer's backward writes each parameter gradient straight into `grad_flat` (through the
        `_vcr_grad_out` hook of diff_gaussian_rasterization) and autograd adopts views of it as `.grad` — no per-step
        memset of the bucket and no accumulate pass.  The price: exactly one backward per `zero_grad()`.
        grad_storage: a flat float32 tensor to use as `grad_flat` (ViewParallel.symmetric_gradient_storage() for the
        NVSwitch-multicast exchange); default is an ordinary allocation.
        names: the keys of `tensors` to pack, in bucket order; default = the rasterizer's inputs (PARAM_LAYOUT).  A trainer
        that keeps the raw parameters in the bucket passes e.g. ("xyz", "f_dc", "f_rest", "opacity", "scaling",
        "rotation") — the reference's optimizer groups (scene/gaussian_model.py:241-248) — and steps them with BucketAdam."""
        self.direct = direct
        names = tuple(names) if names is not None else tuple(n for n, _ in PARAM_LAYOUT)
        self.names = names
        P = tensors[names[0]].shape[0]
        self.P = P
        self.layout = []
        off = 0
        for name in names:
            t = tensors[name]
            n = t.numel()
            self.layout.append((name, off, n, tuple(t.shape)))
            off += _pad4(n)          # the next group starts 16-byte aligned; the (at most 3) pad floats stay zero
        device = device if device is not None else tensors[names[0]].device
        self._allocate(off, device, grad_storage)
        for name, o, n, shape in self.layout:
            self.flat[o:o + n].copy_(tensors[name].reshape(-1))
        self._make_views()

    def _allocate(self, numel, device, grad_storage):
        self.numel = numel
        self.flat = torch.zeros(numel, dtype=torch.float32, device=device)
        if grad_storage is None:
            self.grad_flat = torch.zeros(numel, dtype=torch.float32, device=device)
        else:
            if grad_storage.dtype != torch.float32 or grad_storage.numel() < numel or not grad_storage.is_contiguous():
                raise ValueError("grad_storage must be a contiguous float32 tensor of at least numel_for(tensors) elements")
            if grad_storage.data_ptr() % 16:
                raise ValueError("grad_storage must be 16-byte aligned")
            self.grad_flat = grad_storage.view(-1)[:numel].zero_()

    def _make_views(self):
        self.params = {}
        for name, o, n, shape in self.layout:
            p = self.flat[o:o + n].view(shape).requires_grad_(True)
            if self.direct:
                p._vcr_grad_out = self.grad_flat[o:o + n].view(shape)
            else:
                p.grad = self.grad_flat[o:o + n].view(shape)
            self.params[name] = p

    @staticmethod
    def numel_for(tensors: Dict[str, torch.Tensor], names=None) -> int:
        """Floats of flat storage a bucket of `tensors` needs (every group padded to a multiple of 4 floats)."""
        return sum(_pad4(tensors[name].numel()) for name in (names if names is not None else [n for n, _ in PARAM_LAYOUT]))

    def zero_grad(self):
        if self.direct:
            for p in self.params.values():
                p.grad = None     # the next backward overwrites every element of grad_flat
        else:
            self.grad_flat.zero_()

    @property
    def floats_per_gaussian(self):
        return sum(n for _, _, n, _ in self.layout) // max(self.P, 1)


class BucketAdam:
    """torch.optim.Adam(groups, lr=0.0, eps=1e-15) of the reference (scene/gaussian_model.py:241-258) over a GaussianBucket:
    one learning rate per parameter group = per contiguous range of the flat tensors, the whole step in ONE kernel
    (vcr_adam_step, csrc/adam.cu).  `lrs`: {group name: lr}; change one with set_lr (the xyz schedule,
    gaussian_model.py:264-270).  step(grad_scale) multiplies the gradients by grad_scale on the fly (1/world after a SUM
    all-reduce).  CUDA only."""

    def __init__(self, bucket: GaussianBucket, lrs: Dict[str, float], betas=(0.9, 0.999), eps: float = 1e-15):
        if not bucket.flat.is_cuda:
            raise RuntimeError("BucketAdam needs a CUDA bucket; there is no CPU fallback")
        missing = [n for n in bucket.names if n not in lrs]
        if missing:
            raise KeyError(f"no learning rate for {missing}")
        self.bucket, self.lrs, self.betas, self.eps = bucket, dict(lrs), betas, eps
        self.exp_avg = torch.zeros_like(bucket.flat)
        self.exp_avg_sq = torch.zeros_like(bucket.flat)
        self.steps = 0

    def set_lr(self, name: str, lr: float):
        if name not in self.lrs:
            raise KeyError(name)
        self.lrs[name] = float(lr)

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0):
        import vcr_native as _n
        b = self.bucket
        self.steps += 1
        n = len(b.layout)
        ends = (ctypes.c_longlong * n)(*[off + cnt for _, off, cnt, _ in b.layout])
        lrs = (ctypes.c_float * n)(*[self.lrs[name] for name, _, _, _ in b.layout])
        _n.check(_n.vcr_adam_step(b.numel, b.flat.data_ptr(), b.grad_flat.data_ptr(), self.exp_avg.data_ptr(),
                                  self.exp_avg_sq.data_ptr(), n, ends, lrs, self.betas[0], self.betas[1], self.eps, self.steps,
                                  float(grad_scale), ctypes.c_void_p(torch.cuda.current_stream(b.flat.device).cuda_stream)),
                 "adam_step")


def prune_and_grow(bucket: GaussianBucket, keep: torch.Tensor, n_new: int = 0, optimizer: Optional[BucketAdam] = None,
                   grad_storage: Optional[torch.Tensor] = None, n_kept: Optional[int] = None):
    """Densify / prune on the flat bucket: a new GaussianBucket whose first rows are the rows of `bucket` with keep != 0, in
    order, followed by `n_new` rows for the caller to fill (clones / splits; zero until then), and — when `optimizer` is
    given — a new BucketAdam whose moments are compacted the same way (zeros for the new rows, as
    cat_tensors_to_optimizer does).  Replaces prune_points + densification_postfix of the reference
    (scene/gaussian_model.py:442-531: ~18 boolean-mask indexing operations and as many torch.cat) by one prefix sum and
    one gather launch per 8 arrays (vcr_compact_rows).
    `keep`: bool / uint8 CUDA tensor [P].  One host read of the kept count (the new bucket has to be sized), unless the
    caller already knows it and passes `n_kept`.
    `grad_storage`: flat float32 storage for the new gradient bucket (the symmetric allocation of
    ViewParallel.symmetric_gradient_storage(), which must have been sized for the grown model); default: an ordinary
    allocation.  The new bucket has NONE of the exchange state of the old one: after this call the view-parallel set-up
    (ViewParallel.compact_sh_exchange / overlap_gradients / pick_transport) has to be repeated on the new bucket —
    ViewParallel.exchange_compact() refuses a bucket it was not set up for.
    Returns (new_bucket, new_optimizer or None, n_kept)."""
    import vcr_native as _n
    if not bucket.flat.is_cuda or not keep.is_cuda:
        raise RuntimeError("prune_and_grow needs CUDA tensors; there is no CPU fallback")
    P, dev = bucket.P, bucket.flat.device
    k8 = (keep.to(torch.uint8) if keep.dtype == torch.bool else keep).contiguous()
    if k8.numel() != P or k8.dtype != torch.uint8:
        raise ValueError("keep must have one bool / uint8 entry per Gaussian")
    n_kept = int(k8.sum()) if n_kept is None else int(n_kept)
    P2 = n_kept + int(n_new)
    # group-major layout of the new bucket: every group is compacted into its own 16-byte aligned slice
    new = GaussianBucket.__new__(GaussianBucket)
    new.direct, new.names, new.P, new.layout = bucket.direct, bucket.names, P2, []
    off = 0
    for name, _, n, shape in bucket.layout:
        per = n // max(P, 1)
        new.layout.append((name, off, per * P2, (P2,) + tuple(shape[1:])))
        off += _pad4(per * P2)
    new._allocate(off, dev, grad_storage)
    srcs, dsts, widths = [], [], []
    moments = []
    if optimizer is not None:
        opt2 = BucketAdam.__new__(BucketAdam)
        opt2.bucket, opt2.lrs, opt2.betas, opt2.eps, opt2.steps = new, dict(optimizer.lrs), optimizer.betas, optimizer.eps, optimizer.steps
        opt2.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)          # zeros for the new rows
        opt2.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)
        moments = [(optimizer.exp_avg, opt2.exp_avg), (optimizer.exp_avg_sq, opt2.exp_avg_sq)]
    else:
        opt2 = None
    for (name, o1, n1, _), (_, o2, _, _) in zip(bucket.layout, new.layout):
        per = n1 // max(P, 1)
        for src, dst in [(bucket.flat, new.flat)] + moments:
            srcs.append(src.data_ptr() + 4 * o1); dsts.append(dst.data_ptr() + 4 * o2); widths.append(per)
    scratch = torch.empty(int(_n.vcr_compact_scratch_bytes(P)), dtype=torch.uint8, device=dev)
    n = len(srcs) if n_kept > 0 else 0             # nothing to move when every row is pruned
    if n:
        _n.check(_n.vcr_compact_rows(P, k8.data_ptr(), n, (ctypes.c_void_p * n)(*srcs), (ctypes.c_void_p * n)(*dsts),
                                     (ctypes.c_int * n)(*widths), scratch.data_ptr(), None,
                                     ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)), "compact_rows")
    new._make_views()
    return new, opt2, n_kept


class ViewParallel:
    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.transport = "nccl"        # "multimem" after pick_transport() found the multicast kernel faster
        self._pending = []

    def view_for_step(self, step: int, n_views: int, seed: int = 0) -> int:
        """Rank r takes views r, r+world, ... of a seeded permutation of the camera list, reshuffled every epoch
        (replaces viewpoint_stack.pop(randint(...)), trainer.py:326-328).  Identical on every rank."""
        per_epoch = max(n_views // self.world, 1)
        epoch, k = divmod(step, per_epoch)
        g = torch.Generator().manual_seed(seed * 1_000_003 + epoch)
        perm = torch.randperm(n_views, generator=g)
        return int(perm[(k * self.world + self.rank) % n_views])

    def balanced_view_table(self, costs, seed: int = 0):
        """Cost-balanced replacement for view_for_step: a table of steps, each a list of `world` view ids (entry r is rank
        r's view), such that the views of one step cost about the same and every rank's column has about the same total.
        Every step ends with the gradient exchange, i.e. with a wait for the slowest rank; dealing views of equal cost to
        the ranks of a step removes that wait.  `costs[v]`: a number per view, identical on every rank (e.g. the
        rasterizer's instance count R from the previous epoch, all-gathered).  Views are sorted by cost, cut into groups
        of `world`, each group is rotated by its index (so no rank always gets a group's cheapest view) and the groups
        are visited in a seeded random order shared by all ranks.  len(costs) % world views are left out of the epoch.
        No communication."""
        n, w = len(costs), self.world
        by_cost = sorted(range(n), key=lambda v: (costs[v], v))
        groups = [by_cost[k * w:(k + 1) * w] for k in range(n // w)]
        groups = [g[k % w:] + g[:k % w] for k, g in enumerate(groups)]
        sigma = torch.randperm(len(groups), generator=torch.Generator().manual_seed(seed * 7_919 + 17)).tolist()
        return [groups[j] for j in sigma]

    def allreduce_gradients(self, bucket: GaussianBucket, average: bool = True, async_op: bool = False):
        """SUM (or mean) of the flat gradient bucket over ranks.  Returns the work handle when async_op."""
        if self.world == 1:
            return None
        if getattr(self, "transport", "nccl") == "multimem" and self._in_symmetric_storage(bucket):
            self.multimem_allreduce(bucket)
            if average:
                bucket.grad_flat.div_(self.world)
            return None
        work = dist.all_reduce(bucket.grad_flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        if average:
            if async_op:
                work.wait()
                work = None
            bucket.grad_flat.div_(self.world)
        return work

    def symmetric_gradient_storage(self, numel: int, device):
        """Allocate a float32 gradient bucket in symmetric memory (same offset on every rank) and map the NVLink multicast
        object that spans the ranks' copies.  Collective: every rank calls it with the same numel.  Returns None when the
        GPUs have no multicast support (then use the NCCL transport).  torch's symmetric-memory runtime is used for the
        allocation, the rendezvous and the inter-GPU barrier only; the data movement is this library's kernel."""
        if self.world == 1:
            return None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            group = self.group if self.group is not None else dist.group.WORLD
            t = symm_mem.empty(int(numel), dtype=torch.float32, device=torch.device(device))
            hdl = symm_mem.rendezvous(t, group)
            ok = torch.tensor([1 if getattr(hdl, "multicast_ptr", 0) else 0], device=t.device)
        except Exception:
            t, hdl = None, None
            ok = torch.tensor([0], device=torch.device(device))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)     # every rank takes the same decision
        if int(ok) == 0:
            return None
        self._symm = (t, hdl)
        return t

    # ---- compact exchange of the SH gradient -------------------------------------------------------------------------
    def exchange_floats(self, tensors: Dict[str, torch.Tensor], capacity: Optional[int] = None) -> int:
        """Floats of symmetric storage for a bucket of `tensors` plus the all-gather slices of compact_sh_exchange().
        capacity: the largest number of Gaussians the model may grow to (densification); default = the current count.  The
        symmetric allocation cannot be resized, so a trainer that densifies sizes it for the grown model once."""
        P0 = tensors["means3D"].shape[0]
        P = max(int(capacity or 0), P0)
        return sum(_pad4(tensors[n].numel() // max(P0, 1) * P) for n, _ in PARAM_LAYOUT) + self.world * self._slice_stride(P)

    @staticmethod
    def _slice_stride(P: int) -> int:
        return (3 * P + 3 + 3) // 4 * 4          # [3P colour gradients | campos xyz | pad], a multiple of 16 bytes

    def compact_sh_exchange(self, bucket: GaussianBucket, sh_degree: int = 3):
        """Exchange 3 instead of 48 floats per Gaussian for the SH coefficients.  For ONE view, dL/dSH of a Gaussian is the
        outer product of its clamp-masked colour gradient with the SH basis of the direction camera -> Gaussian, so it is
        enough to all-GATHER the colour gradients (and the ranks' camera positions) and let every rank rebuild and sum the
        outer products itself (vcr_sh_grad_from_colors), in rank order.  The other per-Gaussian gradients (11 floats) are
        all-reduced as before.  Per GPU and step this moves 44 B + world x 12 B per Gaussian instead of 236 B.
        After this call the rasterizer's backward no longer writes the SH gradient into the bucket; call
        exchange_compact(bucket, campos) after the backward in place of allreduce_gradients().  Needs direct=True."""
        self._compact = None
        if not bucket.direct:
            raise ValueError("compact_sh_exchange needs GaussianBucket(direct=True)")
        if "shs" not in bucket.params or "means3D" not in bucket.params:
            raise ValueError("compact_sh_exchange needs a bucket of the rasterizer's inputs (means3D, shs, ...): the rank-one "
                             "structure belongs to the gradient of the `shs` tensor the rasterizer is called with")
        P, dev = bucket.P, bucket.grad_flat.device
        stride = self._slice_stride(P)
        off0 = _pad4(bucket.numel)
        t, _ = getattr(self, "_symm", (None, None))
        if self._in_symmetric_storage(bucket) and t.numel() >= off0 + self.world * stride:
            flat = t.view(-1)[off0:off0 + self.world * stride]
            self._compact_off0 = off0
        else:
            flat = torch.zeros(self.world * stride, dtype=torch.float32, device=dev)
            self._compact_off0 = None            # not reachable through the multicast mapping: NCCL only
        self._slices = flat.view(self.world, stride)
        self._slices.zero_()
        self._compact = (P, stride, int(sh_degree), bucket.grad_flat.data_ptr())
        if self.world > 1:
            bucket.params["shs"]._vcr_sh_compact = self._slices[self.rank, :3 * P].view(P, 3)

    def exchange_compact(self, bucket: GaussianBucket, campos: torch.Tensor, transport: Optional[str] = None):
        """The gradient exchange of compact_sh_exchange(): all-gather of [colour gradients | campos], all-reduce of the
        non-SH groups, reconstruction of the summed SH gradient into the bucket.  On the current stream."""
        if self.world == 1:
            return
        import vcr_native as _n
        if getattr(self, "_compact", None) is None:
            raise RuntimeError("exchange_compact: call compact_sh_exchange(bucket) first")
        P, stride, D, grad_ptr = self._compact
        if bucket.P != P or bucket.grad_flat.data_ptr() != grad_ptr or getattr(bucket.params["shs"], "_vcr_sh_compact", None) is None:
            # e.g. the bucket returned by prune_and_grow(): its gradients are not the ones this exchange was laid out for,
            # and its SH gradient is not being produced in compact form — exchanging would silently diverge the replicas
            raise RuntimeError("exchange_compact: this bucket is not the one compact_sh_exchange() was set up for "
                               "(after prune_and_grow, call compact_sh_exchange / pick_transport again on the new bucket)")
        transport = transport or self.transport
        own = self._slices[self.rank]
        own[3 * P:3 * P + 3].copy_(campos.reshape(-1)[:3])
        ranges = [(off, n) for name, off, n, _ in bucket.layout if name != "shs"]
        stream = ctypes.c_void_p(torch.cuda.current_stream(bucket.grad_flat.device).cuda_stream)
        if transport == "multimem" and self._compact_off0 is not None:
            t, hdl = self._symm
            arr = ctypes.c_longlong * len(ranges)
            hdl.barrier(channel=0, timeout_ms=30000)
            _n.check(_n.vcr_multimem_exchange(ctypes.c_void_p(int(hdl.multicast_ptr)), t.data_ptr(), len(ranges),
                                              arr(*[a for a, _ in ranges]), arr(*[n for _, n in ranges]),
                                              self._compact_off0 + self.rank * stride, stride, self.rank, self.world,
                                              int(os.environ.get("VCR_MM_BLOCKS", "0")), stream), "multimem_exchange")
            hdl.barrier(channel=0, timeout_ms=30000)
        else:
            dist.all_gather_into_tensor(self._slices.view(-1), own, group=self.group)      # in place: own is slice `rank`
            with dist._coalescing_manager(group=self.group, async_ops=False):
                for a, n in ranges:
                    dist.all_reduce(bucket.grad_flat[a:a + n], op=dist.ReduceOp.SUM, group=self.group)
        name, off, n, shape = next(item for item in bucket.layout if item[0] == "shs")
        _n.check(_n.vcr_sh_grad_from_colors(P, D, shape[1], self.world, bucket.params["means3D"].data_ptr(), self._slices.data_ptr(),
                                            stride, bucket.grad_flat[off:off + n].data_ptr(), stream), "sh_grad_from_colors")

    def _in_symmetric_storage(self, bucket: GaussianBucket) -> bool:
        t, _ = getattr(self, "_symm", (None, None))
        return t is not None and bucket.grad_flat.data_ptr() == t.data_ptr()

    def multimem_allreduce(self, bucket: GaussianBucket):
        """SUM of the whole bucket over ranks on the current stream with this library's NVSwitch-multicast kernel
        (vcr_multimem_allreduce_sum): barrier (every rank's gradients are written) -> one launch -> barrier (every rank's
        stores have landed).  The bucket must live in symmetric_gradient_storage()."""
        import vcr_native as _n
        if not self._in_symmetric_storage(bucket):
            raise ValueError("multimem_allreduce needs GaussianBucket(grad_storage=symmetric_gradient_storage(...))")
        _, hdl = self._symm
        one = ctypes.c_longlong * 1
        hdl.barrier(channel=0, timeout_ms=30000)
        _n.check(_n.vcr_multimem_allreduce_sum(ctypes.c_void_p(int(hdl.multicast_ptr)), 1, one(0), one(bucket.numel), self.rank,
                                               self.world, int(os.environ.get("VCR_MM_BLOCKS", "0")),
                                               ctypes.c_void_p(torch.cuda.current_stream(bucket.grad_flat.device).cuda_stream)),
                 "multimem_allreduce_sum")
        hdl.barrier(channel=0, timeout_ms=30000)

    def pick_transport(self, bucket: GaussianBucket, iters: int = 10):
        """Time the exchange of this bucket both ways (NCCL all_reduce; the multicast kernel when the bucket is in symmetric
        storage), keep the faster as `self.transport` and return the measured milliseconds.  Which one wins depends on the
        number of ranks: through the switch every element also loops back to its own GPU, which costs 1/world of the
        traffic — 50 % extra at 2 ranks, 14 % at 8 — while the in-switch reduction saves NCCL's per-peer work.
        Collective; the bucket's gradient contents are clobbered."""
        self.transport = "nccl"
        if self.world == 1:
            return {}
        dev = bucket.grad_flat.device
        if getattr(self, "_compact", None) is not None:      # compact_sh_exchange() is set up: time THAT exchange
            cam = torch.zeros(3, device=dev)
            ways = {"nccl": lambda: self.exchange_compact(bucket, cam, transport="nccl")}
            if self._compact_off0 is not None:
                ways["multimem"] = lambda: self.exchange_compact(bucket, cam, transport="multimem")
        else:
            ways = {"nccl": lambda: dist.all_reduce(bucket.grad_flat, op=dist.ReduceOp.SUM, group=self.group)}
            if self._in_symmetric_storage(bucket):
                ways["multimem"] = lambda: self.multimem_allreduce(bucket)
        ms = {}
        for name, fn in ways.items():
            bucket.grad_flat.zero_()
            for _ in range(3):
                fn()
            torch.cuda.synchronize(dev)
            dist.barrier(group=self.group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)     # same number, hence same choice, on every rank
            ms[name] = float(t)
        self.transport = min(ms, key=ms.get)
        bucket.grad_flat.zero_()
        return ms

    def overlap_gradients(self, bucket: GaussianBucket, n_chunks: int = 4, transport: str = "nccl"):
        """Pipeline the gradient exchange with the tail of the backward: the rasterizer's per-Gaussian backward stage is
        issued as `n_chunks` launches over consecutive Gaussian ranges (vcr_rasterize_backward_chunked) and the all-reduce
        of each range's rows of the bucket is queued as soon as that launch is enqueued, so the exchange of range c runs on
        the communication stream while range c+1 is being computed.  The five row slices of a range (one per parameter
        group) go out as ONE coalesced collective.  Needs a `direct=True` bucket (gradients land in the bucket while the
        backward is still running) and exactly one rasterizer backward per step; call finish_gradients() afterwards in
        place of allreduce_gradients().  A no-op at world size 1."""
        if not bucket.direct:
            raise ValueError("overlap_gradients needs GaussianBucket(direct=True)")
        self._pending = []
        if self.world == 1:
            return
        rows = [(off, n // bucket.P) for _, off, n, _ in bucket.layout]
        self._transport = transport
        if transport == "multimem":
            # transport="multimem": the bucket lives in symmetric_gradient_storage(); each range is exchanged by ONE launch
            # of this library's two-shot kernel (multimem.ld_reduce / multimem.st through the NVSwitch,
            # csrc/view_parallel.cu) on a side stream, bracketed by the symmetric-memory barrier
            import vcr_native as _n
            t, hdl = getattr(self, "_symm", (None, None))
            if t is None or bucket.grad_flat.data_ptr() != t.data_ptr():
                raise ValueError("transport='multimem' needs GaussianBucket(grad_storage=symmetric_gradient_storage(...))")
            dev = bucket.grad_flat.device
            self._comm = torch.cuda.Stream(device=dev)
            mc = ctypes.c_void_p(int(hdl.multicast_ptr))
            nr = len(rows)
            blocks = int(os.environ.get("VCR_MM_BLOCKS", "0"))
            arr = ctypes.c_longlong * nr

            def on_chunk(_chunk, first, count):
                ev = torch.cuda.Event()
                ev.record()                                    # this range's gradient rows are final after this point
                with torch.cuda.stream(self._comm):
                    self._comm.wait_event(ev)
                    hdl.barrier(channel=0, timeout_ms=30000)   # ... on every rank
                    _n.check(_n.vcr_multimem_allreduce_sum(mc, nr, arr(*[off + first * per for off, per in rows]),
                                                           arr(*[count * per for _, per in rows]), self.rank, self.world,
                                                           blocks, ctypes.c_void_p(self._comm.cuda_stream)),
                             "multimem_allreduce_sum")
                self._pending.append(None)

            def finish():
                with torch.cuda.stream(self._comm):
                    hdl.barrier(channel=0, timeout_ms=30000)   # every rank's stores have landed everywhere
                    done = torch.cuda.Event()
                    done.record()
                torch.cuda.current_stream().wait_event(done)
            self._mm_finish = finish
        elif transport == "nccl":
            def on_chunk(_chunk, first, count):
                with dist._coalescing_manager(group=self.group, async_ops=True) as cm:
                    for off, per in rows:
                        dist.all_reduce(bucket.grad_flat[off + first * per: off + (first + count) * per],
                                        op=dist.ReduceOp.SUM, group=self.group)
                self._pending.append(cm)
        else:
            raise ValueError("transport must be 'nccl' or 'multimem'")

        bucket.params["means3D"]._vcr_chunk_hook = (max(int(n_chunks), 1), on_chunk)

    def finish_gradients(self, bucket: GaussianBucket, average: bool = False):
        """Order the current stream after every chunk exchange queued by the overlap_gradients() hook during the last
        backward.  Returns the number of chunk collectives that were pending."""
        pending, self._pending = getattr(self, "_pending", []), []
        if getattr(self, "_transport", "nccl") == "multimem":
            if pending:
                self._mm_finish()
        else:
            for cm in pending:
                cm.wait()
        if average and self.world > 1:
            bucket.grad_flat.div_(self.world)
        return len(pending)

    def accumulate_densification_stats(self, xyz_gradient_accum, denom, max_radii2D, grad_means2D_densify, radii):
        """Per view, LOCAL: add_densification_stats (scene/gaussian_model.py:669-671) + the max_radii2D update
        (trainer.py:345).  No communication: sums and maxima commute, and the statistics are only read when the model
        densifies (every 100 iterations, trainer.py:357-370), so the ranks are combined there, once, by
        sync_densification_stats().  Tensors are updated in place; shapes [P,1], [P,1], [P]."""
        if radii.is_cuda:
            # one fused pass (csrc/view_parallel.cu) instead of ~10 small PyTorch launches
            import vcr_native as _n
            g, r = grad_means2D_densify.contiguous(), radii.contiguous()
            ok = (g.dtype == torch.float32 and r.dtype == torch.int32 and all(
                t.dtype == torch.float32 and t.is_contiguous() and t.is_cuda for t in (xyz_gradient_accum, denom, max_radii2D)))
            if not ok:
                raise TypeError("densification statistics: expected contiguous float32 accumulators, float32 grads, int32 radii")
            _n.check(_n.vcr_densify_stats_accumulate(int(r.numel()), g.data_ptr(), r.data_ptr(), xyz_gradient_accum.data_ptr(),
                                                     denom.data_ptr(), max_radii2D.data_ptr(),
                                                     ctypes.c_void_p(torch.cuda.current_stream(r.device).cuda_stream)),
                     "densify_stats_accumulate")
            return
        # CPU tensors (the gloo tests): the same three updates in PyTorch
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
