#!/usr/bin/env python
"""bench.py — the hot path of VCR-GauS on B200: forward + backward rasterization + D-Normal loss, per view.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|reference-cpu] [--config metric|A|B|C|D|E]

Default workload (BASELINE.json `metric`: "fwd+bwd rasterize ms/view (1M Gauss, 1080p); train views/s at 1/2/4/8 GPU"):
a seeded synthetic scene of 1,000,000 Gaussians (SH degree 3), 1920x1080 views with ray/plane-intersection depth
(`dirs`, the configured default of every reconstruction config, configs/reconstruct.yaml:10), loss = L1(RGB) +
mono-normal + confidence-weighted depth-normal (D-Normal), backward into every per-Gaussian parameter, densification
statistics updated every view.  One "step" = one such view per GPU (fwd + loss + bwd); with N > 1 each rank renders a
different view of the same replicated scene and the per-Gaussian gradients are summed over the ranks (weak scaling).
`value` = views/s over all ranks (so ms/view at N = 1 is `ms_per_step`).  `--config` selects the other BASELINE.json
configurations (bench_common.CONFIGS): A (10k / 256x256), B (300k / 800x600 / 49 views / complete training iteration),
C (1.5M / 1080p / training with densification), D (5M / 1256x828), E (one point of the rasterizer sweep, fwd+bwd only).

Arms (bench_common.py holds what they share; each arm's process imports only its own implementation)
  ours           this repository (GaussianRasterizer -> ctypes -> libvcr_gaus_b200.so, fused loss / prologue / Adam kernels)
  reference      bench_ref.py: the reference's own CUDA extension, Python wrapper, GaussianModel, render() and loss
                 utilities from oracle/_ref — same scene, same views, same loss, same GPU.  (The reference has no CPU
                 implementation of this path; the arm the north star asks to beat is its CUDA rasterizer
                 "on the same B200".)
  reference-cpu  the CPU oracle (oracle/vcr_oracle.c, a port) on all host cores; also timed, on a bounded sample,
                 as `cpu_baseline` inside the default arm at N = 1.
"""
import gc
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench_common as BC  # noqa: E402


def _product_paths():
    for p in (os.path.join(ROOT, "vcr-gaus_b200"),):
        if p not in sys.path:
            sys.path.insert(0, p)


# ------------------------------------------------------------------------------------------------------------------
def measure_view_costs(args, bucket, views, dev):
    """Milliseconds of one rasterizer forward+backward of each view (CUDA events, best of two after one untimed pass) — what
    a trainer knows about a view from the previous epoch.  The instance count R alone predicts it to ~5 % only."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    p = bucket.params
    costs = []
    for v in views:
        i = v["inp"]
        rs = GaussianRasterizationSettings(image_height=args.height, image_width=args.width, tanfovx=i["tanfovx"], tanfovy=i["tanfovy"],
                                           bg=i["bg"].to(dev), scale_modifier=1.0, viewmatrix=i["viewmatrix"].to(dev),
                                           projmatrix=i["projmatrix"].to(dev), sh_degree=3, campos=i["campos"].to(dev),
                                           prefiltered=False, debug=False, f_count=0)
        normals, dirs = i["normals"].to(dev), i["dirs"].to(dev)
        g = torch.ones((10, args.height, args.width), device=dev)
        best = float("inf")
        for rep in range(3):
            bucket.zero_grad()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out, _ = GaussianRasterizer(rs)(means3D=p["means3D"], means2D=None, opacities=p["opacities"], shs=p["shs"],
                                            normals_precomp=normals, scales=p["scales"], rotations=p["rotations"], dirs=dirs)
            out.backward(g)
            e1.record()
            torch.cuda.synchronize(dev)
            if rep > 0:
                best = min(best, e0.elapsed_time(e1))
        costs.append(best)
    bucket.zero_grad()
    return costs


def unpack_u8(u8):
    """uint8 ground truth -> the float image u8 / 255, by the library's kernel, on the current (copy) stream."""
    import vcr_native as N
    out = torch.empty(u8.shape, dtype=torch.float32, device=u8.device)
    N.check(N.vcr_unpack_u8(u8.data_ptr(), out.data_ptr(), u8.numel(), torch.cuda.current_stream(u8.device).cuda_stream), "unpack_u8")
    return out


def wire_unpack(vp):
    """Wire format of the ground-truth image in the e2e region: float32 on one GPU (the copy hides behind the step, an unpack
    kernel only adds work), uint8 + vcr_unpack_u8 when several ranks share the host's memory system.  VCR_BENCH_WIRE overrides."""
    w = os.environ.get("VCR_BENCH_WIRE", "u8" if vp.world > 1 else "f32")
    return unpack_u8 if w == "u8" else None


def clone_params(scene, dev, vp):
    """Rasterizer-input parameters + flat gradient bucket.  With a ViewParallel layer at N > 1 the bucket is placed in
    symmetric memory so that the gradient exchange can go through the NVSwitch multicast kernel (VCR_BENCH_TRANSPORT=nccl
    forces NCCL)."""
    import vcr_view_parallel as VP
    tensors = dict(means3D=scene.means3D.to(dev), shs=scene.shs.to(dev), opacities=scene.opacities.to(dev),
                   scales=scene.scales.to(dev), rotations=scene.rotations.to(dev))
    store = None
    if vp.world > 1 and os.environ.get("VCR_BENCH_TRANSPORT", "multimem") == "multimem":
        # (symmetric storage is also fine for NCCL, so it is allocated whenever the multicast kernel may be chosen)
        store = vp.symmetric_gradient_storage(vp.exchange_floats(tensors), dev)
    b = VP.GaussianBucket(tensors, device=dev, direct=True, grad_storage=store)
    b.transport = "multimem" if store is not None else "nccl"
    b.compact = bool(vp.world > 1 and os.environ.get("VCR_BENCH_COMPACT", "1") == "1")
    if b.compact:
        vp.compact_sh_exchange(b, sh_degree=3)     # exchange 3 instead of 48 floats per Gaussian for the SH coefficients
    return b


class HotPathStepper:
    """One view: (optional H2D of the per-view inputs) -> forward -> loss -> backward -> (N>1: gradient exchange) ->
    densification statistics -> loss readback."""

    def __init__(self, args, scene, views, dev, vp):
        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
        self.GRS, self.GR = GaussianRasterizationSettings, GaussianRasterizer
        self.args, self.dev, self.vp = args, dev, vp
        self.bucket = bucket = clone_params(scene, dev, vp)
        self.views = views
        self.feed = BC.ViewFeeder(views, dev, extra=("dL_dout",) if args.step_kind == "raster" else (), u8_image=wire_unpack(vp))
        P = bucket.P
        self.inside = torch.ones(P, dtype=torch.bool, device=dev)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros((P,), device=dev)
        inp0 = views[0]["inp"]
        # all views share intrinsics and size, so `dirs` is per scene, like scene.dirs (scene/__init__.py:99-102)
        self.dirs = inp0["dirs"].to(dev)
        K = views[0]["cam"].intr
        self.K4 = (float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]))
        self.tanx, self.tany = inp0["tanfovx"], inp0["tanfovy"]
        self.d2h_bytes = 4
        self.loss_host = torch.zeros(1).pin_memory()
        self.diag_events = None
        self.skip = os.environ.get("VCR_BENCH_SKIP", "")   # diagnosis only: "grads", "stats" or "grads,stats"
        # N > 1: pipeline the gradient all-reduce with the per-Gaussian backward stage in this many ranges (0 = one
        # exchange after the backward)
        self.chunks, self.exchange_ms = 0, {}
        if vp.world > 1:
            self.exchange_ms = vp.pick_transport(bucket)          # NCCL or the multicast kernel, whichever is faster here
            forced = os.environ.get("VCR_BENCH_TRANSPORT")
            if forced in ("nccl", "multimem") and (forced == "nccl" or bucket.transport == "multimem"):
                vp.transport = forced
            # pipelining pays with the multicast kernel (a quarter of the SMs, 8 GPUs: -0.04 ms/step), not with NCCL; the
            # compact exchange is too small to be worth pipelining
            self.chunks = 0 if bucket.compact else int(os.environ.get("VCR_BENCH_CHUNKS", "4" if vp.transport == "multimem" else "0"))
        if self.chunks > 0 and "grads" not in self.skip:
            vp.overlap_gradients(bucket, n_chunks=self.chunks, transport=vp.transport)
        # normals_precomp depends on the view (flip towards the camera + w2c rotation): precomputed per view on device
        self.normals = [v["inp"]["normals"].to(dev).requires_grad_(True) for v in views]

    def step(self, i, e2e=False):
        from vcr_dnormal import view_loss
        a, vp = self.args, self.vp
        k = i % self.feed.n
        p = self.bucket.params
        d = self.feed.get(i, e2e)
        view, proj, campos, bg = BC.ViewFeeder.camera(d)
        normals = self.normals[k]
        rs = self.GRS(image_height=a.height, image_width=a.width, tanfovx=self.tanx, tanfovy=self.tany, bg=bg, scale_modifier=1.0,
                      viewmatrix=view, projmatrix=proj, sh_degree=3, campos=campos, prefiltered=False, debug=False, f_count=0)
        means2D = torch.zeros_like(p["means3D"], requires_grad=True)
        means2D_densify = torch.zeros_like(p["means3D"], requires_grad=True)
        out, radii = self.GR(rs)(means3D=p["means3D"], means2D=means2D, means2D_densify=means2D_densify, shs=p["shs"],
                                 normals_precomp=normals, opacities=p["opacities"], scales=p["scales"],
                                 rotations=p["rotations"], dirs=self.dirs, inside=None)
        self.bucket.zero_grad()
        normals.grad = None
        ev = self.diag_events
        if ev is not None:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev.append(e)
            e[0].record()
        if a.step_kind == "raster":
            out.backward(d["dL_dout"])
            loss = out[0, 0, 0].detach()
        else:
            # the three terms through this repository's public loss API (one fused kernel per direction).  Mean over ranks =
            # SUM all-reduce of the gradients of loss / world: the 1/world goes into the loss weights, so there is neither a
            # division pass over the bucket nor an extra scaling kernel
            sc = 1.0 / vp.world
            loss = view_loss(out, d["gt_image"], d["gt_normal"], self.K4, w_l1=sc, w_mono=BC.W_MONO * sc, w_dn=BC.W_DN * sc, exp_t=BC.EXP_T)
            loss.backward()     # N > 1 with chunks: the chunk hook queues the gradient exchange range by range during this call
        if ev is not None:
            e[1].record()
        skip = self.skip
        if vp.world > 1 and "grads" not in skip:
            if getattr(self.bucket, "compact", False):
                vp.exchange_compact(self.bucket, campos)
            elif self.chunks > 0:
                vp.finish_gradients(self.bucket)
            else:
                vp.allreduce_gradients(self.bucket, average=False)
        if ev is not None:
            e[2].record()
        if a.step_kind != "raster" and "stats" not in skip:
            # densification statistics (trainer.py:345-347), every view at every N: one fused local pass; the ranks are
            # combined every 100 steps (the reference densifies every 100 iterations, trainer.py:357-370)
            vp.accumulate_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D, means2D_densify.grad, radii)
            if (i + 1) % 100 == 0:
                vp.sync_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D)
                self.xyz_gradient_accum.zero_(); self.denom.zero_(); self.max_radii2D.zero_()
        if e2e:
            self.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        return loss

    def local_gradient_checksums(self, i):
        """For the exchange self-check: this rank's OWN gradients of view i (no exchange), as float64 sums per group."""
        skip, self.skip = self.skip, "grads,stats"
        p = self.bucket.params
        # the plain backward: dL_dsh written by the kernel itself (no compact colour gradients), no chunk hook queuing exchanges
        compact = p["shs"].__dict__.pop("_vcr_sh_compact", None)
        hook = p["means3D"].__dict__.pop("_vcr_chunk_hook", None)
        try:
            self.step(i, False)
        finally:
            self.skip = skip
            if compact is not None:
                p["shs"]._vcr_sh_compact = compact
            if hook is not None:
                p["means3D"]._vcr_chunk_hook = hook
        return torch.stack([self.bucket.grad_flat[o:o + n].double().sum() for _, o, n, _ in self.bucket.layout])


class TrainStepper:
    """One COMPLETE training iteration per view, from the raw parameters to the updated raw parameters, the way trainer.py
    runs it (trainer.py:323-392): per-Gaussian prologue -> rasterize -> loss (0.8 L1 + 0.2 (1 - SSIM) + mono-normal +
    D-Normal) -> backward -> (N > 1: SUM of the raw gradient bucket) -> densification statistics -> (densify + prune every
    100 iterations) -> Adam.  Fused prologue, fused view loss + fused SSIM, BucketAdam, device-side densification; the
    prologue's and the rasterizer's backward kernels write the raw gradients straight into the flat bucket."""

    def __init__(self, args, scene, views, dev, vp):
        import vcr_view_parallel as VP
        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
        self.GRS, self.GR, self.VP = GaussianRasterizationSettings, GaussianRasterizer, VP
        self.args, self.dev, self.vp = args, dev, vp
        self.feed = BC.ViewFeeder(views, dev, u8_image=wire_unpack(vp))
        self.views = views
        raw = {k: v.to(dev) for k, v in BC.raw_parameters(scene).items()}
        store = None
        self.capacity = int(args.gaussians * (1.6 if args.densify else 1.0))
        if vp.world > 1 and os.environ.get("VCR_BENCH_TRANSPORT", "multimem") == "multimem":
            per = sum(v.numel() // args.gaussians for v in raw.values())
            store = vp.symmetric_gradient_storage((per * self.capacity + 64) // 4 * 4, dev)
        self.store = store
        self.bucket = VP.GaussianBucket(raw, device=dev, names=BC.RAW_NAMES, direct=True, grad_storage=store)
        self.opt = VP.BucketAdam(self.bucket, BC.LRS)
        self.exchange_ms = {}
        if vp.world > 1:
            self.exchange_ms = vp.pick_transport(self.bucket)
            forced = os.environ.get("VCR_BENCH_TRANSPORT")
            if forced == "nccl":
                vp.transport = "nccl"
        self._alloc_stats()
        inp0 = views[0]["inp"]
        self.dirs = inp0["dirs"].to(dev)
        K = views[0]["cam"].intr
        self.K4 = (float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]))
        self.tanx, self.tany = inp0["tanfovx"], inp0["tanfovy"]
        self.extent = scene.extent
        self.d2h_bytes = 4
        self.loss_host = torch.zeros(1).pin_memory()
        self.iteration = 0
        self.P_history = [self.bucket.P]

    def _alloc_stats(self):
        P, dev = self.bucket.P, self.dev
        self.inside = torch.ones(P, dtype=torch.bool, device=dev)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros((P,), device=dev)

    def step(self, i, e2e=False):
        from vcr_dnormal import ssim, view_loss
        from vcr_prologue import gaussian_prologue
        a, vp, p = self.args, self.vp, self.bucket.params
        self.iteration += 1
        d = self.feed.get(i, e2e)
        view, proj, campos, bg = BC.ViewFeeder.camera(d)
        R_w2c = view[:3, :3].t()
        opac, scales, rots, sh, normals = gaussian_prologue(p["xyz"], p["f_dc"], p["f_rest"], p["opacity"], p["scaling"], p["rotation"],
                                                            campos, R_w2c)
        rs = self.GRS(image_height=a.height, image_width=a.width, tanfovx=self.tanx, tanfovy=self.tany, bg=bg, scale_modifier=1.0,
                      viewmatrix=view, projmatrix=proj, sh_degree=3, campos=campos, prefiltered=False, debug=False, f_count=0)
        means2D = torch.zeros_like(p["xyz"], requires_grad=True)
        means2D_densify = torch.zeros_like(p["xyz"], requires_grad=True)
        out, radii = self.GR(rs)(means3D=p["xyz"], means2D=means2D, means2D_densify=means2D_densify, shs=sh, normals_precomp=normals,
                                 opacities=opac, scales=scales, rotations=rots, dirs=self.dirs, inside=None)
        sc = 1.0 / vp.world
        loss = view_loss(out, d["gt_image"], d["gt_normal"], self.K4, w_l1=BC.W_L1 * sc, w_mono=BC.W_MONO * sc, w_dn=BC.W_DN * sc, exp_t=BC.EXP_T)
        loss = loss + (BC.W_SSIM * sc) * (1.0 - ssim(out[:3], d["gt_image"]))
        self.bucket.zero_grad()
        loss.backward()
        if vp.world > 1:
            vp.allreduce_gradients(self.bucket, average=False)
        vp.accumulate_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D, means2D_densify.grad, radii)
        if a.densify and self.iteration % BC.DENSIFY_INTERVAL == 0:
            self.densify()
        self.opt.step()
        if e2e:
            self.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        return loss

    def densify(self):
        """densify_and_prune of the reference (scene/gaussian_model.py:645-659) on the synchronised statistics; identical
        decisions, hence identical replicas, on every rank."""
        from vcr_densify import densify_and_prune
        vp = self.vp
        vp.sync_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D)
        thr = BC.densify_threshold(self.xyz_gradient_accum, self.denom)
        self.bucket, self.opt, info = densify_and_prune(self.bucket, self.opt, self.xyz_gradient_accum, self.denom, self.max_radii2D,
                                                        max_grad=thr, min_opacity=BC.MIN_OPACITY, extent=self.extent,
                                                        max_screen_size=None, percent_dense=BC.PERCENT_DENSE, grad_storage=self.store)
        self._alloc_stats()
        self.P_history.append(self.bucket.P)


def cpu_oracle_views_per_s(args, view, repeat=1):
    """The CPU oracle (a port; the reference has no CPU path) on one full view of the same workload, all host cores."""
    for p in (os.path.join(ROOT, "tests"),):
        if p not in sys.path:
            sys.path.insert(0, p)
    from oracle import oracle as O
    import numpy as np
    inp = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in view["inp"].items()}
    g = np.random.default_rng(0).standard_normal((10, args.height, args.width)).astype(np.float32)
    t0 = time.time()
    for _ in range(repeat):
        f = O.rasterize_forward(inp)
        O.rasterize_backward(inp, f, g)
    dt = (time.time() - t0) / repeat
    return 1.0 / dt, dt, f["binned"]["R"]


def main():
    args = BC.parse()
    BC.claim_stdout()
    if args.impl == "reference":
        import bench_ref
        bench_ref.main(args)
        return
    if args.impl == "reference-cpu":
        if int(os.environ.get("RANK", "0")) == 0:
            scene = BC.build_scene(args)
            views = BC.build_views(args, scene, 1, [0], light=True)
            vps, dt, R = cpu_oracle_views_per_s(args, views[0])
            cores = BC.host_cores()
            BC.emit({"impl": "reference", "metric": BC.METRIC, "value": vps, "unit": "views/s", "n_gpus": 0, "steps": 1, "warmup": 0,
                     "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                     "data": "synthetic", "config": BC.config_dict(args, 1),
                     "cpu_baseline": {"value": vps, "unit": "views/s", "cores": cores, "kind": "port",
                                      "sample": "one full view, rasterize fwd+bwd (no D-Normal), CPU oracle with OpenMP over all host cores"},
                     "e2e": {"value": vps, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        return
    _product_paths()
    import vcr_native as N
    import vcr_view_parallel as VP
    rank, world, local = VP.init_from_env()
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback in the product path)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    vp = VP.ViewParallel()
    all_cores = os.sched_getaffinity(0)
    numa_cores = VP.bind_to_gpu_numa_node(local) if os.environ.get("VCR_BENCH_NUMA", "1") == "1" else 0
    scene = BC.build_scene(args)
    view_ids = [v * world + rank for v in range(args.views)]          # default shard: every world-th camera of the ring
    train = args.step_kind == "train"
    balanced = world > 1 and not train and os.environ.get("VCR_BENCH_BALANCE", "1") == "1"
    if balanced:
        # Cost-balanced sharding: every step ends with the gradient exchange, i.e. with a wait for the slowest rank, so the
        # `world` views of one step should cost the same.  Each rank times its default shard once, the costs are
        # gathered, and ViewParallel.balanced_view_table deals views of equal cost to the ranks of one step.
        tmp = VP.GaussianBucket(dict(means3D=scene.means3D.to(dev), shs=scene.shs.to(dev), opacities=scene.opacities.to(dev),
                                     scales=scene.scales.to(dev), rotations=scene.rotations.to(dev)), device=dev, direct=True)
        mine = torch.tensor(measure_view_costs(args, tmp, BC.build_views(args, scene, world, view_ids, light=True), dev), device=dev)
        del tmp
        allc = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allc, mine)
        costs = [0.0] * (args.views * world)
        for r in range(world):
            for v in range(args.views):
                costs[v * world + r] = float(allc[r][v])
        view_ids = [row[rank] for row in vp.balanced_view_table(costs, seed=0)]
    views = BC.build_views(args, scene, world, view_ids)
    stepper = (TrainStepper if train else HotPathStepper)(args, scene, views, dev, vp)

    if os.environ.get("VCR_RENDER_VERSION"):        # development A/B of the compositing kernels (1 = one pixel per thread)
        N.vcr_debug_render_version(int(os.environ["VCR_RENDER_VERSION"]))
    prof = not args.no_profile and not args.densify
    # ---- device-resident timing (`value`): K steps, launch counters on, per-kernel event brackets OFF ----
    N.vcr_profile_enable(0)
    # warm-up: W steps, but at least three passes over the view set, so that the caching allocator has seen every
    # per-view buffer size (the binning buffer depends on R) and no cudaMalloc lands in the timed region
    # ... and long enough (a fixed count, identical on every rank and in both arms) for clocks and power state to settle
    sampler = BC.ClockSampler(dev.index or 0)
    # no cyclic garbage collection inside the measured regions: a generation-2 pass on ONE rank stalls that rank's launches
    # for milliseconds and, through the per-step exchange, every other rank with it (the step creates no reference cycles)
    gc.collect()
    gc.disable()
    # (VCR_BENCH_WARMUP_FLOOR: profiling runs under ncu lower the floor; bench numbers are always taken with the default)
    floor = int(os.environ.get("VCR_BENCH_WARMUP_FLOOR", "200"))
    n_warm = max(args.warmup, 3 * len(views), floor)
    if args.densify:
        n_warm = max(args.warmup, 3 * len(views))        # a densifying run is timed from a fixed iteration count in both arms
    for i in range(n_warm):
        stepper.step(i, False)
    N.profile_read(reset=True)
    ms, clocks = BC.timed_region(lambda i: stepper.step(i, False), args.steps, 0, dev, world, sampler=sampler, first=n_warm)
    counts = N.profile_read(reset=True)
    nxt = n_warm + args.steps
    # ---- the same K steps once more with every kernel of the library bracketed by CUDA events on its stream ----
    # (the brackets cost ~4 % of a step, so they are kept out of `value`; `ms_per_step_with_kernel_events` reports it)
    kern, ms_prof = {}, None
    if prof:
        N.vcr_profile_enable(1)
        ms_prof, _ = BC.timed_region(lambda i: stepper.step(i, False), args.steps, 1, dev, world, first=nxt)
        kern = N.profile_read(reset=True)
        N.vcr_profile_enable(0)
        nxt += args.steps + 1
    # ---- end-to-end timing (`e2e`): per-view inputs from pinned host memory, loss read back ----
    # (its own warm-up: three passes over the views in e2e mode, so that the caching allocator has seen the prefetch buffers —
    # they are freed late because of record_stream — and no cudaMalloc lands in the timed region)
    w_e2e = max(args.warmup, 3 * len(views))
    ms_e2e, _ = BC.timed_region(lambda i: stepper.step(i, True), args.steps, w_e2e, dev, world, first=nxt)
    nxt += args.steps + w_e2e
    exchange_check = verify_exchange(stepper, vp, nxt, dev) if (world > 1 and not train) else None
    if os.environ.get("VCR_BENCH_DIAG") and not train:
        diagnose(stepper, args, dev, rank, world)
    views_per_s = world * args.steps / (ms / 1e3)
    e2e_vps = world * args.steps / (ms_e2e / 1e3)

    if rank != 0:
        return
    bucket = stepper.bucket
    cfg = BC.config_dict(args, world)
    out = {
        "metric": BC.METRIC, "value": views_per_s, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": cfg,
        "clocks": clocks,
        "host": f"process bound to the {numa_cores} cores next to its GPU" if numa_cores else "no CPU binding",
        "exchange": ({"chosen": vp.transport, "ms_per_exchange": stepper.exchange_ms, "compact_sh": bool(getattr(bucket, "compact", False)),
                      "pipelined_ranges": getattr(stepper, "chunks", 0),
                      "view_sharding": "cost-balanced (balanced_view_table)" if balanced else "every world-th camera",
                      "verified": exchange_check} if world > 1 else None),
        "e2e": {"value": e2e_vps, "unit": "views/s", "h2d_bytes_per_step": stepper.feed.h2d_bytes, "d2h_bytes_per_step": stepper.d2h_bytes,
                "ms_per_step": ms_e2e / args.steps,
                "wire": ("ground-truth image as uint8 (its file format), unpacked to u8/255 on the copy stream" if stepper.feed.u8_image
                         else "ground-truth image f32") + "; ground-truth normal, camera block f32"},
    }
    if hasattr(stepper, "P_history"):
        out["gaussians_after_each_densification"] = stepper.P_history
    # ---- roofline of the dominant kernel + per-kernel table ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    launches = sum(v[2] for k, v in counts.items() if k != "memset")
    out["gpu_launches"] = launches
    if not train:
        stats = scene_stats(stepper, args, clocks)
        per_kernel = kernel_table(kern, stats, args, hbm_peak)
        out["scene"] = stats
        if per_kernel:
            dom = max(per_kernel, key=lambda r: r["ms_per_step"])
            out["roofline"] = roofline_entry(dom, stats, args, hbm_peak, peak_src, clocks)
            out["kernels"] = per_kernel
            out["ms_per_step_with_kernel_events"] = ms_prof / args.steps if ms_prof else None
    else:
        out["kernels_ms_per_step"] = {k: v[0] / max(args.steps, 1) for k, v in kern.items() if v[1]}
        out["ms_per_step_with_kernel_events"] = ms_prof / args.steps if ms_prof else None
    if world == 1 and args.config == "metric" and not args.no_full_step and not os.environ.get("VCR_BENCH_NO_FULL_STEP"):
        try:
            out["full_step"] = time_full_step(args, scene, views, dev, vp, args.steps)
        except Exception as ex:  # a side measurement must never cost the bench line
            out["full_step"] = {"error": f"{type(ex).__name__}: {ex}"}
    if not args.no_cpu_baseline and world == 1:
        os.sched_setaffinity(0, all_cores)        # the CPU baseline gets every host core, not just the GPU's NUMA node
        out["cpu_baseline"] = cpu_baseline(args, scene, views)
    BC.emit(out)


def cpu_baseline(args, scene, views):
    """The CPU oracle (kind "port": the reference has no CPU implementation of the path) on a bounded sample of the
    workload: one full view of THIS config when that takes well under a minute (<= 1.1 M Gaussians), otherwise one view of
    the metric's 1 M / 1080p scene; plus config A (10k Gaussians, one 256x256 view — the configuration the CPU baseline is
    specified on), median of 5."""
    cores = BC.host_cores()
    big = args.gaussians > 1_100_000 or args.width * args.height > 2_100_000
    if big:
        import copy
        a1 = copy.copy(args)
        a1.gaussians, a1.width, a1.height, a1.views = 1_000_000, 1920, 1080, 1
        sc1 = BC.build_scene(a1)
        v1 = BC.build_views(a1, sc1, 1, [0], light=True)[0]
        vps, dt, R = cpu_oracle_views_per_s(a1, v1)
        sample = f"one full 1920x1080 view of the metric's 1M-Gaussian scene (this config is too large for a bounded CPU run), rasterize fwd+bwd ({dt:.1f} s)"
    else:
        vps, dt, R = cpu_oracle_views_per_s(args, views[0])
        sample = f"one full {args.width}x{args.height} view of the same scene, rasterize fwd+bwd, CPU oracle with OpenMP ({dt:.1f} s)"
    res = {"value": vps, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample}
    try:
        import copy
        aA = copy.copy(args)
        aA.gaussians, aA.width, aA.height, aA.views = 10_000, 256, 256, 1
        scA = BC.build_scene(aA)
        vA = BC.build_views(aA, scA, 1, [0], light=True)[0]
        ts = sorted(cpu_oracle_views_per_s(aA, vA)[1] for _ in range(5))
        res["config_A"] = {"ms_per_view": ts[2] * 1e3, "what": "10k Gaussians, one 256x256 view, rasterize fwd+bwd, median of 5"}
    except Exception as ex:
        res["config_A"] = {"error": f"{type(ex).__name__}: {ex}"}
    return res


def time_full_step(args, scene, views, dev, vp, steps):
    """Side measurement at the metric's size: the complete training iteration (TrainStepper) — reported beside the bench
    metric, whose step is the rasterizer + D-Normal hot path only.  `--config B` / `--config C` make it the step."""
    import copy
    a2 = copy.copy(args)
    a2.step_kind, a2.densify = "train", False
    st = TrainStepper(a2, scene, views, dev, vp)
    ms, _ = BC.timed_region(lambda i: st.step(i, False), steps, max(3 * len(views), 24), dev, 1)
    return {"ms_per_step": ms / steps, "views_per_s": 1e3 * steps / ms, "steps": steps,
            "includes": "raw parameters -> prologue -> rasterize fwd -> 0.8 L1 + 0.2 (1-SSIM) + mono-normal + D-Normal -> backward -> Adam (6 groups, eps 1e-15)",
            "reference": "python bench.py --impl reference --config metric reports the hot-path step; --config B the complete iteration on both arms"}


def verify_exchange(stepper, vp, i, dev):
    """After the timed regions, once: is the exchanged gradient of step i the SUM of the ranks' own gradients of their views?
    Every rank first computes its local gradient of view i with the exchange skipped (float64 sums per parameter group,
    all-gathered with NCCL), then the real step; the exchanged bucket's per-group sums must equal the sum over ranks of the
    local ones (1e-4 relative to the group's L1 mass: float-atomic order differs between the two backward runs), and all
    replicas must hold bit-identical buckets."""
    local = stepper.local_gradient_checksums(i)
    l1_local = torch.stack([stepper.bucket.grad_flat[o:o + n].double().abs().sum() for _, o, n, _ in stepper.bucket.layout])
    both = torch.stack([local, l1_local])
    allb = [torch.zeros_like(both) for _ in range(vp.world)]
    dist.all_gather(allb, both)
    want = sum(b[0] for b in allb)
    mass = sum(b[1] for b in allb)
    stepper.step(i, False)
    got = torch.stack([stepper.bucket.grad_flat[o:o + n].double().sum() for _, o, n, _ in stepper.bucket.layout])
    err = float(((got - want).abs() / mass.clamp(min=1e-30)).max())
    # bit-identical replicas: compare a position-weighted integer checksum of the raw bits
    bits = stepper.bucket.grad_flat.view(torch.int32).to(torch.int64)
    chk = torch.stack([bits.sum(), (bits * (torch.arange(bits.numel(), device=dev) % 65521)).sum()])
    allc = [torch.zeros_like(chk) for _ in range(vp.world)]
    dist.all_gather(allc, chk)
    same = all(bool(torch.equal(c, allc[0])) for c in allc)
    return {"sum_of_local_gradients_rel_err": err, "ok": bool(err < 1e-4), "replicas_bit_identical": same,
            "groups": [n for n, _, _, _ in stepper.bucket.layout]}


def diagnose(stepper, args, dev, rank, world):
    """VCR_BENCH_DIAG=1: per-rank split of a step into [start of backward -> gradients written] and [-> exchange done],
    plus the whole step, over 40 steps (printed to stderr by rank 0; not part of the bench line)."""
    n = 40
    stepper.diag_events = []
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    for i in range(n):
        marks[i].record()
        stepper.step(10_000 + i, False)
    marks[n].record()
    torch.cuda.synchronize(dev)
    ev, stepper.diag_events = stepper.diag_events, None
    step = [marks[i].elapsed_time(marks[i + 1]) for i in range(n)]
    bwd = [e[0].elapsed_time(e[1]) for e in ev]
    comm = [e[1].elapsed_time(e[2]) for e in ev]
    fwd = [marks[i].elapsed_time(ev[i][0]) for i in range(n)]
    t = torch.tensor([step, fwd, bwd, comm], device=dev)
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
    else:
        allt = [t]
    if rank == 0:
        a = torch.stack(allt).cpu()            # [world, 4, n]
        names = ["step", "fwd+loss", "bwd", "exchange_tail"]
        rep = {nm: {"mean_per_rank": [round(float(x), 3) for x in a[:, k].mean(dim=1)],
                    "mean_of_max_over_ranks": round(float(a[:, k].max(dim=0).values.mean()), 3),
                    "mean_of_min_over_ranks": round(float(a[:, k].min(dim=0).values.mean()), 3)} for k, nm in enumerate(names)}
        print("DIAG " + json.dumps(rep), file=sys.stderr, flush=True)


def scene_stats(stepper, args, clocks):
    """P, visible P, R, sum n_contrib and the pair counts K_e / K_c of view 0 (what turns kernel times into algorithmic GB/s
    and FLOP/s): K_e = (pixel, Gaussian) pairs whose exponent is evaluated by the forward, K_c = pairs that
    contribute (pass every test).  Counted by the library's instrumented forward (vcr_pair_counters), outside the timed
    regions."""
    import vcr_native as N
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    import diff_gaussian_rasterization as D
    p = stepper.bucket.params
    v = stepper.views[0]["inp"]
    dev = stepper.dev
    rs = GaussianRasterizationSettings(image_height=args.height, image_width=args.width, tanfovx=v["tanfovx"], tanfovy=v["tanfovy"],
                                       bg=v["bg"].to(dev), scale_modifier=1.0, viewmatrix=v["viewmatrix"].to(dev),
                                       projmatrix=v["projmatrix"].to(dev), sh_degree=3, campos=v["campos"].to(dev),
                                       prefiltered=False, debug=False, f_count=0)
    def render():
        with torch.no_grad():
            res = GaussianRasterizer(rs)(means3D=p["means3D"].detach(), means2D=None, opacities=p["opacities"].detach(),
                                         shs=p["shs"].detach(), normals_precomp=stepper.normals[0].detach(),
                                         scales=p["scales"].detach(), rotations=p["rotations"].detach(), dirs=stepper.dirs)
        torch.cuda.synchronize(dev)
        return res

    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    N.check(N.vcr_pair_counters(counters.data_ptr()), "pair_counters")
    try:
        out, radii = render()
    finally:
        N.check(N.vcr_pair_counters(None), "pair_counters")
    timeline = None
    if os.environ.get("VCR_BENCH_TIMELINE"):
        timeline = block_timeline(render, ((args.width + 15) // 16) * ((args.height + 15) // 16), dev, os.environ["VCR_BENCH_TIMELINE"])
    N.profile_read(reset=True)
    c = counters.cpu().tolist()
    P = int(radii.numel())
    Pv = int((radii > 0).sum())
    R = int(getattr(D, "_LAST_NUM_RENDERED", 0))
    return {"P": P, "P_visible": Pv, "R": R, "R_per_P": R / max(P, 1), "pixels": args.width * args.height,
            "tiles": ((args.width + 15) // 16) * ((args.height + 15) // 16),
            "sum_n_contrib": c[2], "K_e": c[0], "K_c": c[1], "warp_candidates": c[3],
            "forward_longest_block_cycles": c[4], "forward_sum_block_cycles": c[5],
            "mean_contributors_per_pixel": c[1] / max(args.width * args.height, 1), "forward_block_timeline": timeline}


def block_timeline(render, blocks, dev, path):
    """Start / end of every block of the forward compositing kernel (vcr_debug_block_timeline): is the kernel as long as its
    longest tile?  Two passes: the production kernel (times), then the instrumented one (candidates per warp).  Writes the raw
    tables to `path` (.npz) and returns the summary."""
    import numpy as np
    import vcr_native as N

    def run(instrumented):
        buf = torch.zeros(16 * blocks, dtype=torch.int64, device=dev)
        cnt = torch.zeros(8, dtype=torch.int64, device=dev)
        render()
        if instrumented:
            N.check(N.vcr_pair_counters(cnt.data_ptr()), "pair_counters")
        N.check(N.vcr_debug_block_timeline(buf.data_ptr()), "block_timeline")
        try:
            render()
        finally:
            N.check(N.vcr_debug_block_timeline(None), "block_timeline")
            N.check(N.vcr_pair_counters(None), "pair_counters")
        t = buf.cpu().numpy().reshape(blocks, 16)
        return t

    t, ti = run(False), run(True)
    np.savez(path, plain=t, instrumented=ti)
    t0 = int(t[:, 0].min())
    start, end = (t[:, 0] - t0) * 1e-3, (t[:, 1] - t0) * 1e-3        # us
    dur = end - start
    order = np.argsort(-dur)[:8]
    ends = np.sort(end)
    n = len(t)
    return {"kernel_span_us": float(end.max()), "blocks": int(n), "sum_block_us": float(dur.sum()),
            "all_blocks_started_by_us": float(start.max()),
            "blocks_done_at_us": {q: float(ends[min(n - 1, int(q * n))]) for q in (0.5, 0.9, 0.99, 0.999)},
            "longest_blocks": [{"block": int(b), "tile": int(t[b, 2] >> 32), "sm": int(t[b, 2] & 0xffffffff), "list": int(t[b, 3]),
                                "warp_candidates": [int(x) for x in ti[b, 4:8]] if b < len(ti) else None,
                                "start_us": float(start[b]), "end_us": float(end[b])} for b in order]}


def roofline_entry(dom, stats, args, hbm_peak, peak_src, clocks):
    """`roofline` of the dominant kernel.  The compositing kernels are bound by FP32 instruction issue, not by HBM (their DRAM
    traffic is below the algorithmic bytes: the per-instance re-reads hit the L2), so for them `bound` is "fp32_issue" and the
    HBM figures the contract asks for are kept beside the FP32 fraction: achieved = (14 K_e + 160 K_c) FLOP (45 K_c
    for the forward) per launch / launch time, peak = 2 x 128 lanes x #SM x SM clock."""
    traffic, issue = None, {}
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tr.get(args.config, {}).get(dom["kernel"]) if isinstance(tr.get(args.config), dict) else None
        same = (args.gaussians, args.width, args.height, args.footprint) == (BC.CONFIGS[args.config]["gaussians"], BC.CONFIGS[args.config]["width"], BC.CONFIGS[args.config]["height"], 2.0)
        if ent and same:
            traffic = int(ent["bytes"])
            issue = {k: ent[k] for k in ("issue_active_pct", "warp_instructions", "dram_pct_of_peak", "source") if k in ent}
    except Exception:
        pass
    compositing = dom["kernel"] in ("render_forward_kernel", "render_backward_kernel")
    entry = {"kernel": dom["kernel"], "bound": "fp32_issue" if compositing else "hbm", "achieved": dom["achieved_gbs"], "peak": hbm_peak,
             "unit": "GB/s", "frac": dom["achieved_gbs"] / hbm_peak, "traffic": traffic, "peak_source": peak_src,
             "algorithmic_bytes_per_launch": dom["bytes_per_launch"], "ncu": issue or None}
    if compositing:
        sms = torch.cuda.get_device_properties(0).multi_processor_count
        mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965
        fp32_peak = 2.0 * 128 * sms * mhz * 1e6 / 1e12
        flop = 14.0 * stats["K_e"] + (160.0 if dom["kernel"] == "render_backward_kernel" else 45.0) * stats["K_c"]
        ach = flop / (dom["avg_launch_ms"] * 1e-3) / 1e12
        entry["fp32"] = {"achieved_tflops": ach, "peak_tflops": fp32_peak, "frac": ach / fp32_peak,
                         "flop_per_launch": flop, "formula": "14*K_e + (160 bwd | 45 fwd)*K_c (SURVEY 8d)", "sm_mhz": mhz, "sms": sms}
        entry["note"] = ("compositing is FP32-issue bound; `achieved/peak/frac` are the HBM figures the contract asks for, `fp32` the "
                         "fraction of the CUDA-core peak the algorithmic flops reach (the flop model counts useful arithmetic only)")
    return entry


def kernel_table(kern, st, args, hbm_peak):
    """Per-kernel average launch time and algorithmic GB/s (byte formulas: SURVEY.md §8d / DESIGN.md)."""
    if not kern:
        return []
    P, Pv, R, N, tiles = st["P"], st["P_visible"], st["R"], st["pixels"], st["tiles"]
    npass_tile = (max(1, math.ceil(math.log2(max(tiles, 2)))) + 7) // 8
    alg = {
        "preprocess_kernel": 336 * Pv + 20 * (P - Pv),
        "scan_block_sums_kernel": 8 * ((P + 255) // 256),
        "duplicate_kernel": 20 * P + 12 * R,
        "radix_pass_kernel": (4 * 24 * P + npass_tile * 16 * R) / (4.0 + npass_tile),   # per launch: 4 passes over the P (u64, u32) depth pairs, the tile passes over R (u32, u32) pairs
        "tile_ranges_kernel": 8 * R + 8 * tiles,
        "render_forward_kernel": 41 * R + 60 * N,
        "render_backward_kernel": 77 * R + 68 * N + 84 * Pv,
        "preprocess_backward_kernel": (92 + 571) * Pv,
        "dnormal_forward_kernel": 29 * N,
        "dnormal_backward_kernel": 33 * N,
        "densify_stats_kernel": 16 * P + 24 * Pv,
    }
    rows = []
    for name, (ms, timed, launches) in kern.items():
        if timed == 0 or name not in alg:
            continue
        avg_ms = ms / timed
        b = alg[name]
        rows.append({"kernel": name, "launches_per_step": launches / max(args.steps, 1), "avg_launch_ms": avg_ms,
                     "ms_per_step": ms / max(args.steps, 1), "bytes_per_launch": b, "achieved_gbs": b / (avg_ms * 1e-3) / 1e9,
                     "frac_of_hbm_peak": b / (avg_ms * 1e-3) / 1e9 / hbm_peak})
    rows.sort(key=lambda r: -r["ms_per_step"])
    return rows


if __name__ == "__main__":
    main()
