#!/usr/bin/env python
"""bench.py — the hot path of VCR-GauS on B200: forward + backward rasterization + D-Normal loss, per view.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|reference-cpu]

Workload (BASELINE.json `metric`: "fwd+bwd rasterize ms/view (1M Gauss, 1080p); train views/s at 1/2/4/8 GPU"):
a seeded synthetic scene of 1,000,000 Gaussians (SH degree 3), 1920x1080 views with ray/plane-intersection depth
(`dirs`, the configured default of every reconstruction config, configs/reconstruct.yaml:10), loss = L1(RGB) +
mono-normal + confidence-weighted depth-normal (D-Normal), backward into every per-Gaussian parameter.
One "step" = one such view per GPU (fwd + loss + bwd); with N > 1 each rank renders a different view of the same
replicated scene and the flat gradient bucket + densification statistics are all-reduced with NCCL (weak scaling).
`value` = views/s over all ranks (so ms/view at N = 1 is `ms_per_step`).  SSIM, the per-Gaussian prologue of render() and
the Adam update are outside the hot path and are not part of a step in either arm; the complete
training iteration including them is timed beside it at N = 1 and reported as `full_step` (both arms).

Arms
  ours           this repository (GaussianRasterizer -> ctypes -> libvcr_gaus_b200.so, fused D-Normal kernels)
  reference      the UNMODIFIED reference CUDA extension (oracle/_ref, built from /root/reference by oracle/build_ref.sh)
                 driven through its `_C` module, with the reference's PyTorch D-Normal op chain — same scene, same
                 views, same loss, same GPU.  (The reference has no CPU implementation of this path, SURVEY.md §8c;
                 the arm the north star asks to beat is its CUDA rasterizer "on the same B200".)
  reference-cpu  the CPU oracle (oracle/vcr_oracle.c, a port) on all host cores; also timed, on a bounded sample,
                 as `cpu_baseline` inside the default arm at N = 1.
"""
import argparse
import gc
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "vcr-gaus_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402

LAMBDA_DN = 0.015   # loss_weight.depth_normal, configs/tnt/base.yaml:12
LAMBDA_MONO = 0.01  # loss_weight.mono_normal order of magnitude (configs/reconstruct.yaml)
EXP_T = 0.5
METRIC = "train views/s (fwd+bwd rasterize + D-Normal loss; 1M Gaussians, 1920x1080)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--gaussians", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--views", type=int, default=8, help="distinct camera poses per rank, cycled")
    ap.add_argument("--footprint", type=float, default=2.0, help="scale multiplier (R/P in [6,10] at the default size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="do not bracket kernels with CUDA events")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region, through NVML from the benchmarking thread itself.

    Polling from a side thread or from an `nvidia-smi -lms` process was tried first and dropped: NVML queries contend
    with this process's CUDA launches for the driver lock and inflated a 90 ms timed region by tens of milliseconds
    (up to 10x with 50 ms polling); even from this thread an uncached query costs 3-10 ms of host time (tools/nvml_probe.py).
    So one sample is taken per timed region: right after the last launch and the closing event have been enqueued, while
    the GPU is still executing the last timed step."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.names = {
                getattr(pynvml, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(pynvml, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            self.ok = True
        except Exception:  # pragma: no cover
            self.ok = False

    def sample(self):
        if not self.ok:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, name in self.names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------------------------
def build_scene(args):
    import vcr_synth as S
    scene = S.make_scene(args.gaussians, seed=0)
    scene.scales.mul_(args.footprint)
    return scene


def build_views(args, scene, world, view_ids, light=False):
    """Views `view_ids` of the ring of args.views * world cameras, on the CPU.  light=True: camera and rasterizer inputs only
    (enough to measure a view's cost), no ground-truth images."""
    import vcr_synth as S
    views = []
    n_total = args.views * world
    for gid in view_ids:
        cam = S.make_camera(args.width, args.height, idx=gid, n_views=n_total)
        view = dict(id=gid, cam=cam, inp=S.raster_inputs(scene, cam, use_dirs=True, use_sem=False, seed=gid // world))
        if not light:
            g = torch.Generator().manual_seed(10_000 + gid)
            H, W = args.height, args.width
            gt_image = torch.rand((3, H, W), generator=g)
            gt_normal = F.normalize(torch.randn((H, W, 3), generator=g) * 0.4 + torch.tensor([0.0, 0.0, -1.0]), dim=-1)
            view.update(gt_image=gt_image.contiguous(), gt_normal=gt_normal.contiguous())
        views.append(view)
    return views


def build_workload(args, rank, world):
    """Scene (identical on every rank) + this rank's default shard of the views (every world-th camera of the ring)."""
    scene = build_scene(args)
    return scene, build_views(args, scene, world, [v * world + rank for v in range(args.views)])


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


def ref_dnormal_chain(depth, K, gt_normal, render_normal, exp_t):
    """The reference's D-Normal chain (tools/normal_utils.py:24-41, tools/graphics_utils.py:111-131,
    tools/loss_utils.py:122-142, trainer.py:266-280), no mask (mask = all ones)."""
    _, H, W = depth.shape
    dev = depth.device
    vx = torch.arange(W, dtype=torch.float32, device=dev).add_(0.5) / (W - 1)
    vy = torch.arange(H, dtype=torch.float32, device=dev).add_(0.5) / (H - 1)
    vy, vx = torch.meshgrid(vy, vx, indexing="ij")
    ndc = torch.stack([vx, vy, depth[0]], dim=-1)
    inv_scale = torch.tensor([[W - 1, H - 1]], device=dev)
    cam_z = ndc[..., 2:3]
    cam = torch.cat([ndc[..., :2] * inv_scale * cam_z, cam_z], dim=-1) @ torch.inverse(K.t())
    dy = torch.gradient(cam, dim=0)[0]
    dx = torch.gradient(cam, dim=1)[0]
    est = F.normalize(torch.cross(dx, dy, dim=-1), p=2, dim=-1)
    rn = render_normal.detach()
    w = torch.exp((torch.sum(rn * gt_normal, dim=-1) - 1) / exp_t).detach()
    mask = torch.ones((H, W), dtype=torch.bool, device=dev)
    e, g, ww = est[mask], gt_normal[mask], w[mask]
    return (ww * torch.abs(e - g).sum(dim=-1)).mean() + (ww * (1.0 - torch.sum(e * g, dim=-1))).mean()


class RefRasterFn(torch.autograd.Function):
    """Autograd shim over the reference `_C` module, argument tuples as in DGR/diff_gaussian_rasterization/__init__.py:186-231,
    393-458 (the reference's Python wrapper is not copied; this builds the same calls)."""

    @staticmethod
    def forward(ctx, C, means3D, shs, normals, opac, scales, rots, dirs, inside, st):
        e = torch.Tensor([])
        R, out, radii, geom, binning, img = C.rasterize_gaussians(
            st["bg"], means3D, e, normals, e, opac, scales, rots, 1.0, e, st["view"], st["proj"], dirs, st["tanx"], st["tany"],
            st["H"], st["W"], shs, st["deg"], st["campos"], False, inside, False)
        ctx.C, ctx.st, ctx.R = C, st, R
        ctx.save_for_backward(means3D, shs, normals, scales, rots, dirs, inside, radii, geom, binning, img, out)
        return out, radii

    @staticmethod
    def backward(ctx, g, _):
        means3D, shs, normals, scales, rots, dirs, inside, radii, geom, binning, img, out = ctx.saved_tensors
        st, e = ctx.st, torch.Tensor([])
        (g2d, g2dd, gcol, gnrm, gobj, gop, g3d, gcov, gsh, gsc, grot) = ctx.C.rasterize_gaussians_backward(
            st["bg"], means3D, radii, e, normals, e, scales, rots, 1.0, e, st["view"], st["proj"], dirs, st["tanx"], st["tany"],
            g.contiguous(), shs, st["deg"], st["campos"], geom, ctx.R, binning, img, out[7:8], out[3:4], inside, False)
        return None, g3d, gsh, gnrm, gop, gsc, grot, None, None, None


def ref_prologue_chain(xyz, fdc, frest, o_raw, s_raw, r_raw, cam_center, R_w2c):
    """What the reference's render() does per Gaussian before the rasterizer call (scene/gaussian_model.py:125-192,
    tools/general_utils.py:98-119, gaussian_renderer/__init__.py:95-101; is_all=True), as the same PyTorch op chain."""
    opacity, scaling, rotation = torch.sigmoid(o_raw), torch.exp(s_raw), F.normalize(r_raw)
    shs = torch.cat((fdc, frest), dim=1)
    r = rotation
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=q.device)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - w * z); R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y); R[:, 2, 1] = 2 * (y * z + w * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    axis = torch.argmin(scaling, dim=-1)
    normal = R.gather(2, axis[:, None, None].expand(-1, 3, -1)).squeeze(-1)
    view_dir = xyz - cam_center
    normal = normal * ((((view_dir * normal).sum(dim=-1) > 0) * 1 - 0.5) * 2)[..., None]
    return opacity, scaling, rotation, shs, normal @ R_w2c.transpose(0, 1)


def time_prologue(args, scene, view, dev, iters=20):
    """render()'s per-Gaussian prologue, forward + backward: this library's fused kernels vs the reference's PyTorch chain.
    Not part of the step (both arms of the bench start from post-activation tensors); reported beside it."""
    from vcr_prologue import gaussian_prologue
    shs = scene.shs.to(dev)
    leaves = [t.requires_grad_(True) for t in (shs[:, :1].clone(), shs[:, 1:].clone(),
                                               torch.logit(scene.opacities.to(dev).clamp(1e-4, 1 - 1e-4)),
                                               torch.log(scene.scales.to(dev)), scene.rotations.to(dev) * 1.3)]
    xyz = scene.means3D.to(dev)
    cc = view["inp"]["campos"].to(dev)
    M = view["inp"]["viewmatrix"].to(dev)[:3, :3].t().contiguous()
    out = {}
    cot = None
    for name, fn in (("fused_ms", gaussian_prologue), ("pytorch_chain_ms", ref_prologue_chain)):
        for k in range(iters + 3):
            if k == 3:
                torch.cuda.synchronize(dev)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            res = fn(xyz, *leaves, cc, M)
            if cot is None:
                cot = [torch.randn_like(t) for t in res]
            torch.autograd.grad(res, leaves, cot)
        e1.record()
        torch.cuda.synchronize(dev)
        out[name] = e0.elapsed_time(e1) / iters
    out["what"] = f"forward+backward of render()'s per-Gaussian prologue, P = {xyz.shape[0]}; not part of the step"
    return out


def ref_ssim_chain(img1, img2):
    """tools/loss_utils.py:47-92 restated (11x11 Gaussian window, grouped conv2d), for the reference arm of `full_step`."""
    gauss = torch.Tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g1 = (gauss / gauss.sum()).unsqueeze(1)
    window = g1.mm(g1.t()).float().unsqueeze(0).unsqueeze(0).expand(3, 1, 11, 11).contiguous().to(img1.device)
    mu1, mu2 = F.conv2d(img1, window, padding=5, groups=3), F.conv2d(img2, window, padding=5, groups=3)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, window, padding=5, groups=3) - mu1_sq
    s2 = F.conv2d(img2 * img2, window, padding=5, groups=3) - mu2_sq
    s12 = F.conv2d(img1 * img2, window, padding=5, groups=3) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()


LAMBDA_DSSIM = 0.2      # configs/reconstruct.yaml lambda_dssim
RAW_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
RAW_LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20, opacity=0.05, scaling=5e-3, rotation=1e-3)


def time_full_step(args, impl, scene, stepper, dev, steps):
    """One COMPLETE training iteration per view, from the raw parameters to the updated raw parameters, the way trainer.py
    runs it: per-Gaussian prologue -> rasterize -> loss (L1 + SSIM + mono-normal + D-Normal) -> backward -> Adam.
    `ours`: fused prologue, fused view loss + fused SSIM, BucketAdam.  `reference`: the reference's PyTorch chains around its
    CUDA rasterizer and torch.optim.Adam(eps=1e-15) with its six groups.  Reported beside the bench metric (whose step is
    the rasterizer + D-Normal hot path only), single GPU."""
    import vcr_view_parallel as VP
    shs = scene.shs.to(dev)
    raw = dict(xyz=scene.means3D.to(dev), f_dc=shs[:, :1].contiguous(), f_rest=shs[:, 1:].contiguous(),
               opacity=torch.logit(scene.opacities.to(dev).clamp(1e-4, 1 - 1e-4)), scaling=torch.log(scene.scales.to(dev)),
               rotation=scene.rotations.to(dev).clone())
    H, W = args.height, args.width
    if impl == "ours":
        from vcr_dnormal import ssim, view_loss
        from vcr_prologue import gaussian_prologue
        bucket = VP.GaussianBucket(raw, device=dev, names=RAW_NAMES, direct=True)   # kernels write the gradients into the bucket
        p, opt = bucket.params, VP.BucketAdam(bucket, RAW_LRS)
    else:
        p = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
        opt = torch.optim.Adam([{"params": [p[k]], "lr": RAW_LRS[k], "name": k} for k in RAW_NAMES], lr=0.0, eps=1e-15)

    def one(i):
        k = i % len(stepper.views)
        d = stepper.devv[k]
        cam = d["cam"]
        view, proj, campos, bg = cam[0:16].view(4, 4), cam[16:32].view(4, 4), cam[32:35], cam[35:38]
        R_w2c = view[:3, :3].t()
        gt_n = d["gt_normal"]
        if impl == "ours":
            opac, scales, rots, sh, normals = gaussian_prologue(p["xyz"], p["f_dc"], p["f_rest"], p["opacity"], p["scaling"], p["rotation"],
                                                                campos, R_w2c)
            rs = stepper.GRS(image_height=H, image_width=W, tanfovx=stepper.tanx, tanfovy=stepper.tany, bg=bg, scale_modifier=1.0,
                             viewmatrix=view, projmatrix=proj, sh_degree=3, campos=campos, prefiltered=False, debug=False, f_count=0)
            means2D = torch.zeros_like(p["xyz"], requires_grad=True)
            out, radii = stepper.GR(rs)(means3D=p["xyz"], means2D=means2D, means2D_densify=torch.zeros_like(p["xyz"], requires_grad=True),
                                        shs=sh, normals_precomp=normals, opacities=opac, scales=scales, rotations=rots,
                                        dirs=stepper.dirs, inside=stepper.inside)
            loss = view_loss(out, d["gt_image"], gt_n, stepper.K4, w_l1=1.0 - LAMBDA_DSSIM, w_mono=LAMBDA_MONO, w_dn=LAMBDA_DN, exp_t=EXP_T)
            loss = loss + LAMBDA_DSSIM * (1.0 - ssim(out[:3], d["gt_image"]))
            bucket.zero_grad()
            loss.backward()
            opt.step()
        else:
            opac, scales, rots, sh, normals = ref_prologue_chain(p["xyz"], p["f_dc"], p["f_rest"], p["opacity"], p["scaling"], p["rotation"],
                                                                 campos, R_w2c)
            st = dict(bg=bg, view=view, proj=proj, campos=campos, tanx=stepper.tanx, tany=stepper.tany, H=H, W=W, deg=3)
            out, radii = RefRasterFn.apply(stepper.C, p["xyz"], sh, normals, opac, scales, rots, stepper.dirs, stepper.inside, st)
            image, depth, normal, _alpha, _sem = out.split([3, 1, 3, 1, 2], dim=0)
            rn = F.normalize(normal.permute(1, 2, 0), dim=-1)
            loss = (1.0 - LAMBDA_DSSIM) * (image - d["gt_image"]).abs().mean() + LAMBDA_DSSIM * (1.0 - ref_ssim_chain(image[None], d["gt_image"][None]))
            loss = loss + LAMBDA_MONO * ((rn - gt_n).abs().sum(-1).mean() + (1.0 - (rn * gt_n).sum(-1)).mean())
            loss = loss + LAMBDA_DN * ref_dnormal_chain(depth, stepper.K, gt_n, rn, EXP_T)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()

    for i in range(max(3 * len(stepper.views), 24)):
        one(i)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        one(i)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    return {"ms_per_step": ms, "views_per_s": 1e3 / ms, "steps": steps,
            "includes": "raw parameters -> prologue -> rasterize fwd -> L1 + SSIM + mono-normal + D-Normal -> backward -> Adam (6 groups, eps 1e-15)"}


def clone_params(scene, dev, direct=False, vp=None):
    """Parameters + flat gradient bucket.  With a ViewParallel layer at N > 1 the bucket is placed in symmetric memory so
    that the gradient exchange can go through the NVSwitch multicast kernel (VCR_BENCH_TRANSPORT=nccl forces NCCL)."""
    import vcr_view_parallel as VP
    tensors = dict(means3D=scene.means3D.to(dev), shs=scene.shs.to(dev), opacities=scene.opacities.to(dev),
                   scales=scene.scales.to(dev), rotations=scene.rotations.to(dev))
    store = None
    if vp is not None and vp.world > 1 and direct and os.environ.get("VCR_BENCH_TRANSPORT", "multimem") == "multimem":
        # (symmetric storage is also fine for NCCL, so it is allocated whenever the multicast kernel may be chosen)
        store = vp.symmetric_gradient_storage(vp.exchange_floats(tensors), dev)
    b = VP.GaussianBucket(tensors, device=dev, direct=direct, grad_storage=store)
    b.transport = "multimem" if store is not None else "nccl"
    b.compact = bool(vp is not None and vp.world > 1 and direct and os.environ.get("VCR_BENCH_COMPACT", "1") == "1")
    if b.compact:
        vp.compact_sh_exchange(b, sh_degree=3)     # exchange 3 instead of 48 floats per Gaussian for the SH coefficients
    return b


class Stepper:
    """One view: (optional H2D of the per-view inputs) -> forward -> loss -> backward -> (N>1: all-reduce) -> loss readback."""

    def __init__(self, args, impl, bucket, views, dev, vp):
        self.args, self.impl, self.bucket, self.dev, self.vp = args, impl, bucket, dev, vp
        self.views = views
        P = bucket.P
        self.inside = torch.ones(P, dtype=torch.bool, device=dev)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=dev)
        self.denom = torch.zeros((P, 1), device=dev)
        self.max_radii2D = torch.zeros((P,), device=dev)
        # per-view host (pinned) and device copies of everything that changes with the view
        self.host, self.devv = [], []
        for v in views:
            inp = v["inp"]
            h = dict(gt_image=v["gt_image"].pin_memory(), gt_normal=v["gt_normal"].pin_memory(),
                     cam=torch.cat([inp["viewmatrix"].reshape(-1), inp["projmatrix"].reshape(-1), inp["campos"].reshape(-1),
                                    inp["bg"].reshape(-1)]).pin_memory())
            self.host.append(h)
            self.devv.append({k: t.to(dev) for k, t in h.items()})
        inp0 = views[0]["inp"]
        # all views share intrinsics and size, so `dirs` is per scene, like scene.dirs (scene/__init__.py:99-102)
        self.dirs = inp0["dirs"].to(dev)
        self.K = views[0]["cam"].intr.to(dev)
        self.K_host = views[0]["cam"].intr
        self.K4 = (float(self.K_host[0, 0]), float(self.K_host[1, 1]), float(self.K_host[0, 2]), float(self.K_host[1, 2]))
        self.tanx, self.tany = inp0["tanfovx"], inp0["tanfovy"]
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host[0].values())
        self.d2h_bytes = 4
        self.loss_host = torch.zeros(1).pin_memory()
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._prefetched = None
        self.diag_events = None
        self.skip = os.environ.get("VCR_BENCH_SKIP", "")   # diagnosis only: "grads", "stats" or "grads,stats"
        # N > 1: pipeline the gradient all-reduce with the per-Gaussian backward stage in this many ranges (0 = one
        # all-reduce after the backward)
        self.chunks, self.exchange_ms = 0, {}
        if impl == "ours" and vp.world > 1:
            self.exchange_ms = vp.pick_transport(bucket)          # NCCL or the multicast kernel, whichever is faster here
            forced = os.environ.get("VCR_BENCH_TRANSPORT")
            if forced in ("nccl", "multimem") and (forced == "nccl" or bucket.transport == "multimem"):
                vp.transport = forced
            # pipelining pays with the multicast kernel (a quarter of the SMs, 8 GPUs: -0.04 ms/step), not with NCCL; the
            # compact exchange is too small to be worth pipelining
            self.chunks = 0 if bucket.compact else int(os.environ.get("VCR_BENCH_CHUNKS", "4" if vp.transport == "multimem" else "0"))
        if self.chunks > 0 and "grads" not in self.skip:
            vp.overlap_gradients(bucket, n_chunks=self.chunks, transport=vp.transport)
        if impl == "reference":
            import ref_ab
            self.C = ref_ab.load()
            if self.C is None:
                raise RuntimeError("oracle/_ref reference extension is not available")
        else:
            from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
            self.GRS, self.GR = GaussianRasterizationSettings, GaussianRasterizer
        # normals_precomp depends on the view (flip towards the camera + w2c rotation): precomputed per view on device
        self.normals = [v["inp"]["normals"].to(dev).requires_grad_(True) for v in views]

    def _issue_prefetch(self, i):
        k = i % len(self.views)
        with torch.cuda.stream(self.copy_stream):
            d = {name: t.to(self.dev, non_blocking=True) for name, t in self.host[k].items()}
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        # keep the tensors alive on the compute stream's timeline
        for t in d.values():
            t.record_stream(torch.cuda.current_stream())
        self._prefetched = (i, d, ev)

    def step(self, i, e2e):
        k = i % len(self.views)
        p = self.bucket.params
        if e2e:
            # per-view inputs come from pinned host memory every step; like a data loader with prefetch depth 1 the
            # copy of view i+1 is issued on a side stream while view i is being rendered (the copies stay inside the
            # timed region, they are just overlapped)
            if self._prefetched is None or self._prefetched[0] != i:
                self._issue_prefetch(i)
            _, d, ev = self._prefetched
            torch.cuda.current_stream().wait_event(ev)
            self._issue_prefetch(i + 1)
        else:
            d = self.devv[k]
        cam = d["cam"]
        view, proj, campos, bg = cam[0:16].view(4, 4), cam[16:32].view(4, 4), cam[32:35], cam[35:38]
        H, W = self.args.height, self.args.width
        normals = self.normals[k]
        if self.impl == "reference":
            st = dict(bg=bg, view=view, proj=proj, campos=campos, tanx=self.tanx, tany=self.tany, H=H, W=W, deg=3)
            out, radii = RefRasterFn.apply(self.C, p["means3D"], p["shs"], normals, p["opacities"], p["scales"], p["rotations"],
                                           self.dirs, self.inside, st)
            means2D_densify = None
        else:
            rs = self.GRS(image_height=H, image_width=W, tanfovx=self.tanx, tanfovy=self.tany, bg=bg, scale_modifier=1.0,
                          viewmatrix=view, projmatrix=proj, sh_degree=3, campos=campos, prefiltered=False, debug=False, f_count=0)
            means2D = torch.zeros_like(p["means3D"], requires_grad=True)
            means2D_densify = torch.zeros_like(p["means3D"], requires_grad=True)
            out, radii = self.GR(rs)(means3D=p["means3D"], means2D=means2D, means2D_densify=means2D_densify, shs=p["shs"],
                                     normals_precomp=normals, opacities=p["opacities"], scales=p["scales"],
                                     rotations=p["rotations"], dirs=self.dirs, inside=self.inside)
        gt_n = d["gt_normal"]
        if self.impl == "reference":
            # channel split exactly like render() (gaussian_renderer/__init__.py:122-123, 133-134), losses as in trainer.py
            image, depth, normal, _alpha, _sem = out.split([3, 1, 3, 1, 2], dim=0)
            rn = F.normalize(normal.permute(1, 2, 0), dim=-1)
            loss = (image - d["gt_image"]).abs().mean()
            loss = loss + LAMBDA_MONO * ((rn - gt_n).abs().sum(-1).mean() + (1.0 - (rn * gt_n).sum(-1)).mean())
            loss = loss + LAMBDA_DN * ref_dnormal_chain(depth, self.K, gt_n, rn, EXP_T)
        else:
            # the same three terms through this repository's public loss API (one fused kernel per direction).  Mean over
            # ranks = SUM all-reduce of the gradients of loss / world: the 1/world goes into the loss weights, so there is
            # neither a division pass over the bucket nor an extra scaling kernel
            from vcr_dnormal import view_loss
            sc = 1.0 / self.vp.world
            loss = view_loss(out, d["gt_image"], gt_n, self.K4, w_l1=sc, w_mono=LAMBDA_MONO * sc, w_dn=LAMBDA_DN * sc, exp_t=EXP_T)
        self.bucket.zero_grad()
        normals.grad = None
        ev = self.diag_events
        if ev is not None:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev.append(e)
            e[0].record()
        if self.impl == "reference" and self.vp.world > 1:
            (loss / self.vp.world).backward()
        else:
            loss.backward()     # ours, N > 1: the chunk hook queues the gradient exchange range by range during this call
        if ev is not None:
            e[1].record()
        skip = self.skip
        if self.vp.world > 1 and "grads" not in skip:
            if getattr(self.bucket, "compact", False):
                self.vp.exchange_compact(self.bucket, campos)
            elif self.chunks > 0:
                self.vp.finish_gradients(self.bucket)
            else:
                self.vp.allreduce_gradients(self.bucket, average=False)
        if ev is not None:
            e[2].record()
        if means2D_densify is not None and "stats" not in skip:
            # densification statistics (trainer.py:345-347), every view at every N: one fused local pass; the ranks are
            # combined every 100 steps (the reference densifies every 100 iterations, trainer.py:357-370)
            self.vp.accumulate_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D,
                                                   means2D_densify.grad, radii)
            if (i + 1) % 100 == 0:
                self.vp.sync_densification_stats(self.xyz_gradient_accum, self.denom, self.max_radii2D)
                self.xyz_gradient_accum.zero_(); self.denom.zero_(); self.max_radii2D.zero_()
        if e2e:
            self.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        return loss


def timed_region(stepper, steps, warmup, e2e, dev, world, sampler=None):
    """`sampler`: a ClockSampler built BEFORE the warm-up (NVML initialisation takes tens of milliseconds of host time; done
    between the warm-up and the first timed step it lets the GPU fall idle and the first timed steps run at ramping clocks)."""
    for i in range(warmup):
        stepper.step(i, e2e)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev0.record()
    for i in range(steps):
        stepper.step(warmup + i, e2e)
    ev1.record()
    if sampler:
        sampler.sample()          # the GPU is still executing the last timed step(s); ev1 is already enqueued, so the
                                  # 3-10 ms an (uncached) NVML query can take on the host does not enter the measurement
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.result() if sampler else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    return ms, clocks


def cpu_oracle_views_per_s(args, view, repeat=1):
    """The CPU oracle (a port; the reference has no CPU path) on one full view of the same workload, all host cores."""
    from oracle import oracle as O
    from util import np_inputs
    import numpy as np
    inp = np_inputs(view["inp"])
    g = np.random.default_rng(0).standard_normal((10, args.height, args.width)).astype(np.float32)
    t0 = time.time()
    for _ in range(repeat):
        f = O.rasterize_forward(inp)
        O.rasterize_backward(inp, f, g)
    dt = (time.time() - t0) / repeat
    return 1.0 / dt, dt, f["binned"]["R"]


_JSON_FD = None


def emit(obj):
    """The ONE JSON line goes to the process's real stdout; everything else written to fd 1 during the run (NCCL prints its
    version banner there) has been redirected to stderr."""
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    global _JSON_FD
    args = parse()
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    import vcr_view_parallel as VP
    rank, world, local = VP.init_from_env()
    if args.impl == "reference-cpu":
        if rank == 0:
            scene, views = build_workload(args, 0, 1)
            vps, dt, R = cpu_oracle_views_per_s(args, views[0])
            cores = os.cpu_count()
            emit(({
                "impl": "reference", "metric": METRIC, "value": vps, "unit": "views/s", "n_gpus": 0, "steps": 1, "warmup": 0,
                "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": {"workload": f"{args.gaussians} Gaussians, {args.width}x{args.height}, SH3, dirs; CPU oracle port"},
                "cpu_baseline": {"value": vps, "unit": "views/s", "cores": cores, "kind": "port",
                                 "sample": "one full view, rasterize fwd+bwd (no D-Normal), OpenMP over all host cores"},
                "e2e": {"value": vps, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback in the product path)"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    vp = VP.ViewParallel()
    all_cores = os.sched_getaffinity(0)
    numa_cores = VP.bind_to_gpu_numa_node(local) if os.environ.get("VCR_BENCH_NUMA", "1") == "1" else 0
    scene = build_scene(args)
    impl = args.impl
    if impl == "reference":
        import ref_ab
        if ref_ab.load() is None:
            if rank == 0:
                emit(({"impl": "reference", "unavailable": "oracle/_ref/_C.so (reference CUDA build) not present on this box"}))
            return
    bucket = clone_params(scene, dev, direct=(impl == "ours"), vp=vp)
    view_ids = [v * world + rank for v in range(args.views)]          # default shard: every world-th camera of the ring
    balanced = world > 1 and impl == "ours" and os.environ.get("VCR_BENCH_BALANCE", "1") == "1"
    if balanced:
        # Cost-balanced sharding: every step ends with the gradient exchange, i.e. with a wait for the slowest rank, so the
        # `world` views of one step should cost the same.  Each rank times its default shard once, the costs are
        # gathered, and ViewParallel.balanced_view_table deals views of equal cost to the ranks of one step.
        mine = torch.tensor(measure_view_costs(args, bucket, build_views(args, scene, world, view_ids, light=True), dev), device=dev)
        allc = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allc, mine)
        costs = [0.0] * (args.views * world)
        for r in range(world):
            for v in range(args.views):
                costs[v * world + r] = float(allc[r][v])
        view_ids = [row[rank] for row in vp.balanced_view_table(costs, seed=0)]
    views = build_views(args, scene, world, view_ids)
    stepper = Stepper(args, impl, bucket, views, dev, vp)

    import vcr_native as N
    prof = (impl == "ours") and not args.no_profile
    # ---- device-resident timing (`value`): K steps, launch counters on, per-kernel event brackets OFF ----
    if impl == "ours":
        N.vcr_profile_enable(0)
    # warm-up: W steps, but at least three passes over the view set, so that the caching allocator has seen every
    # per-view buffer size (the binning buffer depends on R) and no cudaMalloc lands in the timed region
    # ... and long enough (a fixed count, identical on every rank) for clocks and power state to settle under load
    sampler = ClockSampler(dev.index or 0)
    # no cyclic garbage collection inside the measured regions: a generation-2 pass on ONE rank stalls that rank's launches
    # for milliseconds and, through the per-step exchange, every other rank with it (the step creates no reference cycles)
    gc.collect()
    gc.disable()
    # (VCR_BENCH_WARMUP_FLOOR: profiling runs under ncu lower the floor; bench numbers are always taken with the default)
    floor = int(os.environ.get("VCR_BENCH_WARMUP_FLOOR", "200" if impl == "ours" else "60"))
    n_warm = max(args.warmup, 3 * len(views), floor)
    for i in range(n_warm):
        stepper.step(i, False)
    if impl == "ours":
        N.profile_read(reset=True)
    ms, clocks = timed_region(stepper, args.steps, 0, False, dev, world, sampler=sampler)
    counts = N.profile_read(reset=True) if impl == "ours" else {}
    # ---- the same K steps once more with every kernel of the library bracketed by CUDA events on its stream ----
    # (the brackets cost ~5-10 % of a step, so they are kept out of `value`; `ms_per_step_with_kernel_events` reports it)
    kern, ms_prof = {}, None
    if prof:
        N.vcr_profile_enable(1)
        ms_prof, _ = timed_region(stepper, args.steps, 1, False, dev, world)
        kern = N.profile_read(reset=True)
        N.vcr_profile_enable(0)
    # ---- end-to-end timing (`e2e`): per-view inputs from pinned host memory, loss read back ----
    # (its own warm-up: three passes over the views in e2e mode, so that the caching allocator has seen the prefetch buffers —
    # they are freed late because of record_stream — and no cudaMalloc lands in the timed region)
    ms_e2e, _ = timed_region(stepper, args.steps, max(args.warmup, 3 * len(views)), True, dev, world)
    if os.environ.get("VCR_BENCH_DIAG"):
        diagnose(stepper, args, dev, rank, world)
    views_per_s = world * args.steps / (ms / 1e3)
    e2e_vps = world * args.steps / (ms_e2e / 1e3)

    if rank != 0:
        return
    out = {
        "metric": METRIC, "value": views_per_s, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{args.gaussians} Gaussians (SH degree 3), {args.width}x{args.height}, {args.views} views/rank cycled, "
                               f"ray/plane-intersection depth, loss = L1 + mono-normal + D-Normal; rasterize fwd+bwd per view",
                   "parallelism": (f"view-parallel x{world} (replicated Gaussians, all-reduce of the flat gradient bucket via "
                                   + ("own NVSwitch-multicast kernel" if vp.transport == "multimem" else "NCCL")
                                   + (", SH gradient exchanged as 3 colour-gradient floats per Gaussian and rebuilt per rank" if getattr(bucket, "compact", False) else "")
                                   + (f", pipelined with the backward in {stepper.chunks} Gaussian ranges)" if stepper.chunks else ")"))
                   if world > 1 else "single GPU",
                   "l2": "working set per step (parameters + gradients + per-view state > 600 MB) exceeds the 126 MB L2; views rotate every step",
                   "not_in_step": "SSIM, Adam, per-Gaussian prologue (outside the hot path, SURVEY.md 8f; see full_step)",
                   "host": f"process bound to the {numa_cores} cores next to its GPU" if numa_cores else "no CPU binding"},
        "clocks": clocks,
        "exchange": ({"chosen": vp.transport, "ms_per_exchange": stepper.exchange_ms, "compact_sh": bool(getattr(bucket, "compact", False)), "view_sharding": "cost-balanced (balanced_view_table)" if balanced else "every world-th camera"}
                     if world > 1 and impl == "ours" else None),
        "e2e": {"value": e2e_vps, "unit": "views/s", "h2d_bytes_per_step": stepper.h2d_bytes, "d2h_bytes_per_step": stepper.d2h_bytes,
                "ms_per_step": ms_e2e / args.steps},
    }
    if world == 1 and not os.environ.get("VCR_BENCH_NO_FULL_STEP"):
        try:
            out["full_step"] = time_full_step(args, impl, scene, stepper, dev, args.steps)
        except Exception as ex:  # a side measurement must never cost the bench line
            out["full_step"] = {"error": f"{type(ex).__name__}: {ex}"}
    if impl == "reference":
        out["impl"] = "reference"
        out["config"]["reference_build"] = "unmodified reference sources, nvcc 12.9, compute_90 PTX JIT-compiled for sm_100 (see DESIGN.md)"
        out["cpu_baseline"] = {"value": None, "unit": "views/s", "cores": 0, "kind": "reference",
                               "sample": "not a CPU run: the reference's implementation of this path is CUDA-only; this arm is its CUDA rasterizer on this GPU"}
        out["gpu_launches"] = 0
        emit(out)
        return
    # ---- roofline of the dominant kernel + per-kernel table ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    stats = scene_stats(stepper, args)
    per_kernel = kernel_table(kern, stats, args, hbm_peak)
    launches = sum(v[2] for k, v in counts.items() if k != "memset")
    out["gpu_launches"] = launches
    if per_kernel:
        dom = max(per_kernel, key=lambda r: r["ms_per_step"])
        issue = {}
        traffic = None      # DRAM bytes per launch of that kernel, from the committed ncu --set full capture of this workload
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            same = (args.gaussians, args.width, args.height, args.footprint) == (1_000_000, 1920, 1080, 2.0)
            traffic = int(tr[dom["kernel"]]["bytes"]) if same else None
            issue = {k: tr[dom["kernel"]][k] for k in ("issue_active_pct", "warp_instructions", "dram_pct_of_peak") if k in tr[dom["kernel"]]}
        except Exception:
            pass
        out["roofline"] = {"kernel": dom["kernel"], "bound": "hbm", "achieved": dom["achieved_gbs"], "peak": hbm_peak, "unit": "GB/s",
                           "frac": dom["achieved_gbs"] / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": dom["bytes_per_launch"],
                           "note": "compositing kernels are FP32-issue / atomic bound, not HBM bound (DESIGN.md); the HBM fraction is reported as the contract asks",
                           "issue_bound": dict(issue, source=tr[dom["kernel"]]["source"]) if issue else None}
        out["kernels"] = per_kernel
        out["ms_per_step_with_kernel_events"] = ms_prof / args.steps if ms_prof else None
    out["scene"] = stats
    if world == 1:
        try:
            out["render_prologue"] = time_prologue(args, scene, views[0], dev)
        except Exception as ex:  # never lose the bench line over the side measurement
            out["render_prologue"] = {"error": f"{type(ex).__name__}: {ex}"}
    if not args.no_cpu_baseline and world == 1:
        os.sched_setaffinity(0, all_cores)        # the CPU baseline gets every host core, not just the GPU's NUMA node
        vps, dt, R = cpu_oracle_views_per_s(args, views[0])
        out["cpu_baseline"] = {"value": vps, "unit": "views/s", "cores": os.cpu_count(), "kind": "port",
                               "sample": f"one full {args.width}x{args.height} view of the same scene, rasterize fwd+bwd, CPU oracle with OpenMP ({dt:.1f} s)"}
    emit(out)


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
        stepper.step(1000 + i, False)
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


def scene_stats(stepper, args):
    """P, visible P, R for view 0 (needed to turn kernel times into algorithmic GB/s)."""
    p = stepper.bucket.params
    v = stepper.views[0]["inp"]
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = stepper.dev
    rs = GaussianRasterizationSettings(image_height=args.height, image_width=args.width, tanfovx=v["tanfovx"], tanfovy=v["tanfovy"],
                                       bg=v["bg"].to(dev), scale_modifier=1.0, viewmatrix=v["viewmatrix"].to(dev),
                                       projmatrix=v["projmatrix"].to(dev), sh_degree=3, campos=v["campos"].to(dev),
                                       prefiltered=False, debug=False, f_count=0)
    with torch.no_grad():
        out, radii = GaussianRasterizer(rs)(means3D=p["means3D"].detach(), means2D=None, opacities=p["opacities"].detach(),
                                            shs=p["shs"].detach(), normals_precomp=stepper.normals[0].detach(),
                                            scales=p["scales"].detach(), rotations=p["rotations"].detach(), dirs=stepper.dirs)
    import vcr_native as N
    N.profile_read(reset=True)
    P = int(radii.numel())
    Pv = int((radii > 0).sum())
    # R: re-derive from radii/rects is not needed — the autograd node is not kept under no_grad; use the tile formula
    R = int(last_num_rendered())
    return {"P": P, "P_visible": Pv, "R": R, "R_per_P": R / max(P, 1), "pixels": args.width * args.height,
            "tiles": ((args.width + 15) // 16) * ((args.height + 15) // 16)}


def last_num_rendered():
    import diff_gaussian_rasterization as D
    return getattr(D, "_LAST_NUM_RENDERED", 0)


def kernel_table(kern, st, args, hbm_peak):
    """Per-kernel average launch time and algorithmic GB/s (byte formulas: SURVEY.md §8d / DESIGN.md)."""
    if not kern:
        return []
    P, Pv, R, N, tiles = st["P"], st["P_visible"], st["R"], st["pixels"], st["tiles"]
    bits = 32 + max(1, math.ceil(math.log2(tiles + 1)))
    alg = {
        "preprocess_kernel": 336 * Pv + 20 * (P - Pv),
        "scan_block_sums_kernel": 8 * ((P + 255) // 256),
        "duplicate_kernel": 20 * P + 12 * R,
        "radix_pass_kernel": (4 * 24 * P + 2 * 24 * R) / 6.0,      # per launch: 4 passes over the P depth keys, 2 over the R instances
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
