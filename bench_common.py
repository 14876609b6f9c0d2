"""What both arms of bench.py share: the command line, the workload table (BASELINE.json `configs`), the seeded synthetic
scene / cameras / ground truth, clock sampling and the JSON line.  Nothing here imports the product package or the
reference: the scene generator (vcr-gaus_b200/vcr_synth.py, pure PyTorch on the CPU) is loaded by file path, so that the
reference arm's process never has the product package on sys.path and never maps libvcr_gaus_b200.so."""
import argparse
import importlib.util
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.abspath(__file__))

# Loss weights of the reference's reconstruction configs (configs/reconstruct.yaml:21-30, configs/tnt/base.yaml:9-14)
W_L1, W_SSIM = 0.8, 0.2
W_MONO = 0.01
W_DN = 0.015
EXP_T = 5e-3
# learning rates of configs/config_base.yaml:41-50 (position_lr_init * spatial_lr_scale with spatial_lr_scale = 1)
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20, opacity=0.05, scaling=5e-3, rotation=1e-3)
RAW_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
DENSIFY_INTERVAL = 100          # configs/config_base.yaml:53 densification_interval
PERCENT_DENSE = 0.01            # configs/config_base.yaml:52
MIN_OPACITY = 0.005             # trainer.py:370
DENSIFY_QUANTILE = 0.95         # see workload_note("C")

METRIC = "train views/s (fwd+bwd rasterize + D-Normal loss; 1M Gaussians, 1920x1080)"

# BASELINE.json `configs`.  step: "hotpath" = post-activation tensors -> rasterize -> L1 + mono-normal + D-Normal -> backward
# (the metric's step); "train" = one complete training iteration (raw parameters -> prologue -> rasterize -> L1 + SSIM +
# mono-normal + D-Normal -> backward -> densification statistics -> [densify every 100] -> Adam).
CONFIGS = {
    "metric": dict(gaussians=1_000_000, width=1920, height=1080, views=8, step="hotpath", densify=False,
                   what="the configuration BASELINE.json's metric is quoted on"),
    "A": dict(gaussians=10_000, width=256, height=256, views=1, step="hotpath", densify=False,
              what="config A: 10k Gaussians, one 256x256 view (the CPU-runnable case)"),
    "B": dict(gaussians=300_000, width=800, height=600, views=49, step="train", densify=False,
              what="config B: DTU scan24-shaped, 300k Gaussians, 49 views at 800x600, full VCR-GauS loss, 1 GPU"),
    "C": dict(gaussians=1_500_000, width=1920, height=1080, views=8, step="train", densify=True,
              what="config C: TnT Barn-shaped, 1.5M Gaussians, 1920x1080, densification on, view-parallel"),
    "D": dict(gaussians=5_000_000, width=1256, height=828, views=4, step="hotpath", densify=False,
              what="config D: Mip-NeRF 360 garden-shaped, 5M Gaussians, 1256x828, SH degree 3, view-parallel"),
    "E": dict(gaussians=1_000_000, width=1920, height=1080, views=4, step="raster", densify=False,
              what="config E: one point of the rasterizer sweep (--gaussians / --width / --height), rasterize fwd+bwd only, "
                   "upstream gradient N(0,1)"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--config", default="metric", choices=sorted(CONFIGS))
    ap.add_argument("--gaussians", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--views", type=int, default=None, help="distinct camera poses per rank, cycled")
    ap.add_argument("--footprint", type=float, default=2.0, help="scale multiplier of the synthetic scene")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="do not bracket kernels with CUDA events")
    ap.add_argument("--no-full-step", action="store_true", help="skip the side measurement of the complete training iteration")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    for k in ("gaussians", "width", "height", "views"):
        if getattr(args, k) is None:
            setattr(args, k, cfg[k])
    args.step_kind, args.densify = cfg["step"], cfg["densify"]
    return args


def workload(args):
    """`config.workload`: identical string in both arms."""
    kind = {"hotpath": "ray/plane-intersection depth, loss = L1 + mono-normal + D-Normal; rasterize fwd+bwd per view, densification statistics every view",
            "train": "complete training iteration per view: raw parameters -> activations/normals -> rasterize (ray/plane-intersection depth) "
                     "-> 0.8 L1 + 0.2 (1-SSIM) + mono-normal + D-Normal -> backward -> densification statistics -> Adam (6 groups, eps 1e-15)"
                     + (f", densify+prune every {DENSIFY_INTERVAL} steps (grad threshold = the {DENSIFY_QUANTILE:.2f} quantile of the accumulated statistic)" if args.densify else ""),
            "raster": "ray/plane-intersection depth; rasterize fwd+bwd only, upstream gradient N(0,1)"}[args.step_kind]
    return (f"{args.gaussians} Gaussians (SH degree 3), {args.width}x{args.height}, {args.views} views/rank cycled, {kind}")


def config_dict(args, world):
    return {"workload": workload(args), "name": args.config,
            "parallelism": f"view-parallel x{world} (replicated Gaussians, per-step SUM of the per-Gaussian gradients)" if world > 1 else "single GPU",
            "l2": "working set per step (parameters + gradients + per-view state) exceeds the 126 MB L2 for every config but A; views rotate every step"}


def _load_synth():
    name = "vcr_synth"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "vcr-gaus_b200", "vcr_synth.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def build_scene(args):
    S = _load_synth()
    scene = S.make_scene(args.gaussians, seed=0)
    scene.scales.mul_(args.footprint)
    return scene


def build_views(args, scene, world, view_ids, light=False):
    """Views `view_ids` of the ring of args.views * world cameras, on the CPU.  light=True: camera and rasterizer inputs only
    (enough to measure a view's cost), no ground-truth images."""
    S = _load_synth()
    views = []
    n_total = args.views * world
    for gid in view_ids:
        cam = S.make_camera(args.width, args.height, idx=gid, n_views=n_total)
        view = dict(id=gid, cam=cam, inp=S.raster_inputs(scene, cam, use_dirs=True, use_sem=False, seed=gid // world))
        if not light:
            g = torch.Generator().manual_seed(10_000 + gid)
            H, W = args.height, args.width
            # 8-bit ground truth, as image files hold it; the float image both arms train on is u8 / 255 (what the
            # reference's loader computes, utils/general_utils.py PILtoTorch)
            gt_u8 = torch.randint(0, 256, (3, H, W), generator=g, dtype=torch.uint8)
            gt_image = gt_u8.to(torch.float32) / 255.0
            gt_normal = F.normalize(torch.randn((H, W, 3), generator=g) * 0.4 + torch.tensor([0.0, 0.0, -1.0]), dim=-1)
            view.update(gt_image=gt_image.contiguous(), gt_image_u8=gt_u8.contiguous(), gt_normal=gt_normal.contiguous())
            if args.step_kind == "raster":
                view.update(dL_dout=torch.randn((10, H, W), generator=g).contiguous())
        views.append(view)
    return views


def raw_parameters(scene):
    """The raw (pre-activation) parameters of GaussianModel for the synthetic scene: _xyz, _features_dc, _features_rest,
    _opacity (logit), _scaling (log), _rotation."""
    shs = scene.shs
    return dict(xyz=scene.means3D.clone(), f_dc=shs[:, :1].contiguous(), f_rest=shs[:, 1:].contiguous(),
                opacity=torch.logit(scene.opacities.clamp(1e-4, 1 - 1e-4)), scaling=torch.log(scene.scales),
                rotation=scene.rotations.clone())


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


def init_distributed():
    """torch.distributed from the torchrun environment (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*).  (rank, world, local)."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def timed_region(step_fn, steps, warmup, dev, world, sampler=None, first=0):
    """W untimed + exactly K timed calls of step_fn(i), bracketed by barrier + synchronize, CUDA events on the current
    stream, MAX over ranks.  `sampler`: a ClockSampler built BEFORE the warm-up (NVML initialisation takes tens of
    milliseconds of host time; between warm-up and timing it would let the GPU fall idle and the first timed steps would run
    at ramping clocks)."""
    import torch.distributed as dist
    for i in range(warmup):
        step_fn(first + i)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev0.record()
    for i in range(steps):
        step_fn(first + warmup + i)
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


_JSON_FD = None


def claim_stdout():
    """The ONE JSON line goes to the process's real stdout; everything else written to fd 1 during the run (NCCL prints its
    version banner there) is redirected to stderr."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ViewFeeder:
    """The per-view inputs of one rank (ground-truth image, ground-truth normal map, camera block = view matrix | projection
    matrix | camera centre | background), as a data loader hands them over: either already resident on the device
    (`value`) or copied every step from pinned host memory on a side stream, one view ahead of the compute stream (`e2e`;
    the copies stay inside the timed region, they are only overlapped).  Used identically by both arms."""

    def __init__(self, views, dev, extra=(), u8_image=None):
        """u8_image: None, or a function (uint8 device tensor) -> float32 device tensor run on the current stream.  With it the
        ground-truth image crosses the bus as the 8-bit image it is (6.2 MB at 1080p instead of 24.9 MB) and becomes u8 / 255
        on the device, on the copy stream — checked here to be bit-identical to the float image of the other wire format."""
        self.dev, self.n, self.u8_image = dev, len(views), u8_image
        self.host, self.devv = [], []
        for v in views:
            inp = v["inp"]
            h = dict(gt_normal=v["gt_normal"].pin_memory(),
                     cam=torch.cat([inp["viewmatrix"].reshape(-1), inp["projmatrix"].reshape(-1), inp["campos"].reshape(-1),
                                    inp["bg"].reshape(-1)]).pin_memory())
            if u8_image:
                h["gt_image_u8"] = v["gt_image_u8"].pin_memory()
            else:
                h["gt_image"] = v["gt_image"].pin_memory()
            for name in extra:
                h[name] = v[name].pin_memory()
            self.host.append(h)
            d = {k: t.to(dev) for k, t in h.items()}
            if u8_image:
                d["gt_image"] = u8_image(d.pop("gt_image_u8"))
                assert torch.equal(d["gt_image"].cpu(), v["gt_image"]), "u8 / 255 on the device differs from the loader's float image"
            self.devv.append(d)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host[0].values())
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._prefetched = None

    def _issue(self, i):
        k = i % self.n
        with torch.cuda.stream(self.copy_stream):
            d = {name: t.to(self.dev, non_blocking=True) for name, t in self.host[k].items()}
            if self.u8_image:
                d["gt_image"] = self.u8_image(d.pop("gt_image_u8"))
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        for t in d.values():                      # keep the tensors alive on the compute stream's timeline
            t.record_stream(torch.cuda.current_stream())
        self._prefetched = (i, d, ev)

    def get(self, i, e2e):
        if not e2e:
            return self.devv[i % self.n]
        if self._prefetched is None or self._prefetched[0] != i:
            self._issue(i)
        _, d, ev = self._prefetched
        torch.cuda.current_stream().wait_event(ev)
        self._issue(i + 1)
        return d

    @staticmethod
    def camera(d):
        cam = d["cam"]
        return cam[0:16].view(4, 4), cam[16:32].view(4, 4), cam[32:35], cam[35:38]     # view, proj, campos, bg


def densify_threshold(xyz_gradient_accum, denom):
    """The gradient threshold of one densification round, identical in both arms: the DENSIFY_QUANTILE quantile of the
    accumulated statistic grads = xyz_gradient_accum / denom (NaN -> 0, as densify_and_prune forms it,
    scene/gaussian_model.py:646-647).  The reference's fixed 0.0005 (configs/config_base.yaml:56) is tuned to real
    photometric gradients; on random ground truth it would select nearly every Gaussian."""
    g = (xyz_gradient_accum / denom).reshape(-1)
    g = torch.where(torch.isnan(g), torch.zeros_like(g), g)
    k = max(1, min(g.numel(), int(DENSIFY_QUANTILE * g.numel())))
    return float(torch.kthvalue(g, k).values)
