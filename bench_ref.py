"""bench.py --impl reference: the reference's own code for the path, on the same GPU, scene, views and loss.

Everything that computes here is the reference's: its CUDA extension (oracle/_ref/diff_gaussian_rasterization/_C.so, built
from /root/reference by oracle/build_ref.sh: sm_100 SASS; the only deviation from the sources as they lie there are the two
`GLM_FUNC_QUALIFIER`s a temporary copy of the vendored GLM headers gets, without which nvcc 12.9 deletes the colour stores —
DESIGN.md §5), driven through its own Python wrapper `diff_gaussian_rasterization/__init__.py`; its `GaussianModel`
(scene/gaussian_model.py: activations, get_normal, the six-group Adam, densify_and_prune), its `render()`
(gaussian_renderer/__init__.py) and its loss / normal utilities (tools/loss_utils.py, tools/normal_utils.py), all staged
verbatim under oracle/_ref/pyref/.  What this file adds is the loop around them — the loss assembly of
Trainer._compute_loss / _get_total_loss and the statistics updates of Trainer.train_step, whose class cannot be
instantiated without a dataset (cited line by line below) — and, at N > 1, an NCCL all-reduce of the gradients through
the reference's own tools/distributed.py helper (the reference itself never trains on more than one GPU).

No module of the product package is imported and libvcr_gaus_b200.so is never mapped into this process."""
import gc
import os
import sys
import types

import torch
import torch.nn.functional as F

import bench_common as BC

REF_DIR = os.path.join(BC.ROOT, "oracle", "_ref")
REFERENCE_BUILD = ("reference CUDA sources compiled as they lie under /root/reference with nvcc 12.9 for sm_100 (SASS, no PTX); "
                   "the two min/max functors of the vendored GLM 1.0.2 carry GLM_FUNC_QUALIFIER in a temporary header copy "
                   "(without it nvcc deletes the SH colour stores, DESIGN.md 5); Python wrapper, GaussianModel, render() and "
                   "loss utilities are the reference's own files, staged unmodified under oracle/_ref/")


def available():
    need = [os.path.join(REF_DIR, "diff_gaussian_rasterization", "_C.so"), os.path.join(REF_DIR, "diff_gaussian_rasterization", "__init__.py"),
            os.path.join(REF_DIR, "pyref", "tools", "loss_utils.py"), os.path.join(REF_DIR, "pyref", "scene", "gaussian_model.py"),
            os.path.join(REF_DIR, "pyref", "gaussian_renderer", "__init__.py")]
    missing = [p for p in need if not os.path.exists(p)]
    return (not missing), missing


def _import_reference():
    for p in (os.path.join(REF_DIR, "pyref", "_stubs"), os.path.join(REF_DIR, "pyref"), REF_DIR):
        if p not in sys.path:
            sys.path.insert(0, p)
    bad = [p for p in sys.path if p.rstrip("/").endswith("vcr-gaus_b200")]
    assert not bad, "the product package must not be importable from the reference arm"
    import diff_gaussian_rasterization as DGR                         # the reference's wrapper + _C
    from gaussian_renderer import render                              # gaussian_renderer/__init__.py:22-150
    from scene.gaussian_model import GaussianModel                    # scene/gaussian_model.py
    from tools import distributed as rdist                            # tools/distributed.py:110-152
    from tools.loss_utils import cos_weight, l1_loss, monosdf_normal_loss, ssim   # tools/loss_utils.py:36-142
    from tools.normal_utils import compute_normals                    # tools/normal_utils.py:30-41
    assert os.path.realpath(DGR.__file__).startswith(os.path.realpath(REF_DIR)), DGR.__file__
    return types.SimpleNamespace(DGR=DGR, render=render, GaussianModel=GaussianModel, rdist=rdist, cos_weight=cos_weight,
                                 l1_loss=l1_loss, monosdf_normal_loss=monosdf_normal_loss, ssim=ssim, compute_normals=compute_normals)


def _depth_normal_term(R, est_normal, gt_normal, render_normal, mask):
    """Trainer._compute_loss, the 'depth_normal' branch (trainer.py:266-280), statement by statement."""
    render_normal = render_normal.detach()
    with torch.no_grad():
        weights = R.cos_weight(render_normal, gt_normal, BC.EXP_T)
    if mask.sum() != 0:
        est_normal, gt_normal = est_normal[mask], gt_normal[mask]
        render_normal = render_normal[mask]
        weights = weights[mask]
        return R.monosdf_normal_loss(est_normal, gt_normal, weights)
    return 0


class HotPathStepper:
    """The metric's step on the reference: post-activation leaf tensors -> GaussianRasterizer (reference wrapper) -> channel
    split and normalisation of render() (gaussian_renderer/__init__.py:122-134) -> compute_normals -> L1 + mono-normal +
    D-Normal -> backward -> (N > 1: all-reduce) -> densification statistics (trainer.py:345-347)."""

    def __init__(self, args, R, scene, views, dev, world):
        self.args, self.R, self.dev, self.world = args, R, dev, world
        self.feed = BC.ViewFeeder(views, dev, extra=("dL_dout",) if args.step_kind == "raster" else ())
        t = dict(means3D=scene.means3D, shs=scene.shs, opacities=scene.opacities, scales=scene.scales, rotations=scene.rotations)
        self.p = {k: v.to(dev).clone().requires_grad_(True) for k, v in t.items()}      # plain leaf tensors
        P = scene.means3D.shape[0]
        # the three statistics tensors of GaussianModel (training_setup, scene/gaussian_model.py:238-239; create_from_pcd :225)
        self.stats = types.SimpleNamespace(xyz_gradient_accum=torch.zeros((P, 1), device=dev), denom=torch.zeros((P, 1), device=dev),
                                           max_radii2D=torch.zeros((P,), device=dev))
        inp0 = views[0]["inp"]
        self.dirs = inp0["dirs"].to(dev)
        self.K = views[0]["cam"].intr.to(dev)
        self.tanx, self.tany = inp0["tanfovx"], inp0["tanfovy"]
        self.normals = [v["inp"]["normals"].to(dev).requires_grad_(True) for v in views]
        self.mask = torch.ones((args.height, args.width), dtype=torch.bool, device=dev)   # render(): no camera mask, mask_depth_thr = 0
        self.loss_host = torch.zeros(1).pin_memory()
        self.d2h_bytes = 4

    def step(self, i, e2e=False):
        R, a, p = self.R, self.args, self.p
        d = self.feed.get(i, e2e)
        view, proj, campos, bg = BC.ViewFeeder.camera(d)
        k = i % self.feed.n
        # render(): the two screen-space carriers (gaussian_renderer/__init__.py:31-37)
        sp = torch.zeros_like(p["means3D"], dtype=p["means3D"].dtype, requires_grad=True, device="cuda") + 0
        spd = torch.zeros_like(p["means3D"], dtype=p["means3D"].dtype, requires_grad=True, device="cuda") + 0
        sp.retain_grad(); spd.retain_grad()
        rs = R.DGR.GaussianRasterizationSettings(image_height=a.height, image_width=a.width, tanfovx=self.tanx, tanfovy=self.tany, bg=bg,
                                                 scale_modifier=1.0, viewmatrix=view, projmatrix=proj, sh_degree=3, campos=campos,
                                                 prefiltered=False, debug=False, f_count=0)
        out, radii = R.DGR.GaussianRasterizer(raster_settings=rs)(
            means3D=p["means3D"], means2D=sp, means2D_densify=spd, shs=p["shs"], colors_precomp=None, normals_precomp=self.normals[k],
            semantics_precomp=None, opacities=p["opacities"], scales=p["scales"], rotations=p["rotations"], cov3D_precomp=None,
            dirs=self.dirs, inside=None)
        for t in p.values():
            t.grad = None                      # optimizer.zero_grad(set_to_none=True), trainer.py:389
        self.normals[k].grad = None
        if a.step_kind == "raster":
            out.backward(d["dL_dout"])
            loss = out[0, 0, 0].detach()
        else:
            chs = [3, 1, 3, 1]
            image, depth, normal, _alpha = out[:sum(chs)].split(chs, dim=0)              # render() :122-123
            normal = F.normalize(normal.permute(1, 2, 0), dim=-1)                        # :133-134
            est_normal = R.compute_normals(depth, self.K)                                # :136
            gt_n = d["gt_normal"]
            loss = R.l1_loss(image, d["gt_image"])                                       # trainer.py:236
            loss = loss + BC.W_MONO * R.monosdf_normal_loss(normal, gt_n)                # trainer.py:261-264
            loss = loss + BC.W_DN * _depth_normal_term(R, est_normal, gt_n, normal, self.mask)
            if self.world > 1:
                (loss / self.world).backward()
            else:
                loss.backward()
        if self.world > 1:
            for t in p.values():
                R.rdist.dist_all_reduce_tensor(t.grad, reduce="sum")
        if a.step_kind != "raster":
            with torch.no_grad():
                vis, st = radii > 0, self.stats
                st.max_radii2D[vis] = torch.max(st.max_radii2D[vis], radii[vis])          # trainer.py:345
                R.GaussianModel.add_densification_stats(st, spd, vis)                    # scene/gaussian_model.py:669-671
        if e2e:
            self.loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        return loss


class TrainStepper:
    """One complete training iteration on the reference: GaussianModel (raw parameters, its own Adam) -> render() -> the loss
    terms of Trainer._compute_loss (l1, ssim, mono_normal, depth_normal with the weights of configs/reconstruct.yaml +
    configs/tnt/base.yaml) summed as in _get_total_loss -> backward -> statistics -> densify_and_prune every 100
    iterations -> optimizer.step() / zero_grad (trainer.py:323-392)."""

    def __init__(self, args, R, scene, views, dev, world):
        import numpy as np
        from torch import nn
        self.args, self.R, self.dev, self.world = args, R, dev, world
        self.feed = BC.ViewFeeder(views, dev)
        raw = BC.raw_parameters(scene)
        mcfg = types.SimpleNamespace(sh_degree=3, max_mem=1e9, use_decoupled_appearance=False, enable_semantic=False)
        m = R.GaussianModel(mcfg)
        m._xyz = nn.Parameter(raw["xyz"].to(dev).requires_grad_(True))
        m._features_dc = nn.Parameter(raw["f_dc"].to(dev).requires_grad_(True))
        m._features_rest = nn.Parameter(raw["f_rest"].to(dev).requires_grad_(True))
        m._scaling = nn.Parameter(raw["scaling"].to(dev).requires_grad_(True))
        m._rotation = nn.Parameter(raw["rotation"].to(dev).requires_grad_(True))
        m._opacity = nn.Parameter(raw["opacity"].to(dev).requires_grad_(True))
        m.max_radii2D = torch.zeros((m.get_xyz.shape[0]), device="cuda")
        m.active_sh_degree = 3
        m.spatial_lr_scale = 1.0
        m.extent = scene.extent
        opt = types.SimpleNamespace(percent_dense=BC.PERCENT_DENSE, position_lr_init=BC.LRS["xyz"], position_lr_final=1.6e-6,
                                    position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=BC.LRS["f_dc"],
                                    opacity_lr=BC.LRS["opacity"], scaling_lr=BC.LRS["scaling"], rotation_lr=BC.LRS["rotation"])
        m.training_setup(opt)
        self.model = m
        lw = types.SimpleNamespace(semantic=0)
        self.cfg = types.SimpleNamespace(pipline=types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False),
                                         optim=types.SimpleNamespace(loss_weight=lw, mask_depth_thr=0), model=types.SimpleNamespace(ch_sem_feat=2))
        self.weights = dict(l1=BC.W_L1, ssim=BC.W_SSIM, mono_normal=BC.W_MONO, depth_normal=BC.W_DN)
        inp0 = views[0]["inp"]
        self.dirs = inp0["dirs"].to(dev)
        c0 = views[0]["cam"]
        # a Camera as render() reads it (scene/cameras.py): FoV, size, the three matrices, R as numpy (c2w rotation), intr
        self.cams = []
        for v in views:
            c = v["cam"]
            self.cams.append(types.SimpleNamespace(FoVx=c.FoVx, FoVy=c.FoVy, image_height=c.image_height, image_width=c.image_width,
                                                   R=np.asarray(c.R.numpy()), intr=c.intr.to(dev)))
        self.extent = scene.extent
        self.loss_host = torch.zeros(1).pin_memory()
        self.d2h_bytes = 4
        self.iteration = 0
        self.P_history = [int(m.get_xyz.shape[0])]

    def step(self, i, e2e=False):
        R, m = self.R, self.model
        self.iteration += 1
        d = self.feed.get(i, e2e)
        view, proj, campos, bg = BC.ViewFeeder.camera(d)
        cam = self.cams[i % self.feed.n]
        cam.world_view_transform, cam.full_proj_transform, cam.camera_center = view, proj, campos
        pkg = R.render(cam, m, self.cfg, bg, dirs=self.dirs)                             # trainer.py:226
        gt_image, gt_normal = d["gt_image"], d["gt_normal"]
        losses = {}
        losses["l1"] = R.l1_loss(pkg["render"], gt_image)                                # trainer.py:236
        losses["ssim"] = 1.0 - R.ssim(pkg["render"], gt_image)                           # :238
        losses["mono_normal"] = R.monosdf_normal_loss(pkg["normal"], gt_normal)          # :261-264
        losses["depth_normal"] = _depth_normal_term(R, pkg["est_normal"], gt_normal, pkg["normal"], pkg["mask"])   # :266-280
        total = torch.tensor(0., device=torch.device("cuda"))                            # _get_total_loss, :308-318
        for name in self.weights:
            if name in losses:
                total += losses[name] * self.weights[name]
        if self.world > 1:
            (total / self.world).backward()
            for g in m.optimizer.param_groups:
                R.rdist.dist_all_reduce_tensor(g["params"][0].grad, reduce="sum")
        else:
            total.backward()
        vis, radii = pkg["visibility_filter"], pkg["radii"]
        with torch.no_grad():
            m.max_radii2D[vis] = torch.max(m.max_radii2D[vis], radii[vis])               # trainer.py:345
            m.add_densification_stats(pkg["viewspace_points_densify"], vis)              # :346-347
            if self.args.densify and self.iteration % BC.DENSIFY_INTERVAL == 0:          # :357-370
                if self.world > 1:
                    R.rdist.dist_all_reduce_tensor(m.xyz_gradient_accum, reduce="sum")
                    R.rdist.dist_all_reduce_tensor(m.denom, reduce="sum")
                    torch.distributed.all_reduce(m.max_radii2D, op=torch.distributed.ReduceOp.MAX)
                thr = BC.densify_threshold(m.xyz_gradient_accum, m.denom)
                m.densify_and_prune(thr, BC.MIN_OPACITY, self.extent, None, None)
                self.P_history.append(int(m.get_xyz.shape[0]))
            m.optimizer.step()                                                           # :388-389
            m.optimizer.zero_grad(set_to_none=True)
        if e2e:
            self.loss_host.copy_(total.detach().reshape(1), non_blocking=True)
        return total


def main(args):
    ok, missing = available()
    rank, world, local = BC.init_distributed()
    if not ok:
        if rank == 0:
            BC.emit({"impl": "reference", "unavailable": "oracle/_ref is incomplete on this box: " + ", ".join(os.path.relpath(p, BC.ROOT) for p in missing)})
        return
    assert torch.cuda.is_available(), "the reference's implementation of this path is CUDA-only"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    R = _import_reference()
    scene = BC.build_scene(args)
    view_ids = [v * world + rank for v in range(args.views)]
    views = BC.build_views(args, scene, world, view_ids)
    stepper = (TrainStepper if args.step_kind == "train" else HotPathStepper)(args, R, scene, views, dev, world)
    sampler = BC.ClockSampler(dev.index or 0)
    gc.collect()
    gc.disable()
    floor = int(os.environ.get("VCR_BENCH_WARMUP_FLOOR", "200"))
    n_warm = max(args.warmup, 3 * len(views), floor)
    if args.densify:
        n_warm = max(args.warmup, 3 * len(views))        # a densifying run is timed from a fixed iteration count in both arms
    ms, clocks = BC.timed_region(lambda i: stepper.step(i, False), args.steps, n_warm, dev, world, sampler=sampler)
    ms_e2e, _ = BC.timed_region(lambda i: stepper.step(i, True), args.steps, max(args.warmup, 3 * len(views)), dev, world,
                                first=n_warm + args.steps)
    if rank != 0:
        return
    vps, e2e_vps = world * args.steps / (ms / 1e3), world * args.steps / (ms_e2e / 1e3)
    out = {"impl": "reference", "metric": BC.METRIC, "value": vps, "unit": "views/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": BC.config_dict(args, world), "clocks": clocks,
           "e2e": {"value": e2e_vps, "unit": "views/s", "h2d_bytes_per_step": stepper.feed.h2d_bytes, "d2h_bytes_per_step": stepper.d2h_bytes,
                   "ms_per_step": ms_e2e / args.steps},
           "reference_build": REFERENCE_BUILD,
           "cpu_baseline": {"value": None, "unit": "views/s", "cores": 0, "kind": "reference",
                            "sample": "not a CPU run: the reference's implementation of this path is CUDA-only; this arm is its CUDA rasterizer on this GPU"},
           "gpu_launches": 0}
    if hasattr(stepper, "P_history"):
        out["gaussians_after_each_densification"] = stepper.P_history
    BC.emit(out)
