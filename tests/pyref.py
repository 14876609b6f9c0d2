"""The reference's own Python for the path — tools/normal_utils.py, tools/loss_utils.py, tools/general_utils.py,
scene/gaussian_model.py, gaussian_renderer/__init__.py — loaded from the verbatim copies oracle/build_ref.sh stages under
oracle/_ref/pyref/ (git-ignored; they travel to the GPU box, where /root/reference does not exist).

The staged directories go to the END of sys.path: in a test process `diff_gaussian_rasterization` and `simple_knn` must
keep resolving to THIS repository's packages, so the reference's render() and GaussianModel run on the library under test
(the drop-in check), while its loss / normal utilities serve as the pin for the fused kernels.  Test infrastructure only."""
import importlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYREF = os.path.join(ROOT, "oracle", "_ref", "pyref")


def available():
    return os.path.exists(os.path.join(PYREF, "tools", "loss_utils.py")) and os.path.exists(os.path.join(PYREF, "scene", "gaussian_model.py"))


_ns = None


def load():
    """Namespace with the reference's functions, or None when oracle/_ref/pyref has not been staged."""
    global _ns
    if _ns is not None:
        return _ns
    if not available():
        return None
    for p in (PYREF, os.path.join(PYREF, "_stubs")):
        if p not in sys.path:
            sys.path.append(p)
    lu = importlib.import_module("tools.loss_utils")
    nu = importlib.import_module("tools.normal_utils")
    gu = importlib.import_module("tools.general_utils")
    gr = importlib.import_module("tools.graphics_utils")
    assert os.path.realpath(lu.__file__).startswith(os.path.realpath(PYREF)), lu.__file__
    _ns = types.SimpleNamespace(l1_loss=lu.l1_loss, ssim=lu.ssim, monosdf_normal_loss=lu.monosdf_normal_loss, cos_weight=lu.cos_weight,
                                compute_normals=nu.compute_normals, build_rotation=gu.build_rotation, get_all_px_dir=gr.get_all_px_dir,
                                getIntrinsic=gr.getIntrinsic, loss_utils=lu, normal_utils=nu, general_utils=gu, graphics_utils=gr)
    return _ns


def gaussian_model():
    """scene.gaussian_model.GaussianModel of the reference (imports this repository's simple_knn when it is on the path)."""
    load()
    return importlib.import_module("scene.gaussian_model").GaussianModel


def render_fn():
    """gaussian_renderer.render of the reference, bound to whatever `diff_gaussian_rasterization` is first on sys.path."""
    load()
    return importlib.import_module("gaussian_renderer").render


def dnormal_loss(R, est_normal, gt_normal, render_normal, mask, exp_t):
    """The 'depth_normal' branch of Trainer._compute_loss (trainer.py:266-280) with the reference's own functions."""
    import torch
    render_normal = render_normal.detach()
    with torch.no_grad():
        weights = R.cos_weight(render_normal, gt_normal, exp_t)
    if mask.sum() != 0:
        est_normal, gt_normal = est_normal[mask], gt_normal[mask]
        weights = weights[mask]
        return R.monosdf_normal_loss(est_normal, gt_normal, weights)
    return est_normal.sum() * 0


def render_bound(rasterizer_module, tag):
    """A private copy of the reference's gaussian_renderer module whose `from diff_gaussian_rasterization import ...` is
    resolved against `rasterizer_module` (this repository's package, or the reference's own wrapper from
    ref_ab.wrapper()).  Returns its render()."""
    import importlib.util
    load()
    path = os.path.join(PYREF, "gaussian_renderer", "__init__.py")
    spec = importlib.util.spec_from_file_location(f"vcr_pyref_gaussian_renderer_{tag}", path)
    mod = importlib.util.module_from_spec(spec)
    saved = sys.modules.get("diff_gaussian_rasterization")
    sys.modules["diff_gaussian_rasterization"] = rasterizer_module
    try:
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop("diff_gaussian_rasterization", None)
        else:
            sys.modules["diff_gaussian_rasterization"] = saved
    return mod.render


def make_model(raw, device, optim=False):
    """A reference GaussianModel holding the raw parameters `raw` (keys xyz, f_dc, f_rest, opacity, scaling, rotation) as
    nn.Parameters on `device`, SH degree 3 active; with optim=True also its own training_setup() (six-group Adam)."""
    import torch
    from torch import nn
    GaussianModel = gaussian_model()
    m = GaussianModel(types.SimpleNamespace(sh_degree=3, max_mem=1e9, use_decoupled_appearance=False, enable_semantic=False))
    mk = lambda t: nn.Parameter(t.detach().clone().to(device).requires_grad_(True))  # noqa: E731
    m._xyz, m._features_dc, m._features_rest = mk(raw["xyz"]), mk(raw["f_dc"]), mk(raw["f_rest"])
    m._opacity, m._scaling, m._rotation = mk(raw["opacity"]), mk(raw["scaling"]), mk(raw["rotation"])
    m.active_sh_degree = 3
    m.max_radii2D = torch.zeros((m.get_xyz.shape[0]), device=device)
    m.spatial_lr_scale = 1.0
    if optim:
        m.training_setup(types.SimpleNamespace(percent_dense=0.01, position_lr_init=1.6e-4, position_lr_final=1.6e-6, position_lr_delay_mult=0.01,
                                               position_lr_max_steps=30000, feature_lr=2.5e-3, opacity_lr=0.05, scaling_lr=5e-3, rotation_lr=1e-3))
    return m


def render_cfg():
    """The cfg fields render() reads (configs/config_base.yaml pipline.*, optim.loss_weight.semantic, optim.mask_depth_thr)."""
    return types.SimpleNamespace(pipline=types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False),
                                 optim=types.SimpleNamespace(loss_weight=types.SimpleNamespace(semantic=0), mask_depth_thr=0),
                                 model=types.SimpleNamespace(ch_sem_feat=2))


def camera_namespace(cam, device):
    """What render() reads from a scene.cameras.Camera, from a vcr_synth.Camera."""
    import numpy as np
    return types.SimpleNamespace(FoVx=cam.FoVx, FoVy=cam.FoVy, image_height=cam.image_height, image_width=cam.image_width,
                                 world_view_transform=cam.world_view_transform.to(device), full_proj_transform=cam.full_proj_transform.to(device),
                                 camera_center=cam.camera_center.to(device), R=np.asarray(cam.R.numpy()), intr=cam.intr.to(device))
