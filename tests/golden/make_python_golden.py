#!/usr/bin/env python
"""Golden vectors from the reference's own Python (CPU, this container): run once, commit the .npz files.

    python tests/golden/make_python_golden.py

Imports tools.normal_utils / tools.loss_utils / tools.general_utils and scene.gaussian_model of the reference — from the
verbatim copies oracle/build_ref.sh stages under oracle/_ref/pyref (the script stages them from /root/reference first when
that exists) — and writes

  pyref_dnormal.npz   compute_normals, cos_weight, monosdf_normal_loss: est_normal, the D-Normal loss of
                      Trainer._compute_loss and its gradient w.r.t. the depth map, float64 and float32, three cases
                      (confidence on/off, empty mask)
  pyref_image.npz     l1_loss, monosdf_normal_loss(normalize(render normal), gt), ssim and their gradients
  pyref_prologue.npz  GaussianModel.get_opacity / get_scaling / get_rotation / get_features / get_normal(is_all=True) and
                      the flip + rotation of render() (gaussian_renderer/__init__.py:95-101), with the gradients of a
                      random cotangent w.r.t. the raw parameters

The reference hard-codes device='cuda' in build_rotation and get_normal; for this CPU run torch.zeros / torch.ones are
wrapped to drop that keyword (no arithmetic changes)."""
import contextlib
import os
import subprocess
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
if os.path.isdir("/root/reference"):
    subprocess.check_call(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], env=dict(os.environ, STAGE_ONLY="1"))
import pyref  # noqa: E402

R = pyref.load()
assert R is not None, "oracle/_ref/pyref is missing"


@contextlib.contextmanager
def cpu_device():
    z, o = torch.zeros, torch.ones

    def strip(fn):
        def wrapped(*a, **k):
            if k.get("device") in ("cuda", torch.device("cuda")):
                k = dict(k, device="cpu")
            return fn(*a, **k)
        return wrapped
    torch.zeros, torch.ones = strip(z), strip(o)
    try:
        yield
    finally:
        torch.zeros, torch.ones = z, o


def dn_case(W, H, seed, exp_t, empty_mask=False):
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, H), torch.linspace(-1, 1, W), indexing="ij")
    depth = 2.0 + 0.5 * torch.sin(3 * xx) * torch.cos(2 * yy) + 0.3 * xx + 0.02 * torch.randn((H, W), generator=g)
    gt = F.normalize(torch.randn((H, W, 3), generator=g) * 0.3 + torch.tensor([0.0, 0.0, -1.0]), dim=-1)
    rn = (gt + 0.2 * torch.randn((H, W, 3), generator=g)) * (0.2 + torch.rand((H, W, 1), generator=g))
    mask = torch.rand((H, W), generator=g) > (2.0 if empty_mask else 0.2)
    K = torch.tensor([[0.9 * W, 0, W / 2], [0, 0.95 * W, H / 2], [0, 0, 1]], dtype=torch.float32)
    out = dict(depth=depth.numpy(), gt_normal=gt.numpy(), render_normal=rn.numpy(), mask=mask.numpy(), K=K.numpy(), exp_t=np.float64(exp_t))
    for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
        d = depth[None].to(dt).requires_grad_(True)
        est = R.compute_normals(d, K.to(dt))                                       # tools/normal_utils.py:30-41
        rnn = F.normalize(rn.to(dt), dim=-1)                                       # gaussian_renderer/__init__.py:133-134
        loss = pyref.dnormal_loss(R, est, gt.to(dt), rnn, mask, exp_t)             # trainer.py:266-280
        loss.backward()
        out[f"est_normal_{name}"] = est.detach().numpy()
        out[f"loss_{name}"] = np.asarray(loss.detach().numpy())
        out[f"dL_ddepth_{name}"] = d.grad[0].numpy()
    return out


def image_case(W, H, seed):
    g = torch.Generator().manual_seed(seed)
    img = torch.rand((3, H, W), generator=g)
    gt = (img + 0.1 * torch.randn((3, H, W), generator=g)).clamp(0, 1)
    nrm = torch.randn((H, W, 3), generator=g)
    gtn = F.normalize(torch.randn((H, W, 3), generator=g), dim=-1)
    out = dict(image=img.numpy(), gt_image=gt.numpy(), render_normal=nrm.numpy(), gt_normal=gtn.numpy())
    for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
        a = img.to(dt).requires_grad_(True)
        n = nrm.to(dt).requires_grad_(True)
        l1 = R.l1_loss(a, gt.to(dt))                                               # tools/loss_utils.py:36-37
        s = R.ssim(a, gt.to(dt))                                                   # tools/loss_utils.py:54-92
        mono = R.monosdf_normal_loss(F.normalize(n, dim=-1), gtn.to(dt))           # trainer.py:261-264
        gl1, = torch.autograd.grad(l1, a, retain_graph=True)
        gs, = torch.autograd.grad(s, a)
        gm, = torch.autograd.grad(mono, n)
        out.update({f"l1_{name}": l1.detach().numpy(), f"ssim_{name}": s.detach().numpy(), f"mono_{name}": mono.detach().numpy(),
                    f"dl1_{name}": gl1.numpy(), f"dssim_{name}": gs.numpy(), f"dmono_{name}": gm.numpy()})
    return out


def prologue_case(P, seed):
    from torch import nn
    GaussianModel = pyref.gaussian_model()
    g = torch.Generator().manual_seed(seed)
    raw = dict(xyz=torch.randn((P, 3), generator=g), f_dc=torch.randn((P, 1, 3), generator=g), f_rest=0.1 * torch.randn((P, 15, 3), generator=g),
               opacity=2 * torch.randn((P, 1), generator=g), scaling=torch.randn((P, 3), generator=g) - 3, rotation=torch.randn((P, 4), generator=g) * 1.7)
    cam_center = torch.tensor([0.3, -2.0, 1.1])
    A = torch.randn((3, 3), generator=g)
    Q, _ = torch.linalg.qr(A)
    cam_R = Q.numpy()                                                              # viewpoint_camera.R (c2w rotation)
    cot = {k: torch.randn(s, generator=g) for k, s in dict(opacity=(P, 1), scales=(P, 3), rotations=(P, 4), shs=(P, 16, 3), normals=(P, 3)).items()}
    out = {f"raw_{k}": v.numpy() for k, v in raw.items()}
    out.update(cam_center=cam_center.numpy(), cam_R=cam_R, **{f"cot_{k}": v.numpy() for k, v in cot.items()})
    for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
        m = GaussianModel(types.SimpleNamespace(sh_degree=3, max_mem=1e9, use_decoupled_appearance=False, enable_semantic=False))
        m._xyz = nn.Parameter(raw["xyz"].to(dt)); m._features_dc = nn.Parameter(raw["f_dc"].to(dt)); m._features_rest = nn.Parameter(raw["f_rest"].to(dt))
        m._opacity = nn.Parameter(raw["opacity"].to(dt)); m._scaling = nn.Parameter(raw["scaling"].to(dt)); m._rotation = nn.Parameter(raw["rotation"].to(dt))
        with cpu_device():
            opacity, scales, rotations, shs = m.get_opacity, m.get_scaling, m.get_rotation, m.get_features
            normal = m.get_normal(is_all=True)                                      # scene/gaussian_model.py:168-192
            # render(), gaussian_renderer/__init__.py:95-101
            view_dir = m.get_xyz - cam_center.to(dt)
            normal = normal.to(dt) * ((((view_dir * normal).sum(dim=-1) > 0) * 1 - 0.5) * 2)[..., None]
            R_w2c = torch.tensor(cam_R.T).to(dt)
            normals_precomp = normal @ R_w2c.transpose(0, 1)
        res = dict(opacity=opacity, scales=scales, rotations=rotations, shs=shs, normals=normals_precomp)
        total = sum((res[k] * cot[k].to(dt)).sum() for k in res)
        total.backward()
        out.update({f"{k}_{name}": v.detach().numpy() for k, v in res.items()})
        for k, prm in (("f_dc", m._features_dc), ("f_rest", m._features_rest), ("opacity", m._opacity), ("scaling", m._scaling), ("rotation", m._rotation)):
            out[f"grad_{k}_{name}"] = prm.grad.numpy()
    return out


def main():
    dn = {}
    for tag, c in (("a", dn_case(37, 29, 1, 0.5)), ("b", dn_case(64, 48, 2, 0.0)), ("c", dn_case(33, 2, 3, 1.0)), ("d", dn_case(21, 17, 4, 0.005, empty_mask=True))):
        dn.update({f"{tag}_{k}": v for k, v in c.items()})
    np.savez_compressed(os.path.join(HERE, "pyref_dnormal.npz"), **dn)
    np.savez_compressed(os.path.join(HERE, "pyref_image.npz"), **image_case(57, 41, 5))
    np.savez_compressed(os.path.join(HERE, "pyref_prologue.npz"), **prologue_case(257, 6))
    for f in ("pyref_dnormal.npz", "pyref_image.npz", "pyref_prologue.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
