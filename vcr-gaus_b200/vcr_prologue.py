"""Fused per-Gaussian prologue of render(): activations, SH concat and camera-space normals in one kernel per direction.

Replaces, with identical results up to float rounding, what the reference computes with ~40 PyTorch launches per view
before it calls the rasterizer:

    opacity   = pc.get_opacity                 sigmoid(_opacity)                    scene/gaussian_model.py:160-162
    scales    = pc.get_scaling                 exp(_scaling)                        scene/gaussian_model.py:125-128
    rotations = pc.get_rotation                F.normalize(_rotation)               scene/gaussian_model.py:130-132
    shs       = pc.get_features                cat(_features_dc, _features_rest)    scene/gaussian_model.py:138-142
    normal    = pc.get_normal(is_all=...)      column argmin(scaling) of R(q)       scene/gaussian_model.py:168-192
    normals_precomp = flip towards the camera, rotate by R_w2c                      gaussian_renderer/__init__.py:95-101

CUDA tensors only (there is no CPU path)."""
import ctypes

import torch

import vcr_native as _n


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _c(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"gaussian_prologue: {name} must be a CUDA tensor; there is no CPU fallback")
    if t.dtype != torch.float32:
        raise RuntimeError(f"gaussian_prologue: {name} must be float32, got {t.dtype}")
    t = t.contiguous()
    if t.data_ptr() % 16 and t.ndim == 2 and t.shape[1] == 4:
        t = t.clone()       # quaternion arrays are accessed as float4 (include/vcr_gaus_b200.h, "Alignment")
    return t


class _Prologue(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, camera_center, R_w2c, valid):
        xyz, fdc, frest = _c(xyz, "xyz"), _c(features_dc, "features_dc"), _c(features_rest, "features_rest")
        o_raw, s_raw, r_raw = _c(opacity_raw, "opacity"), _c(scaling_raw, "scaling"), _c(rotation_raw, "rotation")
        P = xyz.shape[0]
        n_rest = frest.shape[1] if frest.numel() else 0
        dev = xyz.device
        cc = _c(camera_center.to(dev).reshape(-1), "camera_center")
        M = _c(R_w2c.to(dev).reshape(-1), "R_w2c")
        if cc.numel() != 3 or M.numel() != 9:
            raise RuntimeError("gaussian_prologue: camera_center must have 3 and R_w2c 9 elements")
        v = None
        if valid is not None:
            v = valid.to(torch.uint8).contiguous() if valid.dtype == torch.bool else valid.contiguous()
            if v.numel() != P or not v.is_cuda:
                raise RuntimeError("gaussian_prologue: valid must be a CUDA bool tensor with one entry per Gaussian")
        f = dict(dtype=torch.float32, device=dev)
        opac, scales, rots = torch.empty((P, 1), **f), torch.empty((P, 3), **f), torch.empty((P, 4), **f)
        shs, normals = torch.empty((P, 1 + n_rest, 3), **f), torch.empty((P, 3), **f)
        aux = torch.empty((P,), dtype=torch.uint8, device=dev)
        if P:
            _n.check(_n.vcr_prologue_forward(P, n_rest, xyz.data_ptr(), fdc.data_ptr(), _n.ptr(frest), o_raw.data_ptr(),
                                             s_raw.data_ptr(), r_raw.data_ptr(), _n.ptr(v), cc.data_ptr(), M.data_ptr(),
                                             opac.data_ptr(), scales.data_ptr(), rots.data_ptr(), shs.data_ptr(),
                                             normals.data_ptr(), aux.data_ptr(), _stream(xyz)), "prologue_forward")
        ctx.save_for_backward(r_raw, opac, scales, rots, aux, M)
        ctx.shapes = (features_dc.shape, features_rest.shape, n_rest)
        # optional zero-copy gradient destinations (vcr_view_parallel.GaussianBucket(direct=True)), as in the rasterizer
        ctx.grad_out = tuple(getattr(t, "_vcr_grad_out", None) for t in (features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw))
        return opac, scales, rots, shs, normals

    @staticmethod
    def backward(ctx, g_opac, g_scales, g_rots, g_shs, g_normals):
        r_raw, opac, scales, rots, aux, M = ctx.saved_tensors
        dc_shape, rest_shape, n_rest = ctx.shapes
        P = r_raw.shape[0]
        f = dict(dtype=torch.float32, device=r_raw.device)

        def g(t, shape):
            return torch.zeros(shape, **f) if t is None else _c(t, "gradient")
        g_opac, g_scales, g_rots = g(g_opac, (P, 1)), g(g_scales, (P, 3)), g(g_rots, (P, 4))
        g_shs, g_normals = g(g_shs, (P, 1 + n_rest, 3)), g(g_normals, (P, 3))
        def dest(i, shape):
            buf = ctx.grad_out[i]
            if (buf is not None and buf.is_cuda and buf.dtype == torch.float32 and buf.is_contiguous()
                    and tuple(buf.shape) == tuple(shape) and buf.data_ptr() % 16 == 0):
                return buf.view(shape)       # a fresh view of the bucket: autograd adopts it as .grad without a copy
            return torch.empty(shape, **f)
        d_dc, d_rest = dest(0, dc_shape), dest(1, rest_shape)
        d_o, d_s, d_r = dest(2, (P, 1)), dest(3, (P, 3)), dest(4, (P, 4))
        if P:
            _n.check(_n.vcr_prologue_backward(P, n_rest, r_raw.data_ptr(), opac.data_ptr(), scales.data_ptr(), rots.data_ptr(),
                                              aux.data_ptr(), M.data_ptr(), g_opac.data_ptr(), g_scales.data_ptr(),
                                              g_rots.data_ptr(), g_shs.data_ptr(), g_normals.data_ptr(), d_dc.data_ptr(),
                                              _n.ptr(d_rest), d_o.data_ptr(), d_s.data_ptr(), d_r.data_ptr(), _stream(r_raw)),
                     "prologue_backward")
        return None, d_dc, d_rest, d_o, d_s, d_r, None, None, None


def gaussian_prologue(xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, camera_center, R_w2c, valid=None):
    """(opacity [P,1], scales [P,3], rotations [P,4], shs [P,K,3], normals_precomp [P,3]) from the raw parameters of
    GaussianModel (`_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation`), differentiable w.r.t. all of them
    but xyz (which only decides the flip, as in the reference).  `R_w2c` is `torch.tensor(viewpoint_camera.R.T)`,
    `camera_center` is `viewpoint_camera.camera_center`; `valid` is the inside mask of `get_normal(is_all=False)` (None =
    all, i.e. `is_all=True`).  In render(): means3D = xyz, and these five go straight into the rasterizer call."""
    return _Prologue.apply(xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, camera_center, R_w2c, valid)
