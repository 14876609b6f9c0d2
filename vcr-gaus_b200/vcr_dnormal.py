"""Fused D-Normal regularizer (depth -> normal -> confidence-weighted depth-normal loss).

Replaces, with one kernel per direction, the reference's PyTorch chain
``compute_normals`` (tools/normal_utils.py:30-41) -> ``cos_weight`` + ``monosdf_normal_loss``
(tools/loss_utils.py:122-142) with the masking of trainer.py:266-280 and
gaussian_renderer/__init__.py:125-134.

``compute_normals(depth, K)`` keeps the reference's name and signature (est_normal [H,W,3]) and is differentiable;
``depth_normal_loss(...)`` is the whole regularizer and never materialises est_normal unless asked to.
"""
import ctypes

import torch

import vcr_native as _n


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _intr(K):
    """fx, fy, cx, cy from a 3x3 intrinsic matrix (tools/graphics_utils.py:89-101) or a 4-sequence, on the host."""
    if isinstance(K, torch.Tensor):
        K = K.detach().cpu()
        if K.ndim == 2:
            return float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])
        K = K.tolist()
    fx, fy, cx, cy = K
    return float(fx), float(fy), float(cx), float(cy)


class _DNormal(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth, gt_normal, render_normal, mask, intr, exp_t, mask_depth_max, want_normal):
        if not depth.is_cuda:
            raise RuntimeError("D-Normal (B200 build) needs CUDA tensors; there is no CPU fallback")
        H, W = depth.shape[-2], depth.shape[-1]
        d = depth.contiguous().to(torch.float32)
        g = gt_normal.contiguous().to(torch.float32)
        r = render_normal.contiguous().to(torch.float32) if render_normal is not None else None
        m = mask.contiguous().to(torch.uint8) if mask is not None else None
        fx, fy, cx, cy = intr
        sums = torch.empty(3, dtype=torch.float32, device=d.device)
        est = torch.empty((H, W, 3), dtype=torch.float32, device=d.device) if want_normal else None
        _n.check(_n.vcr_dnormal_forward(d.data_ptr(), g.data_ptr(), _n.ptr(r), _n.ptr(m), W, H, fx, fy, cx, cy, float(exp_t),
                                        float(mask_depth_max), _n.ptr(est), None, sums.data_ptr(), 1, _stream()),
                 "dnormal_forward")
        ctx.save_for_backward(d, g, r if r is not None else torch.empty(0, device=d.device),
                              m if m is not None else torch.empty(0, dtype=torch.uint8, device=d.device), sums)
        ctx.meta = (W, H, fx, fy, cx, cy, float(exp_t), float(mask_depth_max), depth.shape)
        cnt = sums[2]
        loss = torch.where(cnt > 0, (sums[0] + sums[1]) / torch.clamp(cnt, min=1.0), torch.zeros_like(cnt))
        if want_normal:
            return loss, est
        return loss, torch.empty(0, device=d.device)

    @staticmethod
    def backward(ctx, g_loss, g_est):
        d, g, r, m, sums = ctx.saved_tensors
        W, H, fx, fy, cx, cy, exp_t, mdm, shape = ctx.meta
        gl = g_loss.contiguous().to(torch.float32).reshape(1) if g_loss is not None else torch.zeros(1, device=d.device)
        ge = g_est.contiguous().to(torch.float32) if (g_est is not None and g_est.numel() != 0) else None
        out = torch.empty((H, W), dtype=torch.float32, device=d.device)
        _n.check(_n.vcr_dnormal_backward(d.data_ptr(), g.data_ptr(), _n.ptr(r), _n.ptr(m), W, H, fx, fy, cx, cy, exp_t, mdm,
                                         sums.data_ptr(), gl.data_ptr(), _n.ptr(ge), out.data_ptr(), _stream()),
                 "dnormal_backward")
        return out.reshape(shape), None, None, None, None, None, None, None


def depth_normal_loss(depth, intr, gt_normal, render_normal=None, mask=None, exp_t=0.0, mask_depth_max=0.0,
                      return_normal=False):
    """Confidence-weighted depth-normal loss (trainer.py:266-280).

    depth [1,H,W] or [H,W]; intr 3x3 K or (fx,fy,cx,cy); gt_normal [H,W,3]; render_normal [H,W,3] (rasterizer normal map,
    permuted, need not be normalised; only used, detached, for the confidence weight when exp_t > 0); mask bool/uint8 [H,W].
    mask_depth_max > 0 adds ``depth < mask_depth_max`` to the mask (gaussian_renderer/__init__.py:127-131).
    Returns the scalar loss (and est_normal [H,W,3] when return_normal).
    """
    rn = render_normal.detach() if render_normal is not None else None
    loss, est = _DNormal.apply(depth, gt_normal, rn, mask, _intr(intr), exp_t, mask_depth_max, return_normal)
    return (loss, est) if return_normal else loss


def compute_normals(depth_map, K):
    """Same contract as tools/normal_utils.py:30-41: depth [1,H,W] (or [H,W]) -> unit normals [H,W,3], differentiable."""
    H, W = depth_map.shape[-2], depth_map.shape[-1]
    dummy = torch.zeros((H, W, 3), dtype=torch.float32, device=depth_map.device)
    none_mask = torch.zeros((H, W), dtype=torch.uint8, device=depth_map.device)  # empty mask: loss part is identically 0
    _, est = _DNormal.apply(depth_map, dummy, None, none_mask, _intr(K), 0.0, 0.0, True)
    return est


class _ViewLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, outmap, gt_image, gt_normal, mask, intr, exp_t, mask_depth_max, w_l1, w_mono, w_dn):
        if not outmap.is_cuda:
            raise RuntimeError("view_loss (B200 build) needs CUDA tensors; there is no CPU fallback")
        C, H, W = outmap.shape
        if C != 10:
            raise RuntimeError("view_loss expects the rasterizer's [10,H,W] output map")
        o = outmap.contiguous()
        gi = gt_image.contiguous().to(torch.float32)
        gn = gt_normal.contiguous().to(torch.float32)
        m = mask.contiguous().to(torch.uint8) if mask is not None else None
        fx, fy, cx, cy = intr
        sums = torch.empty(8, dtype=torch.float32, device=o.device)
        _n.check(_n.vcr_view_loss_forward(o.data_ptr(), gi.data_ptr(), gn.data_ptr(), _n.ptr(m), W, H, fx, fy, cx, cy,
                                          float(exp_t), float(mask_depth_max), sums.data_ptr(), _stream()), "view_loss_forward")
        ctx.save_for_backward(o, gi, gn, m if m is not None else torch.empty(0, dtype=torch.uint8, device=o.device), sums)
        ctx.meta = (W, H, fx, fy, cx, cy, float(exp_t), float(mask_depth_max), float(w_l1), float(w_mono), float(w_dn))
        loss = torch.empty((), dtype=torch.float32, device=o.device)
        _n.check(_n.vcr_view_loss_total(sums.data_ptr(), W, H, float(w_l1), float(w_mono), float(w_dn), loss.data_ptr(), _stream()),
                 "view_loss_total")
        ctx.mark_non_differentiable(sums)
        return loss, sums

    @staticmethod
    def backward(ctx, g_loss, _g_sums):
        o, gi, gn, m, sums = ctx.saved_tensors
        W, H, fx, fy, cx, cy, exp_t, mdm, w_l1, w_mono, w_dn = ctx.meta
        gl = g_loss.contiguous().to(torch.float32).reshape(1)
        g_out = torch.empty_like(o)
        _n.check(_n.vcr_view_loss_backward(o.data_ptr(), gi.data_ptr(), gn.data_ptr(), _n.ptr(m), W, H, fx, fy, cx, cy, exp_t, mdm,
                                           sums.data_ptr(), gl.data_ptr(), w_l1, w_mono, w_dn, g_out.data_ptr(), _stream()),
                 "view_loss_backward")
        return g_out, None, None, None, None, None, None, None, None, None


def view_loss(outmap, gt_image, gt_normal, intr, mask=None, w_l1=1.0, w_mono=0.0, w_dn=0.0, exp_t=0.0, mask_depth_max=0.0,
              return_terms=False):
    """All image-space loss terms of one training view (except SSIM) from the rasterizer's [10,H,W] output map, fused:

        w_l1 * l1_loss(render, gt_image)                                   (trainer.py:236, tools/loss_utils.py:22-23)
      + w_mono * monosdf_normal_loss(normalize(render_normal), gt_normal)  (trainer.py:261-264)
      + w_dn * depth-normal loss with confidence weights                   (trainer.py:266-280)

    gt_image [3,H,W], gt_normal [H,W,3], intr = 3x3 K or (fx,fy,cx,cy), mask bool/uint8 [H,W] (D-Normal term only).
    Returns the scalar loss; with return_terms also the raw sums (see include/vcr_gaus_b200.h) for logging."""
    loss, sums = _ViewLoss.apply(outmap, gt_image, gt_normal, mask, _intr(intr), exp_t, mask_depth_max, w_l1, w_mono, w_dn)
    return (loss, sums) if return_terms else loss


class _SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2):
        for t, name in ((img1, "img1"), (img2, "img2")):
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError(f"ssim: {name} must be a float32 CUDA tensor; there is no CPU fallback")
        if img1.shape != img2.shape or img1.ndim not in (3, 4):
            raise RuntimeError("ssim: expected two [C,H,W] or [B,C,H,W] tensors of the same shape")
        a, b = img1.contiguous(), img2.contiguous()
        H, W = a.shape[-2], a.shape[-1]
        planes = a.numel() // (H * W)
        dmaps = torch.empty((3,) + tuple(a.shape), dtype=torch.float32, device=a.device)
        total = torch.empty(1, dtype=torch.float64, device=a.device)
        _n.check(_n.vcr_ssim_forward(planes, H, W, a.data_ptr(), b.data_ptr(), dmaps.data_ptr(), total.data_ptr(), _stream()),
                 "ssim_forward")
        ctx.save_for_backward(a, b, dmaps)
        return (total / a.numel()).to(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, g):
        a, b, dmaps = ctx.saved_tensors
        H, W = a.shape[-2], a.shape[-1]
        grad = torch.empty_like(a)
        gg = g.to(torch.float32).contiguous().reshape(1)
        _n.check(_n.vcr_ssim_backward(a.numel() // (H * W), H, W, a.data_ptr(), b.data_ptr(), dmaps.data_ptr(), gg.data_ptr(),
                                      grad.data_ptr(), _stream()), "ssim_backward")
        return grad, None


def ssim(img1, img2, window_size=11, size_average=True):
    """tools/loss_utils.py:54-92 `ssim(img1, img2)` — the mean SSIM of two images with the reference's 11x11 Gaussian window —
    as one fused kernel per direction (the reference: five grouped conv2d + ~15 elementwise launches each way).
    Differentiable w.r.t. img1; img2 is treated as ground truth.  Only the reference's defaults are built."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("fused ssim implements the reference's defaults (window_size=11, size_average=True)")
    return _SSIM.apply(img1, img2)
