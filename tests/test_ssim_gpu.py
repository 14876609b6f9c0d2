"""Fused SSIM (vcr_dnormal.ssim) against the reference's conv2d formulation (tools/loss_utils.py:47-92, restated), value and
gradient, in float64 (the reference's own float32 convolutions run in TF32 on this GPU unless told otherwise)."""
from math import exp

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def reference_ssim(img1, img2, window_size=11):
    gauss = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(window_size)])
    g1 = (gauss / gauss.sum()).unsqueeze(1)
    channel = img1.size(-3)
    window = g1.mm(g1.t()).float().unsqueeze(0).unsqueeze(0).expand(channel, 1, window_size, window_size).contiguous()
    window = window.to(img1.device).type_as(img1)
    pad = window_size // 2
    mu1, mu2 = F.conv2d(img1, window, padding=pad, groups=channel), F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))).mean()


@pytest.mark.parametrize("shape", [(3, 7, 9), (3, 37, 53), (1, 16, 16), (3, 270, 481), (2, 3, 40, 24)])
def test_ssim_value_and_gradient(shape):
    from vcr_dnormal import ssim
    g = torch.Generator().manual_seed(sum(shape))
    gt = torch.rand(shape, generator=g).cuda()
    img = (gt + 0.15 * torch.randn(shape, generator=g).cuda()).clamp(0, 1).requires_grad_(True)
    val = ssim(img, gt)
    (2.5 * (1.0 - val)).backward()
    img64 = img.detach().double().requires_grad_(True)
    ref = reference_ssim(img64 if img64.ndim == 4 else img64[None], gt.double() if gt.ndim == 4 else gt.double()[None])
    (2.5 * (1.0 - ref)).backward()
    assert abs(float(val) - float(ref)) <= 2e-6, (float(val), float(ref))
    err = float((img.grad.double() - img64.grad).norm() / img64.grad.norm())
    assert err <= 1e-5, err


def test_ssim_matches_reference_float32_without_tf32():
    from vcr_dnormal import ssim
    g = torch.Generator().manual_seed(3)
    gt = torch.rand((3, 120, 160), generator=g).cuda()
    img = (gt + 0.1 * torch.randn((3, 120, 160), generator=g).cuda()).requires_grad_(True)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = reference_ssim(img[None], gt[None])
        (g_ref,) = torch.autograd.grad(ref, img)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    val = ssim(img, gt)
    (g_got,) = torch.autograd.grad(val, img)
    assert abs(float(val) - float(ref)) <= 1e-5
    assert float((g_got - g_ref).norm() / g_ref.norm()) <= 1e-4


def test_ssim_identical_images_and_errors():
    from vcr_dnormal import ssim
    x = torch.rand((3, 33, 47)).cuda()
    assert abs(float(ssim(x, x.clone())) - 1.0) <= 1e-6
    with pytest.raises(NotImplementedError):
        ssim(x, x, window_size=7)
    with pytest.raises(RuntimeError):
        ssim(x.cpu(), x.cpu())
