"""Fused SSIM vs the reference's conv2d formulation, forward + backward on one 3x1080x1920 image pair."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vcr-gaus_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_ssim_gpu import reference_ssim  # noqa: E402
from vcr_dnormal import ssim  # noqa: E402

gt = torch.rand((3, 1080, 1920)).cuda()
img = (gt + 0.1 * torch.randn_like(gt)).requires_grad_(True)
for name, fn in (("fused", lambda: ssim(img, gt)), ("pytorch conv2d chain (TF32 allowed, the default)", lambda: reference_ssim(img[None], gt[None]))):
    for k in range(23):
        if k == 3:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        torch.autograd.grad(1.0 - fn(), img)
    e1.record()
    torch.cuda.synchronize()
    print(f"SSIM {name}: {e0.elapsed_time(e1) / 20:.3f} ms per forward+backward", flush=True)
