"""One Adam step over the 59 M-float bucket (1 M Gaussians): BucketAdam (one kernel) vs torch.optim.Adam with the reference's
six groups, default implementation (what the reference runs) and fused=True."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vcr-gaus_b200"))
import vcr_view_parallel as VP  # noqa: E402

P = 1_000_000
dev = torch.device("cuda")
shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20, opacity=0.05, scaling=5e-3, rotation=1e-3)
init = {k: torch.randn(s, device=dev) for k, s in shapes.items()}
bucket = VP.GaussianBucket(init, device=dev, names=tuple(shapes))
bucket.grad_flat.normal_()
ours = VP.BucketAdam(bucket, lrs)


def torch_adam(**kw):
    ps = {k: v.clone().requires_grad_(True) for k, v in init.items()}
    for p in ps.values():
        p.grad = torch.randn_like(p)
    return torch.optim.Adam([{"params": [ps[k]], "lr": lrs[k]} for k in shapes], lr=0.0, eps=1e-15, **kw)


for name, opt in (("BucketAdam (one kernel)", ours), ("torch.optim.Adam, default (reference)", torch_adam()), ("torch.optim.Adam, fused=True", torch_adam(fused=True))):
    for k in range(23):
        if k == 3:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"ADAM {name}: {ms:.3f} ms per step ({59e6 * 28 / ms / 1e6:.0f} GB/s at 28 B per float)", flush=True)
