"""prune_and_grow (one prefix sum + gather launches) vs the reference's way (boolean-mask indexing of every parameter group and
both Adam moments + torch.cat of the new rows), 1 M Gaussians, 20 % pruned, 5 % new."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vcr-gaus_b200"))
import vcr_view_parallel as VP  # noqa: E402

P = 1_000_000
dev = torch.device("cuda")
shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
init = {k: torch.randn(s, device=dev) for k, s in shapes.items()}
bucket = VP.GaussianBucket(init, device=dev, names=tuple(shapes))
opt = VP.BucketAdam(bucket, {k: 1e-3 for k in shapes})
keep = torch.rand(P, device=dev) > 0.2
n_new = P // 20
m = {k: torch.randn_like(v) for k, v in init.items()}
v2 = {k: torch.rand_like(v) for k, v in init.items()}


def ours():
    VP.prune_and_grow(bucket, keep, n_new=n_new, optimizer=opt)


def reference_way():
    out = {}
    for k in shapes:                                   # _prune_optimizer + cat_tensors_to_optimizer
        new_rows = torch.zeros((n_new,) + shapes[k][1:], device=dev)
        out[k] = (torch.cat((init[k][keep], new_rows), dim=0), torch.cat((m[k][keep], torch.zeros_like(new_rows)), dim=0),
                  torch.cat((v2[k][keep], torch.zeros_like(new_rows)), dim=0))
    return out


for name, fn in (("prune_and_grow", ours), ("boolean indexing + cat (reference)", reference_way)):
    for k in range(13):
        if k == 3:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"COMPACT {name}: {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
