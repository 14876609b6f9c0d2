#!/usr/bin/env python
"""Gradient differences on a saturating scene (opaque, large Gaussians): reference vs itself (atomic order), this library's
two-pixel backward and its one-pixel backward vs the reference.   python tools/ab_noise.py [P W H scale opacity seed]"""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vcr-gaus_b200")]
import ref_ab  # noqa: E402
from util import CudaRun, make_case, rel_l2  # noqa: E402
import vcr_native as N  # noqa: E402

a = sys.argv[1:]
P, W, H, scale, opacity, seed = (int(a[0]), int(a[1]), int(a[2]), float(a[3]), float(a[4]), int(a[5])) if len(a) >= 6 else (20000, 517, 299, 8.0, 0.999, 21)
inp, sc, cam = make_case(P, W, H, seed=seed, use_dirs=True, use_sem=True, sh_degree=3)
inp["scales"] = inp["scales"] * scale
inp["opacities"] = torch.full_like(inp["opacities"], opacity)
g = torch.randn((10, H, W), generator=torch.Generator().manual_seed(11))
keys = ("dL_dmeans3D", "dL_dmeans2D", "dL_dsh", "dL_dnormals", "dL_dopacity", "dL_dscales", "dL_drotations")
r1 = ref_ab.RefRun(inp).backward(g)
r2 = ref_ab.RefRun(inp).backward(g)
N.vcr_debug_render_version(2)
c2 = CudaRun(inp).backward(g)
N.vcr_debug_render_version(1)
c1 = CudaRun(inp).backward(g)
N.vcr_debug_render_version(2)
print("%-16s %12s %12s %12s %12s" % ("gradient", "ref vs ref", "2px vs ref", "1px vs ref", "2px vs 1px"))
for k in keys:
    print("%-16s %12.3e %12.3e %12.3e %12.3e" % (k, rel_l2(r2[k], r1[k]), rel_l2(c2[k], r1[k]), rel_l2(c1[k], r1[k]), rel_l2(c2[k], c1[k])))
