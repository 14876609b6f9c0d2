"""Fused render() prologue vs the reference's PyTorch chain, forward + backward, 1 M Gaussians (CUDA events, 20 iterations)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vcr-gaus_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_prologue_gpu import _inputs, reference_chain  # noqa: E402
from vcr_prologue import gaussian_prologue  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
raw, cc, M = _inputs(P, 0)
args = (raw["xyz"], raw["fdc"], raw["frest"], raw["o_raw"], raw["s_raw"], raw["r_raw"], cc, M)
leaves = [raw[k] for k in ("fdc", "frest", "o_raw", "s_raw", "r_raw")]
cot = None


def run(fn):
    global cot
    out = fn(*args) if fn is gaussian_prologue else fn(*args, None)
    if cot is None:
        cot = [torch.randn_like(t) for t in out]
    torch.autograd.grad(out, leaves, cot)


for name, fn in (("fused", gaussian_prologue), ("pytorch chain", reference_chain)):
    for _ in range(5):
        run(fn)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run(fn)
    e1.record()
    torch.cuda.synchronize()
    print(f"PROLOGUE {name}: {e0.elapsed_time(e1) / 20:.3f} ms per forward+backward, P = {P}", flush=True)
