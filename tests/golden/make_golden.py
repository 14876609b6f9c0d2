"""Generate golden vectors from the REFERENCE CUDA extension (oracle/_ref, built from /root/reference by
oracle/build_ref.sh).  Must run on a GPU box:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'

then copy gpurun_out/golden/*.npz into tests/golden/.  The reference ships no golden vectors of its own;
these are outputs of the reference itself on seeded synthetic inputs and pin the CPU oracle (tests/test_oracle_cpu.py).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "vcr-gaus_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import ref_ab  # noqa: E402
from util import make_case, np_inputs  # noqa: E402

CASES = {
    # name: (P, W, H, use_dirs, use_sem, sh_degree, seed)
    "dirs_sem": (2000, 96, 80, True, True, 3, 0),
    "plain": (2000, 100, 72, False, False, 2, 1),
}


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    assert ref_ab.load() is not None, "oracle/_ref is missing"
    for name, (P, W, H, use_dirs, use_sem, deg, seed) in CASES.items():
        inp, sc, cam = make_case(P, W, H, seed=seed, use_dirs=use_dirs, use_sem=use_sem, sh_degree=deg)
        ref = ref_ab.RefRun(inp)
        dec = ref.decode()
        g = torch.randn((10, H, W), generator=torch.Generator().manual_seed(100 + seed))
        if not use_sem:
            g[8:] = 0
        grads = ref.backward(g)
        blob = {"in_" + k: v for k, v in np_inputs(inp).items() if v is not None}
        blob.update({"ref_" + k: v for k, v in dec.items()})
        blob.update({"ref_" + k: v for k, v in grads.items()})
        blob.update(ref_out_color=ref.out.cpu().numpy(), ref_radii=ref.radii.cpu().numpy(), dL_dout=g.numpy(),
                    meta=np.array([P, W, H, int(use_dirs), int(use_sem), deg, seed]))
        if C_knn is not None:
            pts = inp["means3D"].cuda()
            blob["ref_dist2"] = C_knn.distCUDA2(pts).cpu().numpy()
        np.savez_compressed(os.path.join(outdir, f"ref_{name}.npz"), **blob)
        print("wrote", name, "R =", dec["R"])


if __name__ == "__main__":
    C_knn = ref_ab.load("simple_knn")
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
