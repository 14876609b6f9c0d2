import sys, os, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vcr-gaus_b200"), os.path.join(ROOT, "tests")): sys.path.insert(0, p)
import ref_ab
from util import CudaRun, make_case, rel_l2
inp, sc, cam = make_case(10000, 256, 256, seed=0, use_dirs=True, use_sem=False, sh_degree=3)
ref = ref_ab.RefRun(inp); run = CudaRun(inp)
rd, cd = ref.decode(), run.decode()
vis = ref.radii.cpu().numpy() > 0
for k in ("depths","xy","xyz_view","conic_opacity","rgb","cov3D"):
    a, b = cd[k][vis], rd[k][vis]
    print(k, "bit-equal:", np.array_equal(a.view(np.uint32), b.view(np.uint32)), "maxabs", np.abs(a-b).max(), "ref absmean", np.abs(b).mean(), "mine absmean", np.abs(a).mean())
raw = ref.geom.cpu().numpy()
print("geom bytes", raw.size)
out, rout = run.out.detach().cpu().numpy(), ref.out.cpu().numpy()
for c in range(10): print("ch", c, "ref mean", rout[c].mean(), "mine mean", out[c].mean(), "maxabs", np.abs(out[c]-rout[c]).max(), "n!=", (out[c]!=rout[c]).sum())
for k in ("keys_unsorted","vals_unsorted","keys","point_list","ranges","n_contrib","final_T"):
    print(k, np.array_equal(cd[k], rd[k]))
g = torch.randn((10, 256, 256), generator=torch.Generator().manual_seed(11)); g[8:] = 0
rb, cb = ref.backward(g), run.backward(g)
for k in rb:
    if cb.get(k) is not None: print(k, rel_l2(cb[k], rb[k]), np.abs(rb[k]).mean())
print("---- debug rgb ----")
print("ref clamped mean", rd["clamped"][vis].mean(), "mine", cd["clamped"][vis].mean())
print("ref rgb nonzero", (rd["rgb"]!=0).sum(), "any nan", np.isnan(rd["rgb"]).sum())
C = ref_ab.load()
e = torch.Tensor([])
print("e ptr", e.data_ptr())
# degree 0
inp0 = dict(inp); inp0["sh_degree"] = 0
r0 = ref_ab.RefRun(inp0); d0 = r0.decode(); print("deg0 ref rgb absmean", np.abs(d0["rgb"]).mean(), "img mean", r0.out[:3].mean().item())
# with colors_precomp
col = torch.from_numpy(cd["rgb"]).clone()
r1 = ref_ab.RefRun(inp, colors_precomp=col); m1 = CudaRun(inp, colors_precomp=col)
print("precomp: img maxabs", (r1.out - m1.out.detach()).abs().max().item(), "n!=", (r1.out != m1.out.detach()).sum().item())
ecuda = torch.empty(0, device="cuda")
out = C.rasterize_gaussians(ref.bg, ref.means3D, ecuda, ref.normals, ecuda, ref.opac, ref.scales, ref.rots, 1.0, ecuda, ref.view, ref.proj, ref.dirs, ref.tanx, ref.tany, ref.H, ref.W, ref.sh, 3, ref.campos, False, ref.inside, False)
print("cuda-empty: img mean", out[1][:3].mean().item(), "vs mine", run.out[:3].mean().item(), "maxabs", (out[1]-run.out.detach()).abs().max().item())
