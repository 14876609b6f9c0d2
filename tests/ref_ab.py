"""Drive the UNMODIFIED reference CUDA extension (compiled by oracle/build_ref.sh into oracle/_ref/) on the same inputs.

RefRun calls the pybind11 module `_C` of the reference directly (DGR/ext.cpp:15-21) with the argument tuples the
reference's Python wrapper builds (DGR/diff_gaussian_rasterization/__init__.py:186-220 forward, :401-427 backward), because
the parity tests need the opaque buffers it returns; `wrapper()` gives the reference's own Python wrapper (staged beside
`_C.so`, loaded under a private module name so that it cannot shadow this repository's package of the same name) for the
`f_count` modes and for running the reference's render() against the reference build.
The opaque buffers are decoded with the layout of GeometryState / BinningState / ImageState::fromChunk
(DGR/cuda_rasterizer/rasterizer_impl.cu:155-195; obtain() aligns every array to 128 bytes, rasterizer_impl.h:21-27).
Test infrastructure only.
"""
import importlib.util
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
_mods = {}


def wrapper(name="diff_gaussian_rasterization"):
    """The reference's Python package `name` from oracle/_ref (its own __init__.py + _C.so) as module `vcr_ref_<name>`, or
    None when it has not been staged / built."""
    import sys
    key = "pkg:" + name
    if key in _mods:
        return _mods[key]
    d = os.path.join(REF_DIR, name)
    init = os.path.join(d, "__init__.py")
    mod = None
    if os.path.exists(init) and os.path.getsize(init) > 0 and os.path.exists(os.path.join(d, "_C.so")):
        try:
            modname = f"vcr_ref_{name}"
            spec = importlib.util.spec_from_file_location(modname, init, submodule_search_locations=[d])
            mod = importlib.util.module_from_spec(spec)
            sys.modules[modname] = mod                      # `from . import _C` resolves against this entry
            spec.loader.exec_module(mod)
        except Exception as ex:  # pragma: no cover
            print(f"reference package {d} failed to load: {ex}")
            sys.modules.pop(f"vcr_ref_{name}", None)
            mod = None
    _mods[key] = mod
    return mod


def load(name="diff_gaussian_rasterization"):
    """Returns the reference `_C` module or None when oracle/_ref was not built / cannot be loaded."""
    if name in _mods:
        return _mods[name]
    mod = None
    w = wrapper(name)
    if w is not None and hasattr(w, "_C"):
        mod = w._C                                           # one load of the shared object per process
    else:
        path = os.path.join(REF_DIR, name, "_C.so")
        if os.path.exists(path):
            try:
                spec = importlib.util.spec_from_file_location(f"vcr_ref_{name}._C", path)  # PyInit__C; distinct cache key per module
                mod = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(mod)
            except Exception as ex:  # pragma: no cover
                print(f"reference extension {path} failed to load: {ex}")
                mod = None
    _mods[name] = mod
    return mod


def settings(inp, device, f_count=0):
    """GaussianRasterizationSettings of the REFERENCE wrapper for a tests/util.py input dict."""
    w = wrapper()
    return w.GaussianRasterizationSettings(
        image_height=inp["H"], image_width=inp["W"], tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"], bg=inp["bg"].to(device),
        scale_modifier=1.0, viewmatrix=inp["viewmatrix"].to(device), projmatrix=inp["projmatrix"].to(device),
        sh_degree=inp["sh_degree"], campos=inp["campos"].to(device), prefiltered=False, debug=False, f_count=f_count)


def aux(inp, f_count, device="cuda"):
    """f_count = 1: (gaussians_count, important_score, color, radii) of _C.count_gaussians; f_count = 3: (gaussians_count,
    radii) of _C.visi_gaussians_acc — through the reference wrapper's own dispatch (rasterize_gaussians, __init__.py:42-94)."""
    w = wrapper()
    dev = torch.device(device)
    t = lambda x: x.to(dev).contiguous()  # noqa: E731
    means3D = t(inp["means3D"])
    rast = w.GaussianRasterizer(raster_settings=settings(inp, dev, f_count))
    with torch.no_grad():
        out = rast(means3D=means3D, means2D=torch.zeros_like(means3D), means2D_densify=torch.zeros_like(means3D), opacities=t(inp["opacities"]),
                   shs=t(inp["shs"]), scales=t(inp["scales"]), rotations=t(inp["rotations"]))
    torch.cuda.synchronize()
    return out


def _al(x, a=128):
    return (x + a - 1) // a * a


class RefRun:
    def __init__(self, inp, device="cuda", colors_precomp=None, cov3D_precomp=None, inside=None, scale_modifier=1.0):
        C = load()
        assert C is not None
        dev = torch.device(device)
        e = torch.Tensor([])
        t = lambda x: e if x is None else x.to(dev).contiguous()  # noqa: E731
        self.P = P = inp["means3D"].shape[0]
        self.W, self.H = inp["W"], inp["H"]
        self.means3D, self.opac = t(inp["means3D"]), t(inp["opacities"])
        self.sh = t(inp["shs"]) if colors_precomp is None else e
        self.colors = t(colors_precomp)
        self.normals, self.sem = t(inp["normals"]), t(inp["semantics"])
        self.scales = t(inp["scales"]) if cov3D_precomp is None else e
        self.rots = t(inp["rotations"]) if cov3D_precomp is None else e
        self.cov = t(cov3D_precomp)
        self.dirs = t(inp["dirs"])
        self.inside = torch.ones(P, dtype=torch.bool, device=dev) if inside is None else inside.to(dev)
        self.bg, self.view, self.proj, self.campos = t(inp["bg"]), t(inp["viewmatrix"]), t(inp["projmatrix"]), t(inp["campos"])
        self.tanx, self.tany, self.deg, self.mod = inp["tanfovx"], inp["tanfovy"], inp["sh_degree"], scale_modifier
        (self.R, self.out, self.radii, self.geom, self.binning, self.img) = C.rasterize_gaussians(
            self.bg, self.means3D, self.colors, self.normals, self.sem, self.opac, self.scales, self.rots, self.mod, self.cov,
            self.view, self.proj, self.dirs, self.tanx, self.tany, self.H, self.W, self.sh, self.deg, self.campos, False,
            self.inside, False)
        torch.cuda.synchronize()

    def decode(self):
        P, R, N = self.P, self.R, self.W * self.H
        g = self.geom.cpu().numpy()
        base = self.geom.data_ptr()
        off = 0

        def take(buf, count, dtype, itemsize=None):
            nonlocal off
            isz = itemsize or np.dtype(dtype).itemsize
            start = _al(base + off) - base
            arr = buf[start:start + count * isz].view(dtype).copy()
            off = start + count * isz
            return arr

        out = dict(R=R)
        out["depths"] = take(g, P, np.float32)
        out["clamped"] = take(g, 3 * P, np.uint8).reshape(P, 3)
        take(g, P, np.int32)  # internal_radii
        out["xy"] = take(g, 2 * P, np.float32).reshape(P, 2)
        xyz = take(g, 9 * P, np.float32)  # float3[3P] (over-allocated x3); the first P entries are used
        out["xyz_view"] = xyz[:3 * P].reshape(P, 3)
        out["cov3D"] = take(g, 6 * P, np.float32).reshape(P, 6)
        out["conic_opacity"] = take(g, 4 * P, np.float32).reshape(P, 4)
        out["rgb"] = take(g, 3 * P, np.float32).reshape(P, 3)
        out["tiles_touched"] = take(g, P, np.uint32)
        out["point_offsets"] = np.cumsum(out["tiles_touched"], dtype=np.uint64).astype(np.uint32)
        im = self.img.cpu().numpy()
        base, off = self.img.data_ptr(), 0
        out["final_T"] = take(im, N, np.float32)
        out["n_contrib"] = take(im, N, np.uint32)
        tiles = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        out["ranges"] = take(im, 2 * N, np.uint32)[:2 * tiles].reshape(tiles, 2)
        if R > 0:
            b = self.binning.cpu().numpy()
            base, off = self.binning.data_ptr(), 0
            out["point_list"] = take(b, R, np.uint32)
            out["vals_unsorted"] = take(b, R, np.uint32)
            out["keys"] = take(b, R, np.uint64)
            out["keys_unsorted"] = take(b, R, np.uint64)
        return out

    def backward(self, dL_dout):
        C = load()
        g = dL_dout.to(self.out.device).contiguous()
        alpha, depth = self.out[7:8], self.out[3:4]
        (g2d, g2dd, gcol, gnrm, gobj, gop, g3d, gcov, gsh, gsc, grot) = C.rasterize_gaussians_backward(
            self.bg, self.means3D, self.radii, self.colors, self.normals, self.sem, self.scales, self.rots, self.mod, self.cov,
            self.view, self.proj, self.dirs, self.tanx, self.tany, g, self.sh, self.deg, self.campos, self.geom, self.R,
            self.binning, self.img, alpha, depth, self.inside, False)
        torch.cuda.synchronize()
        n = lambda x: x.cpu().numpy()  # noqa: E731
        return dict(dL_dmeans2D=n(g2d), dL_dmeans2D_densify=n(g2dd), dL_dcolors=n(gcol), dL_dnormals=n(gnrm), dL_dobjects=n(gobj),
                    dL_dopacity=n(gop), dL_dmeans3D=n(g3d), dL_dcov3D=n(gcov), dL_dsh=n(gsh), dL_dscales=n(gsc), dL_drotations=n(grot))
