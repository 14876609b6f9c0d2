"""Shared helpers of the parity tests: run the CUDA path through the public API / C ABI and decode its opaque buffers."""
import ctypes

import numpy as np
import torch

import vcr_synth as S


def np_inputs(d):
    return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}


def make_case(P, W, H, seed=0, view=1, use_dirs=True, use_sem=False, sh_degree=3, scale0=None, extent=1.0):
    sc = S.make_scene(P, seed=seed, scale0=scale0, extent=extent)
    cam = S.make_camera(W, H, idx=view, extent=extent, seed=seed)
    return S.raster_inputs(sc, cam, use_dirs=use_dirs, use_sem=use_sem, sh_degree=sh_degree, seed=seed), sc, cam


def settings_from(inp, device, f_count=0, debug=False, scale_modifier=1.0):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=inp["H"], image_width=inp["W"], tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"], bg=inp["bg"].to(device),
        scale_modifier=scale_modifier, viewmatrix=inp["viewmatrix"].to(device), projmatrix=inp["projmatrix"].to(device),
        sh_degree=inp["sh_degree"], campos=inp["campos"].to(device), prefiltered=False, debug=debug, f_count=f_count)


class CudaRun:
    """One forward (+ optional backward) through GaussianRasterizer, keeping the saved opaque buffers for decoding."""

    def __init__(self, inp, device="cuda", colors_precomp=None, cov3D_precomp=None, inside=None, scale_modifier=1.0):
        from diff_gaussian_rasterization import GaussianRasterizer, _RasterizeGaussians
        self.inp = inp
        dev = torch.device(device)
        t = lambda x: None if x is None else x.to(dev).clone().requires_grad_(True)  # noqa: E731
        self.means3D = t(inp["means3D"])
        self.means2D = torch.zeros_like(self.means3D, requires_grad=True)
        self.means2D_densify = torch.zeros_like(self.means3D, requires_grad=True)
        self.opacities = t(inp["opacities"])
        self.shs = t(inp["shs"]) if colors_precomp is None else None
        self.colors_precomp = t(colors_precomp)
        self.normals = t(inp["normals"])
        self.semantics = t(inp["semantics"])
        self.scales = t(inp["scales"]) if cov3D_precomp is None else None
        self.rotations = t(inp["rotations"]) if cov3D_precomp is None else None
        self.cov3D_precomp = t(cov3D_precomp)
        self.dirs = None if inp["dirs"] is None else inp["dirs"].to(dev)
        self.inside = None if inside is None else inside.to(dev)
        self.rs = settings_from(inp, dev, scale_modifier=scale_modifier)
        rast = GaussianRasterizer(self.rs)
        self.out, self.radii = rast(means3D=self.means3D, means2D=self.means2D, means2D_densify=self.means2D_densify,
                                    shs=self.shs, colors_precomp=self.colors_precomp, normals_precomp=self.normals,
                                    semantics_precomp=self.semantics, opacities=self.opacities, scales=self.scales,
                                    rotations=self.rotations, cov3D_precomp=self.cov3D_precomp, dirs=self.dirs, inside=self.inside)
        self.fn = self.out.grad_fn
        torch.cuda.synchronize()

    def saved(self):
        st = self.fn.saved_tensors
        return dict(geom=st[9], binning=st[10], img=st[11], R=self.fn.num_rendered)

    def decode(self):
        import vcr_native as n
        sv = self.saved()
        P = self.means3D.shape[0]
        W, H = self.inp["W"], self.inp["H"]
        R = sv["R"]
        out = dict(R=R)
        if P == 0 or sv["geom"].numel() == 0:
            return out

        def grab(buf, addr, count, dtype):
            off = addr - buf.data_ptr()
            nbytes = count * np.dtype(dtype).itemsize
            return buf[off:off + nbytes].cpu().numpy().view(dtype).copy()

        # The production forward keeps bare tile ids as sort keys and never builds the reference's intermediate lists;
        # this call materialises them (duplicated keys/values in index order, point_offsets, 64-bit sorted keys) so that
        # they can be compared bit for bit.  It does not touch point_list or the tile ranges.
        if R > 0:
            n.check(n.vcr_binning_reference_lists(P, R, W, H, self.radii.data_ptr(), sv["geom"].data_ptr(), sv["binning"].data_ptr(),
                                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "binning_reference_lists")
            torch.cuda.synchronize()
        gv = n.GeomView()
        n.vcr_geom_layout(ctypes.c_void_p(sv["geom"].data_ptr()), P, ctypes.byref(gv))
        g = sv["geom"]
        out.update(depths=grab(g, gv.depths, P, np.float32), clamped=grab(g, gv.clamped, 3 * P, np.uint8).reshape(P, 3),
                   xy=grab(g, gv.means2D, 2 * P, np.float32).reshape(P, 2), xyz_view=grab(g, gv.xyz_view, 3 * P, np.float32).reshape(P, 3),
                   cov3D=grab(g, gv.cov3D, 6 * P, np.float32).reshape(P, 6),
                   conic_opacity=grab(g, gv.conic_opacity, 4 * P, np.float32).reshape(P, 4), rgb=grab(g, gv.rgb, 3 * P, np.float32).reshape(P, 3),
                   tiles_touched=grab(g, gv.tiles_touched, P, np.uint32), point_offsets=grab(g, gv.point_offsets, P, np.uint32))
        iv = n.ImageView()
        n.vcr_image_layout(ctypes.c_void_p(sv["img"].data_ptr()), W, H, ctypes.byref(iv))
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        out.update(final_T=grab(sv["img"], iv.accum_alpha, W * H, np.float32), n_contrib=grab(sv["img"], iv.n_contrib, W * H, np.uint32),
                   ranges=grab(sv["img"], iv.ranges, 2 * tiles, np.uint32).reshape(tiles, 2))
        if R > 0:
            bv = n.BinningView()
            n.vcr_binning_layout(ctypes.c_void_p(sv["binning"].data_ptr()), R, ctypes.byref(bv))
            b = sv["binning"]
            out.update(point_list=grab(b, bv.point_list, R, np.uint32), keys=grab(b, bv.point_list_keys, R, np.uint64),
                       vals_unsorted=grab(b, bv.point_list_unsorted, R, np.uint32),
                       keys_unsorted=grab(b, bv.point_list_keys_unsorted, R, np.uint64))
        return out

    def backward(self, dL_dout):
        self.out.backward(dL_dout.to(self.out.device))
        torch.cuda.synchronize()
        g = lambda x: None if x is None or x.grad is None else x.grad.detach().cpu().numpy()  # noqa: E731
        return dict(dL_dmeans3D=g(self.means3D), dL_dmeans2D=g(self.means2D), dL_dmeans2D_densify=g(self.means2D_densify),
                    dL_dsh=g(self.shs), dL_dcolors=g(self.colors_precomp), dL_dnormals=g(self.normals), dL_dobjects=g(self.semantics),
                    dL_dopacity=g(self.opacities), dL_dscales=g(self.scales), dL_drotations=g(self.rotations),
                    dL_dcov3D=g(self.cov3D_precomp))


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    d = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / d) if d > 0 else float(np.linalg.norm(a - b))
