"""numpy front-end of the CPU oracle (oracle/vcr_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference-cpu legs may import this module; the
product path (vcr-gaus_b200/) never does.  Each function restates one stage of the reference rasterizer; the
file:line citations live beside the C code.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libvcr_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "vcr_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libvcr_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = ctypes.CDLL(_LIB)
        _lib.orc_bin.restype = ctypes.c_int64
        _lib.orc_higher_msb.restype = ctypes.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def higher_msb(n):
    return int(lib().orc_higher_msb(ctypes.c_uint32(n)))


def preprocess(means3D, scales, scale_modifier, rotations, opacities, shs, cov3D_precomp, colors_precomp, viewmatrix,
               projmatrix, campos, W, H, tanfovx, tanfovy, D):
    P = means3D.shape[0]
    M = shs.shape[1] if shs is not None else 0
    out = dict(radii=np.zeros(P, np.int32), xy=np.zeros((P, 2), np.float32), xyz_view=np.zeros((P, 3), np.float32),
               depths=np.zeros(P, np.float32), cov3D=np.zeros((P, 6), np.float32), rgb=np.zeros((P, 3), np.float32),
               conic_opacity=np.zeros((P, 4), np.float32), tiles_touched=np.zeros(P, np.uint32),
               clamped=np.zeros((P, 3), np.uint8))
    args = [_f(means3D), _f(scales), _f(rotations), _f(opacities), _f(shs), _f(cov3D_precomp), _f(colors_precomp),
            _f(viewmatrix), _f(projmatrix), _f(campos)]
    lib().orc_preprocess(ctypes.c_int(P), ctypes.c_int(D), ctypes.c_int(M), _p(args[0]), _p(args[1]),
                         ctypes.c_float(scale_modifier), _p(args[2]), _p(args[3]), _p(args[4]), _p(args[5]), _p(args[6]),
                         _p(args[7]), _p(args[8]), _p(args[9]), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(tanfovx),
                         ctypes.c_float(tanfovy), _p(out["radii"]), _p(out["xy"]), _p(out["xyz_view"]), _p(out["depths"]),
                         _p(out["cov3D"]), _p(out["rgb"]), _p(out["conic_opacity"]), _p(out["tiles_touched"]),
                         _p(out["clamped"]))
    return out


def mark_visible(means3D, viewmatrix):
    P = means3D.shape[0]
    out = np.zeros(P, np.uint8)
    m, v = _f(means3D), _f(viewmatrix)
    lib().orc_mark_visible(ctypes.c_int(P), _p(m), _p(v), _p(out))
    return out.astype(bool)


def bin_instances(pre, W, H):
    P = pre["radii"].shape[0]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    offsets = np.zeros(P, np.uint32)
    common = (ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(pre["radii"]), _p(pre["xy"]), _p(pre["depths"]),
              _p(pre["tiles_touched"]), _p(offsets))
    R = int(lib().orc_bin(*common, None, None, None, None, None))
    n = max(R, 1)
    ku, vu = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    ks, vs = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    lib().orc_bin(*common, _p(ku), _p(vu), _p(ks), _p(vs), _p(ranges))
    return dict(R=R, point_offsets=offsets, keys_unsorted=ku[:R], vals_unsorted=vu[:R], keys=ks[:R], point_list=vs[:R],
                ranges=ranges)


def render_forward(mode, W, H, pre, binned, features, normals, obj, dirs, bg, inside):
    P = pre["radii"].shape[0]
    HW = H * W
    final_T, n_contrib = np.zeros(HW, np.float32), np.zeros(HW, np.uint32)
    out_color = np.zeros((10 if mode == 0 else 3, H, W), np.float32)
    count, score = np.zeros(P, np.int32), np.zeros(P, np.float32)
    a = [_f(features), _f(normals), _f(obj), _f(dirs), _f(bg), None if inside is None else np.ascontiguousarray(inside, np.uint8)]
    pl = binned["point_list"] if binned["R"] > 0 else np.zeros(1, np.uint32)
    lib().orc_render_forward(ctypes.c_int(mode), ctypes.c_int(W), ctypes.c_int(H), _p(binned["ranges"]), _p(pl), _p(pre["xy"]),
                             _p(pre["xyz_view"]), _p(a[0]), _p(pre["depths"]), _p(a[1]), _p(a[2]), _p(a[3]),
                             _p(pre["conic_opacity"]), _p(a[4]), _p(a[5]), _p(final_T), _p(n_contrib), _p(out_color), _p(count),
                             _p(score))
    return dict(out_color=out_color, final_T=final_T, n_contrib=n_contrib, gaussians_count=count, important_score=score)


def rasterize_forward(inp, mode=0):
    """inp: dict of numpy arrays as produced by vcr_synth.raster_inputs (+ optional colors_precomp / cov3D_precomp / inside)."""
    W, H = inp["W"], inp["H"]
    pre = preprocess(inp["means3D"], inp.get("scales"), inp.get("scale_modifier", 1.0), inp.get("rotations"), inp["opacities"],
                     inp.get("shs"), inp.get("cov3D_precomp"), inp.get("colors_precomp"), inp["viewmatrix"], inp["projmatrix"],
                     inp["campos"], W, H, inp["tanfovx"], inp["tanfovy"], inp.get("sh_degree", 3))
    binned = bin_instances(pre, W, H)
    feats = inp["colors_precomp"] if inp.get("colors_precomp") is not None else pre["rgb"]
    P = inp["means3D"].shape[0]
    inside = inp.get("inside")
    if inside is None:
        inside = np.ones(P, np.uint8)
    ren = render_forward(mode, W, H, pre, binned, feats, inp.get("normals"), inp.get("semantics"), inp.get("dirs"), inp["bg"], inside)
    return dict(pre=pre, binned=binned, inside=inside, features=feats, **ren)


def rasterize_backward(inp, fwd, dL_dpix):
    """Hand-derived backward following the reference formulas (not autograd)."""
    W, H = inp["W"], inp["H"]
    P = inp["means3D"].shape[0]
    pre, binned = fwd["pre"], fwd["binned"]
    out = fwd["out_color"]
    g = {k: np.zeros(s, np.float64) for k, s in dict(mean2D=(P, 2), absmean2D=(P, 2), conic=(P, 3), opacity=(P,), color=(P, 3),
                                                      depth=(P,), normal=(P, 3), obj=(P, 2), xyz_view=(P, 3)).items()}
    a = [_f(inp["bg"]), _f(fwd["features"]), _f(inp["normals"]), _f(inp.get("semantics")), _f(inp.get("dirs")),
         _f(out[7]), _f(out[3]), _f(dL_dpix), np.ascontiguousarray(fwd["inside"], np.uint8)]
    pl = binned["point_list"] if binned["R"] > 0 else np.zeros(1, np.uint32)
    lib().orc_render_backward(ctypes.c_int(W), ctypes.c_int(H), _p(binned["ranges"]), _p(pl), _p(a[0]), _p(pre["xy"]),
                              _p(pre["xyz_view"]), _p(pre["conic_opacity"]), _p(a[1]), _p(pre["depths"]), _p(a[2]), _p(a[3]),
                              _p(a[4]), _p(a[5]), _p(a[6]), _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(a[7]), _p(a[8]),
                              _p(g["mean2D"]), _p(g["absmean2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["color"]), _p(g["depth"]),
                              _p(g["normal"]), _p(g["obj"]), _p(g["xyz_view"]))
    shs = _f(inp.get("shs")) if inp.get("colors_precomp") is None else None
    M = shs.shape[1] if shs is not None else 0
    f32 = {k: v.astype(np.float32) for k, v in g.items()}
    res = dict(dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
               dL_dsh=np.zeros((P, max(M, 0), 3), np.float32), dL_dscales=np.zeros((P, 3), np.float32),
               dL_drotations=np.zeros((P, 4), np.float32))
    cov3D = _f(inp["cov3D_precomp"]) if inp.get("cov3D_precomp") is not None else pre["cov3D"]
    scales = _f(inp.get("scales")) if inp.get("cov3D_precomp") is None else None
    rots = _f(inp.get("rotations")) if inp.get("cov3D_precomp") is None else None
    b = [_f(inp["means3D"]), _f(inp["viewmatrix"]), _f(inp["projmatrix"]), _f(inp["campos"])]
    lib().orc_preprocess_backward(ctypes.c_int(P), ctypes.c_int(inp.get("sh_degree", 3)), ctypes.c_int(M), _p(b[0]), _p(pre["radii"]),
                                  _p(shs), _p(pre["clamped"]), _p(scales), _p(rots), ctypes.c_float(inp.get("scale_modifier", 1.0)),
                                  _p(cov3D), _p(b[1]), _p(b[2]), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(inp["tanfovx"]),
                                  ctypes.c_float(inp["tanfovy"]), _p(b[3]), ctypes.c_int(inp.get("dirs") is not None),
                                  _p(f32["mean2D"]), _p(f32["conic"]), _p(f32["xyz_view"]), _p(f32["depth"]), _p(f32["color"]),
                                  _p(res["dL_dmeans3D"]), _p(res["dL_dcov3D"]), _p(res["dL_dsh"]), _p(res["dL_dscales"]),
                                  _p(res["dL_drotations"]))
    pad = np.zeros((P, 1), np.float32)
    res.update(dL_dmeans2D=np.concatenate([f32["mean2D"], pad], 1), dL_dmeans2D_densify=np.concatenate([f32["absmean2D"], pad], 1),
               dL_dcolors=f32["color"], dL_dnormals=f32["normal"], dL_dobjects=f32["obj"], dL_dopacity=f32["opacity"][:, None],
               dL_dconic=f32["conic"], dL_dxyz_view=f32["xyz_view"], dL_ddepth=f32["depth"])
    return res


def dnormal(depth, gt_normal, render_normal, mask, K4, exp_t=0.0, mask_depth_max=0.0, want_grad=True):
    H, W = depth.shape[-2:]
    d = _f(depth.reshape(H, W))
    est = np.zeros((H, W, 3), np.float64)
    sums = np.zeros(3, np.float64)
    gd = np.zeros((H, W), np.float64) if want_grad else None
    g, r = _f(gt_normal), _f(render_normal)
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    fx, fy, cx, cy = K4
    lib().orc_dnormal(_p(d), _p(g), _p(r), _p(m), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(fx), ctypes.c_float(fy),
                      ctypes.c_float(cx), ctypes.c_float(cy), ctypes.c_float(exp_t), ctypes.c_float(mask_depth_max), _p(est),
                      _p(sums), _p(gd))
    loss = (sums[0] + sums[1]) / sums[2] if sums[2] > 0 else 0.0
    return dict(est_normal=est, sums=sums, loss=loss, dL_ddepth=gd)


def knn3(points):
    P = points.shape[0]
    out = np.zeros(P, np.float32)
    p = _f(points)
    lib().orc_knn3(ctypes.c_int(P), _p(p), _p(out))
    return out
