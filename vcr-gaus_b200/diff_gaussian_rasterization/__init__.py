"""Drop-in replacement for the VCR-GauS fork of ``diff_gaussian_rasterization``.

Same public surface as the reference binding (DGR/diff_gaussian_rasterization/__init__.py):
``GaussianRasterizationSettings`` (:461-474), ``GaussianRasterizer`` (:476-540) with ``markVisible`` and the
13-argument ``forward`` returning ``(outmap f32[10,H,W], radii i32[P])``, and ``rasterize_gaussians`` with the
``f_count`` dispatch (:26-94).  Below it, instead of the pybind11 module ``_C`` + torch C++ glue
(DGR/ext.cpp, DGR/rasterize_points.cu), calls go through ctypes into libvcr_gaus_b200.so
(include/vcr_gaus_b200.h).  Python owns every byte of device memory; kernels run on torch's current stream.

There is no CPU path: non-CUDA tensors raise.
"""
import ctypes
from typing import NamedTuple

import torch
import torch.nn as nn

import vcr_native as _n

_LAST_NUM_RENDERED = 0
NUM_CHANNELS = 3
NUM_OBJECTS = 2
NUM_DIST = 0        # DGR/cuda_rasterizer/config.h:17
NUM_DEPTH2 = 0      # DGR/cuda_rasterizer/config.h:18
OUTPUT_CHANNELS = 3 + 1 + 3 + 1 + NUM_OBJECTS + NUM_DIST + NUM_DEPTH2  # DGR/cuda_rasterizer/auxiliary.h:28


def set_variant(num_dist=0, num_depth2=0):
    """What editing NUM_DIST / NUM_DEPTH2 in cuda_rasterizer/config.h and reinstalling does for the reference
    (README.md:150-155), as a process-wide setting: num_dist in (0, 1) adds the depth-distortion channel, num_depth2 in (0, 2)
    the two NDC-depth sums that render() turns into `depth_var` (gaussian_renderer/__init__.py:154-158).  The output map of
    f_count = 0 then has OUTPUT_CHANNELS = 10 + num_dist + num_depth2 channels.  Returns OUTPUT_CHANNELS."""
    global NUM_DIST, NUM_DEPTH2, OUTPUT_CHANNELS
    if num_dist not in (0, 1) or num_depth2 not in (0, 2):
        raise ValueError("NUM_DIST must be 0 or 1 and NUM_DEPTH2 0 or 2")
    NUM_DIST, NUM_DEPTH2 = int(num_dist), int(num_depth2)
    OUTPUT_CHANNELS = int(_n.vcr_variant_channels(NUM_DIST, NUM_DEPTH2))
    return OUTPUT_CHANNELS


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


class _Buffers:
    """The three opaque per-view buffers, grown on request of the library (resizeFunctional, rasterize_points.cu:28-34)."""

    def __init__(self, device):
        self.device = device
        held = self.t = {}

        # the callbacks close over the dict, not over `self`: no reference cycle, so the buffers are released by reference
        # counting as soon as the autograd node drops them (a cycle would keep ~600 MB per view alive until the next pass
        # of the cyclic garbage collector)
        def make(name):
            def alloc(_user, nbytes):
                t = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
                held[name] = t
                return t.data_ptr()
            return alloc
        self.cb = {name: _n.ALLOC_FN(make(name)) for name in ("geom", "binning", "img")}

    def get(self, name):
        return self.t.get(name, torch.empty(0, dtype=torch.uint8, device=self.device))


def _f32c(t, align16=False):
    """contiguous float32 view of an input, the way rasterize_points.cu calls .contiguous() on everything (:119-146).
    align16: the kernels read this array as float4 (rotations, shs); a contiguous view that does not start on a 16-byte
    boundary (a slice of a larger tensor) is copied into a fresh allocation instead of being rejected by the library."""
    if t is None:
        return None
    if t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError("diff_gaussian_rasterization (B200 build) needs CUDA tensors; there is no CPU fallback")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected float32, got {t.dtype}")
    t = t.contiguous()
    if align16 and t.data_ptr() % 16:
        t = t.clone()
    return t


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check_means(means3D):
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:75-77


def _settings_tensors(rs, device):
    bg = _f32c(rs.bg.to(device)) if rs.bg is not None else None
    return bg, _f32c(rs.viewmatrix.to(device)), _f32c(rs.projmatrix.to(device)), _f32c(rs.campos.to(device))


def rasterize_gaussians(means3D, means2D, means2D_densify, sh, colors_precomp, normals_precomp, semantics_precomp,
                        opacities, scales, rotations, cov3Ds_precomp, dirs, inside, raster_settings):
    if raster_settings.f_count == 0:
        return _RasterizeGaussians.apply(means3D, means2D, means2D_densify, sh, colors_precomp, normals_precomp,
                                         semantics_precomp, opacities, scales, rotations, cov3Ds_precomp, dirs, inside,
                                         raster_settings)
    elif raster_settings.f_count == 1:
        return _RasterizeGaussians.forward_count(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                                 cov3Ds_precomp, raster_settings)
    elif raster_settings.f_count == 2:
        return _RasterizeGaussians.forward_visi(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                                cov3Ds_precomp, raster_settings)
    elif raster_settings.f_count == 3:
        return _RasterizeGaussians.forward_visi_acc(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                                    cov3Ds_precomp, raster_settings)
    raise ValueError(f"unknown f_count {raster_settings.f_count}")


def _forward_call(args):
    if NUM_DIST or NUM_DEPTH2:
        return _n.vcr_rasterize_forward_variant(NUM_DIST, NUM_DEPTH2, *args)
    return _n.vcr_rasterize_forward(*args)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, means2D_densify, sh, colors_precomp, normals_precomp, semantics_precomp, opacities,
                scales, rotations, cov3Ds_precomp, dirs, inside, raster_settings):
        rs = raster_settings
        _check_means(means3D)
        device = means3D.device
        P = means3D.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        outmap = torch.empty((OUTPUT_CHANNELS, H, W), dtype=torch.float32, device=device)
        radii = torch.empty((P,), dtype=torch.int32, device=device)
        bufs = _Buffers(device)
        num_rendered = 0
        if P == 0:  # rasterize_points.cu:106
            outmap.zero_()
        else:
            means3D_c, sh_c, colors_c = _f32c(means3D), _f32c(sh, True), _f32c(colors_precomp)
            normals_c, sem_c, opac_c = _f32c(normals_precomp), _f32c(semantics_precomp), _f32c(opacities)
            scales_c, rot_c, cov_c, dirs_c = _f32c(scales), _f32c(rotations, True), _f32c(cov3Ds_precomp), _f32c(dirs)
            inside_c = None
            if inside is not None:              # None = every Gaussian inside (the library's fast path; the reference passes ones)
                inside_c = inside.contiguous()
                if inside_c.dtype != torch.bool:
                    raise RuntimeError("inside must be a bool tensor")
            bg, view, proj, campos = _settings_tensors(rs, device)
            M = sh_c.shape[1] if sh_c is not None and sh_c.numel() != 0 else 0
            n_host = ctypes.c_int(0)
            args = (bufs.cb["geom"], bufs.cb["binning"], bufs.cb["img"], None, P, int(rs.sh_degree), M, _n.ptr(bg), W, H,
                    _n.ptr(means3D_c), _n.ptr(sh_c), _n.ptr(colors_c), _n.ptr(normals_c), _n.ptr(sem_c), _n.ptr(opac_c),
                    _n.ptr(scales_c), float(rs.scale_modifier), _n.ptr(rot_c), _n.ptr(cov_c), _n.ptr(view), _n.ptr(proj),
                    _n.ptr(dirs_c), _n.ptr(campos), float(rs.tanfovx), float(rs.tanfovy), int(bool(rs.prefiltered)),
                    outmap.data_ptr(), radii.data_ptr(), _n.ptr(inside_c), int(bool(rs.debug)), ctypes.byref(n_host),
                    _stream())
            if rs.debug:
                cpu_args = cpu_deep_copy_tuple((rs.bg, means3D, colors_precomp, normals_precomp, semantics_precomp, opacities,
                                                scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
                                                rs.projmatrix, dirs, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
                                                sh, rs.sh_degree, rs.campos, rs.prefiltered, inside, rs.debug))
                try:
                    _n.check(_forward_call(args), "rasterize_gaussians")
                except Exception as ex:
                    torch.save(cpu_args, "snapshot_fw.dump")
                    print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                    raise ex
            else:
                _n.check(_forward_call(args), "rasterize_gaussians")
            num_rendered = n_host.value
            global _LAST_NUM_RENDERED
            _LAST_NUM_RENDERED = num_rendered  # bench.py reads this to report R
        geomBuffer, binningBuffer, imgBuffer = bufs.get("geom"), bufs.get("binning"), bufs.get("img")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.variant = (NUM_DIST, NUM_DEPTH2)
        # Optional zero-copy gradient destinations (vcr_view_parallel.GaussianBucket): a parameter may carry a
        # `_vcr_grad_out` tensor; the backward kernel then writes that gradient straight into it (every element is
        # written) and autograd adopts a view of it as `.grad` instead of allocating + accumulating.
        ctx.grad_out = tuple(getattr(t, "_vcr_grad_out", None) for t in (means3D, sh, opacities, scales, rotations))
        # Optional chunk hook (vcr_view_parallel.ViewParallel.overlap_gradients): `means3D._vcr_chunk_hook = (n_chunks, fn)`
        # makes the backward run its per-Gaussian stage in n_chunks launches and call fn(chunk, first, count) after each.
        ctx.chunk_hook = getattr(means3D, "_vcr_chunk_hook", None)
        # Optional compact SH gradient (vcr_view_parallel.ViewParallel.compact_sh_exchange): `sh._vcr_sh_compact` is a
        # float32 [P,3] destination for the clamp-masked colour gradient; the backward then does NOT write dL_dsh (the
        # returned tensor is filled after the exchange by vcr_sh_grad_from_colors).
        ctx.sh_compact = getattr(sh, "_vcr_sh_compact", None)
        depth, alpha = outmap[3:4], outmap[7:8]
        ctx.save_for_backward(colors_precomp, normals_precomp, semantics_precomp, means3D, scales, rotations, cov3Ds_precomp,
                              radii, sh, geomBuffer, binningBuffer, imgBuffer, alpha, depth, dirs, inside)
        ctx.mark_non_differentiable(radii)
        return outmap, radii

    @staticmethod
    def backward(ctx, grad_color, _):
        rs = ctx.raster_settings
        (colors_precomp, normals_precomp, semantics_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
         binningBuffer, imgBuffer, alpha, depth, dirs, inside) = ctx.saved_tensors
        device = means3D.device
        P = means3D.shape[0]
        H, W = grad_color.shape[1], grad_color.shape[2]
        M = sh.shape[1] if sh.numel() != 0 else 0
        f = dict(dtype=torch.float32, device=device)
        def dest(i, shape):
            buf = ctx.grad_out[i] if hasattr(ctx, "grad_out") else None
            if (buf is not None and buf.is_cuda and buf.dtype == torch.float32 and buf.is_contiguous()
                    and tuple(buf.shape) == tuple(shape) and buf.data_ptr() % 16 == 0):
                return buf.view(shape)   # a fresh view: autograd can adopt it without a copy
            return torch.empty(shape, **f)

        g_means2D = torch.empty((P, 3), **f)
        g_means2D_densify = torch.empty((P, 3), **f)
        # gradients of optional inputs are only produced when the input was given (the library skips the writes otherwise)
        g_colors = torch.empty((P, NUM_CHANNELS), **f) if colors_precomp.numel() != 0 else None
        g_normals = torch.empty((P, 3), **f)
        g_objects = torch.empty((P, NUM_OBJECTS), **f) if semantics_precomp.numel() != 0 else None
        g_opacity = dest(2, (P, 1))
        g_means3D = dest(0, (P, 3))
        g_cov3D = torch.empty((P, 6), **f) if cov3Ds_precomp.numel() != 0 else None
        g_sh = dest(1, (P, M, 3))
        g_scales = dest(3, (P, 3))
        g_rot = dest(4, (P, 4))
        if P != 0:
            scratch = torch.empty(int(_n.vcr_backward_scratch_bytes(P)), dtype=torch.uint8, device=device)
            bg, view, proj, campos = _settings_tensors(rs, device)
            grad_c = _f32c(grad_color)
            alpha_c, depth_c = alpha.contiguous(), depth.contiguous()
            sh_c, rot_c = _f32c(sh, True), _f32c(rotations, True)    # (held until the launch has been enqueued)
            args = (P, int(rs.sh_degree), M, int(ctx.num_rendered), _n.ptr(bg), W, H, _n.ptr(_f32c(means3D)), _n.ptr(sh_c),
                    _n.ptr(_f32c(colors_precomp)), _n.ptr(_f32c(normals_precomp)), _n.ptr(_f32c(semantics_precomp)),
                    alpha_c.data_ptr(), depth_c.data_ptr(), _n.ptr(_f32c(scales)), float(rs.scale_modifier),
                    _n.ptr(rot_c), _n.ptr(_f32c(cov3Ds_precomp)), _n.ptr(view), _n.ptr(proj), _n.ptr(_f32c(dirs)),
                    _n.ptr(campos), float(rs.tanfovx), float(rs.tanfovy), radii.data_ptr(), geomBuffer.data_ptr(),
                    binningBuffer.data_ptr(), imgBuffer.data_ptr(), grad_c.data_ptr(), g_means2D.data_ptr(),
                    g_means2D_densify.data_ptr(), g_opacity.data_ptr(), _n.ptr(g_colors), g_normals.data_ptr(),
                    _n.ptr(g_objects), g_means3D.data_ptr(), _n.ptr(g_cov3D), _n.ptr(g_sh), g_scales.data_ptr(),
                    g_rot.data_ptr(), _n.ptr(inside.contiguous() if inside is not None else None), scratch.data_ptr(), int(bool(rs.debug)), _stream())
            variant = getattr(ctx, "variant", (0, 0))
            if variant != (0, 0):
                if grad_c.shape[0] != _n.vcr_variant_channels(*variant):
                    raise RuntimeError("gradient of the output map has the wrong number of channels for this variant")
                _n.check(_n.vcr_rasterize_backward_variant(variant[0], variant[1], *args), "rasterize_gaussians_backward")
            elif rs.debug:
                cpu_args = cpu_deep_copy_tuple((rs.bg, means3D, radii, colors_precomp, normals_precomp, semantics_precomp, scales,
                                                rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, dirs,
                                                rs.tanfovx, rs.tanfovy, grad_color, sh, rs.sh_degree, rs.campos, geomBuffer,
                                                ctx.num_rendered, binningBuffer, imgBuffer, alpha, depth, inside, rs.debug))
                try:
                    _n.check(_n.vcr_rasterize_backward(*args), "rasterize_gaussians_backward")
                except Exception as ex:
                    torch.save(cpu_args, "snapshot_bw.dump")
                    print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                    raise ex
            elif getattr(ctx, "chunk_hook", None) is not None or getattr(ctx, "sh_compact", None) is not None:
                n_chunks, hook = ctx.chunk_hook if getattr(ctx, "chunk_hook", None) is not None else (1, lambda *a: None)
                compact = getattr(ctx, "sh_compact", None)
                if compact is not None and not (compact.is_cuda and compact.dtype == torch.float32 and compact.is_contiguous()
                                                and compact.numel() == 3 * P):
                    raise RuntimeError("_vcr_sh_compact must be a contiguous float32 CUDA tensor with 3 floats per Gaussian")
                failure = []
                def _cb(_user, chunk, first, count):
                    try:
                        hook(chunk, first, count)
                    except BaseException as ex:      # exceptions cannot cross the C frame; re-raised below
                        failure.append(ex)
                cb = _n.CHUNK_FN(_cb)
                _n.check(_n.vcr_rasterize_backward_chunked(*args, int(n_chunks), cb, None, _n.ptr(compact)),
                         "rasterize_gaussians_backward")
                if failure:
                    raise failure[0]
            else:
                _n.check(_n.vcr_rasterize_backward(*args), "rasterize_gaussians_backward")
        else:
            for t in (g_means2D, g_means2D_densify, g_colors, g_normals, g_objects, g_opacity, g_means3D, g_cov3D, g_sh,
                      g_scales, g_rot):
                if t is not None:
                    t.zero_()

        def match(g, ref):
            # inputs that were absent (zero-sized placeholders) get no gradient
            return g if (g is not None and ref.numel() != 0) else None

        need = ctx.needs_input_grad      # means2D / means2D_densify may have been passed as None
        return (g_means3D, g_means2D if need[1] else None, g_means2D_densify if need[2] else None, match(g_sh, sh), match(g_colors, colors_precomp),
                match(g_normals, normals_precomp), match(g_objects, semantics_precomp), g_opacity, match(g_scales, scales),
                match(g_rot, rotations), match(g_cov3D, cov3Ds_precomp), None, None, None)

    # ---- pruning / visibility variants (no autograd), DGR __init__.py:233-388 ----
    @staticmethod
    def _aux(mode, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
        _check_means(means3D)
        device = means3D.device
        P = means3D.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        radii = torch.zeros((P,), dtype=torch.int32, device=device)
        count = torch.zeros((P,), dtype=torch.int32, device=device)
        score = torch.zeros((P,), dtype=torch.float32, device=device)
        color = torch.zeros((NUM_CHANNELS, H, W), dtype=torch.float32, device=device)
        if P == 0:
            return count, score, color, radii
        bufs = _Buffers(device)
        bg, view, proj, campos = _settings_tensors(rs, device)
        sh_c = _f32c(sh, True)
        M = sh_c.shape[1] if sh_c is not None and sh_c.numel() != 0 else 0
        n_host = ctypes.c_int(0)
        rot_c = _f32c(rotations, True)
        common = (_n.ptr(_f32c(means3D)), _n.ptr(sh_c), _n.ptr(_f32c(colors_precomp)), _n.ptr(_f32c(opacities)),
                  _n.ptr(_f32c(scales)), float(rs.scale_modifier), _n.ptr(rot_c), _n.ptr(_f32c(cov3Ds_precomp)),
                  _n.ptr(view), _n.ptr(proj), _n.ptr(campos), float(rs.tanfovx), float(rs.tanfovy), int(bool(rs.prefiltered)))
        if mode == 1:
            _n.check(_n.vcr_rasterize_count(bufs.cb["geom"], bufs.cb["binning"], bufs.cb["img"], None, P, int(rs.sh_degree), M,
                                            _n.ptr(bg), W, H, *common, color.data_ptr(), count.data_ptr(), score.data_ptr(),
                                            radii.data_ptr(), int(bool(rs.debug)), ctypes.byref(n_host), _stream()),
                     "count_gaussians")
        else:
            _n.check(_n.vcr_rasterize_visi_acc(bufs.cb["geom"], bufs.cb["binning"], bufs.cb["img"], None, P, int(rs.sh_degree), M,
                                               W, H, *common, count.data_ptr(), radii.data_ptr(), int(bool(rs.debug)),
                                               ctypes.byref(n_host), _stream()),
                     "visi_gaussians_acc")
        return count, score, color, radii

    @staticmethod
    def forward_count(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        assert raster_settings.f_count == 1
        with torch.no_grad():
            return _RasterizeGaussians._aux(1, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                            raster_settings)

    @staticmethod
    def forward_visi(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        # f_count == 2 (renderCUDA_visi) has no caller in the reference and is not built.
        raise NotImplementedError("f_count=2 (visi_gaussians) is not part of the B200 build; use f_count=3")

    @staticmethod
    def forward_visi_acc(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        assert raster_settings.f_count == 3
        with torch.no_grad():
            count, _, _, radii = _RasterizeGaussians._aux(3, means3D, sh, colors_precomp, opacities, scales, rotations,
                                                          cov3Ds_precomp, raster_settings)
        return count, radii


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    f_count: int = 0


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: view-space z > 0.2 (checkFrustum, rasterizer_impl.cu:54-66)."""
        with torch.no_grad():
            rs = self.raster_settings
            P = positions.shape[0]
            present = torch.zeros((P,), dtype=torch.bool, device=positions.device)
            if P != 0:
                pos = _f32c(positions)
                view = _f32c(rs.viewmatrix.to(positions.device))
                proj = _f32c(rs.projmatrix.to(positions.device))
                _n.check(_n.vcr_mark_visible(P, pos.data_ptr(), view.data_ptr(), proj.data_ptr(), present.data_ptr(), _stream()),
                         "mark_visible")
        return present

    def forward(self, means3D, means2D, means2D_densify=None, opacities=None, shs=None, colors_precomp=None,
                normals_precomp=None, semantics_precomp=None, scales=None, rotations=None, cov3D_precomp=None, dirs=None,
                inside=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if opacities is None:
            raise Exception('opacities are required')
        dev = means3D.device
        empty = lambda: torch.empty(0, dtype=torch.float32, device=dev)  # noqa: E731
        if means2D_densify is None:
            # plain-3DGS call style (north star: GaussianRasterizer(means3D, means2D, shs, colors_precomp, opacities, ...)):
            # a private carrier keeps the autograd arity of the fork
            means2D_densify = torch.zeros_like(means3D)
        shs = empty() if shs is None else shs
        colors_precomp = empty() if colors_precomp is None else colors_precomp
        semantics_precomp = empty() if semantics_precomp is None else semantics_precomp
        scales = empty() if scales is None else scales
        rotations = empty() if rotations is None else rotations
        cov3D_precomp = empty() if cov3D_precomp is None else cov3D_precomp
        dirs = empty() if dirs is None else dirs
        # inside=None stays None: the library treats a NULL mask as "all inside" (the reference allocates a tensor of ones per
        # call, __init__.py:521-522) and takes its two-pixels-per-thread backward
        if raster_settings.f_count == 0:
            if normals_precomp is None:
                # The reference kernel dereferences `normals` unconditionally (forward.cu:374): with
                # None it would read through a null pointer.  Zero normals give a zero normal map and, in plain depth
                # mode, the same colour/depth/alpha as the original 3DGS rasterizer.
                normals_precomp = torch.zeros_like(means3D)
                if dirs.numel() != 0:
                    raise Exception('normals_precomp is required when dirs (ray/plane intersection depth) is given')
        else:
            normals_precomp = empty() if normals_precomp is None else normals_precomp
        return rasterize_gaussians(means3D, means2D, means2D_densify, shs, colors_precomp, normals_precomp, semantics_precomp,
                                   opacities, scales, rotations, cov3D_precomp, dirs, inside, raster_settings)

    # The reference's forward_count / forward_visi / forward_visi_acc methods pass 9 arguments to the 14-parameter
    # rasterize_gaussians and raise TypeError (DGR __init__.py:543-648); the working route is forward() with
    # raster_settings.f_count set, which is what these do.
    def _with_f_count(self, k):
        return GaussianRasterizer(self.raster_settings._replace(f_count=k))

    def forward_count(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                      cov3D_precomp=None):
        return self._with_f_count(1).forward(means3D, means2D, None, opacities, shs, colors_precomp, None, None, scales,
                                             rotations, cov3D_precomp)

    def forward_visi_acc(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                         cov3D_precomp=None):
        return self._with_f_count(3).forward(means3D, means2D, None, opacities, shs, colors_precomp, None, None, scales,
                                             rotations, cov3D_precomp)
