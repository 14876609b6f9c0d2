"""CPU-only checks of the C-ABI library and the host-side mirror of the reference interface (no compute calls)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    import vcr_native as n
    header = open(os.path.join(ROOT, "include", "vcr_gaus_b200.h")).read()
    declared = set(re.findall(r"\b(vcr_[a-z0-9_]+)\s*\(", header)) - {"vcr_alloc_fn"}
    assert declared, "no declarations found in the header"
    assert declared == set(n.DECLARED_SYMBOLS), declared ^ set(n.DECLARED_SYMBOLS)
    for name in declared:
        assert hasattr(n.lib, name), f"{name} is declared in include/vcr_gaus_b200.h but not exported"
    assert n.vcr_abi_version() == 1


def test_host_side_size_queries_and_msb():
    import vcr_native as n
    from oracle import oracle as O
    for tiles in (1, 2, 255, 256, 257, 1900, 4108, 8160, 32400, 65535, 65536, 1 << 20):
        assert n.vcr_higher_msb(tiles) == O.higher_msb(tiles)
    # documented values: 1080p -> 13, 4K -> 15, 800x600 -> 11, 256^2 -> 9
    assert [n.vcr_higher_msb(t) for t in (8160, 32400, 1900, 256)] == [13, 15, 11, 9]
    g1, g2 = n.vcr_geom_bytes(1000), n.vcr_geom_bytes(2000)
    assert 0 < g1 < g2 and g2 < 2 * g1 + 8192
    assert n.vcr_geom_bytes(1_500_000) < 130 * 1_500_000  # ~116 B per Gaussian (reference: ~115 B)
    assert n.vcr_image_bytes(1920, 1080) >= 8 * 1920 * 1080
    assert n.vcr_binning_bytes(0) > 0 and n.vcr_binning_bytes(1_000_000) >= 36 * 1_000_000
    assert n.vcr_backward_scratch_bytes(1000) >= 1000 * 96
    assert n.vcr_knn_scratch_bytes(1000) > n.vcr_binning_bytes(1000)


def test_layout_accessors_are_consistent():
    import vcr_native as n
    P = 1234
    gv = n.GeomView()
    base = 1 << 20
    n.vcr_geom_layout(ctypes.c_void_p(base), P, ctypes.byref(gv))
    ptrs = [getattr(gv, f) for f, _ in gv._fields_]
    assert all(p % 256 == 0 for p in ptrs) and ptrs == sorted(ptrs)
    assert ptrs[-1] + 4 * P <= base + n.vcr_geom_bytes(P)
    bv = n.BinningView()
    n.vcr_binning_layout(ctypes.c_void_p(base), 5000, ctypes.byref(bv))
    assert bv.point_list == base and bv.point_list_keys >= base + 4 * 5000


def test_api_surface_and_argument_errors():
    import diff_gaussian_rasterization as D
    fields = D.GaussianRasterizationSettings._fields
    assert fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
                      "sh_degree", "campos", "prefiltered", "debug", "f_count")
    rs = D.GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3, torch.zeros(3), False,
                                         False, 0)
    r = D.GaussianRasterizer(rs)
    assert r.raster_settings is rs
    m = torch.zeros((4, 3))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, m, torch.zeros((4, 1)), scales=m, rotations=torch.zeros((4, 4)))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, m, torch.zeros((4, 1)), shs=torch.zeros((4, 16, 3)), colors_precomp=m, scales=m, rotations=torch.zeros((4, 4)))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, m, torch.zeros((4, 1)), shs=torch.zeros((4, 16, 3)))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, m, torch.zeros((4, 1)), shs=torch.zeros((4, 16, 3)), scales=m, rotations=torch.zeros((4, 4)), cov3D_precomp=torch.zeros((4, 6)))
    # no CPU fallback: CPU tensors are refused loudly
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        r(m, m, m, torch.zeros((4, 1)), shs=torch.zeros((4, 16, 3)), normals_precomp=m, scales=m, rotations=torch.zeros((4, 4)))
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(torch.zeros((4, 2)), m, m, torch.zeros((4, 1)), shs=torch.zeros((4, 16, 3)), normals_precomp=m, scales=m,
          rotations=torch.zeros((4, 4)))
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        distCUDA2(torch.zeros((10, 3)))
    from vcr_dnormal import depth_normal_loss
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        depth_normal_loss(torch.ones((1, 8, 8)), (8.0, 8.0, 4.0, 4.0), torch.zeros((8, 8, 3)))


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "vcr-gaus_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                hits = re.findall(r"import\s+oracle|from\s+oracle|libvcr_oracle|vcr_oracle|orc_[a-z_]+\s*\(|oracle[/\\]", src)
                assert not hits, f"{f} reaches into the oracle: {hits}"


def test_synthetic_scene_is_deterministic_and_well_formed():
    import vcr_synth as S
    a, b = S.make_scene(500, seed=3), S.make_scene(500, seed=3)
    for x, y in zip((a.means3D, a.scales, a.rotations, a.opacities, a.shs), (b.means3D, b.scales, b.rotations, b.opacities, b.shs)):
        assert torch.equal(x, y)
    assert torch.allclose(a.rotations.norm(dim=1), torch.ones(500), atol=1e-6)
    assert (a.opacities > 0).all() and (a.opacities < 1).all() and (a.scales > 0).all()
    cam = S.make_camera(64, 48, idx=2)
    w2c = cam.world_view_transform.t()
    assert torch.allclose(w2c[:3, :3] @ w2c[:3, :3].t(), torch.eye(3), atol=1e-5)
    # camera looks at the origin: the origin projects to the image centre with positive depth
    o = torch.tensor([0.0, 0.0, 0.0, 1.0]) @ cam.full_proj_transform
    assert abs(float(o[0] / o[3])) < 1e-4 and abs(float(o[1] / o[3])) < 1e-4 and float(o[3]) > 0
    d = S.pixel_dirs(cam)
    assert d.shape == (3, 48, 64) and torch.allclose(d.norm(dim=0), torch.ones(48, 64), atol=1e-5)
