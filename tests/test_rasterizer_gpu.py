"""GPU parity of the CUDA rasterizer (through GaussianRasterizer -> ctypes -> C ABI) against the CPU oracle.

Gates (BASELINE.json north_star):
  * radii, tiles_touched, point offsets, duplicated keys, sorted keys / point list, tile ranges: BIT-EXACT
  * forward maps (colour, depth, normal, alpha, semantics): <= 1e-5 absolute.  exp() on the GPU (MUFU.EX2) is not
    reproducible on a CPU, so a handful of pixels may flip one of the threshold tests (alpha < 1/255, T < 1e-4,
    depth < 0.2) relative to the oracle; at most FLIP_FRAC of the pixels may exceed 1e-5, and those by <= FLIP_ABS.
    (Against the reference CUDA build, test_reference_ab_gpu.py, no such allowance is made.)
  * gradients: <= 1e-3 relative L2 (float atomics reorder the sums).
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from util import CudaRun, make_case, np_inputs, rel_l2, settings_from

pytestmark = pytest.mark.gpu

ABS_TOL = 1e-5
FLIP_FRAC = 2e-4
FLIP_ABS = 2e-2
GRAD_TOL = 1e-3

CASES = [
    # P, W, H, use_dirs, use_sem, sh_degree
    (10000, 256, 256, True, False, 3),     # BASELINE config A
    (10000, 256, 256, False, True, 3),     # plain depth + semantics
    (3000, 200, 120, True, True, 2),       # ragged tiles (200 = 12.5 tiles, 120 = 7.5 tiles), lower SH degree
    (1, 64, 64, True, False, 0),           # single Gaussian
    (40000, 320, 240, False, False, 1),
]


def check_images(out, ref, what):
    diff = np.abs(out.astype(np.float64) - ref.astype(np.float64))
    scale = np.maximum(1.0, np.abs(ref))          # depth / normal channels divide by a weight that can be ~1e-10
    bad = diff > ABS_TOL * scale
    frac = bad.mean()
    assert frac <= FLIP_FRAC, f"{what}: {bad.sum()} of {bad.size} values differ by more than {ABS_TOL} (max {diff.max()})"
    return frac


def forward_checks(inp, run, orc):
    dec = run.decode()
    pre, binned = orc["pre"], orc["binned"]
    radii = run.radii.cpu().numpy()
    assert np.array_equal(radii, pre["radii"]), f"radii mismatch at {np.nonzero(radii != pre['radii'])[0][:10]}"
    vis = radii > 0
    assert np.array_equal(dec["tiles_touched"], pre["tiles_touched"])
    assert np.array_equal(dec["point_offsets"], binned["point_offsets"])
    assert dec["R"] == binned["R"]
    assert np.array_equal(dec["depths"][vis].view(np.uint32), pre["depths"][vis].view(np.uint32)), "depth bits"
    assert np.array_equal(dec["xy"][vis].view(np.uint32), pre["xy"][vis].view(np.uint32)), "pixel centre bits"
    np.testing.assert_allclose(dec["conic_opacity"][vis], pre["conic_opacity"][vis], rtol=2e-5, atol=1e-7)
    if run.colors_precomp is None:  # with precomputed colours the SH -> RGB arrays are not produced
        np.testing.assert_allclose(dec["rgb"][vis], pre["rgb"][vis], rtol=0, atol=2e-6)
        assert (dec["clamped"][vis] != pre["clamped"][vis]).mean() < 1e-3
    if binned["R"] > 0:
        assert np.array_equal(dec["keys_unsorted"], binned["keys_unsorted"]), "duplicated keys"
        assert np.array_equal(dec["vals_unsorted"], binned["vals_unsorted"]), "duplicated values"
        assert np.array_equal(dec["keys"], binned["keys"]), "sorted keys"
        assert np.array_equal(dec["point_list"], binned["point_list"]), "sorted point list"
    assert np.array_equal(dec["ranges"], binned["ranges"]), "tile ranges"
    out = run.out.detach().cpu().numpy()
    ref = orc["out_color"]
    # depth and normal are ratios over the accumulated weight: compare where the weight is not vanishing
    H, W = inp["H"], inp["W"]
    check_images(out[:3], ref[:3], "rgb")
    check_images(out[7], ref[7], "alpha")
    check_images(out[8:10], ref[8:10], "semantics")
    solid = ref[7] > 1e-3
    check_images(out[3][solid], ref[3][solid], "depth")
    for c in range(4, 7):
        check_images(out[c][solid], ref[c][solid], "normal")
    assert np.abs(out - ref)[:, solid].max() <= FLIP_ABS
    nc = dec["n_contrib"].reshape(H, W)
    assert (nc != orc["n_contrib"].reshape(H, W)).mean() <= FLIP_FRAC
    return dec


@pytest.mark.parametrize("P,W,H,use_dirs,use_sem,deg", CASES)
def test_forward_backward_vs_oracle(P, W, H, use_dirs, use_sem, deg):
    inp, sc, cam = make_case(P, W, H, seed=0, use_dirs=use_dirs, use_sem=use_sem, sh_degree=deg)
    npi = np_inputs(inp)
    orc = O.rasterize_forward(npi)
    run = CudaRun(inp)
    forward_checks(inp, run, orc)
    g = torch.randn((10, H, W), generator=torch.Generator().manual_seed(7))
    if not use_sem:
        g[8:] = 0
    ob = O.rasterize_backward(npi, orc, g.numpy())
    cb = run.backward(g)
    for k in ("dL_dmeans3D", "dL_dmeans2D", "dL_dmeans2D_densify", "dL_dsh", "dL_dnormals", "dL_dopacity", "dL_dscales",
              "dL_drotations"):
        e = rel_l2(cb[k], ob[k])
        assert e <= GRAD_TOL, f"{k}: relative L2 error {e}"
    if use_sem:
        assert rel_l2(cb["dL_dobjects"], ob["dL_dobjects"]) <= GRAD_TOL


def test_colors_and_cov_precomp():
    P, W, H = 5000, 160, 160
    inp, sc, cam = make_case(P, W, H, seed=3, use_dirs=True, use_sem=False)
    npi = np_inputs(inp)
    pre = O.preprocess(npi["means3D"], npi["scales"], 1.0, npi["rotations"], npi["opacities"], npi["shs"], None, None,
                       npi["viewmatrix"], npi["projmatrix"], npi["campos"], W, H, npi["tanfovx"], npi["tanfovy"], 3)
    colors = torch.rand((P, 3), generator=torch.Generator().manual_seed(5))
    # covariance of every Gaussian (also the culled ones) from the oracle's own scale/rotation path
    far = dict(npi); far["viewmatrix"] = np.eye(4, dtype=np.float32); far["viewmatrix"][3, 2] = 10.0
    far["projmatrix"] = far["viewmatrix"]
    cov = torch.from_numpy(O.preprocess(npi["means3D"], npi["scales"], 1.0, npi["rotations"], npi["opacities"], npi["shs"], None, None,
                                        far["viewmatrix"], far["projmatrix"], npi["campos"], W, H, 1.0, 1.0, 3)["cov3D"])
    npi2 = dict(npi, colors_precomp=colors.numpy(), cov3D_precomp=cov.numpy())
    orc = O.rasterize_forward(npi2)
    run = CudaRun(inp, colors_precomp=colors, cov3D_precomp=cov)
    forward_checks(inp, run, orc)
    g = torch.randn((10, H, W), generator=torch.Generator().manual_seed(8)); g[8:] = 0
    ob = O.rasterize_backward(npi2, orc, g.numpy())
    cb = run.backward(g)
    for k in ("dL_dmeans3D", "dL_dcolors", "dL_dcov3D", "dL_dopacity", "dL_dnormals"):
        assert rel_l2(cb[k], ob[k]) <= GRAD_TOL, k
    assert cb["dL_dsh"] is None and cb["dL_dscales"] is None


def test_inside_mask_and_scale_modifier():
    P, W, H = 4000, 128, 128
    inp, sc, cam = make_case(P, W, H, seed=4, use_dirs=True)
    inside = torch.rand(P, generator=torch.Generator().manual_seed(2)) > 0.3
    npi = dict(np_inputs(inp), inside=inside.numpy().astype(np.uint8), scale_modifier=0.7)
    orc = O.rasterize_forward(npi)
    run = CudaRun(inp, inside=inside, scale_modifier=0.7)
    forward_checks(inp, run, orc)
    g = torch.randn((10, H, W), generator=torch.Generator().manual_seed(9)); g[8:] = 0
    ob = O.rasterize_backward(npi, orc, g.numpy())
    cb = run.backward(g)
    for k in ("dL_dmeans3D", "dL_dsh", "dL_dnormals", "dL_dopacity", "dL_dscales", "dL_drotations"):
        assert rel_l2(cb[k], ob[k]) <= GRAD_TOL, k


def test_everything_culled_and_empty():
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import settings_from
    inp, sc, cam = make_case(500, 96, 64, seed=1)
    inp = dict(inp)
    inp["means3D"] = inp["means3D"] + torch.tensor([1000.0, 0.0, 0.0])   # far outside the frustum / behind
    run = CudaRun(inp)
    out = run.out.detach().cpu().numpy()
    orc = O.rasterize_forward(np_inputs(inp))
    assert orc["binned"]["R"] == 0, "the case is built so that nothing is rendered"
    assert run.saved()["R"] == 0 and bool((run.radii == 0).all())
    np.testing.assert_allclose(out, orc["out_color"], atol=1e-6)
    cb = run.backward(torch.ones_like(run.out))
    for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations", "dL_dnormals"):
        assert np.all(cb[k] == 0), k
    # P == 0 short-circuit (rasterize_points.cu:106)
    dev = torch.device("cuda")
    rs = settings_from(inp, dev)
    z = torch.zeros((0, 3), device=dev)
    o, r = GaussianRasterizer(rs)(means3D=z, means2D=z, means2D_densify=z, opacities=torch.zeros((0, 1), device=dev),
                                  shs=torch.zeros((0, 16, 3), device=dev), normals_precomp=z, scales=z,
                                  rotations=torch.zeros((0, 4), device=dev))
    assert o.shape == (10, 64, 96) and float(o.abs().max()) == 0 and r.numel() == 0


def test_mark_visible_and_aux_modes():
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import settings_from
    P, W, H = 6000, 160, 128
    inp, sc, cam = make_case(P, W, H, seed=6, use_dirs=False)
    npi = np_inputs(inp)
    dev = torch.device("cuda")
    rast = GaussianRasterizer(settings_from(inp, dev))
    vis = rast.markVisible(inp["means3D"].to(dev)).cpu().numpy()
    assert np.array_equal(vis, O.mark_visible(npi["means3D"], npi["viewmatrix"]))
    kw = dict(means3D=inp["means3D"].to(dev), means2D=None, opacities=inp["opacities"].to(dev), shs=inp["shs"].to(dev),
              scales=inp["scales"].to(dev), rotations=inp["rotations"].to(dev))
    # f_count = 1
    cnt, score, color, radii = GaussianRasterizer(settings_from(inp, dev, f_count=1))(**kw)
    oc = O.rasterize_forward(npi, mode=1)
    assert np.array_equal(radii.cpu().numpy(), oc["pre"]["radii"])
    # The CPU oracle cannot reproduce MUFU.EX2: a pair whose alpha lies within an ulp of 1/255 (or whose T lies within an
    # ulp of 1e-4) is counted by one side and not by the other — the FLIP_FRAC allowance of the pixels below, times 5 because
    # there are ~3.4 pixels per Gaussian here and a flipped pair changes one Gaussian's count (and, through T, possibly the
    # next one's).  The exact gate for these modes is the A/B against the reference build (test_reference_ab_gpu.py).
    assert (cnt.cpu().numpy() != oc["gaussians_count"]).mean() <= 5 * FLIP_FRAC
    np.testing.assert_allclose(score.cpu().numpy(), oc["important_score"], rtol=1e-3, atol=1.0)
    d = np.abs(color.cpu().numpy() - oc["out_color"][:3])
    assert (d > ABS_TOL).mean() <= FLIP_FRAC
    # f_count = 3
    cnt3, radii3 = GaussianRasterizer(settings_from(inp, dev, f_count=3))(**kw)
    o3 = O.rasterize_forward(npi, mode=3)
    assert ((cnt3.cpu().numpy() > 0) != (o3["gaussians_count"] > 0)).mean() <= 5 * FLIP_FRAC
    assert abs(int(cnt3.sum()) - int(o3["gaussians_count"].sum())) <= 4


def test_no_reference_cycles_per_view():
    """A forward+backward must release its per-view buffers by reference counting alone: with the cyclic garbage collector
    switched off (as training loops do to avoid its pauses) device memory must not grow from view to view."""
    import gc
    from diff_gaussian_rasterization import GaussianRasterizer
    inp, _, _ = make_case(20000, 400, 300, seed=2)
    dev = torch.device("cuda")
    params = {k: inp[k].to(dev).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    normals, dirs = inp["normals"].to(dev), inp["dirs"].to(dev)

    def one_view():
        out, radii = GaussianRasterizer(settings_from(inp, dev))(
            means3D=params["means3D"], means2D=torch.zeros_like(params["means3D"], requires_grad=True), shs=params["shs"],
            normals_precomp=normals, opacities=params["opacities"], scales=params["scales"], rotations=params["rotations"], dirs=dirs)
        out.sum().backward()
        for p in params.values():
            p.grad = None

    gc.collect()
    gc.disable()
    try:
        one_view()
        torch.cuda.synchronize()
        base = torch.cuda.memory_allocated()
        for _ in range(5):
            one_view()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == base, "per-view buffers are kept alive by a reference cycle"
    finally:
        gc.enable()


def test_radix_timeout_is_reported_not_swallowed(tmp_path, monkeypatch):
    """The sort's bounded look-back: when it gives up, the call that contains it cannot know yet (nothing is read back after
    the sort), but it must not fault, and the next forward on the device — or the same one in debug mode — must fail loudly."""
    import vcr_native as N
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import settings_from
    monkeypatch.chdir(tmp_path)                                  # a failing debug forward writes snapshot_fw.dump into the cwd
    inp, sc, cam = make_case(20_000, 320, 240, seed=3)           # 5 radix blocks over P, > 2 over R: every pass has a look-back
    dev = torch.device("cuda")
    kw = dict(means3D=inp["means3D"].to(dev), means2D=None, opacities=inp["opacities"].to(dev), shs=inp["shs"].to(dev),
              normals_precomp=inp["normals"].to(dev), scales=inp["scales"].to(dev), rotations=inp["rotations"].to(dev), dirs=inp["dirs"].to(dev))
    good, _ = GaussianRasterizer(settings_from(inp, dev))(**kw)
    torch.cuda.synchronize()
    def forward(debug=False):
        out, _ = GaussianRasterizer(settings_from(inp, dev, debug=debug))(**kw)
        torch.cuda.synchronize()
        return out

    N.vcr_debug_radix_fault(1)
    try:
        forward()                                                # poisoned: returns, its outputs are unwritten, no fault
        with pytest.raises(RuntimeError, match="look-back timed out in an earlier call"):
            forward()                                            # the next forward reads the flag back together with R
    finally:
        N.vcr_debug_radix_fault(0)
    with pytest.raises(RuntimeError, match="look-back timed out in an earlier call"):
        forward()                     # (the failed call above had already queued its own, sabotaged, depth sort: drained here)
    assert torch.equal(forward(), good)                          # the flag has been consumed and the library works again
    N.vcr_debug_radix_fault(1)
    try:
        with pytest.raises(RuntimeError, match="instance lists of this call are invalid"):
            forward(debug=True)                                  # debug mode: reported by the failing call itself
    finally:
        N.vcr_debug_radix_fault(0)
    assert torch.equal(forward(), good)


def test_packed_expf_is_the_library_expf_for_every_argument():
    """render_forward2_kernel evaluates expf for its two pixels with packed FMAs (vcr_expf2, csrc/common.cuh); the images stay
    bit-identical to the reference build only if that is expf for every input.  tools/expf2_check.cu compares the two over all
    2^32 arguments on both lanes (built by __graft_entry__.build(); about two seconds on a B200)."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "bin", "expf2_check")
    if not os.path.exists(exe):
        pytest.skip("tools/bin/expf2_check not built (run __graft_entry__.build())")
    res = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and " 0 mismatching arguments" in res.stdout, res.stdout + res.stderr

