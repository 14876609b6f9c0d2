"""GPU tests of the view-parallel pieces that run on one GPU: the chunked backward (vcr_rasterize_backward_chunked, the hook
the gradient exchange is pipelined through) and the fused densification-statistics pass."""
import numpy as np
import pytest
import torch

from util import CudaRun, make_case, rel_l2, settings_from

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P", [3000, 3001, 10_007])
def test_chunked_backward_equals_single_launch(P):
    """The per-Gaussian backward stage in 5 launches writes the same gradients as in one, straight into the flat bucket,
    and reports consecutive ranges starting at multiples of 256 that cover every Gaussian.  P = 3001 / 10007: the bucket's
    groups would start at float offsets 3P, 51P, ... — 16-byte aligned only through the bucket's padding."""
    import vcr_view_parallel as VP
    from diff_gaussian_rasterization import GaussianRasterizer
    W, H = 320, 200                               # 12+ blocks of 256 -> ragged split into 5 ranges, last block partial
    inp, _, _ = make_case(P, W, H, seed=3, use_dirs=True, use_sem=False)
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(5)
    dL = torch.randn((10, H, W), generator=g)

    plain = CudaRun(inp)
    ref = plain.backward(dL)

    tensors = {k: inp[k].to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    bucket = VP.GaussianBucket(tensors, device=dev, direct=True)
    assert all(off % 4 == 0 and bucket.params[name].data_ptr() % 16 == 0 for name, off, _, _ in bucket.layout)
    for _, off, n, _ in bucket.layout:
        bucket.grad_flat[off:off + n].fill_(float("nan"))      # every element of every group must be overwritten
    calls = []
    bucket.params["means3D"]._vcr_chunk_hook = (5, lambda c, first, count: calls.append((c, first, count)))
    p = bucket.params
    normals = inp["normals"].to(dev).requires_grad_(True)
    out, radii = GaussianRasterizer(settings_from(inp, dev))(
        means3D=p["means3D"], means2D=torch.zeros_like(p["means3D"], requires_grad=True), shs=p["shs"], normals_precomp=normals,
        opacities=p["opacities"], scales=p["scales"], rotations=p["rotations"], dirs=inp["dirs"].to(dev))
    out.backward(dL.to(dev))
    torch.cuda.synchronize()

    assert [c for c, _, _ in calls] == [0, 1, 2, 3, 4]
    assert calls[0][1] == 0 and all(first % 256 == 0 for _, first, _ in calls)
    assert all(calls[i][1] + calls[i][2] == calls[i + 1][1] for i in range(4)) and calls[-1][1] + calls[-1][2] == P
    assert p["means3D"].grad.data_ptr() == bucket.grad_flat.data_ptr(), "gradient was not written in place"
    assert torch.isfinite(bucket.grad_flat).all()
    names = dict(means3D="dL_dmeans3D", shs="dL_dsh", opacities="dL_dopacity", scales="dL_dscales", rotations="dL_drotations")
    for name, off, n, shape in bucket.layout:
        got = bucket.grad_flat[off:off + n].view(shape).cpu().numpy()
        # two runs of the compositing backward differ by float-atomic ordering only
        assert rel_l2(got, ref[names[name]]) <= 2e-4, name
    assert rel_l2(normals.grad.cpu().numpy(), ref["dL_dnormals"]) <= 2e-4


def test_chunk_hook_exception_surfaces():
    import vcr_view_parallel as VP
    from diff_gaussian_rasterization import GaussianRasterizer
    inp, _, _ = make_case(600, 96, 64, seed=1)
    dev = torch.device("cuda")
    bucket = VP.GaussianBucket({k: inp[k].to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}, device=dev, direct=True)

    def boom(c, first, count):
        raise KeyError("from the hook")
    bucket.params["means3D"]._vcr_chunk_hook = (2, boom)
    p = bucket.params
    out, _ = GaussianRasterizer(settings_from(inp, dev))(
        means3D=p["means3D"], means2D=None, shs=p["shs"], normals_precomp=inp["normals"].to(dev), opacities=p["opacities"],
        scales=p["scales"], rotations=p["rotations"], dirs=inp["dirs"].to(dev))
    with pytest.raises(KeyError):
        out.sum().backward()
    torch.cuda.synchronize()


def test_densify_stats_kernel_matches_reference_updates():
    """vcr_densify_stats_accumulate == trainer.py:345 + scene/gaussian_model.py:669-671, two views in a row."""
    import vcr_view_parallel as VP
    vp = VP.ViewParallel()
    P = 100_003
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(11)
    accum, denom, max_r = torch.zeros((P, 1), device=dev), torch.zeros((P, 1), device=dev), torch.zeros(P, device=dev)
    e_accum, e_denom, e_max = accum.clone(), denom.clone(), max_r.clone()
    for _ in range(2):
        grad = torch.randn((P, 3), generator=g).to(dev)
        radii = (torch.randint(0, 60, (P,), generator=g, dtype=torch.int32) * (torch.rand(P, generator=g) > 0.4)).to(torch.int32).to(dev)
        vp.accumulate_densification_stats(accum, denom, max_r, grad, radii)
        vis = radii > 0                                                           # the reference's three updates
        e_max[vis] = torch.max(e_max[vis], radii[vis].float())
        e_accum[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
        e_denom[vis] += 1
    torch.cuda.synchronize()
    assert torch.equal(denom, e_denom) and torch.equal(max_r, e_max)
    torch.testing.assert_close(accum, e_accum, rtol=1e-6, atol=1e-7)              # sqrt(x*x+y*y): contraction may differ by 1 ulp
    assert float(denom.max()) == 2.0 and float((denom == 0).sum()) > 0


def test_bucket_adam_matches_torch_adam():
    """BucketAdam == torch.optim.Adam with the reference's six groups (scene/gaussian_model.py:241-258, eps=1e-15), three
    steps with a learning-rate change in between and a gradient scale."""
    import vcr_view_parallel as VP
    P = 10_007
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(1)
    shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
    names = tuple(shapes)
    init = {k: torch.randn(s, generator=g).to(dev) for k, s in shapes.items()}
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20, opacity=0.05, scaling=5e-3, rotation=1e-3)
    bucket = VP.GaussianBucket(init, device=dev, names=names)
    assert bucket.floats_per_gaussian == 59 and tuple(bucket.params) == names
    opt = VP.BucketAdam(bucket, lrs)
    ref_p = {k: v.clone().requires_grad_(True) for k, v in init.items()}
    ref = torch.optim.Adam([{"params": [ref_p[k]], "lr": lrs[k], "name": k} for k in names], lr=0.0, eps=1e-15)
    for step in range(3):
        grads = {k: torch.randn(s, generator=g).to(dev) * (10.0 ** (step - 1)) for k, s in shapes.items()}
        scale = 0.5 if step == 1 else 1.0
        for k in names:
            bucket.params[k].grad.copy_(grads[k])
            ref_p[k].grad = grads[k] * scale
        if step == 2:
            opt.set_lr("xyz", 8e-5)
            ref.param_groups[0]["lr"] = 8e-5
        opt.step(grad_scale=scale)
        ref.step()
        for k in names:
            torch.testing.assert_close(bucket.params[k].detach(), ref_p[k].detach(), rtol=2e-6, atol=1e-7, msg=lambda m, k=k: f"{k} step {step}: {m}")
    with pytest.raises(KeyError):
        VP.BucketAdam(bucket, {"xyz": 1.0})


@pytest.mark.parametrize("deg,P", [(3, 4000), (1, 4000), (0, 4000), (3, 3001), (3, 10_007)])
def test_compact_sh_gradient_reconstruction(deg, P):
    """dL_dsh rebuilt from the clamp-masked colour gradient + camera position (vcr_sh_grad_from_colors) equals the SH gradient
    the backward writes itself; every other gradient is unchanged by the compact mode."""
    import ctypes
    import vcr_native as N
    from diff_gaussian_rasterization import GaussianRasterizer
    W, H = 200, 160
    inp, _, _ = make_case(P, W, H, seed=6, sh_degree=deg)
    dev = torch.device("cuda")
    dL = torch.randn((10, H, W), generator=torch.Generator().manual_seed(2)).to(dev)
    plain = CudaRun(inp)
    ref = plain.backward(dL.cpu())
    assert np.abs(ref["dL_dsh"]).max() > 0

    leaves = {k: inp[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    stride = (3 * P + 3 + 3) // 4 * 4
    slices = torch.zeros((2, stride), device=dev)                   # "two ranks": the second one contributes nothing
    leaves["shs"]._vcr_sh_compact = slices[0, :3 * P].view(P, 3)
    out, _ = GaussianRasterizer(settings_from(inp, dev))(
        means3D=leaves["means3D"], means2D=None, shs=leaves["shs"], normals_precomp=inp["normals"].to(dev), opacities=leaves["opacities"],
        scales=leaves["scales"], rotations=leaves["rotations"], dirs=inp["dirs"].to(dev))
    out.backward(dL)
    slices[0, 3 * P:3 * P + 3] = inp["campos"].to(dev)
    slices[1, 3 * P:3 * P + 3] = torch.tensor([9.0, 9.0, 9.0], device=dev)
    M = inp["shs"].shape[1]
    rebuilt = torch.full((P, M, 3), float("nan"), device=dev)
    N.check(N.vcr_sh_grad_from_colors(P, deg, M, 2, leaves["means3D"].data_ptr(), slices.data_ptr(), stride, rebuilt.data_ptr(),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "sh_grad_from_colors")
    torch.cuda.synchronize()
    assert torch.isfinite(rebuilt).all()
    # two runs of the float-atomic backward: the order in which the warps' partial sums land differs from run to run, and
    # the scale / rotation gradients are differences of large terms (measured up to 9e-5; the gate for gradients is 1e-3)
    assert rel_l2(rebuilt.cpu().numpy(), ref["dL_dsh"]) <= 2e-4
    for k, name in (("means3D", "dL_dmeans3D"), ("opacities", "dL_dopacity"), ("scales", "dL_dscales"), ("rotations", "dL_drotations")):
        assert rel_l2(leaves[k].grad.cpu().numpy(), ref[name]) <= 2e-4, k


def test_prune_and_grow_matches_boolean_indexing():
    """prune_and_grow == the reference's prune_points + densification_postfix on parameters and both Adam moments
    (scene/gaussian_model.py:442-531): kept rows in order, new rows appended with zero moments."""
    import vcr_view_parallel as VP
    P, n_new = 70_001, 1234
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(8)
    shapes = dict(xyz=(P, 3), f_dc=(P, 1, 3), f_rest=(P, 15, 3), opacity=(P, 1), scaling=(P, 3), rotation=(P, 4))
    names = tuple(shapes)
    init = {k: torch.randn(s, generator=g).to(dev) for k, s in shapes.items()}
    bucket = VP.GaussianBucket(init, device=dev, names=names)
    opt = VP.BucketAdam(bucket, {k: 1e-3 for k in names})
    opt.exp_avg.normal_(generator=None); opt.exp_avg_sq.uniform_()
    opt.steps = 7
    keep = (torch.rand(P, generator=g) > 0.37).to(dev)
    keep[:300] = False; keep[-5:] = True                                # a fully pruned block and a kept tail
    old_m = {k: opt.exp_avg[o:o + n].view(s).clone() for k, o, n, s in bucket.layout}
    old_v = {k: opt.exp_avg_sq[o:o + n].view(s).clone() for k, o, n, s in bucket.layout}
    new, opt2, n_kept = VP.prune_and_grow(bucket, keep, n_new=n_new, optimizer=opt)
    torch.cuda.synchronize()
    assert n_kept == int(keep.sum()) and new.P == n_kept + n_new and new.floats_per_gaussian == 59 and opt2.steps == 7
    assert all(o % 4 == 0 for _, o, _, _ in new.layout)
    for k, o, n, s in new.layout:
        assert s[0] == new.P and tuple(s[1:]) == tuple(shapes[k][1:])
        assert torch.equal(new.params[k].detach()[:n_kept], init[k][keep]), k          # param[mask]
        m2, v2 = opt2.exp_avg[o:o + n].view(s), opt2.exp_avg_sq[o:o + n].view(s)
        assert torch.equal(m2[:n_kept], old_m[k][keep]) and torch.equal(v2[:n_kept], old_v[k][keep]), k
        assert float(m2[n_kept:].abs().max()) == 0.0 and float(v2[n_kept:].abs().max()) == 0.0   # cat(zeros)
        assert new.params[k].grad.data_ptr() == new.grad_flat[o:].data_ptr()
    # nothing kept / everything kept
    none, _, z = VP.prune_and_grow(bucket, torch.zeros(P, dtype=torch.bool, device=dev), n_new=3)
    assert z == 0 and none.P == 3
    full, _, a = VP.prune_and_grow(bucket, torch.ones(P, dtype=torch.bool, device=dev))
    torch.cuda.synchronize()
    assert a == P and torch.equal(full.flat, bucket.flat)


def test_rasterize_through_bucket_after_prune_and_grow():
    """A densification leaves a Gaussian count that is not a multiple of 4.  The rasterizer's float4 accesses (shs, rotations,
    dL_dsh, dL_drot) must still find 16-byte aligned groups in the new bucket, and forward + backward through it must give
    what plain leaf tensors of the same values give."""
    import vcr_view_parallel as VP
    from diff_gaussian_rasterization import GaussianRasterizer
    P, W, H = 5000, 240, 176
    inp, _, _ = make_case(P, W, H, seed=4, use_dirs=True, use_sem=False)
    dev = torch.device("cuda")
    tensors = {k: inp[k].to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    bucket = VP.GaussianBucket(tensors, device=dev, direct=True)
    keep = torch.ones(P, dtype=torch.bool, device=dev)
    keep[torch.arange(0, P, 7, device=dev)] = False                  # 4285 kept
    n_new = 2
    new, _, n_kept = VP.prune_and_grow(bucket, keep, n_new=n_new)
    P2 = n_kept + n_new
    assert P2 % 4 != 0 and new.P == P2
    with torch.no_grad():                                            # the new rows: clones of the first two kept Gaussians
        for name in new.names:
            new.params[name][n_kept:] = new.params[name][:n_new]
    assert all(new.params[name].data_ptr() % 16 == 0 for name in new.names)
    sel = torch.cat([keep.nonzero(as_tuple=True)[0], keep.nonzero(as_tuple=True)[0][:n_new]]).cpu()
    inp2 = dict(inp)
    for k in ("means3D", "shs", "opacities", "scales", "rotations", "normals"):
        inp2[k] = inp[k][sel].contiguous()
    for name in new.names:
        assert torch.equal(new.params[name].detach().cpu(), inp2[name])
    dL = torch.randn((10, H, W), generator=torch.Generator().manual_seed(9))
    plain = CudaRun(inp2)
    ref = plain.backward(dL)
    p = new.params
    out, radii = GaussianRasterizer(settings_from(inp2, dev))(
        means3D=p["means3D"], means2D=None, shs=p["shs"], normals_precomp=inp2["normals"].to(dev), opacities=p["opacities"],
        scales=p["scales"], rotations=p["rotations"], dirs=inp2["dirs"].to(dev))
    assert torch.equal(out, plain.out) and torch.equal(radii, plain.radii)
    out.backward(dL.to(dev))
    torch.cuda.synchronize()
    assert p["shs"].grad.data_ptr() == new.grad_flat[new.layout[1][1]:].data_ptr(), "gradient was not written in place"
    names = dict(means3D="dL_dmeans3D", shs="dL_dsh", opacities="dL_dopacity", scales="dL_dscales", rotations="dL_drotations")
    for name, off, n, shape in new.layout:
        assert rel_l2(new.grad_flat[off:off + n].view(shape).cpu().numpy(), ref[names[name]]) <= 2e-4, name


def test_misaligned_float4_arrays():
    """Through the Python API a rotations / shs view that does not start on a 16-byte boundary is copied and works; through the
    C ABI the same pointer is refused with a message instead of faulting inside a kernel."""
    import ctypes
    import vcr_native as N
    from diff_gaussian_rasterization import GaussianRasterizer, _Buffers
    P, W, H = 1001, 96, 64
    inp, _, _ = make_case(P, W, H, seed=2)
    dev = torch.device("cuda")
    plain = CudaRun(inp)
    # views at float offset 1 of a larger allocation: contiguous, 4-byte aligned, not 16-byte aligned
    rot_store = torch.zeros(4 * P + 1, device=dev); rot_store[1:] = inp["rotations"].to(dev).reshape(-1)
    sh_store = torch.zeros(48 * P + 1, device=dev); sh_store[1:] = inp["shs"].to(dev).reshape(-1)
    rot, sh = rot_store[1:].view(P, 4).requires_grad_(True), sh_store[1:].view(P, 16, 3).requires_grad_(True)
    assert rot.data_ptr() % 16 != 0 and sh.data_ptr() % 16 != 0
    means = inp["means3D"].to(dev).requires_grad_(True)
    out, radii = GaussianRasterizer(settings_from(inp, dev))(
        means3D=means, means2D=None, shs=sh, normals_precomp=inp["normals"].to(dev), opacities=inp["opacities"].to(dev),
        scales=inp["scales"].to(dev), rotations=rot, dirs=inp["dirs"].to(dev))
    assert torch.equal(out, plain.out) and torch.equal(radii, plain.radii)
    out.sum().backward()
    torch.cuda.synchronize()
    assert torch.isfinite(rot.grad).all() and torch.isfinite(sh.grad).all()
    # C ABI: refused before any launch
    bufs = _Buffers(dev)
    o = torch.empty((10, H, W), device=dev); r = torch.empty(P, dtype=torch.int32, device=dev)
    ins = torch.ones(P, dtype=torch.bool, device=dev)
    t = {k: inp[k].to(dev).contiguous() for k in ("bg", "means3D", "normals", "opacities", "scales", "viewmatrix", "projmatrix", "dirs", "campos", "shs")}
    n_host = ctypes.c_int(0)
    st = N.vcr_rasterize_forward(bufs.cb["geom"], bufs.cb["binning"], bufs.cb["img"], None, P, 3, 16, t["bg"].data_ptr(), W, H,
                                 t["means3D"].data_ptr(), t["shs"].data_ptr(), None, t["normals"].data_ptr(), None, t["opacities"].data_ptr(),
                                 t["scales"].data_ptr(), 1.0, rot.data_ptr(), None, t["viewmatrix"].data_ptr(), t["projmatrix"].data_ptr(),
                                 t["dirs"].data_ptr(), t["campos"].data_ptr(), float(inp["tanfovx"]), float(inp["tanfovy"]), 0,
                                 o.data_ptr(), r.data_ptr(), ins.data_ptr(), 0, ctypes.byref(n_host),
                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert st != 0 and b"16-byte aligned" in N.vcr_last_error()
    torch.cuda.synchronize()
