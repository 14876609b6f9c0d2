"""Fused render() prologue (vcr_prologue.gaussian_prologue) against the reference's PyTorch chain, forward and gradients."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def build_rotation(r):
    """tools/general_utils.py:98-119 (restated)."""
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device, dtype=r.dtype)
    r_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r_ * z); R[:, 0, 2] = 2 * (x * z + r_ * y)
    R[:, 1, 0] = 2 * (x * y + r_ * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r_ * x)
    R[:, 2, 0] = 2 * (x * z - r_ * y); R[:, 2, 1] = 2 * (y * z + r_ * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def reference_chain(xyz, fdc, frest, o_raw, s_raw, r_raw, cam_center, R_w2c, valid):
    """scene/gaussian_model.py:125-192 + gaussian_renderer/__init__.py:95-101 (restated)."""
    opacity, scaling, rotation = torch.sigmoid(o_raw), torch.exp(s_raw), F.normalize(r_raw)
    shs = torch.cat((fdc, frest), dim=1)
    if valid is None:
        valid = torch.ones(xyz.shape[0], dtype=torch.bool, device=xyz.device)
        normal = None
    else:
        normal = torch.zeros_like(xyz)
    rots = build_rotation(rotation[valid])
    axis = torch.argmin(scaling[valid], dim=-1)
    normals = rots.gather(2, axis[:, None, None].expand(-1, 3, -1)).squeeze(-1)
    if normal is not None:
        normal[valid] = normals
        normals = normal
    view_dir = xyz - cam_center
    normals = normals * ((((view_dir * normals).sum(dim=-1) > 0) * 1 - 0.5) * 2)[..., None]
    return opacity, scaling, rotation, shs, normals @ R_w2c.transpose(0, 1)


def _inputs(P, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    dev = torch.device("cuda")
    mk = lambda *s, scale=1.0: (torch.randn(*s, generator=g) * scale).to(dev).to(dtype).requires_grad_(True)  # noqa: E731
    raw = dict(xyz=mk(P, 3), fdc=mk(P, 1, 3), frest=mk(P, 15, 3, scale=0.1), o_raw=mk(P, 1, scale=2.0), s_raw=mk(P, 3, scale=0.7) ,
               r_raw=mk(P, 4))
    cam_center = torch.tensor([0.3, -2.0, 1.5], device=dev, dtype=dtype)
    A = torch.randn(3, 3, generator=g, dtype=torch.float64)
    Q, _ = torch.linalg.qr(A)
    return raw, cam_center, Q.to(dev).to(dtype)


@pytest.mark.parametrize("P,use_valid", [(1, False), (257, True), (100_003, False), (100_003, True)])
def test_prologue_matches_reference_chain(P, use_valid):
    from vcr_prologue import gaussian_prologue
    raw, cc, M = _inputs(P, 4)
    valid = None
    if use_valid:
        valid = (torch.rand(P, generator=torch.Generator().manual_seed(9)) > 0.3).cuda()
    got = gaussian_prologue(raw["xyz"], raw["fdc"], raw["frest"], raw["o_raw"], raw["s_raw"], raw["r_raw"], cc, M, valid)
    exp = reference_chain(raw["xyz"], raw["fdc"], raw["frest"], raw["o_raw"], raw["s_raw"], raw["r_raw"], cc, M, valid)
    names = ("opacity", "scales", "rotations", "shs", "normals")
    for n, a, b in zip(names, got, exp):
        assert a.shape == b.shape, n
        torch.testing.assert_close(a, b, rtol=2e-6, atol=2e-6, msg=lambda m, n=n: f"{n}: {m}")
    assert torch.equal(got[3], exp[3])                                   # the concat is a copy: bit-exact
    # gradients: one random cotangent per output, through both graphs
    g = torch.Generator().manual_seed(12)
    cot = [torch.randn(t.shape, generator=g).cuda() for t in exp]
    leaves = [raw[k] for k in ("fdc", "frest", "o_raw", "s_raw", "r_raw")]
    g_got = torch.autograd.grad(got, leaves, cot, allow_unused=True)
    g_exp = torch.autograd.grad(exp, leaves, cot, allow_unused=True)
    for k, a, b in zip(("fdc", "frest", "o_raw", "s_raw", "r_raw"), g_got, g_exp):
        err = float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
        assert err <= 1e-5, (k, err)
    assert torch.equal(g_got[0], cot[3][:, :1]) and torch.equal(g_got[1], cot[3][:, 1:])


def test_prologue_gradient_against_float64_chain():
    """Independent of torch's float32 kernels: the same chain in float64 is the judge (<= 1e-5 relative)."""
    from vcr_prologue import gaussian_prologue
    P = 4099
    raw, cc, M = _inputs(P, 21)
    got = gaussian_prologue(raw["xyz"], raw["fdc"], raw["frest"], raw["o_raw"], raw["s_raw"], raw["r_raw"], cc, M)
    raw64 = {k: v.detach().double().requires_grad_(True) for k, v in raw.items()}
    exp = reference_chain(raw64["xyz"], raw64["fdc"], raw64["frest"], raw64["o_raw"], raw64["s_raw"], raw64["r_raw"], cc.double(), M.double(), None)
    # skip Gaussians whose flip test or argmin is decided within float32 rounding
    same = ((got[4].double() - exp[4]).abs().max(dim=1).values < 1e-4)
    assert float(same.float().mean()) > 0.999
    g = torch.Generator().manual_seed(2)
    cot = [torch.randn(t.shape, generator=g).cuda() for t in got]
    cot[4] = cot[4] * same[:, None]
    keys = ("o_raw", "s_raw", "r_raw")
    g_got = torch.autograd.grad(got, [raw[k] for k in keys], cot)
    g_exp = torch.autograd.grad(exp, [raw64[k] for k in keys], [c.double() for c in cot])
    for k, a, b in zip(keys, g_got, g_exp):
        err = float((a.double() - b).norm() / b.norm())
        assert err <= 1e-5, (k, err)


def test_prologue_feeds_the_rasterizer():
    """End to end: raw parameters -> prologue -> rasterizer -> loss; every raw parameter receives a finite gradient."""
    from vcr_prologue import gaussian_prologue
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import make_case, settings_from
    inp, _, cam = make_case(3000, 160, 120, seed=5)
    dev = torch.device("cuda")
    shs = inp["shs"].to(dev)
    raw = dict(xyz=inp["means3D"].to(dev).requires_grad_(True), fdc=shs[:, :1].clone().requires_grad_(True),
               frest=shs[:, 1:].clone().requires_grad_(True),
               o_raw=torch.logit(inp["opacities"].to(dev).clamp(1e-4, 1 - 1e-4)).requires_grad_(True),
               s_raw=torch.log(inp["scales"].to(dev)).requires_grad_(True), r_raw=(inp["rotations"].to(dev) * 1.7).requires_grad_(True))
    view = inp["viewmatrix"].to(dev)                     # world_view_transform: [R | 0; t | 1] (transposed convention)
    R_w2c = view[:3, :3].t().contiguous()
    opac, scales, rots, sh, normals = gaussian_prologue(raw["xyz"], raw["fdc"], raw["frest"], raw["o_raw"], raw["s_raw"], raw["r_raw"],
                                                        inp["campos"].to(dev), R_w2c)
    out, radii = GaussianRasterizer(settings_from(inp, dev))(
        means3D=raw["xyz"], means2D=torch.zeros_like(raw["xyz"], requires_grad=True), shs=sh, normals_precomp=normals, opacities=opac,
        scales=scales, rotations=rots, dirs=inp["dirs"].to(dev))
    out.square().mean().backward()
    for k, v in raw.items():
        assert v.grad is not None and torch.isfinite(v.grad).all() and float(v.grad.abs().sum()) > 0, k


def test_prologue_writes_gradients_into_a_direct_bucket():
    """With GaussianBucket(direct=True, names=raw groups) the prologue's and the rasterizer's backward kernels write every raw
    gradient straight into the flat bucket; same numbers as the ordinary autograd path."""
    import vcr_view_parallel as VP
    from vcr_prologue import gaussian_prologue
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import make_case, settings_from
    inp, _, _ = make_case(2500, 128, 96, seed=8)
    dev = torch.device("cuda")
    shs = inp["shs"].to(dev)
    raw = dict(xyz=inp["means3D"].to(dev), f_dc=shs[:, :1].contiguous(), f_rest=shs[:, 1:].contiguous(),
               opacity=torch.logit(inp["opacities"].to(dev).clamp(1e-4, 1 - 1e-4)), scaling=torch.log(inp["scales"].to(dev)),
               rotation=inp["rotations"].to(dev) * 0.6)
    names = tuple(raw)
    view = inp["viewmatrix"].to(dev)
    dL = torch.randn((10, 96, 128), generator=torch.Generator().manual_seed(4)).to(dev)

    def run(p):
        opac, scales, rots, sh, normals = gaussian_prologue(p["xyz"], p["f_dc"], p["f_rest"], p["opacity"], p["scaling"], p["rotation"],
                                                            inp["campos"].to(dev), view[:3, :3].t())
        out, _ = GaussianRasterizer(settings_from(inp, dev))(
            means3D=p["xyz"], means2D=None, shs=sh, normals_precomp=normals, opacities=opac, scales=scales, rotations=rots,
            dirs=inp["dirs"].to(dev))
        out.backward(dL)

    bucket = VP.GaussianBucket(raw, device=dev, names=names, direct=True)
    bucket.grad_flat.fill_(float("nan"))
    run(bucket.params)
    torch.cuda.synchronize()
    assert torch.isfinite(bucket.grad_flat).all(), "some gradient rows were not written into the bucket"
    for k in names:
        assert bucket.params[k].grad.data_ptr() == bucket.grad_flat[bucket.layout[names.index(k)][1]:].data_ptr(), k
    plain = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    run(plain)
    for k, off, n, shape in bucket.layout:
        a, b = bucket.grad_flat[off:off + n].view(shape).double(), plain[k].grad.double()
        assert float((a - b).norm() / b.norm().clamp_min(1e-30)) <= 5e-5, k   # two runs of the float-atomic backward
