"""World-size-2 gloo tests (CPU) of the view-parallel layer: gradient bucket all-reduce and densification statistics.

The check is the property the layer promises: after the exchange every rank holds exactly what ONE
process would have accumulated by running the `world` views one after the other, the way the reference trainer does
(trainer.py:339-347, scene/gaussian_model.py:669-671)."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vcr-gaus_b200")


def _fake_view_grads(P, view):
    g = torch.Generator().manual_seed(100 + view)
    grads = {"means3D": torch.randn((P, 3), generator=g), "shs": torch.randn((P, 16, 3), generator=g),
             "opacities": torch.randn((P, 1), generator=g), "scales": torch.randn((P, 3), generator=g),
             "rotations": torch.randn((P, 4), generator=g)}
    radii = torch.randint(0, 40, (P,), generator=g, dtype=torch.int32) * (torch.rand(P, generator=g) > 0.3)
    g2d = torch.randn((P, 3), generator=g).abs()
    return grads, radii.to(torch.int32), g2d


def _groups(bucket):
    """The gradient bucket without the alignment padding between the groups: the groups' rows, concatenated in bucket order."""
    for _, off, _, _ in bucket.layout:
        assert off % 4 == 0, "every group must start on a 16-byte boundary"
    return torch.cat([bucket.grad_flat[off:off + n] for _, off, n, _ in bucket.layout]).clone()


def _worker(rank, world, port, P, out_dir):
    sys.path.insert(0, PKG)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import vcr_view_parallel as VP
    r, w, _ = VP.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    vp = VP.ViewParallel()
    tensors = {"means3D": torch.zeros((P, 3)), "shs": torch.zeros((P, 16, 3)), "opacities": torch.zeros((P, 1)),
               "scales": torch.ones((P, 3)), "rotations": torch.zeros((P, 4))}
    if rank == 1:  # replicas start different on purpose; broadcast must fix that
        tensors["means3D"] += 5.0
    bucket = VP.GaussianBucket(tensors)
    vp.broadcast_parameters(bucket, src=0)
    assert bucket.floats_per_gaussian == 59
    accum, denom, max_r = torch.zeros((P, 1)), torch.zeros((P, 1)), torch.zeros(P)
    lazy = [torch.zeros((P, 1)), torch.zeros((P, 1)), torch.zeros(P)]   # local accumulation, combined once at the end
    for step in range(2):
        view = vp.view_for_step(step, n_views=8, seed=1)
        grads, radii, g2d = _fake_view_grads(P, view)
        bucket.zero_grad()
        loss = sum((bucket.params[k] * grads[k]).sum() for k in grads)  # d loss / d param = grads[k]
        loss.backward()
        vp.allreduce_gradients(bucket, average=False)
        vp.reduce_densification_stats(accum, denom, max_r, g2d, radii)
        vp.accumulate_densification_stats(lazy[0], lazy[1], lazy[2], g2d, radii)
        torch.save({"view": view, "grad": _groups(bucket), "accum": accum.clone(), "denom": denom.clone(),
                    "max_r": max_r.clone(), "flat": bucket.flat.clone()}, os.path.join(out_dir, f"r{rank}_s{step}.pt"))
    vp.sync_densification_stats(*lazy)
    # chunk-pipelined exchange: the rasterizer's backward calls the hook once per Gaussian range; here the hook is driven
    # by hand over ragged ranges, with the gradient rows of a range written just before its call
    dbucket = VP.GaussianBucket(tensors, direct=True)
    vp.overlap_gradients(dbucket, n_chunks=3)
    n_chunks, hook = dbucket.params["means3D"]._vcr_chunk_hook
    grads, _, _ = _fake_view_grads(P, 40 + rank)
    bounds = [0, 100, 101, P]
    for c in range(n_chunks):
        a, b = bounds[c], bounds[c + 1]
        for name, off, n, shape in dbucket.layout:
            per = n // P
            dbucket.grad_flat[off + a * per: off + b * per] = grads[name].reshape(-1)[a * per: b * per]
        hook(c, a, b - a)
    assert vp.finish_gradients(dbucket) == n_chunks
    torch.save({"grad": _groups(dbucket)}, os.path.join(out_dir, f"r{rank}_chunked.pt"))
    torch.save({"accum": lazy[0], "denom": lazy[1], "max_r": lazy[2]}, os.path.join(out_dir, f"r{rank}_lazy.pt"))
    dist.destroy_process_group()


def test_world2_gloo_matches_sequential(tmp_path):
    P, world = 257, 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, P, str(tmp_path)), nprocs=world, join=True)
    accum, denom, max_r = torch.zeros((P, 1)), torch.zeros((P, 1)), torch.zeros(P)
    seen = set()
    for step in range(2):
        r0 = torch.load(tmp_path / f"r0_s{step}.pt")
        r1 = torch.load(tmp_path / f"r1_s{step}.pt")
        assert r0["view"] != r1["view"], "ranks must render different views"
        seen |= {r0["view"], r1["view"]}
        assert torch.equal(r0["flat"], r1["flat"]), "replicas differ after broadcast"
        # sequential reference: the same two views on one process
        total = None
        for v in (r0["view"], r1["view"]):
            grads, radii, g2d = _fake_view_grads(P, v)
            flat = torch.cat([grads[k].reshape(-1) for k in ("means3D", "shs", "opacities", "scales", "rotations")])
            total = flat if total is None else total + flat
            vis = radii > 0
            accum[vis] += torch.norm(g2d[vis, :2], dim=-1, keepdim=True)   # add_densification_stats
            denom[vis] += 1
            max_r[vis] = torch.max(max_r[vis], radii[vis].float())          # trainer.py:345
        for r in (r0, r1):
            torch.testing.assert_close(r["grad"], total, rtol=1e-6, atol=1e-6)
            torch.testing.assert_close(r["accum"], accum, rtol=1e-6, atol=1e-6)
            assert torch.equal(r["denom"], denom) and torch.equal(r["max_r"], max_r)
    total = sum(torch.cat([_fake_view_grads(P, 40 + r)[0][k].reshape(-1) for k in ("means3D", "shs", "opacities", "scales", "rotations")])
                for r in (0, 1))
    for r in (0, 1):
        torch.testing.assert_close(torch.load(tmp_path / f"r{r}_chunked.pt")["grad"], total, rtol=1e-6, atol=1e-6)
    assert len(seen) == 4, "two steps x two ranks must cover four distinct views of the epoch"
    # the lazily combined statistics (local accumulation + one sync) equal the per-step ones and the sequential ones
    for r in (0, 1):
        lz = torch.load(tmp_path / f"r{r}_lazy.pt")
        torch.testing.assert_close(lz["accum"], accum, rtol=1e-6, atol=1e-6)
        assert torch.equal(lz["denom"], denom) and torch.equal(lz["max_r"], max_r)


def test_single_process_is_a_no_op():
    sys.path.insert(0, PKG)
    import vcr_view_parallel as VP
    vp = VP.ViewParallel()
    assert vp.world == 1 and vp.rank == 0
    t = {"means3D": torch.randn((5, 3)), "shs": torch.randn((5, 16, 3)), "opacities": torch.rand((5, 1)),
         "scales": torch.rand((5, 3)), "rotations": torch.randn((5, 4))}
    b = VP.GaussianBucket(t)
    (b.params["means3D"].sum() * 2 + b.params["shs"].sum()).backward()
    off_sh = b.layout[1][1]
    assert off_sh == 16 and off_sh % 4 == 0                                   # 15 floats of means3D + 1 pad float
    assert torch.equal(b.grad_flat[:15], torch.full((15,), 2.0)) and torch.equal(b.grad_flat[off_sh:off_sh + 240], torch.ones(240))
    assert float(b.grad_flat[15]) == 0.0 and float(b.flat[15]) == 0.0         # the pad stays zero
    assert vp.allreduce_gradients(b) is None
    views = [vp.view_for_step(s, 6, seed=0) for s in range(6)]
    assert sorted(views) == list(range(6))


def test_balanced_view_table():
    """Views of one step cost about the same, nobody gets a view twice, and the ranks' totals are level."""
    sys.path.insert(0, PKG)
    import vcr_view_parallel as VP
    vp = VP.ViewParallel()
    vp.world, vp.rank = 4, 1
    g = torch.Generator().manual_seed(3)
    costs = (1.0 + torch.rand(38, generator=g)).tolist()            # 38 views, 4 ranks -> 9 steps, 2 views left out
    table = vp.balanced_view_table(costs, seed=5)
    assert len(table) == 9 and all(len(row) == 4 for row in table)
    flat = [v for row in table for v in row]
    assert len(set(flat)) == 36
    left_out = set(range(38)) - set(flat)
    assert all(costs[v] >= max(costs[u] for u in flat) for v in left_out)   # the two most expensive ones
    srt = sorted(costs)
    worst_gap = max(srt[i + 3] - srt[i] for i in range(len(srt) - 3))
    assert all(max(costs[v] for v in row) - min(costs[v] for v in row) <= worst_gap + 1e-12 for row in table)
    totals = [sum(costs[row[r]] for row in table) for r in range(4)]
    assert max(totals) - min(totals) < 0.02 * sum(totals) / 4
    assert table != vp.balanced_view_table(costs, seed=6), "the visiting order must change with the seed"
    assert sorted(map(sorted, table)) == sorted(map(sorted, vp.balanced_view_table(costs, seed=6))), "... but not the groups"


def test_bucket_layouts_and_cuda_only_pieces():
    """Host-side bookkeeping of the bucket (custom group names, exchange sizing) and the loud failure of the CUDA-only
    pieces on CPU tensors (there is no CPU fallback for kernels)."""
    import pytest
    sys.path.insert(0, PKG)
    import vcr_view_parallel as VP
    P = 10
    raw = dict(xyz=torch.randn(P, 3), f_dc=torch.randn(P, 1, 3), f_rest=torch.randn(P, 15, 3), opacity=torch.randn(P, 1),
               scaling=torch.randn(P, 3), rotation=torch.randn(P, 4))
    names = tuple(raw)
    b = VP.GaussianBucket(raw, names=names)
    pad4 = lambda n: (n + 3) // 4 * 4  # noqa: E731
    padded = sum(pad4(t.numel()) for t in raw.values())
    assert b.numel == padded and b.floats_per_gaussian == 59 and tuple(b.params) == names
    assert all(off % 4 == 0 for _, off, _, _ in b.layout), "groups must start 16-byte aligned for every P"
    assert [n for n, _, _, _ in b.layout] == list(names)
    assert torch.equal(b.params["f_rest"].detach(), raw["f_rest"]) and b.params["f_rest"].grad.shape == (P, 15, 3)
    assert VP.GaussianBucket.numel_for(raw, names) == padded
    vp = VP.ViewParallel()
    std = dict(means3D=raw["xyz"], shs=torch.randn(P, 16, 3), opacities=raw["opacity"], scales=raw["scaling"], rotations=raw["rotation"])
    stride = VP.ViewParallel._slice_stride(P)
    assert stride % 4 == 0 and stride >= 3 * P + 3
    assert vp.exchange_floats(std) == sum(pad4(t.numel()) for t in std.values()) + vp.world * stride
    grown = vp.exchange_floats(std, capacity=3 * P + 1)
    assert grown >= VP.GaussianBucket.numel_for({k: torch.zeros((3 * P + 1,) + t.shape[1:]) for k, t in std.items()}) + vp.world * VP.ViewParallel._slice_stride(3 * P + 1)
    with pytest.raises(RuntimeError):
        VP.BucketAdam(b, {k: 1e-3 for k in names})
    with pytest.raises(RuntimeError):
        VP.prune_and_grow(b, torch.ones(P, dtype=torch.bool))
    with pytest.raises(ValueError):
        vp.compact_sh_exchange(VP.GaussianBucket(std))                       # needs direct=True
    with pytest.raises(ValueError):
        vp.compact_sh_exchange(VP.GaussianBucket(raw, names=names, direct=True))   # needs the rasterizer's inputs
    with pytest.raises(ValueError):
        vp.overlap_gradients(VP.GaussianBucket(std))
    assert VP.bind_to_gpu_numa_node(0) >= 0
