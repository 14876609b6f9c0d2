"""Two-GPU tests of the NVSwitch-multicast gradient exchange (csrc/view_parallel.cu, vcr_multimem_allreduce_sum) and of the
chunk-pipelined backward that feeds it.  Skipped on boxes with fewer than two GPUs or without NVLink multicast."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vcr-gaus_b200")
GROUPS = ("means3D", "shs", "opacities", "scales", "rotations")


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, PKG)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import ctypes
    import vcr_native as N
    import vcr_view_parallel as VP
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import make_case, settings_from
    VP.init_from_env(backend="nccl")
    dev = torch.device("cuda", rank)
    vp = VP.ViewParallel()
    result = {"multicast": False}
    # ---- scene: the same Gaussians on both ranks, a different view per rank ----
    P, W, H = 5000, 320, 200
    inp, _, _ = make_case(P, W, H, seed=3, view=1 + rank, use_dirs=True, use_sem=False)
    tensors = {k: inp[k].to(dev) for k in GROUPS}
    numel = VP.GaussianBucket.numel_for(tensors)
    store = vp.symmetric_gradient_storage(numel + 64, dev)
    if store is not None:
        result["multicast"] = True
        t, hdl = vp._symm
        # ---- 1. the kernel alone: ragged, unaligned ranges; untouched elements stay as they were ----
        g = torch.Generator().manual_seed(7 + rank)
        mine = torch.randn(numel + 64, generator=g).to(dev)
        t.copy_(mine)
        other = torch.randn(numel + 64, generator=torch.Generator().manual_seed(7 + (1 - rank))).to(dev)
        ranges = [(1, 10), (13, 1), (16, 4096), (5001, 12345), (20000, 3), (numel - 7, 71)]
        torch.cuda.synchronize(dev)
        hdl.barrier(channel=0, timeout_ms=30000)
        arr = ctypes.c_longlong * len(ranges)
        N.check(N.vcr_multimem_allreduce_sum(ctypes.c_void_p(int(hdl.multicast_ptr)), len(ranges), arr(*[a for a, _ in ranges]),
                                             arr(*[n for _, n in ranges]), rank, world, 8,
                                             ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)), "mm")
        hdl.barrier(channel=0, timeout_ms=30000)
        torch.cuda.synchronize(dev)
        expect = mine.clone()
        for a, n in ranges:
            expect[a:a + n] = mine[a:a + n] + other[a:a + n]
        result["kernel_equal"] = bool(torch.equal(t, expect))      # two addends: the sum is exact either way round
        result["kernel_maxdiff"] = float((t - expect).abs().max())
        # ---- 2. chunk-pipelined backward + multicast exchange == sum of the two ranks' gradients ----
        bucket = VP.GaussianBucket(tensors, device=dev, direct=True, grad_storage=store)
        result["pick"] = vp.pick_transport(bucket, iters=2)           # times both ways; leaves the gradients zeroed
        result["picked"] = vp.transport
        # whole-bucket exchange through allreduce_gradients() with the multicast transport
        vp.transport = "multimem"
        bucket.grad_flat.copy_(mine[:numel])
        vp.allreduce_gradients(bucket, average=False)
        torch.cuda.synchronize(dev)
        result["whole_equal"] = bool(torch.equal(bucket.grad_flat, mine[:numel] + other[:numel]))
        vp.overlap_gradients(bucket, n_chunks=3, transport="multimem")
        p = bucket.params
        for step in range(2):           # twice: the second step must not see anything left over from the first
            bucket.zero_grad()
            out, _ = GaussianRasterizer(settings_from(inp, dev))(
                means3D=p["means3D"], means2D=None, shs=p["shs"], normals_precomp=inp["normals"].to(dev), opacities=p["opacities"],
                scales=p["scales"], rotations=p["rotations"], dirs=inp["dirs"].to(dev))
            dL = torch.randn((10, H, W), generator=torch.Generator().manual_seed(50 + rank)).to(dev)
            out.backward(dL)
            n_pending = vp.finish_gradients(bucket)
            torch.cuda.synchronize(dev)
        result["pending"] = n_pending
        result["exchanged"] = bucket.grad_flat.clone().cpu()
        # the same view without any exchange
        plain = VP.GaussianBucket(tensors, device=dev, direct=True)
        q = plain.params
        out, _ = GaussianRasterizer(settings_from(inp, dev))(
            means3D=q["means3D"], means2D=None, shs=q["shs"], normals_precomp=inp["normals"].to(dev), opacities=q["opacities"],
            scales=q["scales"], rotations=q["rotations"], dirs=inp["dirs"].to(dev))
        out.backward(dL)
        torch.cuda.synchronize(dev)
        result["local"] = plain.grad_flat.clone().cpu()
    # ---- 3. compact SH exchange (3 instead of 48 floats per Gaussian for the SH coefficients), both transports ----
    for transport in (("multimem", "nccl") if store is not None else ("nccl",)):
        vp2 = VP.ViewParallel()
        store2 = vp2.symmetric_gradient_storage(vp2.exchange_floats(tensors), dev) if transport == "multimem" else None
        if transport == "multimem" and store2 is None:
            continue
        b2 = VP.GaussianBucket(tensors, device=dev, direct=True, grad_storage=store2)
        vp2.compact_sh_exchange(b2, sh_degree=3)
        vp2.transport = transport
        q = b2.params
        for step in range(2):
            b2.zero_grad()
            b2.grad_flat.fill_(float("nan"))
            out, _ = GaussianRasterizer(settings_from(inp, dev))(
                means3D=q["means3D"], means2D=None, shs=q["shs"], normals_precomp=inp["normals"].to(dev), opacities=q["opacities"],
                scales=q["scales"], rotations=q["rotations"], dirs=inp["dirs"].to(dev))
            dL = torch.randn((10, H, W), generator=torch.Generator().manual_seed(50 + rank)).to(dev)
            out.backward(dL)
            vp2.exchange_compact(b2, inp["campos"].to(dev))
            torch.cuda.synchronize(dev)
        result[f"compact_{transport}"] = b2.grad_flat.clone().cpu()
    if "local" not in result:
        plain = VP.GaussianBucket(tensors, device=dev, direct=True)
        q = plain.params
        out, _ = GaussianRasterizer(settings_from(inp, dev))(
            means3D=q["means3D"], means2D=None, shs=q["shs"], normals_precomp=inp["normals"].to(dev), opacities=q["opacities"],
            scales=q["scales"], rotations=q["rotations"], dirs=inp["dirs"].to(dev))
        out.backward(torch.randn((10, H, W), generator=torch.Generator().manual_seed(50 + rank)).to(dev))
        torch.cuda.synchronize(dev)
        result["local"] = plain.grad_flat.clone().cpu()
    torch.save(result, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_multicast_exchange_two_gpus(tmp_path):
    port = 29700 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r = [torch.load(tmp_path / f"r{k}.pt") for k in range(2)]
    total = r[0]["local"].double() + r[1]["local"].double()
    for key in [k for k in r[0] if k.startswith("compact_")]:
        assert torch.equal(r[0][key], r[1][key]), f"{key}: replicas must hold identical sums"
        err = float((r[0][key].double() - total).norm() / total.norm())
        assert err <= 2e-4, (key, err)      # run-to-run float-atomic order of the compositing backward (gate: 1e-3)
    assert "compact_nccl" in r[0]
    if not r[0]["multicast"]:
        pytest.skip("no NVLink multicast on this box")
    assert "compact_multimem" in r[0]
    for k in range(2):
        assert r[k]["kernel_equal"], r[k]["kernel_maxdiff"]
        assert r[k]["pending"] == 3
        assert r[k]["whole_equal"]
        assert set(r[k]["pick"]) == {"nccl", "multimem"} and r[k]["picked"] in r[k]["pick"]
    assert r[0]["picked"] == r[1]["picked"], "ranks must agree on the transport"
    for k in range(2):
        pass
    assert torch.equal(r[0]["exchanged"], r[1]["exchanged"]), "replicas must receive bit-identical sums"
    err = float((r[0]["exchanged"].double() - total).norm() / total.norm())
    assert err <= 2e-4, err          # float-atomic ordering of the compositing backward differs run to run (gate: 1e-3)
