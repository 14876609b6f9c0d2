"""Microbenchmark of the gradient exchange alone: this library's NVSwitch-multicast kernel vs NCCL all_reduce on a bucket of
the bench's size (59 floats x 1M Gaussians = 236 MB).  torchrun --nproc-per-node N tools/mm_bench.py"""
import ctypes
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vcr-gaus_b200"))
import vcr_native as N  # noqa: E402
import vcr_view_parallel as VP  # noqa: E402

rank, world, local = VP.init_from_env()
dev = torch.device("cuda", local)
vp = VP.ViewParallel()
numel = 59_000_000
store = vp.symmetric_gradient_storage(numel, dev)
plain = torch.ones(numel, device=dev)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


res = {"world": world, "MB": numel * 4 / 1e6}
res["nccl_ms"] = timed(lambda: dist.all_reduce(plain))
if store is not None:
    t, hdl = vp._symm
    t.fill_(1.0)
    mc = ctypes.c_void_p(int(hdl.multicast_ptr))
    arr = ctypes.c_longlong * 1
    for blocks in (18, 37, 74, 148, 296):
        def mm(blocks=blocks):
            s = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            hdl.barrier(channel=0, timeout_ms=30000)
            N.check(N.vcr_multimem_allreduce_sum(mc, 1, arr(0), arr(numel), rank, world, blocks, s), "mm")
            hdl.barrier(channel=0, timeout_ms=30000)
        res[f"multimem_{blocks}_ms"] = timed(mm)
    res["barrier_pair_ms"] = timed(lambda: (hdl.barrier(channel=0, timeout_ms=30000), hdl.barrier(channel=0, timeout_ms=30000)))
    res["nccl_on_symmetric_ms"] = timed(lambda: dist.all_reduce(t))
else:
    res["multimem"] = "no multicast support"
if rank == 0:
    print("MMBENCH " + str(res), flush=True)
dist.barrier()
dist.destroy_process_group()
