"""Probe: does this box expose NVLink multicast (NVLS) / peer mappings through torch symmetric memory?
torchrun --nproc-per-node 2 tools/symm_probe.py"""
import os, time, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm_mem
    n = 64 * 1024 * 1024
    t = symm_mem.empty(n, dtype=torch.float32, device=torch.device("cuda", local))
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    mc = getattr(hdl, "multicast_ptr", None)
    print(f"rank {rank}: symm_mem ok, buffer_ptrs={[hex(p) for p in hdl.buffer_ptrs]}, multicast_ptr={hex(mc) if mc else mc}, "
          f"signal_pad_ptrs={len(hdl.signal_pad_ptrs)}", flush=True)
    t.fill_(float(rank + 1))
    hdl.barrier()
    if mc:
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(3):
            ev0.record()
            torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
            ev1.record(); torch.cuda.synchronize()
            print(f"rank {rank}: multimem_all_reduce_ 256 MB: {ev0.elapsed_time(ev1):.3f} ms, t[0]={float(t[0])}", flush=True)
    x = torch.ones(59_000_000, device="cuda")
    for it in range(3):
        torch.cuda.synchronize(); ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
        ev0.record(); dist.all_reduce(x); ev1.record(); torch.cuda.synchronize()
        print(f"rank {rank}: nccl all_reduce 236 MB: {ev0.elapsed_time(ev1):.3f} ms", flush=True)
except Exception as ex:
    print(f"rank {rank}: symmetric memory probe failed: {type(ex).__name__}: {ex}", flush=True)
dist.destroy_process_group()
