#!/usr/bin/env python
"""Record a torch.profiler (CUPTI) trace of a few bench steps and print the GPU timeline of one step:
kernel busy time, idle gaps, and the largest gaps with the kernels around them.  usage: trace_step.py [ours|reference]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

def main(impl):
    sys.argv = ["bench.py", "--steps", "3", "--warmup", "3", "--no-cpu-baseline", "--impl", impl]
    import bench_common as BC
    args = BC.parse()
    dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
    scene = BC.build_scene(args)
    views = BC.build_views(args, scene, 1, list(range(args.views)))
    if impl == "ours":
        bench._product_paths()
        import vcr_view_parallel as VP
        st = bench.HotPathStepper(args, scene, views, dev, VP.ViewParallel())
    else:
        import bench_ref
        st = bench_ref.HotPathStepper(args, bench_ref._import_reference(), scene, views, dev, 1)
    for i in range(4): st.step(i, False)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(4, 7): st.step(i, False)
        torch.cuda.synchronize()
    out = os.path.join(ROOT, "gpurun_out", f"trace_{impl}.json")
    prof.export_chrome_trace(out)
    ev = json.load(open(out))["traceEvents"]
    k = sorted([e for e in ev if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")], key=lambda e: e["ts"])
    t0, t1 = k[0]["ts"], max(e["ts"] + e["dur"] for e in k)
    busy = sum(e["dur"] for e in k)
    print(f"{impl}: 3 steps span {(t1 - t0) / 1e3:.3f} ms, GPU busy {busy / 1e3:.3f} ms, idle {(t1 - t0 - busy) / 1e3:.3f} ms, {len(k)} GPU ops")
    gaps = []
    end = k[0]["ts"] + k[0]["dur"]
    for a, b in zip(k, k[1:]):
        end = max(end, a["ts"] + a["dur"])
        g = b["ts"] - end
        if g > 3: gaps.append((g, a["name"][:50], b["name"][:50]))
    gaps.sort(reverse=True)
    print("largest idle gaps (us): before <- after")
    for g, a, b in gaps[:25]: print(f"  {g:8.1f}  {a:50s} -> {b}")
    print("sum of gaps > 3us:", sum(g for g, _, _ in gaps) / 1e3, "ms")
    agg = {}
    for e in k: agg[e["name"][:60]] = agg.get(e["name"][:60], 0) + e["dur"]
    for n, d in sorted(agg.items(), key=lambda kv: -kv[1])[:25]: print(f"  {d / 3:9.1f} us/step  {n}")

if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "ours")
