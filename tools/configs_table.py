#!/usr/bin/env python
"""Collect the bench JSON lines of a round into profiles/rNN_configs.md (one row per config / sweep point, both arms).
usage: python tools/configs_table.py OUT.md LABEL=path.json[,ref.json] ...   and  --sweep gpurun_out  for the config E points"""
import glob
import json
import os
import re
import sys


def load(p):
    try:
        return json.loads(open(p).read().strip().splitlines()[-1])
    except Exception:
        return None


def row(label, ours, ref):
    sc = ours.get("scene", {}) if ours else {}
    k = {r["kernel"]: r for r in (ours or {}).get("kernels", [])}
    def us(n):
        return f"{k[n]['avg_launch_ms'] * 1e3:.0f}" if n in k else "-"
    def gbs(n):
        return f"{k[n]['achieved_gbs']:.0f}" if n in k else "-"
    o, r = (ours or {}).get("ms_per_step"), (ref or {}).get("ms_per_step")
    oe, re_ = (ours or {}).get("e2e", {}).get("ms_per_step"), (ref or {}).get("e2e", {}).get("ms_per_step")
    clk = (ours or {}).get("clocks") or {}
    f = lambda x, d=3: "-" if x is None else f"{x:.{d}f}"  # noqa: E731
    return (f"| {label} | {f(o)} | {f(r)} | {f(r / o, 2) if o and r else '-'} | {f(oe)} | {f(re_)} | {sc.get('P', '-')} | {sc.get('R', '-')} | "
            f"{f(sc.get('R_per_P'), 2)} | {sc.get('sum_n_contrib', '-')} | {sc.get('K_e', '-')} | {sc.get('K_c', '-')} | {us('render_forward_kernel')} / {us('render_backward_kernel')} | "
            f"{gbs('preprocess_kernel')} / {gbs('radix_pass_kernel')} / {gbs('preprocess_backward_kernel')} | {clk.get('sm_mhz', '-')} {','.join(clk.get('reasons', [])) or 'none'} |")


def main():
    out = sys.argv[1]
    lines = ["| config | ours ms/step | reference ms/step | ratio | ours e2e ms | reference e2e ms | P | R | R/P | sum n_contrib | K_e | K_c | fwd / bwd compositing us | alg. GB/s preprocess / radix pass / per-Gaussian bwd | SM MHz, throttle reasons |",
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for a in sys.argv[2:]:
        if a.startswith("--sweep="):
            d = a.split("=", 1)[1]
            pts = sorted(glob.glob(os.path.join(d, "sweep_ours_*.json")), key=lambda p: (int(re.findall(r"_(\d+)x", p)[0]), int(re.findall(r"ours_(\d+)_", p)[0])))
            for p in pts:
                P, res = re.findall(r"ours_(\d+)_(\d+x\d+)", p)[0]
                lines.append(row(f"E: {int(P):,} @ {res}", load(p), load(p.replace("sweep_ours_", "sweep_reference_"))))
            continue
        label, paths = a.split("=", 1)
        ps = paths.split(",")
        lines.append(row(label, load(ps[0]), load(ps[1]) if len(ps) > 1 else None))
    open(out, "a").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
