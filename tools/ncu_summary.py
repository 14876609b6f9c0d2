#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, without a GPU) into a small text file for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/NAME.txt

Per kernel launch: duration, registers, occupancy, issue utilisation, instruction counts, DRAM bytes (traffic), L2/L1
throughput, top stall reasons, and the SASS regions grouped by execution count (needs -lineinfo / --import-source on).
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
    "sm__inst_executed_pipe_xu.sum", "smsp__inst_executed_pipe_lsu.sum", "smsp__inst_executed_op_global_red.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main(rep):
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu summary of {rep}\n")
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        print(f"## {name[:100]}")
        for m in METRICS:
            if m in idx and r[idx[m]] not in ("", "n/a"):
                print(f"  {m:72s} {r[idx[m]]:>20s} {units[idx[m]]}")
        try:
            rd = float(r[idx["dram__bytes_read.sum"]].replace(",", ""))
            wr = float(r[idx["dram__bytes_write.sum"]].replace(",", ""))
            ur, uw = units[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_write.sum"]]
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            print(f"  {'traffic = dram read + write':72s} {(rd * scale.get(ur, 1) + wr * scale.get(uw, 1)) / 1e6:>20.3f} MB")
        except Exception:
            pass
        st = []
        for h, i in idx.items():
            if "smsp__average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio") and r[i] not in ("", "n/a"):
                st.append((float(r[i].replace(",", "")), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
        st.sort(reverse=True)
        print("  top stall reasons (warps per issue-active cycle): " + ", ".join(f"{n} {v:.2f}" for v, n in st[:6]))
        print()
    # SASS regions by execution count, per kernel
    src = ncu_csv(rep, "source")
    cur, block = None, []

    def flush():
        if not block:
            return
        h = block[0]
        try:
            i_ex, i_s = h.index("Instructions Executed"), h.index("# Samples")
        except ValueError:
            return
        data = [(int(r[i_ex] or 0), int(r[i_s] or 0)) for r in block[1:] if len(r) > max(i_ex, i_s)]
        tot, tots = sum(d[0] for d in data) or 1, sum(d[1] for d in data) or 1
        print(f"## SASS regions of {cur[:90]}  (total warp-instructions {tot:,})")
        lvl, first, acc, accs, n = None, 0, 0, 0, 0
        for k, (ex, s) in enumerate(data + [(-1, 0)]):
            if lvl is None or ex < 0 or abs(ex - lvl) > 0.15 * max(lvl, 1):
                if n and acc / tot > 0.01:
                    print(f"  instr {first:4d}-{k - 1:4d} ({n:3d} instr)  executed {lvl / 1e6:9.2f} M times each  {acc / tot * 100:5.1f}% of instructions  {accs / tots * 100:5.1f}% of stall samples")
                lvl, first, acc, accs, n = ex, k, 0, 0, 0
            acc += max(ex, 0); accs += s; n += 1
        print()

    for r in src:
        if len(r) >= 2 and r[0] == "Kernel Name":
            flush()
            cur, block = r[1], []
        elif cur is not None:
            block.append(r)
    flush()


if __name__ == "__main__":
    main(sys.argv[1])
