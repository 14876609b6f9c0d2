#!/usr/bin/env python
"""Summarise `-Xptxas -v` logs: kernel, registers, smem, spills."""
import re, sys, glob, subprocess
for f in sorted(glob.glob(sys.argv[1] if len(sys.argv) > 1 else "vcr-gaus_b200/build/*.ptxas.log")):
    txt = open(f).read()
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?(\d+) bytes spill stores, (\d+) bytes spill loads.*?Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", txt, re.S):
        name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
        print(f"{name[:70]:70s} regs={m.group(4):>3s} smem={m.group(5) or 0:>6} spill_st={m.group(2)} spill_ld={m.group(3)}")
