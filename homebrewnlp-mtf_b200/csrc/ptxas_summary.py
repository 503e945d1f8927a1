#!/usr/bin/env python
"""Summarise build/*.ptxas.log: registers, spills, shared memory per kernel (reads nvcc -Xptxas -v output)."""
import glob, re, subprocess, sys
rows = []
for f in sorted(glob.glob("build/*.ptxas.log")):
    txt = open(f).read()
    for m in re.finditer(r"Compiling entry function '(\S+)'.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(.*)", txt):
        name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
        rows.append((name[:70], int(m.group(5)), int(m.group(2)), int(m.group(3))))
for r in rows:
    if len(sys.argv) < 2 or r[2] or r[3] or sys.argv[1] in r[0]:
        print(f"{r[0]:70s} regs={r[1]:3d} stack={r[2]:5d} spill_st={r[3]}")
