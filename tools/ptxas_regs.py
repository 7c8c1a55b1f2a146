#!/usr/bin/env python
"""Registers / spills / stack per kernel from csrc/build/*.ptxas.log (nvcc -Xptxas -v).  usage: ptxas_regs.py [filter]"""
import glob, os, re, subprocess, sys
here = os.path.dirname(os.path.abspath(__file__))
flt = sys.argv[1] if len(sys.argv) > 1 else ""
for f in sorted(glob.glob(os.path.join(here, "..", "monte_carlo_path_tracing_b200", "csrc", "build", "*.ptxas.log"))):
    name = None
    for line in open(f):
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::|mcpt::", "", name).split("(")[0][:60]
            stack = ""
            continue
        if "stack frame" in line:
            stack = line.strip()
        m = re.search(r"Used (\d+) registers", line)
        if m and name and flt in name:
            print("%-62s %4s regs  %s" % (name, m.group(1), stack))
