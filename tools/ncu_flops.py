#!/usr/bin/env python
"""FP32 flops executed per kernel launch, from an ncu report captured with `--set full --import-source on`:
sums `Predicated-On Thread Instructions Executed` of the SASS page by opcode.  Weights: FFMA 2, FFMA2 4, FADD / FMUL 1,
FADD2 / FMUL2 2, MUFU 1 (one special-function result), F2F/I2F etc. 0.  Also prints the opcode mix of the FP pipes.
usage: ncu_flops.py report.ncu-rep [--json out.json --sweeps-lightsel N --Lpad N --elements-shade N ...]"""
import csv
import io
import json
import re
import subprocess
import sys

W = {"FFMA": 2, "FFMA2": 4, "FADD": 1, "FMUL": 1, "FADD2": 2, "FMUL2": 2, "MUFU": 1, "FMNMX": 0, "FSEL": 0, "FSETP": 0,
     "DFMA": 0, "DADD": 0, "DMUL": 0}

def launches(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    cur = None
    out = []
    hdr = None
    for row in csv.reader(io.StringIO(txt)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            cur = {"kernel": re.sub(r"mcpt::|<unnamed>::|\(anonymous namespace\)::", "", row[1]).split("(")[0].replace("void ", ""), "ops": {}, "thread_inst": 0, "warp_inst": 0}
            out.append(cur); hdr = None; continue
        if row[0] == "Address":
            hdr = row; iS = hdr.index("Source"); iP = hdr.index("Predicated-On Thread Instructions Executed"); iI = hdr.index("Instructions Executed"); continue
        if cur is None or hdr is None:
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", row[iS])
        if not m:
            continue
        op = m.group(2)
        try:
            n = int(row[iP]); wi = int(row[iI])
        except ValueError:
            continue
        cur["ops"][op] = cur["ops"].get(op, 0) + n
        cur["thread_inst"] += n; cur["warp_inst"] += wi
    # the page lists every launch twice (CUDA-C view and SASS view share the table): drop exact repeats in sequence
    ded = []
    for l in out:
        if ded and ded[-1]["kernel"] == l["kernel"] and ded[-1]["ops"] == l["ops"]:
            continue
        ded.append(l)
    for l in ded:
        l["flops"] = sum(W.get(op, 0) * n for op, n in l["ops"].items())
    return ded

if __name__ == "__main__":
    rep = sys.argv[1]
    L = launches(rep)
    for i, l in enumerate(L):
        fp = {k: v for k, v in l["ops"].items() if k in W and v}
        print("%2d %-28s thread-inst %.4g  warp-inst %.4g  flops %.4g  " % (i, l["kernel"], l["thread_inst"], l["warp_inst"], l["flops"]),
              {k: "%.3g" % v for k, v in sorted(fp.items(), key=lambda kv: -kv[1])})
    if "--json" in sys.argv:
        json.dump(L, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
