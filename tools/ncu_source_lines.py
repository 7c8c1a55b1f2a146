#!/usr/bin/env python
"""Per-source-line hot spots from an ncu report captured with --import-source on (code built with -lineinfo).
usage: ncu_source_lines.py report.ncu-rep [launch-index] [top-n]
Prints file:line, warp-instructions executed, average active threads, stall samples."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
# the output is a sequence of (File Path, Function Name, header, rows...) groups per launch and per file
groups = []
cur = None
fpath = fn = None
for row in csv.reader(io.StringIO(txt)):
    if not row:
        continue
    if row[0] == "File Path":
        fpath = row[1]; continue
    if row[0] == "Function Name":
        fn = row[1]; continue
    if row[0] == "Line No":
        cur = {"file": fpath, "fn": fn, "hdr": row, "rows": []}
        groups.append(cur); continue
    if cur is not None:
        cur["rows"].append(row)
# launches: a new launch starts when a (file) repeats for the same function sequence; split by seeing fpath repeat
launches = []
seen = set()
for g in groups:
    key = g["file"]
    if key in seen:
        launches.append([]); seen = set()
    if not launches:
        launches.append([])
    seen.add(key)
    launches[-1].append(g)
L = launches[which]
print("launch %d of %d: %s" % (which, len(launches), L[0]["fn"]))
items = []
tot_inst = tot_thr = tot_samp = 0
for g in L:
    h = g["hdr"]
    iI = h.index("Instructions Executed"); iT = h.index("Thread Instructions Executed"); iS = h.index("# Samples")
    for r in g["rows"]:
        if r[0] == "":
            continue  # SASS row
        try:
            inst = int(r[iI]); thr = int(r[iT]); samp = int(r[iS])
        except ValueError:
            continue
        items.append((inst, thr, samp, "%s:%s" % (g["file"].split("/")[-1], r[0]), r[1].strip()[:90]))
        tot_inst += inst; tot_thr += thr; tot_samp += samp
print("total warp-inst %d, avg active threads %.2f, samples %d" % (tot_inst, tot_thr / max(tot_inst, 1), tot_samp))
by = sorted(items, key=lambda x: -x[2])[:topn]
print("%7s %7s %6s  %-22s %s" % ("samp%", "inst%", "thr", "where", "source"))
for inst, thr, samp, where, src in by:
    print("%6.2f%% %6.2f%% %6.1f  %-22s %s" % (100 * samp / max(tot_samp, 1), 100 * inst / max(tot_inst, 1), thr / max(inst, 1), where, src))
