#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump of the blind-rotation kernel per barrier-delimited segment:
instruction mix (per warp-iteration) and stall-sample shares.  Usage: ncu_segments.py file.csv [warp_iters]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
segs, cur = [], dict(n=0, samples=0, mix=collections.Counter(), stalls=collections.Counter(), execd=0)
def cat(op):
    if op.startswith(("DFMA", "DADD", "DMUL", "DSETP")): return "fp64"
    if op.startswith(("LDS", "STS")): return "smem"
    if op.startswith(("LDG", "STG", "LD.", "ST.")): return "gmem"
    if op.startswith(("LDL", "STL")): return "local"
    if op.startswith(("F2I", "I2F", "F2F", "FRND", "MUFU")): return "xu"
    if op.startswith(("BAR", "WARPSYNC", "BRA", "EXIT", "NOP", "BSSY", "BSYNC")): return "ctrl"
    if op.startswith(("IMAD", "FFMA", "FMUL", "FADD")): return "fma"
    return "alu"
loop_exec = max(int(r[ix["Instructions Executed"]] or 0) for r in data)
for r in data:
    src = r[ix["Source"]].split()
    if not src: continue
    op = src[1] if src[0].startswith("@") else src[0]
    ex = int(r[ix["Instructions Executed"]] or 0)
    s = int(r[ix["# Samples"]] or 0)
    if ex >= loop_exec // 2:   # inside the main loop
        cur["n"] += 1; cur["samples"] += s; cur["mix"][cat(op)] += ex / loop_exec
        for c in stall_cols: cur["stalls"][c] += int(r[ix[c]] or 0)
    if op.startswith("BAR") and ex >= loop_exec // 2 or (op == "BRA" and ex >= loop_exec // 2 and cur["n"] > 500):
        segs.append(cur); cur = dict(n=0, samples=0, mix=collections.Counter(), stalls=collections.Counter(), execd=0)
if cur["n"]: segs.append(cur)
tot = sum(s["samples"] for s in segs)
for i, s in enumerate(segs):
    print(f"segment {i}: {s['n']} SASS instr, {100*s['samples']/tot:.1f}% of samples; per-warp-iter mix:",
          {k: round(v) for k, v in s["mix"].most_common()})
    top = s["stalls"].most_common(6)
    print("     stalls:", {k.replace('stall_', ''): f"{100*v/max(1,s['samples']):.0f}%" for k, v in top})
