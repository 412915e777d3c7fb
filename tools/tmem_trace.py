#!/usr/bin/env python
"""Print the phase timeline recorded by a -DB200TFHE_TMEM_TRACE build of blind_rotate_tmem.cu (file named by
B200TFHE_TRACE_FILE): per warp, the start offset of each iteration relative to warp 0 and the duration of each phase."""
import sys
NAMES = ["ring-wait", "decompose", "fwd-pass1", "transpose", "pass2+own", "ct-barrier", "mac", "inv-pass1", "transpose",
         "inv-pass2", "accumulate"]
rows = {}
for ln in open(sys.argv[1]):
    v = [int(x) for x in ln.split()]
    rows[(v[0], v[1])] = v[2:]
t0 = rows[(0, 0)][0]
its = sorted({k[1] for k in rows})
print("warp(q,r) iter  start  | " + " ".join(f"{n[:9]:>9}" for n in NAMES) + " |  total")
for w in range(8):
    for it in its:
        s = rows[(w, it)]
        d = [s[k + 1] - s[k] for k in range(11)]
        print(f"w{w} ({w & 3},{w >> 2})  {it}  {s[0] - t0:7d} | " + " ".join(f"{x:9d}" for x in d) + f" | {s[11] - s[0]:6d}")
