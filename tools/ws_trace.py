#!/usr/bin/env python
"""Phase timeline of a -DB200TFHE_WS_TRACE build of blind_rotate_ws.cu (file named by B200TFHE_TRACE_FILE)."""
import sys
rows = {}
for ln in open(sys.argv[1]):
    v = [int(x) for x in ln.split()]
    rows[(v[0], v[1])] = v[2:]
t0 = rows[(0, 1)][0]
print("Fourier warps: start | top | waitX0 | ldX0 | fwd0 | waitX1 | ldX1 | fwd1 | mac0 | inv0 | mac1 | inv1 | iteration")
for w in range(4):
    for it in (1, 2):
        s = rows[(w, it)]
        d = [s[k + 1] - s[k] for k in range(11)]
        print(f"F{w} it{it} {s[0] - t0:7d} |", " ".join(f"{x:6d}" for x in d), "|", rows[(w, it + 1)][0] - s[0])
print("torus warps: start | waitY0 | acc0 | dec0 | gap | waitY1 | acc1 | dec1 | iteration")
for w in range(4, 8):
    for it in (1, 2):
        s = rows[(w, it)]
        d = [s[k + 1] - s[k] for k in range(7)]
        print(f"T{w - 4} it{it} {s[0] - t0:7d} |", " ".join(f"{x:6d}" for x in d), "|", rows[(w, it + 1)][0] - s[0])
