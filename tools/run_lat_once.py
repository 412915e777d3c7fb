#!/usr/bin/env python
"""One device-resident blind rotation of ONE ciphertext at PARAM_MESSAGE_2_CARRY_2 (the latency kernel): the target of an ncu capture.
   python tools/run_lat_once.py [batch] [lat_cluster_max]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402

ctx = t.Context(0)
p = t.parameters.get("PARAM_MESSAGE_2_CARRY_2")
rng = np.random.default_rng(1)
n, k, N, l = p.lwe_dimension, p.glwe_dimension, p.polynomial_size, p.pbs_level
bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, 2 ** 63, n * l * (k + 1) ** 2 * N, dtype=np.uint64))
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
if len(sys.argv) > 2:
    ctx.set_option("lat_cluster_max", int(sys.argv[2]))
d_in = torch.randint(-2 ** 63, 2 ** 63 - 1, (B, n + 1), dtype=torch.int64, device="cuda")
d_lut = torch.randint(-2 ** 63, 2 ** 63 - 1, ((k + 1) * N,), dtype=torch.int64, device="cuda")
d_out = torch.zeros((B, k * N + 1), dtype=torch.int64, device="cuda")
torch.cuda.synchronize()
stream = torch.cuda.ExternalStream(ctx.stream)
lat = []
for it in range(3 if os.environ.get("NCU") else 260):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ctx.check(ctx._lib.b200tfhe_pbs_batch_dev(ctx._h, bsk._h, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), C.c_void_p(d_lut.data_ptr()), None, B))
    e1.record(stream)
    torch.cuda.synchronize()
    if it >= 200:
        lat.append(e0.elapsed_time(e1))
lat.sort()
print("ok", f"batch {B}: blind rotation p50 {lat[len(lat) // 2]:.4f} ms" if lat else "")
