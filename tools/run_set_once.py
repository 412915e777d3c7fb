#!/usr/bin/env python
"""One device-resident blind rotation of a batch at a named parameter set (random key and inputs): the target of an ncu capture.
   python tools/run_set_once.py PARAM_MESSAGE_1_CARRY_1 16384 [br_kernel]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402

ctx = t.Context(0)
p = t.parameters.get(sys.argv[1])
B = int(sys.argv[2])
if len(sys.argv) > 3:
    ctx.set_option("br_kernel", int(sys.argv[3]))
rng = np.random.default_rng(1)
n, k, N, l = p.lwe_dimension, p.glwe_dimension, p.polynomial_size, p.pbs_level
dt, tdt, lo, hi = (np.uint64, torch.int64, -2 ** 63, 2 ** 63 - 1) if p.torus_bits == 64 else (np.uint32, torch.int32, -2 ** 31, 2 ** 31 - 1)
bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, 2 ** 32, n * l * (k + 1) ** 2 * N, dtype=np.uint64).astype(dt))
d_in = torch.randint(lo, hi, (B, n + 1), dtype=tdt, device="cuda")
d_lut = torch.randint(lo, hi, ((k + 1) * N,), dtype=tdt, device="cuda")
d_out = torch.zeros((B, k * N + 1), dtype=tdt, device="cuda")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
stream = torch.cuda.ExternalStream(ctx.stream)
for it in range(3):
    if it == 2:
        e0.record(stream)
    ctx.check(ctx._lib.b200tfhe_pbs_batch_dev(ctx._h, bsk._h, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), C.c_void_p(d_lut.data_ptr()), None, B))
e1.record(stream)
ctx.synchronize()
torch.cuda.synchronize()
print(f"ok {sys.argv[1]} batch {B}: {e0.elapsed_time(e1):.3f} ms -> {B / e0.elapsed_time(e1):.1f} k PBS/s")
