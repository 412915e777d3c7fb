#!/usr/bin/env python
"""A/B of the experimental variants of blind_rotate_x16_kernel (context option "x16_variant") against the generic kernel at boolean
DEFAULT_PARAMETERS: device-resident PBS launches timed with CUDA events on the library's stream, decrypt-checked."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

o = Oracle("BOOLEAN_DEFAULT").keygen(fourier=False)
ctx = t.Context(0)
lib = ctx._lib
p = t.parameters.get("BOOLEAN_DEFAULT_PARAMETERS")
bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk.astype(np.uint32))
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
bits = (np.arange(B) % 2).astype(bool)
cts = np.tile(o.encrypt(np.where(bits[:64], 1 << 29, 7 << 29).astype(np.uint64), "small", seed=7).astype(np.uint32), (B // 64, 1))
lut = o.boolean_lut().astype(np.uint32)
d_in = torch.from_numpy(cts.view(np.int32)).cuda()
d_lut = torch.from_numpy(lut.view(np.int32)).cuda()
d_out = torch.zeros((B, o.big + 1), dtype=torch.int32, device="cuda")
stream = torch.cuda.ExternalStream(ctx.stream)
VP = C.c_void_p
res = {}
variants = [("generic", 1, 0), ("x16_v0", 6, 0), ("x16_v1_tma", 6, 1), ("x16b_all_transforms_first", 6, 4), ("x16b_tma_ring", 6, 12), ("generic_again", 1, 0)]
for name, kern, var in variants:
    ctx.set_option("br_kernel", kern)
    ctx.set_option("x16_variant", var)
    run = lambda: ctx.check(lib.b200tfhe_pbs_batch_dev(ctx._h, bsk._h, VP(d_in.data_ptr()), VP(d_out.data_ptr()), VP(d_lut.data_ptr()), None, B))
    d_out.zero_()
    torch.cuda.synchronize()
    run(); run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(3):
        run()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    out = d_out[:64].cpu().numpy().view(np.uint32)
    ok = bool(np.array_equal(o.phase(out.astype(np.uint64), "big") < np.uint64(1 << 31), bits[:64]))
    res[name] = {"ms": ms, "pbs_per_s": B / ms * 1e3, "decrypt_ok": ok}
ctx.set_option("x16_variant", -1)
print(json.dumps(res, indent=1))
