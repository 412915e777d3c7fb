#!/usr/bin/env python
"""Measured error variance of the bootstrapped phase (BASELINE north_star) at PARAM_MESSAGE_2_CARRY_2 with real keys and
noise: GPU (through the C ABI) vs the CPU oracle vs the analytic law.  Prints one JSON object (committed copy:
profiles/r01_noise.json).  Uses oracle/ as the checker only.   python tools/measure_noise.py [--batch 4096] [--ref 256]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--ref", type=int, default=256)
    args = ap.parse_args()
    out = {}
    ctx = t.Context(0)
    for name in ["MESSAGE_2_CARRY_2", "MESSAGE_1_CARRY_1"]:
        o = Oracle(name).keygen()
        p = t.core_crypto.make_params(o.n, o.k, o.N, o.b, o.l, o.b_ks, o.l_ks, o.bits)
        bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
        rng = np.random.default_rng(3)
        msgs = rng.integers(0, o.pmod, args.batch)
        f = lambda x: (5 * x + 3) % o.pmod
        lut = o.lut(f)
        cts = o.encrypt(o.encode(msgs), "small")
        want = o.encode(np.array([f(int(m)) for m in msgs]))
        sd = lambda a, b: (a - b).astype(np.int64).astype(np.float64) / 2.0 ** 64
        e_gpu = sd(o.phase(t.programmable_bootstrap_lwe_ciphertext(cts, lut, bsk)), want)
        e_ref = sd(o.phase(o.pbs(cts[:args.ref], lut)), want[:args.ref])
        B2 = 2.0 ** (2 * o.b)
        theory = o.n / 2 * (1 + o.k * o.N / 2) / (12 * B2 ** o.l) + o.n * o.l * (o.k + 1) * o.N * (B2 + 2) / 12 * o.p.glwe_std ** 2
        out[name] = {"gpu_samples": args.batch, "oracle_samples": args.ref,
                     "gpu_log2_variance": float(np.log2(e_gpu.var())), "oracle_log2_variance": float(np.log2(e_ref.var())),
                     "analytic_log2_variance": float(np.log2(theory)), "gpu_mean": float(e_gpu.mean()),
                     "gpu_log2_max_abs_error": float(np.log2(np.abs(e_gpu).max())),
                     "decoding_margin_log2": float(np.log2(1.0 / (4 * o.pmod))),
                     "all_decrypt_correctly": bool((o.decode(o.phase(t.programmable_bootstrap_lwe_ciphertext(cts, lut, bsk))) % o.pmod == np.array([f(int(m)) for m in msgs])).all())}
        bsk.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
