#!/usr/bin/env python
"""End-to-end throughput of the other BASELINE.json configs (parity-test cases, not bench lines) through the host-pointer
C ABI with synthetic uniform keys (timing is data-independent): boolean NAND batch, 1_1 / 2_2 / 3_3 apply_lookup_table.
Prints one JSON object; the committed copy is profiles/r01_configs.json.   python tools/bench_configs.py [--reps 3]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402
from tfhe_rs_odd_b200 import boolean  # noqa: E402

# name: (n, k, N, pbs_base_log, pbs_level, ks_base_log, ks_level, bits, batch)
CONFIGS = {
    "boolean DEFAULT_PARAMETERS NAND": (722, 2, 512, 6, 3, 3, 4, 32, 8192),
    "shortint PARAM_MESSAGE_1_CARRY_1": (684, 3, 512, 18, 1, 4, 3, 64, 16384),
    "shortint PARAM_MESSAGE_2_CARRY_2": (742, 1, 2048, 23, 1, 3, 5, 64, 16384),
    "shortint PARAM_MESSAGE_3_CARRY_3": (864, 1, 8192, 15, 2, 3, 6, 64, 2048),
    # multi-bit bootstrap keys: trailing element = LweBskGroupingFactor
    "shortint PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_2": (818, 1, 2048, 22, 1, 5, 3, 64, 16384, 2),
    "shortint PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3": (888, 1, 2048, 21, 1, 7, 2, 64, 16384, 3),
}


def flop_per_pbs(n, k, N, l):
    M = N // 2
    return n * ((k + 1) * (l + 1) * 5 * M * (M.bit_length() - 1) + (k + 1) ** 2 * l * M * 8 + (k + 1) * (l + 1) * M * 6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only", default="", help="substring of the config names to run")
    args = ap.parse_args()
    ctx = t.Context(0)
    rng = np.random.default_rng(1)
    out = {}
    for name, cfg in CONFIGS.items():
        if args.only not in name:
            continue
        (n, k, N, b, l, bks, lks, bits, B), G = cfg[:9], (cfg[9] if len(cfg) > 9 else 0)
        dt = np.uint64 if bits == 64 else np.uint32
        p = t.core_crypto.make_params(n, k, N, b, l, bks, lks, bits)
        big = k * N
        hi = 2 ** bits
        if G:
            bsk = t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(
                ctx, p, rng.integers(0, hi, (n // G << G) * l * (k + 1) ** 2 * N, dtype=np.uint64), G)
        else:
            bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, hi, n * l * (k + 1) ** 2 * N, dtype=np.uint64).astype(dt))
        ksk = t.LweKeyswitchKey.upload(ctx, p, rng.integers(0, hi, big * lks * (n + 1), dtype=np.uint64).astype(dt))
        if bits == 32:
            a = rng.integers(0, hi, (B, n + 1), dtype=np.uint64).astype(dt)
            c = rng.integers(0, hi, (B, n + 1), dtype=np.uint64).astype(dt)
            fn = lambda: boolean.gate_batch(ksk, bsk, 1, a, c, pbs_first=True)
        else:
            cts = rng.integers(0, hi, (B, big + 1), dtype=np.uint64)
            lut = rng.integers(0, hi, (k + 1) * N, dtype=np.uint64)
            fn = lambda: t.keyswitch_programmable_bootstrap(ksk, bsk, cts, lut)
        fn()
        ts = []
        for _ in range(args.reps):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        dtm = sorted(ts)[len(ts) // 2]
        out[name] = {"batch": B, "ms": dtm * 1e3, "ks_pbs_per_s_e2e": B / dtm,
                     "timing": "host wall clock around the host-pointer call (pageable numpy buffers, H2D + D2H inside)"}
        if not G:   # the classic flop model does not describe the multi-bit loop
            out[name].update(flop_per_pbs=flop_per_pbs(n, k, N, l), algorithmic_tflops=B / dtm * flop_per_pbs(n, k, N, l) / 1e12)
        else:
            out[name]["grouping_factor"] = G
        if bits == 64:   # single KS+PBS latency through the same host call
            one = cts[:1].copy()
            t.keyswitch_programmable_bootstrap(ksk, bsk, one, lut)
            l1 = []
            for _ in range(20):
                t0 = time.perf_counter(); t.keyswitch_programmable_bootstrap(ksk, bsk, one, lut); l1.append(time.perf_counter() - t0)
            out[name]["single_ks_pbs_ms_host_call_p50"] = sorted(l1)[10] * 1e3
        bsk.close(); ksk.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
