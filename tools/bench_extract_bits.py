"""Throughput of the batched bit extraction (b200tfhe_extract_bits_batch_u64) at PARAM_MESSAGE_2_CARRY_2, host buffers in and out.
Prints one JSON line.  Usage: python tools/bench_extract_bits.py [--batch 4096] [--nbits 5] [--reps 5]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--nbits", type=int, default=5)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    import tfhe_rs_odd_b200 as t
    ctx = t.Context(0)
    n, k, N, b, l, bks, lks = 742, 1, 2048, 23, 1, 3, 5
    p = t.core_crypto.make_params(n, k, N, b, l, bks, lks, 64)
    rng = np.random.default_rng(1)
    # synthetic uniform keys: timing does not depend on key values
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, 2 ** 64, n * l * (k + 1) ** 2 * N, dtype=np.uint64))
    ksk = t.LweKeyswitchKey.upload(ctx, p, rng.integers(0, 2 ** 64, k * N * lks * (n + 1), dtype=np.uint64))
    cts = rng.integers(0, 2 ** 64, size=(a.batch, p.glwe_dimension * p.polynomial_size + 1), dtype=np.uint64)
    f = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized
    f(cts, bsk, ksk, 64 - a.nbits, a.nbits)
    ts = []
    for _ in range(a.reps):
        t0 = time.perf_counter()
        f(cts, bsk, ksk, 64 - a.nbits, a.nbits)
        ts.append(time.perf_counter() - t0)
    dt = float(np.median(ts))
    print(json.dumps({"op": "extract_bits", "params": "MESSAGE_2_CARRY_2", "batch": a.batch, "nbits": a.nbits,
                      "ms_per_call": dt * 1e3, "ciphertexts_per_s": a.batch / dt,
                      "ks_pbs_rounds_per_s": a.batch * (a.nbits - 1) / dt}))


if __name__ == "__main__":
    main()
