"""Throughput of the whole WoP-PBS of one shortint block at WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS (n 769, N 2048, pbs 15/2, ks 6/2,
pfks 15/2, cbs 5/3): extract 4 bits, circuit bootstrap + vertical packing with `--luts` tables; host buffers in and out, synthetic
uniform keys (timing is data independent).  Prints one JSON line.  Usage: python tools/bench_wopbs.py [--batch 256] [--reps 3]"""
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
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--nbits", type=int, default=4)
    ap.add_argument("--luts", type=int, default=1)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    import tfhe_rs_odd_b200 as t
    cc = t.core_crypto
    ctx = t.Context(0)
    n, k, N, b, l, bks, lks = 769, 1, 2048, 15, 2, 6, 2
    pfb, pfl, cb, cl = 15, 2, 5, 3
    p = cc.make_params(n, k, N, b, l, bks, lks, 64)
    rng = np.random.default_rng(1)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, 2 ** 64, n * l * (k + 1) ** 2 * N, dtype=np.uint64))
    ksk = t.LweKeyswitchKey.upload(ctx, p, rng.integers(0, 2 ** 64, k * N * lks * (n + 1), dtype=np.uint64))
    pf = cc.LwePrivateFunctionalPackingKeyswitchKeyList.upload(
        ctx, p, pfb, pfl, rng.integers(0, 2 ** 64, (k + 1) * (k * N + 1) * pfl * (k + 1) * N, dtype=np.uint64))
    cts = rng.integers(0, 2 ** 64, size=(a.batch, k * N + 1), dtype=np.uint64)
    luts = rng.integers(0, 2 ** 64, size=(a.luts, N), dtype=np.uint64)

    def run():
        t0 = time.perf_counter()
        bits = cc.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 64 - a.nbits, a.nbits)
        t1 = time.perf_counter()
        cc.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(bits, luts, bsk, pf, cb, cl)
        return t1 - t0, time.perf_counter() - t1

    run()
    ts = np.array([run() for _ in range(a.reps)])
    eb, cv = np.median(ts[:, 0]), np.median(ts[:, 1])
    print(json.dumps({"op": "wopbs", "params": "WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS", "batch": a.batch, "nbits": a.nbits, "luts": a.luts,
                      "extract_bits_ms": eb * 1e3, "cbs_vertical_packing_ms": cv * 1e3, "blocks_per_s": a.batch / (eb + cv),
                      "bootstraps_per_block": (a.nbits - 1) + a.nbits * cl}))


if __name__ == "__main__":
    main()
