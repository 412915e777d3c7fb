#!/usr/bin/env python
"""Throughput of two device-resident circuits (BASELINE.json configs[3] and [4]) with synthetic uniform keys (timing is
data-independent; correctness of the same circuits is checked in tests/test_waves.py with real keys):
  * M independent FheUint32 multiplications (16 blocks at PARAM_MESSAGE_2_CARRY_2) -> multiplications / s
  * S independent Trivium streams at PARAM_MESSAGE_1_CARRY_1 -> keystream bits / s
Wall clock around OpGraph.run + a stream synchronise; includes the Python scheduling of the operation graph."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfhe_rs_odd_b200 as t  # noqa: E402
from tfhe_rs_odd_b200 import parameters, radix, shortint  # noqa: E402
from tfhe_rs_odd_b200.trivium import TriviumStreamShortint  # noqa: E402
from tfhe_rs_odd_b200.waves import CiphertextPool, OpGraph  # noqa: E402


def keys(ctx, name, rng):
    p = parameters.get(name)
    n, k, N, l, lks = p.lwe_dimension, p.glwe_dimension, p.polynomial_size, p.pbs_level, p.ks_level
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, rng.integers(0, 2 ** 64, n * l * (k + 1) ** 2 * N, dtype=np.uint64))
    ksk = t.LweKeyswitchKey.upload(ctx, p, rng.integers(0, 2 ** 64, k * N * lks * (n + 1), dtype=np.uint64))
    mm = int(round(parameters.plaintext_modulus(name) ** 0.5))
    return p, ksk, bsk, shortint.ServerKey(ksk, bsk, mm, mm, "ks_pbs")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--muls", type=int, default=64)
    ap.add_argument("--streams", type=int, default=32)
    args = ap.parse_args()
    ctx = t.Context(0)
    rng = np.random.default_rng(5)
    out = {}

    p, ksk, bsk, sk = keys(ctx, "PARAM_MESSAGE_2_CARRY_2_KS_PBS", rng)
    M, NB = args.muls, 16
    per = 1200
    pool = CiphertextPool(ctx, p, capacity=M * per, big=True)
    pool.set_luts(np.stack(radix.radix_luts(sk)))
    pool.write(0, rng.integers(0, 2 ** 64, (2 * M * NB, pool.len), dtype=np.uint64))
    def mul_once():
        g = OpGraph(pool)
        alloc = radix.SlotAllocator(3 * M * NB, pool.capacity)
        t0 = time.perf_counter()
        for m in range(M):
            radix.mul(g, alloc, m * NB, M * NB + m * NB, 2 * M * NB + m * NB, NB)
        t1 = time.perf_counter()
        waves = g.run(ksk, bsk)
        ctx.synchronize()
        return t1 - t0, time.perf_counter() - t1, waves
    mul_once()
    build_s, run_s, waves = mul_once()
    pbs = sum(w[1] for w in waves)
    out["FheUint32 multiplication (16 blocks, PARAM_MESSAGE_2_CARRY_2)"] = {
        "independent_multiplications": M, "waves": len(waves), "bootstraps": pbs, "first_wave_bootstraps": waves[0][1],
        "graph_build_s": build_s, "run_s": run_s, "multiplications_per_s": M / (build_s + run_s), "ks_pbs_per_s": pbs / run_s,
        "reference_published": "333 ms per multiplication, all cores of m6i.metal (BASELINE.md)"}
    pool.close(); bsk.close(); ksk.close()

    p, ksk, bsk, sk = keys(ctx, "PARAM_MESSAGE_1_CARRY_1_KS_PBS", rng)
    S = args.streams
    luts = np.stack([sk.generate_lookup_table(lambda z: (z >> 1) & z & 1).acc, sk.generate_lookup_table(lambda z: z % 2).acc])
    key_cts = rng.integers(0, 2 ** 64, (S, 80, p.glwe_dimension * p.polynomial_size + 1), dtype=np.uint64)
    t0 = time.perf_counter()
    tr = TriviumStreamShortint(ctx, p, ksk, bsk, luts, key_cts, rng.integers(0, 2, (S, 80)), encode_bit=lambda b: b << 61)
    ctx.synchronize()
    init_s = time.perf_counter() - t0
    ts = []
    for _ in range(3):
        t0 = time.perf_counter(); tr.next_64(read=False); ctx.synchronize(); ts.append(time.perf_counter() - t0)
    dt = sorted(ts)[1]
    out["Trivium keystream (PARAM_MESSAGE_1_CARRY_1)"] = {
        "streams": S, "warmup_1152_steps_s": init_s, "s_per_64_bits_per_stream_batch": dt, "keystream_bits_per_s": 64 * S / dt,
        "bootstraps_per_batch": 64 * S * 8, "ks_pbs_per_s": 64 * S * 8 / dt, "reference_published": "none (BASELINE.md)"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
