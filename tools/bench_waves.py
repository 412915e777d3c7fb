#!/usr/bin/env python
"""Throughput of two device-resident circuits (BASELINE.json configs[3] and [4]) with synthetic uniform keys (timing is
data-independent; correctness of the same circuits is checked in tests/test_waves.py with real keys):
  * M independent FheUint32 multiplications (16 blocks at PARAM_MESSAGE_2_CARRY_2) -> multiplications / s
  * S independent Trivium streams at PARAM_MESSAGE_1_CARRY_1 -> keystream bits / s
Wall clock around OpGraph.run + a stream synchronise; includes the Python scheduling of the operation graph.
Under torchrun (one process per GPU) every rank runs its own M multiplications / S streams on its own device -- independent
circuits stay on one GPU, so intermediate ciphertexts never cross GPUs -- with a barrier on both sides of each timed
region; the reported time is the maximum over ranks and the totals are whole-job (weak scaling)."""
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
    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        v = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        return float(v.item())

    ctx = t.Context(local)
    rng = np.random.default_rng(5)     # same synthetic keys on every rank (bench.py shows the NCCL key broadcast)
    out = {"n_gpus": world}

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
        barrier()
        t1 = time.perf_counter()
        waves = g.run(ksk, bsk)
        ctx.synchronize()
        barrier()
        return max_over_ranks(t1 - t0), max_over_ranks(time.perf_counter() - t1), waves
    mul_once()
    build_s, run_s, waves = mul_once()
    pbs = sum(w[1] for w in waves)
    out["FheUint32 multiplication (16 blocks, PARAM_MESSAGE_2_CARRY_2)"] = {
        "independent_multiplications": M * world, "per_gpu": M, "waves": len(waves), "bootstraps": pbs * world, "first_wave_bootstraps": waves[0][1],
        "graph_build_s": build_s, "run_s": run_s, "multiplications_per_s": M * world / (build_s + run_s), "ks_pbs_per_s": pbs * world / run_s,
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
        barrier()
        t0 = time.perf_counter(); tr.next_64(read=False); ctx.synchronize(); barrier(); ts.append(max_over_ranks(time.perf_counter() - t0))
    dt = sorted(ts)[1]
    out["Trivium keystream (PARAM_MESSAGE_1_CARRY_1)"] = {
        "streams": S * world, "per_gpu": S, "warmup_1152_steps_s": init_s, "s_per_64_bits_per_stream_batch": dt, "keystream_bits_per_s": 64 * S * world / dt,
        "bootstraps_per_batch": 64 * S * 8 * world, "ks_pbs_per_s": 64 * S * 8 * world / dt, "reference_published": "none (BASELINE.md)"}
    if rank == 0:
        print(json.dumps(out, indent=1))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
