#!/usr/bin/env python
"""Generates tests/golden/next_rows_golden.npz: keys, inputs and outputs for the SURVEY 8f rows built after the KS+PBS path --
multi-bit PBS on a tiny set (n = 4, grouping factor 2) and the gadget engine's integer operations (packing keyswitch, GLWE x
plaintext polynomial + extraction, pack_into_new_accumulator) on a tiny polynomial size (N = 64) so the stored packing key stays
small.  Same provenance and purpose as make_golden.py: frozen outputs of the pinned oracle, checked against the oracle in the
`-m "not gpu"` suite and against the CUDA path in the `-m gpu` suite (integer operations bit for bit).
        python tests/golden/make_golden_next.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "next_rows_golden.npz")


def multibit_set(z):
    name = "GOLD_MB"
    o = Oracle(name, seed=0x601D0003).keygen()
    rng = np.random.default_rng(0x601F)
    B = 6
    msgs = rng.integers(0, o.pmod, B)
    cts = o.encrypt(o.encode(msgs), "big", seed=0x601D0400)
    luts = np.stack([o.lut(lambda x: x), o.lut(lambda x: (3 * x + 1) % o.pmod)])
    idx = (np.arange(B) % 2).astype(np.uint32)
    out = o.ks_pbs(cts, luts, idx)
    z.update({f"{name}.small_key": o.small_key.astype(np.uint8), f"{name}.glwe_key": o.glwe_key.astype(np.uint8),
              f"{name}.ksk": o.ksk, f"{name}.bsk": o.bsk, f"{name}.cts": cts, f"{name}.msgs": msgs, f"{name}.luts": luts,
              f"{name}.lut_idx": idx, f"{name}.ks_pbs_phase": o.phase(out, "big"),
              f"{name}.ks_pbs_msgs": o.decode(o.phase(out, "big")) % o.pmod})
    assert (z[f"{name}.ks_pbs_msgs"] == np.where(idx == 0, msgs, (3 * msgs + 1) % o.pmod)).all()


def gadget_set(z):
    name = "GOLD_PK"
    o = Oracle(name, seed=0x601D0004)
    o.keygen(fourier=False)
    pksk = o.gen_pksk()
    rng = np.random.default_rng(0x6020)
    lwes = rng.integers(0, 2 ** 64, (5, o.big + 1), dtype=np.uint64)
    lwes[0, :] = 0
    lwes[1, :] = np.uint64(0xFFFFFFFFFFFFFFFF)
    glwes = o.packing_keyswitch(pksk, lwes)
    polys = np.zeros((2, o.N), np.uint64)
    polys[0, [1, 17, 40]] = [3, 2 ** 64 - 1, 5]
    polys[1] = rng.integers(0, 2 ** 64, o.N, dtype=np.uint64)
    ext = o.glwe_mul_plain_extract(glwes, polys)
    packed = np.stack([o.pack_into_accumulator(pksk, lwes[2:4], 3), o.pack_into_accumulator(pksk, lwes[:5], 5)])
    z.update({f"{name}.glwe_key": o.glwe_key.astype(np.uint8), f"{name}.pksk": pksk, f"{name}.lwes": lwes, f"{name}.pks_out": glwes,
              f"{name}.polys": polys, f"{name}.mul_extract_out": ext, f"{name}.packed_p3_count2": packed[0],
              f"{name}.packed_p5_count5": packed[1]})


def main():
    z = {}
    multibit_set(z)
    gadget_set(z)
    np.savez_compressed(OUT, **z)
    print(OUT, os.path.getsize(OUT), "bytes;", len(z), "arrays")


if __name__ == "__main__":
    main()
