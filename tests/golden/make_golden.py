#!/usr/bin/env python
"""Generates tests/golden/ks_pbs_golden.npz: keys, inputs and outputs of the keyswitch + PBS path on three tiny parameter
sets (the polynomial sizes / decompositions of 2_2, of the reference's FFT_U64 test set and of the boolean set, with n = 4).

The Rust reference cannot run in this image (no rustc) and holds no ciphertext-level fixtures, so these vectors
come from the CPU restatement in oracle/ after it was pinned against the reference's own known-answer tests
(tests/test_oracle_kat.py).  They freeze that state: the `-m "not gpu"` suite checks the oracle still reproduces them, the
`-m gpu` suite checks the CUDA path against them (integer stages bit-exact, decrypted messages equal, phase within the
analytic bound).        python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, lib  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ks_pbs_golden.npz")
B = 6


def shortint_set(name, z):
    o = Oracle(name, seed=0x601D0001).keygen()
    dt = np.uint64
    rng = np.random.default_rng(0x601D)
    msgs = rng.integers(0, o.pmod, B)
    cts = o.encrypt(o.encode(msgs), "big", seed=0x601D0100)
    cts[0, :-1] = 0                       # degenerate rows: zero mask, all-ones mask
    cts[1, :-1] = np.uint64(0xFFFFFFFFFFFFFFFF)
    ks = o.keyswitch(cts)
    ms = np.array([[lib().orc_modswitch(int(v), int(np.log2(o.N)), o.bits) for v in row] for row in ks], np.uint32)
    luts = np.stack([o.lut(lambda x: x), o.lut(lambda x: (3 * x + 1) % o.pmod)])
    idx = (np.arange(B) % 2).astype(np.uint32)
    out = o.ks_pbs(cts, luts, idx)
    z.update({f"{name}.small_key": o.small_key.astype(np.uint8), f"{name}.glwe_key": o.glwe_key.astype(np.uint8),
              f"{name}.ksk": o.ksk.astype(dt), f"{name}.bsk": o.bsk.astype(dt), f"{name}.cts": cts, f"{name}.msgs": msgs,
              f"{name}.ks_out": ks, f"{name}.modswitch": ms, f"{name}.luts": luts, f"{name}.lut_idx": idx,
              f"{name}.ks_pbs_out": out, f"{name}.ks_pbs_phase": o.phase(out, "big"),
              f"{name}.ks_pbs_msgs": o.decode(o.phase(out, "big")) % o.pmod})
    exp = np.where(idx == 0, msgs, (3 * msgs + 1) % o.pmod)
    assert (z[f"{name}.ks_pbs_msgs"][2:] == exp[2:]).all(), name


def boolean_set(name, z):
    o = Oracle(name, seed=0x601D0002).keygen()
    rng = np.random.default_rng(0x601E)
    a_bits, b_bits = rng.integers(0, 2, B), rng.integers(0, 2, B)
    enc = lambda bits, seed: o.encrypt(np.where(bits == 1, 1 << 29, (1 << 32) - (1 << 29)).astype(np.uint64), "small", seed=seed)
    a, b = enc(a_bits, 0x601D0200), enc(b_bits, 0x601D0300)
    GATE_NAND = 1
    pro = o.gate_prologue(GATE_NAND, a, b)
    lut = o.boolean_lut()
    out = o.pbs_ks(pro, lut)
    z.update({f"{name}.small_key": o.small_key.astype(np.uint8), f"{name}.glwe_key": o.glwe_key.astype(np.uint8),
              f"{name}.ksk": o.ksk.astype(np.uint32), f"{name}.bsk": o.bsk.astype(np.uint32), f"{name}.a": a.astype(np.uint32),
              f"{name}.b": b.astype(np.uint32), f"{name}.a_bits": a_bits, f"{name}.b_bits": b_bits,
              f"{name}.nand_prologue": pro.astype(np.uint32), f"{name}.nand_out": out.astype(np.uint32),
              f"{name}.nand_bits": o.decode_bool(o.phase(out, "small"))})
    assert (z[f"{name}.nand_bits"] == (1 - (a_bits & b_bits)).astype(bool)).all()


def main():
    z = {}
    shortint_set("GOLD_2_2", z)
    shortint_set("GOLD_U64", z)
    boolean_set("GOLD_U32", z)
    np.savez_compressed(OUT, **z)
    print(OUT, os.path.getsize(OUT), "bytes;", len(z), "arrays")


if __name__ == "__main__":
    main()
