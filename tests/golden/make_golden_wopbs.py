#!/usr/bin/env python
"""Generates tests/golden/wopbs_golden.npz: frozen inputs / outputs of the pinned oracle for the integer stage of the WoP-PBS row
(SURVEY 8f rank 4): the private functional packing keyswitch with the circuit-bootstrap key list, on a tiny polynomial size
(N = 64, pfks level 2 / base 2^15 as in WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS) so the stored key list stays small.  Checked against
the oracle in the `-m "not gpu"` suite and bit for bit against the CUDA path in the `-m gpu` suite.
        python tests/golden/make_golden_wopbs.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "wopbs_golden.npz")
PF_BASE_LOG, PF_LEVEL = 15, 2


def main():
    name = "GOLD_PK"
    o = Oracle(name, seed=0x601D0005)
    o.keygen(fourier=False)
    pf = o.gen_cbs_pfpksk_list(PF_BASE_LOG, PF_LEVEL)
    rng = np.random.default_rng(0x6021)
    lwes = rng.integers(0, 2 ** 64, (6, o.big + 1), dtype=np.uint64)
    lwes[0, :] = 0
    lwes[1, :] = np.uint64(0xFFFFFFFFFFFFFFFF)
    lwes[2, :] = np.uint64(1 << 33)            # half of the last kept digit of a 30-bit decomposition: rounding boundary
    lwes[3, :] = np.uint64((1 << 63) - 1)
    key_len = (o.big + 1) * PF_LEVEL * o.glwe_len
    outs = np.stack([o.pfpks(pf[r * key_len:(r + 1) * key_len], PF_BASE_LOG, PF_LEVEL, lwes) for r in range(o.k + 1)])
    # semantic pin: an encryption of m goes to -m * S_r (r < k) and to +m (r = k)
    m = np.uint64(5 << 60)
    ct = o.encrypt(np.array([m], dtype=np.uint64), "big", seed=0x601D0500)
    dec = np.stack([o.glwe_phase(o.pfpks(pf[r * key_len:(r + 1) * key_len], PF_BASE_LOG, PF_LEVEL, ct)[0]) for r in range(o.k + 1)])
    z = {f"{name}.glwe_key": o.glwe_key.astype(np.uint8), f"{name}.pfpksk_list": pf, f"{name}.lwes": lwes, f"{name}.pfpks_out": outs,
         f"{name}.enc5": ct, f"{name}.enc5_phases": dec, "pf_base_log": np.int64(PF_BASE_LOG), "pf_level": np.int64(PF_LEVEL)}
    np.savez_compressed(OUT, **z)
    print(OUT, os.path.getsize(OUT), "bytes;", len(z), "arrays")


if __name__ == "__main__":
    main()
