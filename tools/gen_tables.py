#!/usr/bin/env python
"""Generates (and checks) the correctly rounded constant table of tfhe-rs-odd_b200/csrc/fft_reg.cuh:
COS128_TAB[q] = cos(2 pi q / 128), q = 0..32, evaluated with mpmath at 200 bits and rounded to the nearest double.

    python tools/gen_tables.py            # print the table as C++
    python tools/gen_tables.py --check    # compare with the table in fft_reg.cuh, exit 1 on any difference
"""
import os
import re
import sys

import mpmath

mpmath.mp.prec = 200


def table():
    # q = 32 is a quarter turn: exactly 0 (the 200-bit pi leaves 6e-62 there)
    return [0.0 if q == 32 else float(mpmath.cos(2 * mpmath.pi * q / 128)) for q in range(33)]


def main():
    tab = table()
    if "--check" in sys.argv:
        src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tfhe-rs-odd_b200", "csrc", "fft_reg.cuh")).read()
        body = re.search(r"COS128_TAB\[33\]\s*=\s*\{([^}]*)\}", src).group(1)
        have = [float(x) for x in body.replace("\n", " ").split(",") if x.strip()]
        bad = [(q, a, b) for q, (a, b) in enumerate(zip(have, tab)) if a != b]
        if len(have) != 33 or bad:
            print("COS128_TAB differs from the correctly rounded values:", bad or len(have))
            sys.exit(1)
        print("COS128_TAB: 33 entries, all correctly rounded")
        return
    print("constexpr double COS128_TAB[33] = {")
    for i in range(0, 33, 6):
        print("    " + ", ".join(repr(v) for v in tab[i:i + 6]) + ("," if i + 6 < 33 else "};"))


if __name__ == "__main__":
    main()
