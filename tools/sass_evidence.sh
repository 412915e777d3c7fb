#!/bin/bash
# Counts, per in-tree object built for sm_100a, the SASS mnemonics that prove the Blackwell-only paths (tcgen05 = UTC*, LDTM/STTM; TMA =
# UTMALDG / UBLKCP; mbarrier = SYNCS.*) and the FP64 / shared-memory / shuffle mix of the blind-rotation kernels.
#   tools/sass_evidence.sh > profiles/rNN_sass_evidence.txt
cd "$(dirname "$0")/../tfhe-rs-odd_b200/csrc" || exit 1
echo "# SASS evidence of the final build (cuobjdump -sass of the in-tree objects, sm_100a; counts per object, all kernels of the object together)"
for o in keyswitch_tc blind_rotate_tmem blind_rotate_big blind_rotate_lat blind_rotate_x16 blind_rotate_pair blind_rotate_mb blind_rotate external_product; do
    echo "## $o.o"
    cuobjdump -sass $o.o | grep -E "^\s+/\*[0-9a-f]{4}\*/" | awk '{print $2}' | sed 's/;$//' |
        grep -E "^(UTC|UTMA|UBLKCP|LDTM|STTM|SYNCS|DFMA|DADD|DMUL|DMMA|LDS|STS|LDG|SHFL|BAR|F2I|I2F)" | sed -E 's/^(DFMA|DADD|DMUL|LDS|STS|LDG|SHFL|BAR|F2I|I2F)\..*/\1/' | sort | uniq -c | sort -rn
done
