"""Radix integers on device-resident waves: the block structure of integer::RadixCiphertext (blocks of `message_modulus`,
least significant first) evaluated through `waves.OpGraph`.

  * unchecked block-wise addition + the reference's sequential carry propagation (full_propagate:
    integer/server_key/radix/mod.rs, message_extract / carry_extract per block);
  * schoolbook multiplication in the structure of unchecked_mul_parallelized (integer/server_key/radix_parallel/mul.rs:329-424):
    one wave of bivariate bootstraps producing the lsb / msb block products (256 of them for 16 blocks), then column sums
    reduced in waves of message_extract / carry_extract over chunks of five terms until every column holds one clean block.
Slots are plain pool indices; the caller owns the layout.
"""
from collections import defaultdict

LUT_MSG, LUT_CARRY, LUT_MUL_LSB, LUT_MUL_MSB = 0, 1, 2, 3


def radix_luts(server_key):
    """Accumulators [message_extract, carry_extract, bivariate product lsb, bivariate product msb] of a shortint ServerKey."""
    m = server_key.message_modulus
    return [server_key.generate_lookup_table(lambda z: z % m).acc,
            server_key.generate_lookup_table(lambda z: z // m).acc,
            server_key.generate_lookup_table(lambda z: ((z // m) * (z % m)) % m).acc,
            server_key.generate_lookup_table(lambda z: ((z // m) * (z % m)) // m).acc]


class SlotAllocator:
    def __init__(self, first, capacity):
        self.next, self.end = first, capacity

    def take(self, n=1):
        s = self.next
        self.next += n
        if self.next > self.end:
            raise MemoryError("ciphertext pool exhausted")
        return s


def add_propagate(g, x, y, out, carry, nb):
    """out[i] <- block i of x + y (mod m^nb); x, y, out, carry: first slots of nb consecutive blocks (carry: scratch)."""
    for i in range(nb):
        terms = [(x + i, 1), (y + i, 1)] + ([(carry + i - 1, 1)] if i else [])
        g.op(out + i, terms, lut=LUT_MSG)
        if i + 1 < nb:
            g.op(carry + i, terms, lut=LUT_CARRY)


def mul(g, alloc, x, y, out, nb, message_modulus=4):
    """out[i] <- block i of x * y (mod m^nb) for clean inputs (every block < m).  Scratch slots come from `alloc`."""
    m = message_modulus
    cols = defaultdict(list)
    for i in range(nb):
        for j in range(nb - i):
            lo = alloc.take()
            g.unchecked_apply_lookup_table_bivariate(lo, x + i, y + j, m, LUT_MUL_LSB)
            cols[i + j].append(lo)
            if i + j + 1 < nb:
                hi = alloc.take()
                g.unchecked_apply_lookup_table_bivariate(hi, x + i, y + j, m, LUT_MUL_MSB)
                cols[i + j + 1].append(hi)
    # Every term is < m and a block holds m*m - 1, so (m*m - 1) // (m - 1) terms add without overflow: 5 at m = 4, the chunk size of
    # the reference's column sums (radix_parallel/sum.rs).  A column with more terms than one chunk reduces its FULL chunks only and
    # lets the remainder wait for the carries that arrive next round: 547 bootstraps per 16-block product instead of the 614 of
    # eager 4-term groups, in the same 16 waves.
    chunk = min((m * m - 1) // (m - 1), g.MAX_TERMS)
    while any(len(v) > 1 for v in cols.values()):
        nxt = defaultdict(list)
        for c in range(nb):
            terms = cols.get(c, [])
            if len(terms) <= 1:
                nxt[c].extend(terms)
                continue
            full = len(terms) > chunk
            for k in range(0, len(terms), chunk):
                grp = terms[k:k + chunk]
                if len(grp) == 1 or (full and len(grp) < chunk):
                    nxt[c].extend(grp)
                    continue
                msg = alloc.take()
                g.op(msg, [(s, 1) for s in grp], lut=LUT_MSG)
                nxt[c].append(msg)
                if c + 1 < nb:
                    car = alloc.take()
                    g.op(car, [(s, 1) for s in grp], lut=LUT_CARRY)
                    nxt[c + 1].append(car)
        cols = nxt
    for c in range(nb):
        if cols.get(c):
            g.op(out + c, [(cols[c][0], 1)])
        else:
            g.op(out + c, [])                      # no term: the block is zero
