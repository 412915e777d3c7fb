"""Device-resident waves of shortint / boolean operations.

The reference runs a circuit as rayon-parallel layers of "leveled prologue + apply_lookup_table"
(integer/server_key/radix_parallel/mul.rs:329-424, apps/trivium/src/trivium/trivium_shortint.rs:81-168), moving every
ciphertext through host memory.  Here the ciphertexts of a circuit live in a `CiphertextPool` in HBM and an `OpGraph`
turns a list of operations

    pool[dst] <- LUT( sum_k scale_k * pool[src_k] + plaintext )           (LUT optional)

into dependency levels (ASAP list scheduling with read-after-write, write-after-read and write-after-write hazards on
pool slots) and executes one `b200tfhe_wave_run` per level: gather + batched KS+PBS + scatter, nothing leaves the device.
The operation shape is the reference's: unchecked_add / unchecked_scalar_mul (lwe_linear_algebra.rs:191-215, 556-572),
the bivariate packing lhs*factor + rhs (shortint/server_key/bivariate_pbs.rs:167-182), message / carry extraction
(shortint/server_key/mod.rs:539-545, 619-623).
"""
import ctypes as C
from collections import defaultdict

import numpy as np

from . import _native
from .core_crypto import B200TfheError, _dtype, _ptr


class WaveOp(C.Structure):
    """Layout of b200tfhe_wave_op (include/b200tfhe.h)."""
    _fields_ = [("dst", C.c_uint32), ("src", C.c_uint32 * 4), ("src4", C.c_uint32), ("scale", C.c_int64 * 4),
                ("plaintext", C.c_uint64), ("lut", C.c_int32), ("scale4", C.c_int32)]


class CiphertextPool:
    """`capacity` LWE ciphertexts resident in HBM.  big=True: kN+1 scalars per slot, bootstrap = keyswitch -> PBS (shortint
    *_KS_PBS parameter sets); big=False: n+1 scalars, PBS -> keyswitch (boolean DEFAULT_PARAMETERS)."""

    def __init__(self, ctx, params, capacity, big=True):
        self.ctx, self.params, self.capacity, self.big = ctx, params, int(capacity), bool(big)
        self.bits = params.torus_bits
        self.len = (params.glwe_dimension * params.polynomial_size if big else params.lwe_dimension) + 1
        h = C.c_void_p()
        ctx.check(ctx._lib.b200tfhe_pool_create(ctx._h, C.byref(params), 1 if big else 0, self.capacity, C.byref(h)))
        self._h = h
        self.n_luts = 0

    def write(self, first, cts):
        a = np.ascontiguousarray(cts, dtype=_dtype(self.bits)).reshape(-1, self.len)
        self.ctx.check(self.ctx._lib.b200tfhe_pool_write(self.ctx._h, self._h, int(first), a.shape[0], _ptr(a)))

    def read(self, first, count):
        out = np.empty((int(count), self.len), _dtype(self.bits))
        self.ctx.check(self.ctx._lib.b200tfhe_pool_read(self.ctx._h, self._h, int(first), int(count), _ptr(out)))
        return out

    def set_luts(self, luts):
        k1n = (self.params.glwe_dimension + 1) * self.params.polynomial_size
        a = np.ascontiguousarray(luts, dtype=_dtype(self.bits)).reshape(-1, k1n)
        self.ctx.check(self.ctx._lib.b200tfhe_pool_set_luts(self.ctx._h, self._h, _ptr(a), a.shape[0]))
        self.n_luts = a.shape[0]

    def close(self):
        if self._h:
            self.ctx._lib.b200tfhe_pool_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OpGraph:
    """Records operations on pool slots, assigns each to the earliest wave its hazards allow, runs the waves in order.

    An operation recorded at wave L reads its sources as they are BEFORE wave L and writes its destination at the END of wave
    L, so:  RAW  L > wave of the last write of every source;  WAR  L >= wave of every earlier reader of dst;
    WAW  L > wave of the last write of dst."""

    MAX_TERMS = 5          # the fifth term's multiplier is 32 bits wide (b200tfhe_wave_op.scale4)

    def __init__(self, pool):
        self.pool = pool
        self.ops = []                            # (wave, dst, terms, plaintext, lut)
        self._written = defaultdict(int)         # slot -> wave of its last write (0: before the first wave)
        self._read = defaultdict(int)            # slot -> latest wave that reads it

    def op(self, dst, terms, plaintext=0, lut=None):
        """pool[dst] <- LUT_lut(sum scale * pool[src] + plaintext); terms: iterable of (src_slot, scale).  Returns the wave."""
        merged = defaultdict(int)
        for s, c in terms:
            merged[int(s)] += int(c)
        terms = [(s, c) for s, c in merged.items() if c != 0]
        if len(terms) > self.MAX_TERMS:
            raise B200TfheError(f"an operation takes at most {self.MAX_TERMS} source terms, got {len(terms)}")
        if len(terms) == self.MAX_TERMS:         # the narrow multiplier goes last
            terms.sort(key=lambda sc: -(2 ** 31) <= sc[1] < 2 ** 31)
            if not -(2 ** 31) <= terms[-1][1] < 2 ** 31:
                raise B200TfheError("the fifth term of an operation takes a 32-bit multiplier")
        cap = self.pool.capacity
        if not (0 <= dst < cap) or any(not (0 <= s < cap) for s, _ in terms):
            raise B200TfheError("pool slot out of range")
        if lut is not None and not (0 <= lut < max(1, self.pool.n_luts)):
            raise B200TfheError("accumulator index out of range")
        wave = 1 + max([self._written[s] for s, _ in terms] + [0])
        wave = max(wave, self._read[dst], self._written[dst] + 1)
        for s, _ in terms:
            self._read[s] = max(self._read[s], wave)
        self._written[dst] = wave
        self.ops.append((wave, int(dst), terms, int(plaintext) & (2 ** 64 - 1), lut))
        return wave

    # the reference's operator names, on slots
    def unchecked_add(self, dst, a, b):
        return self.op(dst, [(a, 1), (b, 1)])

    def unchecked_scalar_mul(self, dst, a, scalar):
        return self.op(dst, [(a, scalar)])

    def apply_lookup_table(self, dst, a, lut):
        return self.op(dst, [(a, 1)], lut=lut)

    def unchecked_apply_lookup_table_bivariate(self, dst, lhs, rhs, factor, lut):
        """bivariate_pbs.rs:167-182: lhs * factor + rhs, then the bivariate accumulator."""
        return self.op(dst, [(lhs, factor), (rhs, 1)], lut=lut)

    def waves(self):
        by = defaultdict(list)
        for o in self.ops:
            by[o[0]].append(o)
        return [by[w] for w in sorted(by)]

    def compile(self):
        """Freezes the recorded operations into a `WavePlan` (ctypes operation arrays, one block per wave) and clears the graph.
        The plan can be replayed any number of times: the scheduling work is done once, off the critical path."""
        waves = []
        for wave in self.waves():
            wave = sorted(wave, key=lambda o: o[4] is None)        # bootstrapped operations first
            waves.append((wave, sum(1 for o in wave if o[4] is not None)))
        total = sum(len(w) for w, _ in waves)
        arr = (WaveOp * max(1, total))()
        i = 0
        for wave, _ in waves:
            for (_, dst, terms, pt, lut) in wave:
                arr[i].dst = dst
                for k, (s, c) in enumerate(terms):
                    if k < 4:
                        arr[i].src[k] = s
                        arr[i].scale[k] = c
                    else:
                        arr[i].src4 = s
                        arr[i].scale4 = c
                arr[i].plaintext = pt
                arr[i].lut = -1 if lut is None else lut
                i += 1
        plan = WavePlan(self.pool, arr, [len(w) for w, _ in waves], [n for _, n in waves])
        self.ops = []
        self._written.clear()
        self._read.clear()
        return plan

    def run(self, ksk, bsk):
        """Executes and clears the recorded operations.  Returns [(n_ops, n_pbs)] per wave."""
        return self.compile().run(ksk, bsk)


class WavePlan:
    """A compiled circuit: every wave's operations in one array, replayed by ONE C call (b200tfhe_wave_sequence_run)."""

    def __init__(self, pool, ops, n_ops, n_pbs):
        self.pool, self.ops = pool, ops
        self.n_ops = np.ascontiguousarray(n_ops, dtype=np.uint32)
        self.n_pbs = np.ascontiguousarray(n_pbs, dtype=np.uint32)

    @property
    def stats(self):
        return [(int(a), int(b)) for a, b in zip(self.n_ops, self.n_pbs)]

    def run(self, ksk, bsk):
        """Asynchronous on the context's stream.  Returns [(n_ops, n_pbs)] per wave."""
        ctx = self.pool.ctx
        ctx.check(ctx._lib.b200tfhe_wave_sequence_run(ctx._h, ksk._h if ksk else None, bsk._h if bsk else None, self.pool._h, self.ops,
                                                      _ptr(self.n_ops), _ptr(self.n_pbs), len(self.n_ops)))
        return self.stats


_native  # noqa: B018  (imported for its side effect of declaring the prototypes)
