"""Host-side mirror of the boolean gate engine for the bootstrapped path.

Reference (relative to /root/reference/tfhe/src/boolean): engine/mod.rs:607-798 (BinaryGatesEngine: trivial /
encrypted dispatch + linear combinations), engine/bootstrapping.rs:257-401 (Bootstrapper::{bootstrap,
bootstrap_keyswitch, keyswitch_bootstrap, apply_bootstrapping_pattern}), mod.rs:74-80 (PLAINTEXT_TRUE/FALSE).
Ciphertexts are batches: `Ciphertext.Encrypted` = uint32 array [B][dim+1]; `Ciphertext.Trivial` = bool array [B].
"""
import ctypes as C
import numpy as np

from .core_crypto import B200TfheError, _ptr

PLAINTEXT_TRUE = 1 << 29
PLAINTEXT_FALSE = 7 << 29
GATES = {"and": 0, "nand": 1, "or": 2, "nor": 3, "xor": 4, "xnor": 5}


def gate_batch(ksk, bsk, gate, a, b, pbs_first=True):
    """One binary gate over a list of encrypted pairs: linear combination + PBS/KS in apply_bootstrapping_pattern order."""
    ctx, p = bsk.ctx, bsk.params
    gate = GATES[gate] if isinstance(gate, str) else int(gate)
    dim = p.lwe_dimension if pbs_first else p.glwe_dimension * p.polynomial_size
    a = np.ascontiguousarray(a, dtype=np.uint32).reshape(-1, dim + 1)
    b = np.ascontiguousarray(b, dtype=np.uint32).reshape(-1, dim + 1)
    if a.shape != b.shape:
        raise B200TfheError("gate operands must have the same batch size")
    out = np.empty_like(a)
    ctx.check(ctx._lib.b200tfhe_gate_batch_u32(ctx._h, ksk._h, bsk._h, gate, 1 if pbs_first else 0,
                                               _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
    return out


class ServerKey:
    """boolean::ServerKey{bootstrapping_key, key_switching_key, pbs_order} (boolean/engine/bootstrapping.rs:93-97)."""

    def __init__(self, ksk, bsk, pbs_first=True):
        self.ksk, self.bsk, self.pbs_first = ksk, bsk, pbs_first

    def _gate(self, name, a, b):
        ta, tb = np.asarray(a).dtype == np.bool_, np.asarray(b).dtype == np.bool_
        clear = {"and": np.logical_and, "nand": lambda x, y: ~(x & y), "or": np.logical_or,
                 "nor": lambda x, y: ~(x | y), "xor": np.logical_xor, "xnor": lambda x, y: ~(x ^ y)}[name]
        if ta and tb:                                   # Trivial/Trivial arm: clear logic (engine/mod.rs:614-616)
            return clear(np.asarray(a), np.asarray(b))
        if ta or tb:                                    # Encrypted/Trivial arms (engine/mod.rs:980-993 pattern)
            enc, triv = (b, np.asarray(a)) if ta else (a, np.asarray(b))
            enc = np.ascontiguousarray(enc, dtype=np.uint32)
            out = enc.copy()
            neg = self.not_(enc)
            dim = enc.shape[-1] - 1
            def trivial(v):
                t = np.zeros_like(enc)
                t[..., dim] = PLAINTEXT_TRUE if v else PLAINTEXT_FALSE
                return t
            for i, m in enumerate(triv.reshape(-1)):
                res = {"and": enc[i] if m else trivial(False)[i], "nand": neg[i] if m else trivial(True)[i],
                       "or": trivial(True)[i] if m else enc[i], "nor": trivial(False)[i] if m else neg[i],
                       "xor": neg[i] if m else enc[i], "xnor": enc[i] if m else neg[i]}[name]
                out[i] = res
            return out
        return gate_batch(self.ksk, self.bsk, name, a, b, self.pbs_first)

    def and_(self, a, b): return self._gate("and", a, b)
    def nand(self, a, b): return self._gate("nand", a, b)
    def or_(self, a, b): return self._gate("or", a, b)
    def nor(self, a, b): return self._gate("nor", a, b)
    def xor(self, a, b): return self._gate("xor", a, b)
    def xnor(self, a, b): return self._gate("xnor", a, b)

    @staticmethod
    def not_(a):
        """NOT is a negation, no bootstrap (engine/mod.rs:377-399)."""
        a = np.asarray(a)
        if a.dtype == np.bool_:
            return ~a
        return (np.uint32(0) - a.astype(np.uint32)).astype(np.uint32)
