"""Host-side mirror of shortint::ServerKey for the lookup-table path, batch-first.

Reference (relative to /root/reference/tfhe/src/shortint):
  ServerKey::{apply_lookup_table, apply_lookup_table_assign}            server_key/mod.rs:457-476
  keyswitch_programmable_bootstrap_assign / programmable_bootstrap_keyswitch_assign   :783-933
  trivial_pbs_assign                                                     :763-781
  generate_lookup_table -> fill_accumulator                              :383-399, engine/mod.rs:72-128
  Ciphertext{ct, degree, noise_level, message_modulus, carry_modulus, pbs_order}      ciphertext/mod.rs:263-270
The reference handles one ciphertext per call and parallelises with rayon over a Vec; here a `CiphertextList`
is the unit: trivial ciphertexts are resolved on the host exactly like trivial_pbs_assign, the encrypted ones are
gathered into one structure-of-arrays batch for the GPU, and degree / noise_level are updated as the reference does.
"""
import numpy as np

from . import core_crypto as cc
from .core_crypto import B200TfheError

NOISE_ZERO = 0
NOISE_NOMINAL = 1


class LookupTable:
    """LookupTableOwned{acc: GlweCiphertext, degree} (server_key/mod.rs:322-331)."""

    def __init__(self, acc, degree):
        self.acc, self.degree = acc, degree


class CiphertextList:
    """A list of shortint ciphertexts: data [B][kN+1] (or [B][n+1] for PBS->KS keys) + per-ciphertext metadata."""

    def __init__(self, data, degree, noise_level):
        self.data = np.ascontiguousarray(data, dtype=np.uint64)
        B = self.data.shape[0]
        self.degree = np.broadcast_to(np.asarray(degree, dtype=np.uint64), (B,)).copy()
        self.noise_level = np.broadcast_to(np.asarray(noise_level, dtype=np.uint64), (B,)).copy()

    def __len__(self):
        return self.data.shape[0]

    def is_trivial(self):
        """noise_level == ZERO and an all-zero mask (ciphertext/mod.rs:371-374)."""
        return (self.noise_level == NOISE_ZERO) & ~self.data[:, :-1].any(axis=1)

    def copy(self):
        return CiphertextList(self.data.copy(), self.degree.copy(), self.noise_level.copy())


def _check_fourier_key(fourier, N, polys=4):
    """Sanity check of a deserialised Fourier bootstrap key before it is uploaded verbatim (EXPERIMENTAL entry point: no
    reference-produced bytes exist in this repository to pin the element order and scaling of concrete-fft's serialised buffer).
    A Fourier polynomial produced by forward_as_torus (fft/mod.rs:197-218) in natural frequency order inverts to torus coefficients
    in [-1/2, 1/2); a different order or scaling (missing 2^-BITS, a factor M) does not -- raise instead of bootstrapping garbage."""
    M = N // 2
    f = fourier.reshape(-1)[: 2 * M * polys].reshape(-1, M, 2)
    j = np.arange(M)
    untwist = np.exp(-1j * np.pi * j / N)
    for q in range(f.shape[0]):
        c = np.fft.ifft(f[q, :, 0] + 1j * f[q, :, 1]) * untwist
        hi = max(np.abs(c.real).max(), np.abs(c.imag).max())
        if not np.isfinite(hi) or hi > 0.5 + 1e-6:
            raise B200TfheError(f"Fourier bootstrap key polynomial {q} does not invert to torus coefficients in [-1/2, 1/2) (max {hi:.3g}): "
                                "unexpected element order or scaling of the serialised key")


class ServerKey:
    """shortint::ServerKey{key_switching_key, bootstrapping_key (Classic or MultiBit: the device key handle carries the
    grouping factor), message_modulus, carry_modulus, pbs_order}."""

    def __init__(self, ksk, bsk, message_modulus, carry_modulus, pbs_order="ks_pbs"):
        if bsk.params.torus_bits != 64:
            raise B200TfheError("shortint ciphertexts live on the 2^64 torus")
        self.key_switching_key, self.bootstrapping_key = ksk, bsk
        self.message_modulus, self.carry_modulus, self.pbs_order = message_modulus, carry_modulus, pbs_order
        self.params = bsk.params

    @classmethod
    def from_bincode(cls, ctx, data):
        """Load `bincode::serialize(&shortint::ServerKey)` bytes (serialization.py; server_key/mod.rs:283-297): the keyswitch key
        goes to the device as is, the Fourier bootstrap key (Classic or MultiBit) is uploaded in its serialised order."""
        from . import serialization as ser
        d = ser.deserialize_shortint_server_key(data)
        k, b = d["key_switching_key"], d["bootstrapping_key"]
        glwe_dim, N = b["glwe_size"] - 1, b["polynomial_size"]
        if k["input_lwe_dimension"] != glwe_dim * N or k["output_lwe_size"] != b["input_lwe_dimension"] + 1:
            raise B200TfheError("server key: keyswitch key and bootstrap key dimensions differ")
        params = cc.make_params(b["input_lwe_dimension"], glwe_dim, N, b["decomp_base_log"], b["decomp_level_count"],
                                k["decomp_base_log"], k["decomp_level_count"], 64)
        ksk = cc.LweKeyswitchKey.upload(ctx, params, k["data"])
        _check_fourier_key(np.asarray(b["fourier"], dtype=np.float64), params.polynomial_size)
        bsk = cc.FourierLweBootstrapKey.from_fourier(ctx, params, b["fourier"], b["grouping_factor"])
        sk = cls(ksk, bsk, d["message_modulus"], d["carry_modulus"], d["pbs_order"])
        sk.max_degree, sk.max_noise_level = d["max_degree"], d["max_noise_level"]
        return sk

    # ---- LUT generation: engine/mod.rs:72-128 -----------------------------------------------------------
    def generate_lookup_table(self, f):
        p = self.params
        N, k = p.polynomial_size, p.glwe_dimension
        modulus_sup = self.message_modulus * self.carry_modulus
        box = N // modulus_sup
        delta = (1 << 63) // modulus_sup
        vals = np.array([int(f(i)) for i in range(modulus_sup)], dtype=np.uint64)
        body = np.repeat(vals * np.uint64(delta), box)          # wrapping u64 multiply
        half = box // 2
        body[:half] = np.uint64(0) - body[:half]
        body = np.roll(body, -half)                             # rotate_left(half_box_size)
        acc = np.zeros((k + 1) * N, np.uint64)
        acc[k * N:] = body
        return LookupTable(acc, int(vals.max()))

    # ---- apply_lookup_table(_assign): server_key/mod.rs:457-476 -------------------------------------------
    def apply_lookup_table(self, cts, luts, lut_idx=None):
        out = cts.copy()
        self.apply_lookup_table_assign(out, luts, lut_idx)
        return out

    def apply_lookup_table_assign(self, cts, luts, lut_idx=None):
        """luts: one LookupTable or a list; lut_idx[b] picks ciphertext b's table (None: table 0 for all)."""
        if isinstance(luts, LookupTable):
            luts = [luts]
        B = len(cts)
        idx = np.zeros(B, np.uint32) if lut_idx is None else np.ascontiguousarray(lut_idx, dtype=np.uint32)
        if idx.shape != (B,) or (B and idx.max() >= len(luts)):
            raise B200TfheError("lut_idx must hold one valid table index per ciphertext")
        accs = np.stack([l.acc for l in luts])
        degs = np.array([l.degree for l in luts], dtype=np.uint64)
        triv = cts.is_trivial()
        if triv.any():
            self._trivial_pbs_assign(cts, accs, idx, np.nonzero(triv)[0])
        enc = np.nonzero(~triv)[0]
        if enc.size:
            fn = cc.keyswitch_programmable_bootstrap if self.pbs_order == "ks_pbs" else cc.programmable_bootstrap_keyswitch
            res = fn(self.key_switching_key, self.bootstrapping_key, cts.data[enc], accs, lut_idx=idx[enc])
            cts.data[enc] = res
            cts.noise_level[enc] = NOISE_NOMINAL               # server_key/mod.rs:855-856
        cts.degree[:] = degs[idx]
        return cts

    def _trivial_pbs_assign(self, cts, accs, idx, rows):
        """server_key/mod.rs:763-781: plain table look-up on the body, padding bit negates."""
        p = self.params
        N, k = p.polynomial_size, p.glwe_dimension
        modulus_sup = self.message_modulus * self.carry_modulus
        delta = np.uint64((1 << 63) // modulus_sup)
        box = N // modulus_sup
        body = cts.data[rows, -1]
        v = body // delta
        neg = v >= modulus_sup
        pos = (v % np.uint64(modulus_sup)).astype(np.int64) * box
        res = accs[idx[rows], k * N + pos]
        cts.data[rows, -1] = np.where(neg, np.uint64(0) - res, res)

    # ---- bivariate helper used by integer mul / trivium callers: bivariate_pbs.rs:167-182 ---------------------
    def unchecked_apply_lookup_table_bivariate(self, left, right, lut, factor=None):
        """left*factor + right (wrapping, no PBS) then apply_lookup_table; factor defaults to message_modulus."""
        factor = np.uint64(self.message_modulus if factor is None else factor)
        # noise of scalar_mul + add as the reference tracks it (bivariate_pbs.rs:167-182 -> unchecked_scalar_mul_assign + unchecked_add_assign)
        tmp = CiphertextList(left.data * factor + right.data, left.degree * factor + right.degree,
                             left.noise_level * factor + right.noise_level)
        return self.apply_lookup_table_assign(tmp, lut)
