"""Key / ciphertext wire formats of the path: what `bincode::serialize(&key)` (bincode 1.3.3, default
options, tfhe/Cargo.toml:36) writes for the reference's serde-derived types, so a serialised key set can be loaded without a
Rust re-export step.

bincode 1.x default encoding: little-endian fixed-width integers, usize as u64, u128 as 16 bytes, f64 as 8 bytes, bool as one
byte, Vec<T> / seq as a u64 length followed by the elements, struct fields in declaration order with no framing, newtype
structs transparent, enum variants as a u32 index.  Type layouts restated from (relative to /root/reference/tfhe/src/):
  CiphertextModulus            core_crypto/commons/ciphertext_modulus.rs:41-64   {modulus: u128 (0 = native), scalar_bits: usize}
  LweCiphertext                core_crypto/entities/lwe_ciphertext.rs:500-507    {data, ciphertext_modulus}
  LweCiphertextList            entities/lwe_ciphertext_list.rs:9-17              {data, lwe_size, ciphertext_modulus}
  GlweCiphertext               entities/glwe_ciphertext.rs:285-293               {data, polynomial_size, ciphertext_modulus}
  LweKeyswitchKey              entities/lwe_keyswitch_key.rs:76-86               {data, base_log, level, output_lwe_size, modulus}
  LwePackingKeyswitchKey       entities/lwe_packing_keyswitch_key.rs:13-24       {data, base_log, level, output_glwe_size, output_polynomial_size, modulus}
  GgswCiphertextList           entities/ggsw_ciphertext_list.rs:9-20             {data, glwe_size, polynomial_size, base_log, level, modulus}
  LweBootstrapKey              entities/lwe_bootstrap_key.rs:98-106              {ggsw_list}
  LweMultiBitBootstrapKey      entities/lwe_multi_bit_bootstrap_key.rs:10-20     {ggsw_list, grouping_factor}
  FourierPolynomialList        fft_impl/fft64/math/fft/mod.rs:588-717            seq of 2 + count: polynomial_size, count, then per polynomial
                                                                                 concrete-fft's serialize_fourier_buffer
  FourierLweBootstrapKey       fft_impl/fft64/crypto/bootstrap.rs:25-33          {fourier, input_lwe_dimension, glwe_size, base_log, level}
  FourierLweMultiBitBootstrapKey  entities/lwe_multi_bit_bootstrap_key.rs:316-325   {... , grouping_factor}
  ShortintBootstrappingKey     shortint/server_key/mod.rs:112-120               Classic(..) = 0 | MultiBit{fourier_bsk, deterministic_execution} = 1
  shortint ServerKey           shortint/server_key/mod.rs:283-297
  shortint Ciphertext          shortint/ciphertext/mod.rs:261-270
PINNING: the reference holds no serialised fixtures and its tests only round-trip through bincode itself, so the byte layouts
above are derived from the type definitions ("parity unpinned" for this row).  One element comes from an un-vendored
dependency: concrete-fft 0.3.0 `Plan::serialize_fourier_buffer` writes one Fourier polynomial as a seq of N/2 c64
(num_complex::Complex<f64> = two f64) in the plan-independent NATURAL frequency order; that order is what this library's
device layout is for N <= 2048 (pinned against numpy's FFT in tests/test_serialization.py) and a fixed permutation of it
for N = 8192.
"""
import struct

import numpy as np

from .core_crypto import B200TfheError

PBS_ORDER = {0: "ks_pbs", 1: "pbs_ks"}          # PBSOrder::{KeyswitchBootstrap = 0, BootstrapKeyswitch = 1}
PBS_ORDER_INV = {v: k for k, v in PBS_ORDER.items()}


class Reader:
    def __init__(self, data):
        self.b, self.o = memoryview(bytes(data)) if not isinstance(data, (bytes, bytearray, memoryview)) else memoryview(data), 0

    def _take(self, n):
        if self.o + n > len(self.b):
            raise B200TfheError("bincode: unexpected end of input")
        v = self.b[self.o:self.o + n]
        self.o += n
        return v

    def u8(self):
        return self._take(1)[0]

    def u32(self):
        return struct.unpack("<I", self._take(4))[0]

    def u64(self):
        return struct.unpack("<Q", self._take(8))[0]

    def u128(self):
        lo, hi = struct.unpack("<QQ", self._take(16))
        return lo | (hi << 64)

    def bool(self):
        v = self.u8()
        if v > 1:
            raise B200TfheError("bincode: invalid bool")
        return bool(v)

    def array(self, dtype, count):
        dt = np.dtype(dtype).newbyteorder("<")
        return np.frombuffer(self._take(count * dt.itemsize), dtype=dt, count=count).astype(dtype, copy=True)

    def vec(self, dtype):
        return self.array(dtype, self.u64())

    def done(self):
        if self.o != len(self.b):
            raise B200TfheError(f"bincode: {len(self.b) - self.o} trailing bytes")


class Writer:
    def __init__(self):
        self.parts = []

    def u8(self, v):
        self.parts.append(struct.pack("<B", v))

    def u32(self, v):
        self.parts.append(struct.pack("<I", v))

    def u64(self, v):
        self.parts.append(struct.pack("<Q", int(v)))

    def u128(self, v):
        self.parts.append(struct.pack("<QQ", int(v) & (2 ** 64 - 1), int(v) >> 64))

    def array(self, a, dtype):
        self.parts.append(np.ascontiguousarray(a, dtype=np.dtype(dtype).newbyteorder("<")).tobytes())

    def vec(self, a, dtype):
        a = np.asarray(a).reshape(-1)
        self.u64(a.size)
        self.array(a, dtype)

    def bytes(self):
        return b"".join(self.parts)


def _dtype(bits):
    return {32: np.uint32, 64: np.uint64}[bits]


# ---- CiphertextModulus ---------------------------------------------------------------------------------------------------
def _read_modulus(r, bits):
    modulus, scalar_bits = r.u128(), r.u64()
    if scalar_bits != bits:
        raise B200TfheError(f"Expected an unsigned integer with {bits} bits, found {scalar_bits} bits")   # ciphertext_modulus.rs:74-81
    if modulus != 0:
        raise B200TfheError("only the native ciphertext modulus is supported on this path")
    return modulus


def _write_modulus(w, bits):
    w.u128(0)
    w.u64(bits)


# ---- ciphertexts -----------------------------------------------------------------------------------------------------------
def deserialize_lwe_ciphertext(data, bits=64):
    r = Reader(data)
    ct = r.vec(_dtype(bits))
    _read_modulus(r, bits)
    r.done()
    return ct


def serialize_lwe_ciphertext(ct, bits=64):
    w = Writer()
    w.vec(ct, _dtype(bits))
    _write_modulus(w, bits)
    return w.bytes()


def deserialize_lwe_ciphertext_list(data, bits=64):
    r = Reader(data)
    flat, lwe_size = r.vec(_dtype(bits)), r.u64()
    _read_modulus(r, bits)
    r.done()
    if lwe_size == 0 or flat.size % lwe_size:
        raise B200TfheError("LweCiphertextList: data length is not a multiple of lwe_size")
    return flat.reshape(-1, lwe_size)


def serialize_lwe_ciphertext_list(cts, bits=64):
    cts = np.asarray(cts)
    w = Writer()
    w.vec(cts, _dtype(bits))
    w.u64(cts.shape[-1])
    _write_modulus(w, bits)
    return w.bytes()


def deserialize_glwe_ciphertext(data, bits=64):
    r = Reader(data)
    flat, poly = r.vec(_dtype(bits)), r.u64()
    _read_modulus(r, bits)
    r.done()
    return flat, poly


def serialize_glwe_ciphertext(glwe, polynomial_size, bits=64):
    w = Writer()
    w.vec(glwe, _dtype(bits))
    w.u64(polynomial_size)
    _write_modulus(w, bits)
    return w.bytes()


# ---- keys, standard domain ---------------------------------------------------------------------------------------------
def _read_ksk(r, bits):
    data, base_log, level, out_size = r.vec(_dtype(bits)), r.u64(), r.u64(), r.u64()
    _read_modulus(r, bits)
    if level == 0 or out_size == 0 or data.size % (level * out_size):
        raise B200TfheError("LweKeyswitchKey: inconsistent sizes")
    return {"data": data, "decomp_base_log": base_log, "decomp_level_count": level, "output_lwe_size": out_size,
            "input_lwe_dimension": data.size // (level * out_size)}


def deserialize_lwe_keyswitch_key(data, bits=64):
    r = Reader(data)
    k = _read_ksk(r, bits)
    r.done()
    return k


def _write_ksk(w, data, base_log, level, output_lwe_size, bits):
    w.vec(data, _dtype(bits))
    w.u64(base_log); w.u64(level); w.u64(output_lwe_size)
    _write_modulus(w, bits)


def serialize_lwe_keyswitch_key(data, base_log, level, output_lwe_size, bits=64):
    w = Writer()
    _write_ksk(w, data, base_log, level, output_lwe_size, bits)
    return w.bytes()


def deserialize_lwe_packing_keyswitch_key(data, bits=64):
    r = Reader(data)
    d, base_log, level, glwe_size, poly = r.vec(_dtype(bits)), r.u64(), r.u64(), r.u64(), r.u64()
    _read_modulus(r, bits)
    r.done()
    return {"data": d, "decomp_base_log": base_log, "decomp_level_count": level, "output_glwe_size": glwe_size,
            "output_polynomial_size": poly, "input_lwe_dimension": d.size // (level * glwe_size * poly)}


def serialize_lwe_packing_keyswitch_key(data, base_log, level, glwe_size, polynomial_size, bits=64):
    w = Writer()
    w.vec(data, _dtype(bits))
    for v in (base_log, level, glwe_size, polynomial_size):
        w.u64(v)
    _write_modulus(w, bits)
    return w.bytes()


def _read_ggsw_list(r, bits):
    data, glwe_size, poly, base_log, level = r.vec(_dtype(bits)), r.u64(), r.u64(), r.u64(), r.u64()
    _read_modulus(r, bits)
    per = level * glwe_size * glwe_size * poly
    if per == 0 or data.size % per:
        raise B200TfheError("GgswCiphertextList: inconsistent sizes")
    return {"data": data, "glwe_size": glwe_size, "polynomial_size": poly, "decomp_base_log": base_log,
            "decomp_level_count": level, "ggsw_count": data.size // per}


def deserialize_lwe_bootstrap_key(data, bits=64):
    r = Reader(data)
    k = _read_ggsw_list(r, bits)
    r.done()
    k["input_lwe_dimension"] = k["ggsw_count"]
    return k


def deserialize_lwe_multi_bit_bootstrap_key(data, bits=64):
    r = Reader(data)
    k = _read_ggsw_list(r, bits)
    k["grouping_factor"] = G = r.u64()
    r.done()
    if G == 0 or k["ggsw_count"] % (1 << G):
        raise B200TfheError("LweMultiBitBootstrapKey: GGSW count is not a multiple of 2^grouping_factor")
    k["input_lwe_dimension"] = k["ggsw_count"] // (1 << G) * G
    return k


def serialize_lwe_bootstrap_key(data, glwe_size, polynomial_size, base_log, level, bits=64, grouping_factor=None):
    w = Writer()
    w.vec(data, _dtype(bits))
    for v in (glwe_size, polynomial_size, base_log, level):
        w.u64(v)
    _write_modulus(w, bits)
    if grouping_factor is not None:
        w.u64(grouping_factor)
    return w.bytes()


# ---- keys, Fourier domain --------------------------------------------------------------------------------------------------
def _read_fourier_list(r):
    total = r.u64()
    poly, count = r.u64(), r.u64()
    if total != 2 + count:
        raise B200TfheError("FourierPolynomialList: sequence of two fields and Fourier polynomials")   # fft/mod.rs:659
    M = poly // 2
    out = np.empty((count, M, 2), np.float64)
    for i in range(count):
        if r.u64() != M:
            raise B200TfheError("FourierPolynomialList: Fourier polynomial of the wrong size")
        out[i] = r.array(np.float64, 2 * M).reshape(M, 2)
    return out, poly


def _write_fourier_list(w, fourier, polynomial_size):
    f = np.asarray(fourier, dtype=np.float64).reshape(-1, polynomial_size // 2, 2)
    w.u64(2 + f.shape[0]); w.u64(polynomial_size); w.u64(f.shape[0])
    for poly in f:
        w.u64(poly.shape[0])
        w.array(poly, np.float64)


def _read_fourier_bsk(r, multi_bit):
    fourier, poly = _read_fourier_list(r)
    k = {"fourier": fourier, "polynomial_size": poly, "input_lwe_dimension": r.u64(), "glwe_size": r.u64(),
         "decomp_base_log": r.u64(), "decomp_level_count": r.u64(), "grouping_factor": r.u64() if multi_bit else 1}
    G = k["grouping_factor"]
    ggsw = k["input_lwe_dimension"] if not multi_bit else k["input_lwe_dimension"] // G * (1 << G)
    if fourier.shape[0] != ggsw * k["decomp_level_count"] * k["glwe_size"] ** 2:
        raise B200TfheError("Fourier bootstrap key: polynomial count does not match the dimensions")
    return k


def deserialize_fourier_lwe_bootstrap_key(data, multi_bit=False):
    r = Reader(data)
    k = _read_fourier_bsk(r, multi_bit)
    r.done()
    return k


def _write_fourier_bsk(w, fourier, polynomial_size, input_lwe_dimension, glwe_size, base_log, level, grouping_factor=None):
    _write_fourier_list(w, fourier, polynomial_size)
    for v in (input_lwe_dimension, glwe_size, base_log, level):
        w.u64(v)
    if grouping_factor is not None:
        w.u64(grouping_factor)


def serialize_fourier_lwe_bootstrap_key(fourier, polynomial_size, input_lwe_dimension, glwe_size, base_log, level, grouping_factor=None):
    w = Writer()
    _write_fourier_bsk(w, fourier, polynomial_size, input_lwe_dimension, glwe_size, base_log, level, grouping_factor)
    return w.bytes()


# ---- shortint ---------------------------------------------------------------------------------------------------------------
def deserialize_shortint_server_key(data):
    """-> dict with key_switching_key, bootstrapping_key (Fourier, natural order), moduli, max_degree, max_noise_level, pbs_order."""
    r = Reader(data)
    ksk = _read_ksk(r, 64)
    variant = r.u32()
    if variant == 0:
        bsk = _read_fourier_bsk(r, False)
        bsk["deterministic_execution"] = None
    elif variant == 1:
        bsk = _read_fourier_bsk(r, True)
        bsk["deterministic_execution"] = r.bool()
    else:
        raise B200TfheError(f"ShortintBootstrappingKey: unknown variant {variant}")
    out = {"key_switching_key": ksk, "bootstrapping_key": bsk, "message_modulus": r.u64(), "carry_modulus": r.u64(),
           "max_degree": r.u64(), "max_noise_level": r.u64()}
    _read_modulus(r, 64)
    order = r.u32()
    if order not in PBS_ORDER:
        raise B200TfheError(f"PBSOrder: unknown variant {order}")
    out["pbs_order"] = PBS_ORDER[order]
    r.done()
    return out


def serialize_shortint_server_key(ksk, ks_base_log, ks_level, output_lwe_size, fourier, polynomial_size, input_lwe_dimension, glwe_size,
                                  pbs_base_log, pbs_level, message_modulus, carry_modulus, max_degree, max_noise_level, pbs_order,
                                  grouping_factor=None, deterministic_execution=False):
    w = Writer()
    _write_ksk(w, ksk, ks_base_log, ks_level, output_lwe_size, 64)
    w.u32(0 if grouping_factor is None else 1)
    _write_fourier_bsk(w, fourier, polynomial_size, input_lwe_dimension, glwe_size, pbs_base_log, pbs_level, grouping_factor)
    if grouping_factor is not None:
        w.u8(1 if deterministic_execution else 0)
    for v in (message_modulus, carry_modulus, max_degree, max_noise_level):
        w.u64(v)
    _write_modulus(w, 64)
    w.u32(PBS_ORDER_INV[pbs_order])
    return w.bytes()


def deserialize_shortint_ciphertext(data):
    r = Reader(data)
    ct = r.vec(np.uint64)
    _read_modulus(r, 64)
    out = {"ct": ct, "degree": r.u64(), "noise_level": r.u64(), "message_modulus": r.u64(), "carry_modulus": r.u64()}
    order = r.u32()
    if order not in PBS_ORDER:
        raise B200TfheError(f"PBSOrder: unknown variant {order}")
    out["pbs_order"] = PBS_ORDER[order]
    r.done()
    return out


def serialize_shortint_ciphertext(ct, degree, noise_level, message_modulus, carry_modulus, pbs_order):
    w = Writer()
    w.vec(ct, np.uint64)
    _write_modulus(w, 64)
    for v in (degree, noise_level, message_modulus, carry_modulus):
        w.u64(v)
    w.u32(PBS_ORDER_INV[pbs_order])
    return w.bytes()
