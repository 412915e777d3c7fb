"""Host-side mirror of the fork's gadget engine for the operations built around the blind rotation.

Reference (relative to /root/reference/tfhe/src/gadget):
  Encoding (odd plaintext modulus p, parts of Z_p per element of Z_o)       ciphertext/mod.rs:24-255
  Memory::create_accumulator / as_buffers / as_buffers_common_factor        engine/bootstrapping.rs:41-71,146-205,214-240
  Bootstrapper::bootstrap (blind rotation without sample extract + extract) engine/bootstrapping.rs:389-447
  Bootstrapper::create_vi_for_mvb / mvb_bootstrap                           engine/bootstrapping.rs:503-543,567-622
  Bootstrapper::pack_into_new_accumulator                                   engine/bootstrapping.rs:690-773
  keyswitch_lwe_ciphertext_into_glwe_ciphertext                             core_crypto/algorithms/lwe_packing_keyswitch.rs:100-187
Everything numeric runs on the GPU through the C ABI (include/b200tfhe.h); this file only builds the small plaintext
tables (accumulators, v_i polynomials) exactly as the reference does, batch-first.  The reference's debugging side effects
(println!, SerializableCiphertext::append_to_file("mvb")) are not reproduced.
"""
import ctypes as C

import numpy as np

from . import core_crypto as cc
from .core_crypto import B200TfheError, _ptr

U64 = (1 << 64)


class Encoding:
    """Encoding{origin_modulus o, parts: for each i in Z_o the set of Z_p elements that encode it, modulus_p}."""

    def __init__(self, origin_modulus, parts, modulus_p):
        parts = [frozenset(int(x) for x in part) for part in parts]
        if len(parts) != origin_modulus or any(x >= modulus_p for part in parts for x in part):
            raise B200TfheError("This Arithmetic Encoding is not correct !")
        self.origin_modulus, self.parts, self.modulus_p = int(origin_modulus), parts, int(modulus_p)

    @classmethod
    def new_canonical(cls, origin_modulus, values_for_singletons, modulus_p):
        return cls(origin_modulus, [{v} for v in values_for_singletons], modulus_p)

    def get_modulus(self):
        return self.modulus_p

    def get_origin_modulus(self):
        return self.origin_modulus

    def is_canonical(self):
        return all(len(part) == 1 for part in self.parts)

    def get_part_single_value_if_canonical(self, i):
        if not self.is_canonical():
            raise B200TfheError("encoding is not canonical")
        return next(iter(self.parts[i]))

    def inverse_encoding(self, x):
        for i, part in enumerate(self.parts):
            if x in part:
                return i
        return None

    def negative_on_p_ring(self, x):
        return (self.modulus_p - x) % self.modulus_p

    def apply_lut_to_encoding(self, f):
        """ciphertext/mod.rs:225-254: element f(i) of the output inherits the Z_p values of i."""
        parts = [set() for _ in range(self.origin_modulus)]
        for i in range(self.origin_modulus):
            parts[f(i)] |= self.parts[i]
        return Encoding(self.origin_modulus, parts, self.modulus_p)

    def encode(self, message):
        """engine/mod.rs:126-134 encode_message_into_plaintext: floor(2^64 / p) * z."""
        return (U64 // self.modulus_p * self.get_part_single_value_if_canonical(message)) % U64

    def decode_phase(self, phase):
        """nearest multiple of 2^64/p -> Z_p element -> Z_o element (None when no part contains it)."""
        z = int(round(int(phase) * self.modulus_p / U64)) % self.modulus_p
        return self.inverse_encoding(z)


def create_accumulator(encoding_in, encoding_out):
    """Memory::create_accumulator (bootstrapping.rs:41-71): the Z_p' value of each of the p windows (odd p)."""
    p = encoding_in.get_modulus()
    if p % 2 != 1 or not encoding_out.is_canonical():
        raise B200TfheError("create_accumulator needs an odd input modulus and a canonical output encoding")
    accu = [0] * p
    for k in range(p):
        if k % 2 == 0:
            i = encoding_in.inverse_encoding(k // 2)
            accu[k] = 0 if i is None else encoding_out.get_part_single_value_if_canonical(i)
        else:
            i = encoding_in.inverse_encoding((p + 1) // 2 + (k - 1) // 2)
            accu[k] = 0 if i is None else encoding_out.negative_on_p_ring(encoding_out.get_part_single_value_if_canonical(i))
    return accu


def accumulator_glwe(params, encoding_in, encoding_out):
    """Memory::as_buffers, p != 2 branch (bootstrapping.rs:146-186): trivial GLWE whose body holds the windows."""
    N, k = params.polynomial_size, params.glwe_dimension
    p, new_p = encoding_in.get_modulus(), encoding_out.get_modulus()
    accu = create_accumulator(encoding_in, encoding_out)
    unit = U64 // new_p
    body = np.zeros(N, np.uint64)
    shift = N // (2 * p)
    body[:shift] = np.uint64(unit * accu[0] % U64)
    for kk in range(1, p):
        body[shift + (kk - 1) * N // p: shift + kk * N // p] = np.uint64(unit * accu[kk] % U64)
    body[N - shift:] = np.uint64(unit * ((new_p - accu[0]) % new_p) % U64)
    out = np.zeros((k + 1) * N, np.uint64)
    out[k * N:] = body
    return out


def common_factor_accumulator(params, encoding_out):
    """Memory::as_buffers_common_factor (bootstrapping.rs:214-240): body filled with 2^64/p' (2^63/p' for an even p')."""
    N, k = params.polynomial_size, params.glwe_dimension
    constant = (1 << 63) if encoding_out.get_modulus() % 2 == 0 else (1 << 64)
    out = np.zeros((k + 1) * N, np.uint64)
    out[k * N:] = np.uint64(constant // encoding_out.get_modulus())
    return out


def create_vi_for_mvb(params, enc_in, enc_out):
    """Bootstrapper::create_vi_for_mvb (bootstrapping.rs:503-543): plaintext polynomial with p non-zero coefficients whose
    product with the rotated common factor rebuilds the accumulator of (enc_in -> enc_out)."""
    N = params.polynomial_size
    accu = create_accumulator(enc_in, enc_out)
    p, new_p = enc_in.get_modulus(), enc_out.get_modulus()
    if new_p % 2 == 1:
        inv2 = (new_p + 1) // 2
        accu = [x * inv2 % new_p for x in accu]
    elif new_p == 2:
        new_p = 4
    v = np.zeros(N, np.uint64)
    for i in range(p - 1):
        v[N // (2 * p) + i * N // p] = (accu[i + 1] - accu[i]) % new_p
    v[N // (2 * p) + (p - 1) * N // p] = (new_p - accu[0] - accu[p - 1]) % new_p
    return v


class LwePackingKeyswitchKey:
    """Device-resident LwePackingKeyswitchKey (entities/lwe_packing_keyswitch_key.rs): [kN][l_ks][(k+1)N]."""

    def __init__(self, ctx, params, handle):
        self.ctx, self.params, self._h = ctx, params, handle

    @classmethod
    def upload(cls, ctx, params, pksk):
        a = np.ascontiguousarray(pksk, dtype=np.uint64).reshape(-1)
        kN = params.glwe_dimension * params.polynomial_size
        expect = kN * params.ks_level * (params.glwe_dimension + 1) * params.polynomial_size
        if a.size != expect:
            raise B200TfheError(f"packing keyswitch key has {a.size} scalars, parameters imply {expect}")
        h = C.c_void_p()
        ctx.check(ctx._lib.b200tfhe_pksk_upload_u64(ctx._h, C.byref(params), _ptr(a), C.byref(h)))
        return cls(ctx, params, h)

    def close(self):
        if getattr(self, "_h", None):
            self.ctx._lib.b200tfhe_pksk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def keyswitch_lwe_ciphertext_into_glwe_ciphertext(pksk, input_lwe):
    """in: [kN+1] or [B][kN+1] -> GLWE [(k+1)N] per ciphertext.  Bit-exact."""
    p, ctx = pksk.params, pksk.ctx
    a, single = cc._as_list(input_lwe, p.glwe_dimension * p.polynomial_size + 1, 64, "packing keyswitch input")
    out = np.empty((a.shape[0], (p.glwe_dimension + 1) * p.polynomial_size), np.uint64)
    ctx.check(ctx._lib.b200tfhe_packing_keyswitch_batch_u64(ctx._h, pksk._h, _ptr(a), _ptr(out), a.shape[0]))
    return out[0] if single else out


def glwe_mul_plain_extract(ctx, params, glwe, polys):
    """out[b][i] = sample_extract_0(glwe[b] * polys[i]) (mvb_bootstrap post-processing).  Bit-exact."""
    glwe_len = (params.glwe_dimension + 1) * params.polynomial_size
    g, _ = cc._as_list(glwe, glwe_len, 64, "GLWE input")
    pl = np.ascontiguousarray(polys, dtype=np.uint64).reshape(-1, params.polynomial_size)
    out = np.empty((g.shape[0], pl.shape[0], params.glwe_dimension * params.polynomial_size + 1), np.uint64)
    ctx.check(ctx._lib.b200tfhe_glwe_mul_plain_extract_batch_u64(ctx._h, C.byref(params), _ptr(g), _ptr(pl), pl.shape[0], _ptr(out), g.shape[0]))
    return out


class Bootstrapper:
    """gadget Bootstrapper + ServerKey{bootstrapping_key, key_switching_key, lwe_packing_keyswitch_key, pbs_order}."""

    def __init__(self, bsk, ksk=None, pksk=None, pbs_order="ks_pbs"):
        self.bootstrapping_key, self.key_switching_key, self.lwe_packing_keyswitch_key = bsk, ksk, pksk
        self.pbs_order, self.params, self.ctx = pbs_order, bsk.params, bsk.ctx

    def bootstrap(self, input_lwe, enc_in, enc_out):
        """bootstrapping.rs:389-447 behind apply_bootstrapping_pattern (:887-901): accumulator of (enc_in -> enc_out)."""
        acc = accumulator_glwe(self.params, enc_in, enc_out)
        if self.pbs_order == "ks_pbs":
            return cc.keyswitch_programmable_bootstrap(self.key_switching_key, self.bootstrapping_key, input_lwe, acc)
        return cc.programmable_bootstrap_keyswitch(self.key_switching_key, self.bootstrapping_key, input_lwe, acc)

    def mvb_bootstrap(self, ciphertexts, enc_in, encs_out, lut_fis, keyswitch_first=None):
        """bootstrapping.rs:567-622 over a list: returns [B][len(lut_fis)][kN+1].  keyswitch_first defaults to the key's
        pbs order (Big key choice: the inputs live under the big key and are keyswitched first)."""
        p, ctx = self.params, self.ctx
        ks = (self.pbs_order == "ks_pbs") if keyswitch_first is None else keyswitch_first
        if len({e.get_modulus() for e in encs_out}) != 1:
            raise B200TfheError("mvb_bootstrap: all output encodings must share one modulus")   # the reference's open "problème"
        polys = np.stack([create_vi_for_mvb(p, enc_in.apply_lut_to_encoding(lambda x, t=tuple(lut): t[x]), enc_out)
                          for enc_out, lut in zip(encs_out, lut_fis)])
        acc = common_factor_accumulator(p, encs_out[0])
        width = p.glwe_dimension * p.polynomial_size + 1 if ks else p.lwe_dimension + 1
        a, _ = cc._as_list(ciphertexts, width, 64, "MVB input")
        out = np.empty((a.shape[0], polys.shape[0], p.glwe_dimension * p.polynomial_size + 1), np.uint64)
        ksk_h = self.key_switching_key._h if ks else None
        ctx.check(ctx._lib.b200tfhe_mvb_batch_u64(ctx._h, ksk_h, self.bootstrapping_key._h, _ptr(a), _ptr(acc), _ptr(polys),
                                                  polys.shape[0], _ptr(out), a.shape[0]))
        return out

    def pack_into_new_accumulator(self, lwe_ciphertexts, p):
        """bootstrapping.rs:690-773: [count][kN+1] (or [B][count][kN+1]) -> GLWE accumulator [(k+1)N] (per batch entry)."""
        prm, ctx = self.params, self.ctx
        width = prm.glwe_dimension * prm.polynomial_size + 1
        a = np.ascontiguousarray(lwe_ciphertexts, dtype=np.uint64)
        single = a.ndim == 2
        a = a.reshape(-1, a.shape[-2], a.shape[-1])
        if a.shape[2] != width:
            raise B200TfheError(f"pack_into_new_accumulator: ciphertext length {a.shape[2]} != expected {width}")
        out = np.empty((a.shape[0], (prm.glwe_dimension + 1) * prm.polynomial_size), np.uint64)
        ctx.check(ctx._lib.b200tfhe_pack_into_accumulator_batch_u64(ctx._h, self.lwe_packing_keyswitch_key._h, _ptr(a), a.shape[1],
                                                                    int(p), _ptr(out), a.shape[0]))
        return out[0] if single else out
