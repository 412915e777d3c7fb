"""Host-side mirror of tfhe::core_crypto for the KS+PBS path, over the C ABI (include/b200tfhe.h).

Names and argument meaning follow the reference (paths relative to /root/reference/tfhe/src/core_crypto):
  keyswitch_lwe_ciphertext                     algorithms/lwe_keyswitch.rs:96-170
  programmable_bootstrap_lwe_ciphertext        algorithms/lwe_programmable_bootstrapping.rs:1017-1113
  blind_rotate_assign                          algorithms/lwe_programmable_bootstrapping.rs:232-289
  programmable_bootstrap_..._without_sample_extract   :1115-1162 (fork)
  convert_standard_lwe_bootstrap_key_to_fourier       algorithms/lwe_bootstrap_key_conversion.rs:21-152
  extract_lwe_sample_from_glwe_ciphertext      algorithms/glwe_sample_extraction.rs:91-147
  fast_pbs_modulus_switch                      fft_impl/common.rs:26-43
  convert_standard_lwe_multi_bit_bootstrap_key_to_fourier   algorithms/lwe_multi_bit_bootstrap_key_conversion.rs
  multi_bit_programmable_bootstrap_lwe_ciphertext, multi_bit_blind_rotate_assign
                                               algorithms/lwe_multi_bit_programmable_bootstrapping.rs:295-546,1035-1127
The reference API is one ciphertext per call; every function here also accepts a list (2-D array), which is
the unit the GPU works on.  Errors: the reference panics on dimension / modulus mismatch; here the C ABI
returns non-zero and this layer raises B200TfheError with the library's message.
"""
import ctypes as C
import numpy as np

from . import _native
from ._native import Params


class B200TfheError(RuntimeError):
    pass


def _dtype(bits):
    return np.uint64 if bits == 64 else np.uint32


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Context:
    """One CUDA device + stream + workspace (the thread-local engine of shortint/engine/mod.rs:23-69)."""

    def __init__(self, device=0):
        self._lib = _native.lib()
        h = C.c_void_p()
        rc = self._lib.b200tfhe_context_create(int(device), C.byref(h))
        if rc != 0:
            raise B200TfheError(self._lib.b200tfhe_last_error(None).decode())
        self._h = h
        self.device = device

    def check(self, rc):
        if rc != 0:
            raise B200TfheError(self._lib.b200tfhe_last_error(self._h).decode())

    def synchronize(self):
        self.check(self._lib.b200tfhe_synchronize(self._h))

    def set_option(self, key, value):
        """Tuning / test knob of this context (b200tfhe_context_set_option): "br_kernel", "lat_max", "lat_cluster_max", "mb_lat_max", "ks_tensor_core", "ks_tc_min", "chunk_min"."""
        self.check(self._lib.b200tfhe_context_set_option(self._h, key.encode(), int(value)))

    @property
    def stream(self):
        return self._lib.b200tfhe_context_stream(self._h)

    @property
    def launch_count(self):
        return int(self._lib.b200tfhe_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.b200tfhe_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_params(n, k, N, pbs_base_log, pbs_level, ks_base_log, ks_level, bits):
    return Params(n, k, N, pbs_base_log, pbs_level, ks_base_log, ks_level, bits)


class FourierLweBootstrapKey:
    """Device-resident Fourier bootstrap key (fft_impl/fft64/crypto/bootstrap.rs:27-33)."""

    def __init__(self, ctx, params, handle):
        self.ctx, self.params, self._h = ctx, params, handle

    @property
    def bits(self):
        return self.params.torus_bits

    def device_buffer(self):
        p, n = C.c_void_p(), C.c_size_t()
        self.ctx.check(self.ctx._lib.b200tfhe_bsk_device_buffer(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    @property
    def grouping_factor(self):
        """1 for a classic key, LweBskGroupingFactor for a multi-bit key."""
        return int(self.ctx._lib.b200tfhe_bsk_grouping_factor(self._h))

    @classmethod
    def alloc(cls, ctx, params, grouping_factor=1):
        h = C.c_void_p()
        if grouping_factor > 1:
            ctx.check(ctx._lib.b200tfhe_multibit_bsk_alloc(ctx._h, C.byref(params), int(grouping_factor), C.byref(h)))
        else:
            ctx.check(ctx._lib.b200tfhe_bsk_alloc(ctx._h, C.byref(params), C.byref(h)))
        return cls(ctx, params, h)

    @classmethod
    def from_fourier(cls, ctx, params, fourier, grouping_factor=1):
        """Key already in the Fourier domain, the reference's serialised order: [ggsw][l][k+1][k+1][N/2] complex f64 in natural
        frequency order (what bincode-deserialising a FourierLweBootstrapKey yields)."""
        f = np.ascontiguousarray(fourier, dtype=np.float64).reshape(-1)
        G = int(grouping_factor)
        ggsw = params.lwe_dimension if G <= 1 else (params.lwe_dimension // G << G)
        expect = ggsw * params.pbs_level * (params.glwe_dimension + 1) ** 2 * params.polynomial_size
        if f.size != expect:
            raise B200TfheError(f"Fourier bootstrap key has {f.size} f64 values, parameters imply {expect}")
        h = C.c_void_p()
        ctx.check(ctx._lib.b200tfhe_bsk_upload_fourier(ctx._h, C.byref(params), G, _ptr(f), C.byref(h)))
        return cls(ctx, params, h)

    def close(self):
        if getattr(self, "_h", None):
            self.ctx._lib.b200tfhe_bsk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def convert_standard_lwe_bootstrap_key_to_fourier(ctx, params, std_bsk):
    """std_bsk: flat [n][l][k+1][k+1][N] scalars (entities/lwe_bootstrap_key.rs:99-106) -> FourierLweBootstrapKey."""
    dt = _dtype(params.torus_bits)
    a = np.ascontiguousarray(std_bsk, dtype=dt).reshape(-1)
    expect = params.lwe_dimension * params.pbs_level * (params.glwe_dimension + 1) ** 2 * params.polynomial_size
    if a.size != expect:
        raise B200TfheError(f"bootstrap key has {a.size} scalars, parameters imply {expect}")
    h = C.c_void_p()
    fn = ctx._lib.b200tfhe_bsk_upload_u64 if params.torus_bits == 64 else ctx._lib.b200tfhe_bsk_upload_u32
    ctx.check(fn(ctx._h, C.byref(params), _ptr(a), C.byref(h)))
    return FourierLweBootstrapKey(ctx, params, h)


def convert_standard_lwe_bootstrap_key_to_fourier_dev(ctx, params, d_std_bsk_ptr):
    h = C.c_void_p()
    ctx.check(ctx._lib.b200tfhe_bsk_convert_dev(ctx._h, C.byref(params), C.c_void_p(d_std_bsk_ptr), C.byref(h)))
    return FourierLweBootstrapKey(ctx, params, h)


# FourierLweMultiBitBootstrapKey (fft_impl/fft64/crypto/bootstrap ... entities/lwe_multi_bit_bootstrap_key.rs): the same
# device handle type with grouping_factor > 1
FourierLweMultiBitBootstrapKey = FourierLweBootstrapKey


def convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, params, std_bsk, grouping_factor):
    """std_bsk: flat [n/G][2^G][l][k+1][k+1][N] scalars (entities/lwe_multi_bit_bootstrap_key.rs) -> device Fourier key."""
    if params.torus_bits != 64:
        raise B200TfheError("multi-bit bootstrap keys exist for 64-bit parameter sets only")
    G = int(grouping_factor)
    a = np.ascontiguousarray(std_bsk, dtype=np.uint64).reshape(-1)
    if G < 2 or params.lwe_dimension % G:
        raise B200TfheError(f"grouping factor {G} must be >= 2 and divide the LWE dimension {params.lwe_dimension}")
    expect = (params.lwe_dimension // G << G) * params.pbs_level * (params.glwe_dimension + 1) ** 2 * params.polynomial_size
    if a.size != expect:
        raise B200TfheError(f"multi-bit bootstrap key has {a.size} scalars, parameters imply {expect}")
    h = C.c_void_p()
    ctx.check(ctx._lib.b200tfhe_multibit_bsk_upload_u64(ctx._h, C.byref(params), G, _ptr(a), C.byref(h)))
    return FourierLweBootstrapKey(ctx, params, h)


class LweKeyswitchKey:
    """Device-resident keyswitch key in the reference layout (entities/lwe_keyswitch_key.rs:77-108)."""

    def __init__(self, ctx, params, handle):
        self.ctx, self.params, self._h = ctx, params, handle

    @property
    def bits(self):
        return self.params.torus_bits

    @classmethod
    def upload(cls, ctx, params, ksk):
        dt = _dtype(params.torus_bits)
        a = np.ascontiguousarray(ksk, dtype=dt).reshape(-1)
        expect = params.glwe_dimension * params.polynomial_size * params.ks_level * (params.lwe_dimension + 1)
        if a.size != expect:
            raise B200TfheError(f"keyswitch key has {a.size} scalars, parameters imply {expect}")
        h = C.c_void_p()
        fn = ctx._lib.b200tfhe_ksk_upload_u64 if params.torus_bits == 64 else ctx._lib.b200tfhe_ksk_upload_u32
        ctx.check(fn(ctx._h, C.byref(params), _ptr(a), C.byref(h)))
        return cls(ctx, params, h)

    @classmethod
    def alloc(cls, ctx, params):
        h = C.c_void_p()
        ctx.check(ctx._lib.b200tfhe_ksk_alloc(ctx._h, C.byref(params), C.byref(h)))
        return cls(ctx, params, h)

    def device_buffer(self):
        p, n = C.c_void_p(), C.c_size_t()
        self.ctx.check(self.ctx._lib.b200tfhe_ksk_device_buffer(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def finalize(self):
        self.ctx.check(self.ctx._lib.b200tfhe_ksk_finalize(self.ctx._h, self._h))

    def close(self):
        if getattr(self, "_h", None):
            self.ctx._lib.b200tfhe_ksk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- batched operations (host arrays in, host arrays out) ------------------------------------------------------
def _suf(bits):
    return "u64" if bits == 64 else "u32"


def _as_list(a, width, bits, what):
    a = np.ascontiguousarray(a, dtype=_dtype(bits))
    single = a.ndim == 1
    a = a.reshape(-1, a.shape[-1]) if a.size else a.reshape(0, width)
    if a.shape[1] != width:
        raise B200TfheError(f"{what}: ciphertext length {a.shape[1]} != expected {width}")
    return a, single


def _luts(acc, idx, p, batch):
    glwe = (p.glwe_dimension + 1) * p.polynomial_size
    luts = np.ascontiguousarray(acc, dtype=_dtype(p.torus_bits)).reshape(-1, glwe)
    li = None
    if idx is not None:
        li = np.ascontiguousarray(idx, dtype=np.uint32).reshape(-1)
        if li.size != batch:
            raise B200TfheError("lut_idx must have one entry per ciphertext")
    elif luts.shape[0] > 1 and batch > 0:
        # several accumulators and no index: one per ciphertext, or an error -- never "silently accumulator 0 for everybody"
        if luts.shape[0] != batch:
            raise B200TfheError(f"{luts.shape[0]} accumulators for {batch} ciphertexts and no lut_idx: pass one accumulator, one per "
                                "ciphertext, or an index")
        li = np.arange(batch, dtype=np.uint32)
    return luts, li


def keyswitch_lwe_ciphertext(ksk, input_lwe, output_lwe=None):
    """out = keyswitch(in).  in: [kN+1] or [B][kN+1]; returns / fills out: [n+1] or [B][n+1].  Bit-exact."""
    p, ctx = ksk.params, ksk.ctx
    a, single = _as_list(input_lwe, p.glwe_dimension * p.polynomial_size + 1, p.torus_bits, "keyswitch input")
    out = np.empty((a.shape[0], p.lwe_dimension + 1), _dtype(p.torus_bits))
    fn = getattr(ctx._lib, f"b200tfhe_keyswitch_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, ksk._h, _ptr(a), _ptr(out), a.shape[0]))
    res = out[0] if single else out
    if output_lwe is not None:
        np.copyto(output_lwe, res)
        return output_lwe
    return res


def programmable_bootstrap_lwe_ciphertext(input_lwe, accumulator, fourier_bsk, output_lwe=None, lut_idx=None):
    """out = PBS(in, accumulator).  in: [n+1] or [B][n+1]; accumulator: GLWE [(k+1)N] or a list of them
    selected per ciphertext by lut_idx; out: [kN+1] per ciphertext."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    a, single = _as_list(input_lwe, p.lwe_dimension + 1, p.torus_bits, "PBS input")
    luts, li = _luts(accumulator, lut_idx, p, a.shape[0])
    out = np.empty((a.shape[0], p.glwe_dimension * p.polynomial_size + 1), _dtype(p.torus_bits))
    fn = getattr(ctx._lib, f"b200tfhe_pbs_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, fourier_bsk._h, _ptr(a), _ptr(out), _ptr(luts), luts.shape[0], _ptr(li), a.shape[0]))
    res = out[0] if single else out
    if output_lwe is not None:
        np.copyto(output_lwe, res)
        return output_lwe
    return res


def programmable_bootstrap_lwe_ciphertext_without_sample_extract(input_lwe, accumulator, fourier_bsk, lut_idx=None):
    """Fork entry (lwe_programmable_bootstrapping.rs:1115-1162): returns the rotated GLWE [(k+1)N] per ciphertext."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    a, single = _as_list(input_lwe, p.lwe_dimension + 1, p.torus_bits, "blind-rotate input")
    luts, li = _luts(accumulator, lut_idx, p, a.shape[0])
    out = np.empty((a.shape[0], (p.glwe_dimension + 1) * p.polynomial_size), _dtype(p.torus_bits))
    fn = getattr(ctx._lib, f"b200tfhe_blind_rotate_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, fourier_bsk._h, _ptr(a), _ptr(out), _ptr(luts), luts.shape[0], _ptr(li), a.shape[0]))
    return out[0] if single else out


def blind_rotate_assign(input_lwe, accumulator_glwe, fourier_bsk):
    """acc <- blind rotation of acc by the (modulus-switched) input, in place (one accumulator per ciphertext)."""
    p = fourier_bsk.params
    glwe = (p.glwe_dimension + 1) * p.polynomial_size
    acc = np.asarray(accumulator_glwe)
    a, _ = _as_list(input_lwe, p.lwe_dimension + 1, p.torus_bits, "blind-rotate input")
    luts = np.ascontiguousarray(acc, dtype=_dtype(p.torus_bits)).reshape(-1, glwe)
    if luts.shape[0] != a.shape[0]:
        raise B200TfheError("blind_rotate_assign needs one accumulator per ciphertext")
    res = programmable_bootstrap_lwe_ciphertext_without_sample_extract(a, luts, fourier_bsk, np.arange(a.shape[0], dtype=np.uint32))
    acc[...] = res.reshape(acc.shape)
    return acc


def _ggsw_idx(ggsw_idx, batch):
    if ggsw_idx is None:
        return None
    return np.ascontiguousarray(np.broadcast_to(np.asarray(ggsw_idx, dtype=np.uint32), (batch,)))


def add_external_product_assign(out, ggsw, glwe, ggsw_idx=None):
    """out <- out + ggsw (x) glwe (algorithms/lwe_programmable_bootstrapping.rs:309-475), in place over a list.

    ggsw: a FourierLweBootstrapKey used as a list of lwe_dimension Fourier GGSW ciphertexts; ggsw_idx selects one per
    ciphertext (scalar or array; None = GGSW 0)."""
    p, ctx = ggsw.params, ggsw.ctx
    w = (p.glwe_dimension + 1) * p.polynomial_size
    dt = _dtype(p.torus_bits)
    if not (isinstance(out, np.ndarray) and out.dtype == dt and out.flags.c_contiguous and out.size % w == 0):
        raise B200TfheError(f"external product output must be a contiguous {dt.__name__} array of whole GLWE ciphertexts")
    g, _ = _as_list(glwe, w, p.torus_bits, "external-product input")
    batch = out.size // w
    if g.shape[0] != batch:
        raise B200TfheError("external product: out and glwe hold different numbers of ciphertexts")
    fn = getattr(ctx._lib, f"b200tfhe_external_product_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, ggsw._h, _ptr(_ggsw_idx(ggsw_idx, batch)), _ptr(out), _ptr(g), batch))
    return out


def cmux_assign(ct0, ct1, ggsw, ggsw_idx=None):
    """cmux_assign (lwe_programmable_bootstrapping.rs:552-765 -> ggsw.rs:766-777), in place over a list, both operands updated as
    the reference updates them: ct1 <- ct1 - ct0, ct0 <- ct0 + ggsw (x) ct1."""
    p, ctx = ggsw.params, ggsw.ctx
    w = (p.glwe_dimension + 1) * p.polynomial_size
    dt = _dtype(p.torus_bits)
    for a in (ct0, ct1):
        if not (isinstance(a, np.ndarray) and a.dtype == dt and a.flags.c_contiguous and a.size % w == 0):
            raise B200TfheError(f"cmux operands must be contiguous {dt.__name__} arrays of whole GLWE ciphertexts")
    if ct0.size != ct1.size:
        raise B200TfheError("cmux: ct0 and ct1 hold different numbers of ciphertexts")
    batch = ct0.size // w
    fn = getattr(ctx._lib, f"b200tfhe_cmux_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, ggsw._h, _ptr(_ggsw_idx(ggsw_idx, batch)), _ptr(ct0), _ptr(ct1), batch))
    return ct0, ct1


def keyswitch_programmable_bootstrap(ksk, fourier_bsk, input_lwe, accumulator, lut_idx=None):
    """KS then PBS (shortint keyswitch_programmable_bootstrap_assign core, shortint/server_key/mod.rs:783-857)."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    w = p.glwe_dimension * p.polynomial_size + 1
    a, single = _as_list(input_lwe, w, p.torus_bits, "KS+PBS input")
    luts, li = _luts(accumulator, lut_idx, p, a.shape[0])
    out = np.empty((a.shape[0], w), _dtype(p.torus_bits))
    fn = getattr(ctx._lib, f"b200tfhe_ks_pbs_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, ksk._h, fourier_bsk._h, _ptr(a), _ptr(out), _ptr(luts), luts.shape[0], _ptr(li), a.shape[0]))
    return out[0] if single else out


def programmable_bootstrap_keyswitch(ksk, fourier_bsk, input_lwe, accumulator, lut_idx=None):
    """PBS then KS (shortint :859-933; boolean bootstrap_keyswitch, boolean/engine/bootstrapping.rs:299-344)."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    w = p.lwe_dimension + 1
    a, single = _as_list(input_lwe, w, p.torus_bits, "PBS+KS input")
    luts, li = _luts(accumulator, lut_idx, p, a.shape[0])
    out = np.empty((a.shape[0], w), _dtype(p.torus_bits))
    fn = getattr(ctx._lib, f"b200tfhe_pbs_ks_batch_{_suf(p.torus_bits)}")
    ctx.check(fn(ctx._h, ksk._h, fourier_bsk._h, _ptr(a), _ptr(out), _ptr(luts), luts.shape[0], _ptr(li), a.shape[0]))
    return out[0] if single else out


def fast_pbs_modulus_switch(ctx, scalars, polynomial_size, bits):
    a = np.ascontiguousarray(scalars, dtype=_dtype(bits)).reshape(-1)
    out = np.empty(a.size, np.uint32)
    log2n = int(polynomial_size).bit_length() - 1
    fn = getattr(ctx._lib, f"b200tfhe_modulus_switch_{_suf(bits)}")
    ctx.check(fn(ctx._h, _ptr(a), _ptr(out), a.size, log2n))
    return out


def extract_lwe_sample_from_glwe_ciphertext(ctx, glwe, glwe_dimension, polynomial_size, bits):
    """nth = MonomialDegree(0).  glwe: [(k+1)N] or [B][(k+1)N] -> [kN+1] per ciphertext.  Bit-exact."""
    a, single = _as_list(glwe, (glwe_dimension + 1) * polynomial_size, bits, "sample-extract input")
    out = np.empty((a.shape[0], glwe_dimension * polynomial_size + 1), _dtype(bits))
    fn = getattr(ctx._lib, f"b200tfhe_sample_extract_batch_{_suf(bits)}")
    ctx.check(fn(ctx._h, _ptr(a), _ptr(out), glwe_dimension, polynomial_size, a.shape[0]))
    return out[0] if single else out


# ---- multi-bit entry points: same kernels behind the same C calls, the key handle carries the grouping factor ---------
def multi_bit_programmable_bootstrap_lwe_ciphertext(input_lwe, accumulator, multi_bit_bsk, output_lwe=None, lut_idx=None,
                                                    thread_count=None):
    """algorithms/lwe_multi_bit_programmable_bootstrapping.rs:1035-1127.  thread_count (the reference's CPU producer
    threads per ciphertext) is accepted and ignored: the batch is the unit of parallelism here."""
    if multi_bit_bsk.grouping_factor < 2:
        raise B200TfheError("multi_bit_programmable_bootstrap_lwe_ciphertext needs a multi-bit bootstrap key")
    return programmable_bootstrap_lwe_ciphertext(input_lwe, accumulator, multi_bit_bsk, output_lwe, lut_idx)


multi_bit_deterministic_programmable_bootstrap_lwe_ciphertext = multi_bit_programmable_bootstrap_lwe_ciphertext


def multi_bit_blind_rotate_assign(input_lwe, accumulator_glwe, multi_bit_bsk, thread_count=None):
    """:295-546 (and the deterministic variant :548-...): acc <- multi-bit blind rotation, in place."""
    if multi_bit_bsk.grouping_factor < 2:
        raise B200TfheError("multi_bit_blind_rotate_assign needs a multi-bit bootstrap key")
    return blind_rotate_assign(input_lwe, accumulator_glwe, multi_bit_bsk)


# ---- bit extraction (first half of the WoP-PBS) -------------------------------------------------------------------------
def extract_bits_from_lwe_ciphertext_mem_optimized(lwe_in, fourier_bsk, ksk, delta_log, number_of_bits_to_extract, lwe_list_out=None):
    """algorithms/lwe_wopbs.rs:321-364 -> fft64/crypto/wop_pbs/mod.rs:66-225 (argument order of the reference; fft and stack
    have no counterpart).  lwe_in: [kN+1] or [B][kN+1] under the big key; result [number_of_bits][n+1] per ciphertext under the
    small key, most significant extracted bit first, every bit scaled by q/2."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    if p.torus_bits != 64:
        raise B200TfheError("extract_bits: 64-bit torus only")
    a, single = _as_list(lwe_in, p.glwe_dimension * p.polynomial_size + 1, 64, "extract_bits input")
    nb = int(number_of_bits_to_extract)
    if nb < 0 or int(delta_log) < 0:
        raise B200TfheError("extract_bits: negative delta_log / bit count")
    out = np.zeros((a.shape[0], nb, p.lwe_dimension + 1), np.uint64)
    ctx.check(ctx._lib.b200tfhe_extract_bits_batch_u64(ctx._h, ksk._h, fourier_bsk._h, _ptr(a), int(delta_log), nb, _ptr(out), a.shape[0]))
    res = out[0] if single else out
    if lwe_list_out is not None:
        np.copyto(lwe_list_out, res)
        return lwe_list_out
    return res


# ---- circuit bootstrap + vertical packing (second half of the WoP-PBS) ---------------------------------------------------
class LwePrivateFunctionalPackingKeyswitchKeyList:
    """Device-resident key list of the circuit bootstrap (entities/lwe_private_functional_packing_keyswitch_key_list.rs,
    generated by algorithms/lwe_wopbs.rs:80-160): [k+1][kN+1][level 1..L][(k+1)N]."""

    def __init__(self, ctx, params, handle, decomp_base_log, decomp_level_count):
        self.ctx, self.params, self._h = ctx, params, handle
        self.decomp_base_log, self.decomp_level_count = decomp_base_log, decomp_level_count

    @classmethod
    def upload(cls, ctx, params, decomp_base_log, decomp_level_count, key_list):
        a = np.ascontiguousarray(key_list, dtype=np.uint64).reshape(-1)
        k, N = params.glwe_dimension, params.polynomial_size
        expect = (k + 1) * (k * N + 1) * int(decomp_level_count) * (k + 1) * N
        if a.size != expect:
            raise B200TfheError(f"pfpksk list has {a.size} scalars, parameters imply {expect}")
        h = C.c_void_p()
        ctx.check(ctx._lib.b200tfhe_cbs_pfpksk_upload_u64(ctx._h, C.byref(params), int(decomp_base_log), int(decomp_level_count), _ptr(a), C.byref(h)))
        return cls(ctx, params, h, int(decomp_base_log), int(decomp_level_count))

    def close(self):
        if getattr(self, "_h", None):
            self.ctx._lib.b200tfhe_cbs_pfpksk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pfpksk_list, key_index, input_lwe):
    """algorithms/lwe_private_functional_packing_keyswitch.rs:8-89 with key `key_index` of the list: [kN+1] or [B][kN+1] ->
    GLWE [(k+1)N] per ciphertext.  Bit-exact."""
    p, ctx = pfpksk_list.params, pfpksk_list.ctx
    a, single = _as_list(input_lwe, p.glwe_dimension * p.polynomial_size + 1, 64, "pfpks input")
    out = np.empty((a.shape[0], (p.glwe_dimension + 1) * p.polynomial_size), np.uint64)
    ctx.check(ctx._lib.b200tfhe_pfpks_batch_u64(ctx._h, pfpksk_list._h, int(key_index), _ptr(a), _ptr(out), a.shape[0]))
    return out[0] if single else out


def circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(lwe_list_in, big_lut_as_polynomial_list, fourier_bsk,
                                                                                 pfpksk_list, base_log_cbs, level_cbs,
                                                                                 number_of_luts=None, lwe_list_out=None):
    """algorithms/lwe_wopbs.rs:651-704 -> fft64/crypto/wop_pbs/mod.rs:647-853.  lwe_list_in: [nbits][n+1] or [B][nbits][n+1]
    (extract_bits output); big_lut_as_polynomial_list: [number_of_luts * polys_per_lut][N]; the reference deduces
    number_of_luts from lwe_list_out (:736) -- here it is an argument (default: one polynomial per table); result
    [number_of_luts][kN+1] per input."""
    p, ctx = fourier_bsk.params, fourier_bsk.ctx
    if p.torus_bits != 64:
        raise B200TfheError("circuit bootstrap: 64-bit torus only")
    a = np.ascontiguousarray(lwe_list_in, dtype=np.uint64)
    if a.ndim not in (2, 3) or a.shape[-1] != p.lwe_dimension + 1:
        raise B200TfheError(f"circuit bootstrap input: ciphertext length {a.shape[-1] if a.ndim else 0} != expected {p.lwe_dimension + 1}")
    single = a.ndim == 2
    a = a.reshape(-1, a.shape[-2], a.shape[-1])
    luts = np.ascontiguousarray(big_lut_as_polynomial_list, dtype=np.uint64).reshape(-1, p.polynomial_size)
    n_luts = luts.shape[0] if number_of_luts is None else int(number_of_luts)
    if n_luts < 1 or luts.shape[0] % n_luts:
        raise B200TfheError("circuit bootstrap: the polynomial list does not split into number_of_luts tables")   # chunks_exact (:741)
    out = np.zeros((a.shape[0], n_luts, p.glwe_dimension * p.polynomial_size + 1), np.uint64)
    ctx.check(ctx._lib.b200tfhe_cbs_vertical_packing_batch_u64(ctx._h, fourier_bsk._h, pfpksk_list._h, _ptr(a), a.shape[1], _ptr(luts), n_luts,
                                                               luts.shape[0] // n_luts, int(base_log_cbs), int(level_cbs), _ptr(out), a.shape[0]))
    res = out[0] if single else out
    if lwe_list_out is not None:
        np.copyto(lwe_list_out, res)
        return lwe_list_out
    return res
