"""ctypes binding of libb200tfhe.so (C ABI: include/b200tfhe.h).  No CPU fallback: a missing library is an error."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200TFHE_LIB") or os.path.join(_HERE, "libb200tfhe.so")  # env override: A/B builds of the kernels


class Params(C.Structure):
    """b200tfhe_params (core_crypto/commons/parameters.rs:40-236 newtypes, flattened)."""
    _fields_ = [("lwe_dimension", C.c_int32), ("glwe_dimension", C.c_int32), ("polynomial_size", C.c_int32),
                ("pbs_base_log", C.c_int32), ("pbs_level", C.c_int32), ("ks_base_log", C.c_int32),
                ("ks_level", C.c_int32), ("torus_bits", C.c_int32)]


class NativeLibraryMissing(RuntimeError):
    pass


_lib = None
_VP, _U32, _SZ, _I = C.c_void_p, C.c_uint32, C.c_size_t, C.c_int
_PP = C.POINTER(C.c_void_p)

# name -> argtypes (all return int unless listed in _RESTYPES)
_PROTOS = {
    "b200tfhe_context_create": [_I, _PP],
    "b200tfhe_context_destroy": [_VP],
    "b200tfhe_synchronize": [_VP],
    "b200tfhe_last_error": [_VP],
    "b200tfhe_context_stream": [_VP],
    "b200tfhe_context_set_option": [_VP, C.c_char_p, C.c_int64],
    "b200tfhe_params_supported": [C.POINTER(Params)],
    "b200tfhe_bsk_upload_u64": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_bsk_upload_u32": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_bsk_convert_dev": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_bsk_alloc": [_VP, C.POINTER(Params), _PP],
    "b200tfhe_multibit_bsk_upload_u64": [_VP, C.POINTER(Params), _U32, _VP, _PP],
    "b200tfhe_multibit_bsk_convert_dev": [_VP, C.POINTER(Params), _U32, _VP, _PP],
    "b200tfhe_multibit_bsk_alloc": [_VP, C.POINTER(Params), _U32, _PP],
    "b200tfhe_bsk_grouping_factor": [_VP],
    "b200tfhe_bsk_upload_fourier": [_VP, C.POINTER(Params), _U32, _VP, _PP],
    "b200tfhe_multibit_params_supported": [C.POINTER(Params), _U32],
    "b200tfhe_bsk_device_buffer": [_VP, _PP, C.POINTER(_SZ)],
    "b200tfhe_bsk_destroy": [_VP],
    "b200tfhe_ksk_upload_u64": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_ksk_upload_u32": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_ksk_alloc": [_VP, C.POINTER(Params), _PP],
    "b200tfhe_ksk_device_buffer": [_VP, _PP, C.POINTER(_SZ)],
    "b200tfhe_ksk_finalize": [_VP, _VP],
    "b200tfhe_ksk_destroy": [_VP],
    "b200tfhe_gate_batch_u32": [_VP, _VP, _VP, _I, _I, _VP, _VP, _VP, _U32],
    "b200tfhe_keyswitch_batch_dev": [_VP, _VP, _VP, _VP, _U32],
    "b200tfhe_keyswitch_batch_noalloc_dev": [_VP, _VP, _VP, _VP, _U32],
    "b200tfhe_pbs_batch_dev": [_VP, _VP, _VP, _VP, _VP, _VP, _U32],
    "b200tfhe_blind_rotate_batch_dev": [_VP, _VP, _VP, _VP, _VP, _VP, _U32],
    "b200tfhe_ks_pbs_batch_dev": [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _U32],
    "b200tfhe_pbs_ks_batch_dev": [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _U32],
    "b200tfhe_launch_count": [_VP],
    "b200tfhe_probe_fp64_tflops": [_VP, C.c_double, C.POINTER(C.c_double)],
    "b200tfhe_pool_create": [_VP, C.POINTER(Params), _I, _U32, _PP],
    "b200tfhe_pool_destroy": [_VP],
    "b200tfhe_pool_write": [_VP, _VP, _U32, _U32, _VP],
    "b200tfhe_pool_read": [_VP, _VP, _U32, _U32, _VP],
    "b200tfhe_pool_device_buffer": [_VP, _PP, C.POINTER(_SZ)],
    "b200tfhe_pool_set_luts": [_VP, _VP, _VP, _U32],
    "b200tfhe_pksk_upload_u64": [_VP, C.POINTER(Params), _VP, _PP],
    "b200tfhe_pksk_destroy": [_VP],
    "b200tfhe_packing_keyswitch_batch_u64": [_VP, _VP, _VP, _VP, _U32],
    "b200tfhe_glwe_mul_plain_extract_batch_u64": [_VP, C.POINTER(Params), _VP, _VP, _U32, _VP, _U32],
    "b200tfhe_mvb_batch_u64": [_VP, _VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32],
    "b200tfhe_pack_into_accumulator_batch_u64": [_VP, _VP, _VP, _U32, _U32, _VP, _U32],
    "b200tfhe_extract_bits_batch_u64": [_VP, _VP, _VP, _VP, _U32, _U32, _VP, _U32],
    "b200tfhe_cbs_pfpksk_upload_u64": [_VP, C.POINTER(Params), _U32, _U32, _VP, _PP],
    "b200tfhe_cbs_pfpksk_destroy": [_VP],
    "b200tfhe_pfpks_batch_u64": [_VP, _VP, _U32, _VP, _VP, _U32],
    "b200tfhe_cbs_vertical_packing_batch_u64": [_VP, _VP, _VP, _VP, _U32, _VP, _U32, _U32, _U32, _U32, _VP, _U32],
    "b200tfhe_wave_run": [_VP, _VP, _VP, _VP, _VP, _U32, _U32],
    "b200tfhe_wave_sequence_run": [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _U32],
}
_PROTOS.update({
    "b200tfhe_multi_context_create": [_I, _PP],
    "b200tfhe_multi_context_destroy": [_VP],
    "b200tfhe_multi_context_gpus": [_VP],
    "b200tfhe_multi_context_get": [_VP, _I],
    "b200tfhe_multi_last_error": [_VP],
    "b200tfhe_multi_keys_upload_u64": [_VP, C.POINTER(Params), _VP, _VP, _PP],
    "b200tfhe_multi_keys_upload_u32": [_VP, C.POINTER(Params), _VP, _VP, _PP],
    "b200tfhe_multi_keys_destroy": [_VP],
})
for _s in ("u64", "u32"):
    _PROTOS[f"b200tfhe_multi_keyswitch_batch_{_s}"] = [_VP, _VP, _VP, _VP, _U32]
    _PROTOS[f"b200tfhe_multi_pbs_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_multi_ks_pbs_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_multi_pbs_ks_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_keyswitch_batch_{_s}"] = [_VP, _VP, _VP, _VP, _U32]
    _PROTOS[f"b200tfhe_pbs_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_blind_rotate_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_ks_pbs_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_pbs_ks_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _VP, _U32, _VP, _U32]
    _PROTOS[f"b200tfhe_external_product_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32]
    _PROTOS[f"b200tfhe_cmux_batch_{_s}"] = [_VP, _VP, _VP, _VP, _VP, _U32]
    _PROTOS[f"b200tfhe_modulus_switch_{_s}"] = [_VP, _VP, _VP, _SZ, _U32]
    _PROTOS[f"b200tfhe_sample_extract_batch_{_s}"] = [_VP, _VP, _VP, _U32, _U32, _U32]
_RESTYPES = {"b200tfhe_last_error": C.c_char_p, "b200tfhe_multi_last_error": C.c_char_p, "b200tfhe_multi_context_get": C.c_void_p, "b200tfhe_context_stream": C.c_void_p, "b200tfhe_launch_count": C.c_uint64,
             "b200tfhe_bsk_grouping_factor": C.c_uint32}

EXPORTED_SYMBOLS = tuple(sorted(_PROTOS))


def lib():
    """Load libb200tfhe.so; raises NativeLibraryMissing (never falls back to a CPU path)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryMissing(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                f"or `make -C tfhe-rs-odd_b200/csrc` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, args in _PROTOS.items():
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = _RESTYPES.get(name, C.c_int)
        _lib = L
    return _lib
