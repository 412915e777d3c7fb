"""b200tfhe: B200-native batched keyswitch + programmable bootstrap for TFHE (sm_100a only, no CPU fallback).

Layout: csrc/ (CUDA kernels + C ABI -> libb200tfhe.so), _native.py (ctypes binding), core_crypto.py /
shortint.py / boolean.py (host-side mirrors of the reference's interfaces for this path), parameters.py.
The directory name carries a hyphen (repo convention); import it as `tfhe_rs_odd_b200` via the shim at the
repository root.
"""
from . import _native  # noqa: F401
from ._native import Params, NativeLibraryMissing, LIB_PATH  # noqa: F401
from .core_crypto import *  # noqa: F401,F403
from .core_crypto import B200TfheError, Context  # noqa: F401
from . import parameters  # noqa: F401
