"""Multi-GPU plumbing for the path: independent ciphertexts shard statically over ranks, keys are broadcast once.

The reference's only parallelism is rayon over independent ciphertexts (tfhe/benches/core_crypto/pbs_bench.rs:517-532,
tfhe/src/integer/server_key/radix_parallel/mul.rs:355-404).  Here: one process per GPU (torch.distributed), a contiguous
static partition of the batch (remainder to the low ranks), NO per-batch collective; the only communication is a
one-time broadcast of the keys (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import ctypes as C

import numpy as np


def shard_range(batch, world_size, rank):
    """[begin, end) of the ciphertext indices owned by `rank`: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    base, rem = divmod(int(batch), int(world_size))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard(array, world_size, rank):
    b, e = shard_range(len(array), world_size, rank)
    return array[b:e]


def broadcast_key_arrays(arrays, src=0, group=None, device=None):
    """Broadcast standard-domain key arrays (numpy, uint64/uint32) from rank `src`; every rank passes arrays of the right
    shape/dtype (contents ignored on non-source ranks).  Returns the received numpy arrays."""
    import torch
    import torch.distributed as dist
    out = []
    for a in arrays:
        signed = {np.dtype(np.uint64): np.int64, np.dtype(np.uint32): np.int32}[np.dtype(a.dtype)]
        t = torch.from_numpy(np.ascontiguousarray(a).view(signed).copy())
        if device is not None:
            t = t.to(device)
        dist.broadcast(t, src, group=group)
        out.append(t.cpu().numpy().view(a.dtype))
    return out


def gather_shards(local_out, batch, group=None):
    """All ranks end with the full [batch][...] result, assembled in index order (host-side gather)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    parts = [None] * world
    dist.all_gather_object(parts, np.ascontiguousarray(local_out), group=group)
    full = np.concatenate(parts, axis=0)
    assert full.shape[0] == batch
    return full


class MultiGpu:
    """Several GPUs of one node behind one handle (b200tfhe_multi_*): ONE host batch is split over the devices inside the library
    (one host thread + stream per GPU, results at the caller's indices); keys are converted on device 0 and copied device-to-device.
    The in-process alternative to one-process-per-GPU sharding for a caller that holds one Vec<Ciphertext>."""

    def __init__(self, n_gpus=0):
        from . import _native
        from .core_crypto import B200TfheError
        self._lib, self._err = _native.lib(), B200TfheError
        h = C.c_void_p()
        if self._lib.b200tfhe_multi_context_create(int(n_gpus), C.byref(h)):
            raise B200TfheError(self._lib.b200tfhe_multi_last_error(None).decode())
        self._h, self._keys, self.params = h, None, None
        self.n_gpus = int(self._lib.b200tfhe_multi_context_gpus(h))

    def _check(self, rc):
        if rc:
            raise self._err(self._lib.b200tfhe_multi_last_error(self._h).decode())

    def upload_keys(self, params, std_bsk=None, ksk=None):
        dt = np.uint64 if params.torus_bits == 64 else np.uint32
        b = None if std_bsk is None else np.ascontiguousarray(std_bsk, dtype=dt).reshape(-1)
        k = None if ksk is None else np.ascontiguousarray(ksk, dtype=dt).reshape(-1)
        h = C.c_void_p()
        fn = self._lib.b200tfhe_multi_keys_upload_u64 if params.torus_bits == 64 else self._lib.b200tfhe_multi_keys_upload_u32
        self._check(fn(self._h, C.byref(params), b.ctypes.data_as(C.c_void_p) if b is not None else None,
                       k.ctypes.data_as(C.c_void_p) if k is not None else None, C.byref(h)))
        if self._keys:
            self._lib.b200tfhe_multi_keys_destroy(self._keys)
        self._keys, self.params = h, params

    def _run(self, name, cts, in_len, out_len, luts=None, lut_idx=None, out=None):
        p = self.params
        dt = np.uint64 if p.torus_bits == 64 else np.uint32
        a = np.ascontiguousarray(cts, dtype=dt).reshape(-1, in_len)
        if out is None:
            out = np.empty((a.shape[0], out_len), dt)
        fn = getattr(self._lib, f"b200tfhe_multi_{name}_batch_{'u64' if p.torus_bits == 64 else 'u32'}")
        vp = lambda x: x.ctypes.data_as(C.c_void_p) if x is not None else None
        if luts is None:
            self._check(fn(self._h, self._keys, vp(a), vp(out), a.shape[0]))
        else:
            l = np.ascontiguousarray(luts, dtype=dt).reshape(-1, (p.glwe_dimension + 1) * p.polynomial_size)
            li = None if lut_idx is None else np.ascontiguousarray(lut_idx, dtype=np.uint32)
            self._check(fn(self._h, self._keys, vp(a), vp(out), vp(l), l.shape[0], vp(li), a.shape[0]))
        return out

    def keyswitch(self, cts, out=None):
        p = self.params
        return self._run("keyswitch", cts, p.glwe_dimension * p.polynomial_size + 1, p.lwe_dimension + 1, out=out)

    def keyswitch_programmable_bootstrap(self, cts, luts, lut_idx=None, out=None):
        w = self.params.glwe_dimension * self.params.polynomial_size + 1
        return self._run("ks_pbs", cts, w, w, luts, lut_idx, out)

    def programmable_bootstrap_keyswitch(self, cts, luts, lut_idx=None, out=None):
        w = self.params.lwe_dimension + 1
        return self._run("pbs_ks", cts, w, w, luts, lut_idx, out)

    def programmable_bootstrap(self, cts, luts, lut_idx=None, out=None):
        p = self.params
        return self._run("pbs", cts, p.lwe_dimension + 1, p.glwe_dimension * p.polynomial_size + 1, luts, lut_idx, out)

    def close(self):
        if getattr(self, "_keys", None):
            self._lib.b200tfhe_multi_keys_destroy(self._keys)
            self._keys = None
        if getattr(self, "_h", None):
            self._lib.b200tfhe_multi_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
