"""Multi-GPU plumbing for the path: independent ciphertexts shard statically over ranks, keys are broadcast once.

The reference's only parallelism is rayon over independent ciphertexts (tfhe/benches/core_crypto/pbs_bench.rs:517-532,
tfhe/src/integer/server_key/radix_parallel/mul.rs:355-404).  Here: one process per GPU (torch.distributed), a contiguous
static partition of the batch (remainder to the low ranks), NO per-batch collective; the only communication is a
one-time broadcast of the keys (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
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
