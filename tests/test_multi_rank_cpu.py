"""World-size-2 gloo test (CPU) of the multi-GPU host logic: static sharding + one-time key broadcast + index-ordered
gather.  The per-rank compute is the oracle here (no GPU in this container): the test checks the plumbing, i.e. that the
sharded result equals the single-process result bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.oracle import Oracle
    from tfhe_rs_odd_b200 import parallel
    o = Oracle("TOY_U64", nthreads=1).keygen()
    if rank != 0:                       # non-source ranks start from garbage keys and must receive rank 0's
        o.ksk = np.zeros_like(o.ksk); o.bsk = np.ones_like(o.bsk)
    o.ksk, o.bsk = parallel.broadcast_key_arrays([o.ksk, o.bsk], src=0)
    o.keygen()                           # keys are cached by name; restore and re-derive the Fourier key
    B = 11
    msgs = np.arange(B) % o.pmod
    cts = o.encrypt(o.encode(msgs), "big")
    b, e = parallel.shard_range(B, world, rank)
    local = o.ks_pbs(cts[b:e], o.lut(lambda x: x))
    full = parallel.gather_shards(local, B)
    if rank == 0:
        ref = o.ks_pbs(cts, o.lut(lambda x: x))
        ret["equal"] = bool(np.array_equal(full, ref))
        ret["dec"] = bool(np.array_equal(o.decode(o.phase(full, "big")) % o.pmod, msgs))
    dist.destroy_process_group()


def test_shard_range_partitions():
    from tfhe_rs_odd_b200.parallel import shard_range
    for batch in [0, 1, 7, 8, 16384, 16385]:
        for world in [1, 2, 4, 8]:
            spans = [shard_range(batch, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_two_rank_gloo_sharded_ks_pbs_equals_single_process():
    import random
    port = 29600 + random.randint(0, 300)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret["equal"] and ret["dec"]
