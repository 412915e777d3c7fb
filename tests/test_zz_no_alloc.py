"""b200tfhe_keyswitch_batch_noalloc_dev: the allocation-free keyswitch (digit scratch in a fixed __device__ array) against the
ordinary b200tfhe_keyswitch_batch_dev and the oracle, bit for bit, on two calls in a row (the shared array keeps the first
call's contents).  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from conftest import get_oracle, params_of

pytestmark = pytest.mark.gpu


def _dev_call(ctx, fn, ksk, cts, n_out):
    """Run one *_dev keyswitch entry point on device buffers (torch only owns the memory)."""
    import torch
    d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    d_out = torch.full((cts.shape[0], n_out), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx.check(fn(ctx._h, ksk._h, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), cts.shape[0]))
    ctx.synchronize()
    return d_out.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("name", ["MESSAGE_2_CARRY_2", "TOY_KSW10"])
def test_keyswitch_noalloc_matches_ordinary_entry_point_twice(ctx, name):
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    ksk = t.LweKeyswitchKey.upload(ctx, params_of(o), o.ksk.astype(np.uint64))
    lib = ctx._lib
    rng = np.random.default_rng(0xA110C)
    # first call: a ragged batch (vs the 128-row tile) with boundary rows; second call: a SMALLER batch of other inputs, so
    # stale digits of the first call sit right behind the rows the second one reads
    for B in (150, 37):
        cts = rng.integers(0, 2 ** 64, size=(B, o.big + 1), dtype=np.uint64)
        cts[0, :] = 0
        cts[1, :] = np.uint64(2 ** 64 - 1)
        cts[2, :] = np.uint64(1 << (64 - o.b_ks * o.l_ks - 1))   # rounding boundary of the decomposition
        ref_dev = _dev_call(ctx, lib.b200tfhe_keyswitch_batch_dev, ksk, cts, o.n + 1)
        got = _dev_call(ctx, lib.b200tfhe_keyswitch_batch_noalloc_dev, ksk, cts, o.n + 1)
        assert np.array_equal(got, ref_dev), f"noalloc keyswitch differs from keyswitch_batch_dev (batch {B})"
        assert np.array_equal(got, o.keyswitch(cts)), f"noalloc keyswitch differs from the oracle (batch {B})"


def test_keyswitch_noalloc_rejects_batches_beyond_the_fixed_scratch(ctx):
    import torch
    import tfhe_rs_odd_b200 as t
    o = get_oracle("MESSAGE_2_CARRY_2")
    ksk = t.LweKeyswitchKey.upload(ctx, params_of(o), o.ksk.astype(np.uint64))
    kp = o.big * o.l_ks                                   # 10240 digit bytes per ciphertext (already a multiple of 16)
    too_many = (4 << 20) // kp + 1
    d_in = torch.zeros((too_many, o.big + 1), dtype=torch.int64, device="cuda")
    d_out = torch.zeros((too_many, o.n + 1), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = ctx._lib.b200tfhe_keyswitch_batch_noalloc_dev(ctx._h, ksk._h, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), too_many)
    assert rc != 0
    msg = ctx._lib.b200tfhe_last_error(ctx._h).decode()
    assert "fixed device array" in msg and str(too_many - 1) in msg, msg
    # the largest admissible batch still runs
    rc = ctx._lib.b200tfhe_keyswitch_batch_noalloc_dev(ctx._h, ksk._h, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), too_many - 1)
    assert rc == 0, ctx._lib.b200tfhe_last_error(ctx._h).decode()
    ctx.synchronize()
