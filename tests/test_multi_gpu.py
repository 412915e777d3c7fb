"""Several GPUs behind one library handle (b200tfhe_multi_*): one host batch is split over every visible GPU inside the library;
results must be bit-identical to the single-GPU call (same kernels, same operation order per ciphertext, only the batch is cut)
and decrypt like the oracle.  With one visible GPU the same code path runs with a single shard."""
import numpy as np
import pytest

from conftest import get_oracle, params_of

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["TOY_2_2", "TOY_U32"])
def test_multi_gpu_matches_single_gpu_bit_for_bit(ctx, name):
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200.parallel import MultiGpu
    o = get_oracle(name)
    p = params_of(o)
    dt = np.uint64 if o.bits == 64 else np.uint32
    mg = MultiGpu(0)                                   # every visible device
    assert mg.n_gpus >= 1
    mg.upload_keys(p, o.bsk.astype(dt), o.ksk.astype(dt))
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk.astype(dt))
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk.astype(dt))
    B = 37                                             # ragged against any GPU count
    if o.order == "ks_pbs":
        msgs = np.arange(B) % o.pmod
        cts = o.encrypt(o.encode(msgs), "big", seed=0x3617)
        luts = np.stack([o.lut(lambda x: x), o.lut(lambda x: (3 * x + 1) % o.pmod)])
        idx = (np.arange(B) % 2).astype(np.uint32)
        got = mg.keyswitch_programmable_bootstrap(cts, luts, idx)
        one = t.keyswitch_programmable_bootstrap(ksk, bsk, cts, luts, idx)
        assert np.array_equal(got, one)
        want = np.where(idx == 0, msgs, (3 * msgs + 1) % o.pmod)
        assert np.array_equal(o.decode(o.phase(got, "big")) % o.pmod, want)
        ks = mg.keyswitch(cts)
        assert np.array_equal(ks, o.keyswitch(cts))    # integer stage: bit-exact against the oracle too
    else:
        bits = (np.arange(B) % 2).astype(bool)
        cts = o.encrypt(np.where(bits, 1 << 29, 7 << 29).astype(np.uint64), "small", seed=0x3618).astype(np.uint32)
        lut = o.boolean_lut().astype(np.uint32)
        got = mg.programmable_bootstrap_keyswitch(cts, lut)
        one = t.programmable_bootstrap_keyswitch(ksk, bsk, cts, lut)
        assert np.array_equal(got, one)
        assert np.array_equal(o.decode_bool(o.phase(got.astype(np.uint64), "small")), bits)
    # a batch smaller than the GPU count and an empty batch
    small = cts[:1]
    if o.order == "ks_pbs":
        assert np.array_equal(mg.keyswitch_programmable_bootstrap(small, luts[0]), t.keyswitch_programmable_bootstrap(ksk, bsk, small, luts[0]))
        assert mg.keyswitch_programmable_bootstrap(cts[:0], luts[0]).shape[0] == 0
    mg.close()


def test_multi_gpu_argument_errors():
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200.parallel import MultiGpu
    with pytest.raises(t.B200TfheError, match="visible"):
        MultiGpu(4096)
