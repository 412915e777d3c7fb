"""Multi-bit PBS (reference algorithms/lwe_multi_bit_programmable_bootstrapping.rs).

Acceptance criterion of the reference's own test (core_crypto/algorithms/test/lwe_multi_bit_programmable_bootstrapping.rs:
every message encrypted -> multi-bit PBS with an accumulator -> decrypts to f(message)), applied to
  * the oracle restatement (CPU),
  * the CUDA stage functions executed lane by lane on the CPU (tests/native/emul.cpp, same headers as the kernel),
  * the CUDA kernels through the C ABI (-m gpu), together with decrypted / phase parity against the oracle.
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT, get_oracle, params_of, phase_bound
from oracle.oracle import _p

TOY = ["TOY_MB_G2", "TOY_MB_G3", "TOY_MB22_G2", "TOY_MB22_G3"]


def _f(o):
    return lambda x: (3 * x + 1) % o.pmod


def _phase_diff(a, b):
    return np.abs((a - b).astype(np.int64).astype(np.float64)) / 2.0 ** 64


@pytest.mark.parametrize("name", TOY)
def test_oracle_multibit_pbs_decrypts_every_message(name):
    o = get_oracle(name)
    msgs = np.arange(2 * o.pmod) % o.pmod
    cts = o.encrypt(o.encode(msgs), under="small", std=1e-9)
    out = o.pbs(cts, o.lut(_f(o)))
    assert np.array_equal(o.decode(o.phase(out)), np.array([_f(o)(int(m)) for m in msgs], np.uint64))
    big = o.encrypt(o.encode(msgs), under="big")
    out = o.ks_pbs(big, o.lut(lambda x: x))
    assert np.array_equal(o.decode(o.phase(out)), msgs.astype(np.uint64))


def test_oracle_multibit_key_layout():
    """GGSW idx of a group encrypts prod_b (s_b XOR NOT bit_{G-1-b}(idx)) (combine_key_bits,
    lwe_multi_bit_bootstrap_key_generation.rs:401-427): exactly one GGSW per group encrypts 1, the others 0, and its
    index spells the group's key bits, first key element in the most significant position."""
    o = get_oracle("TOY_MB_G3")
    G, NG, N, k1, l = 3, 8, o.N, o.k + 1, o.l
    bsk = o.bsk.reshape(o.n // G, NG, l, k1, k1, N)
    for gi in range(o.n // G):
        bits = o.small_key[gi * G:(gi + 1) * G]
        want = int(bits[0]) * 4 + int(bits[1]) * 2 + int(bits[2])
        for idx in range(NG):
            # last row of level 1 is a GLWE encryption of m * q/B (ggsw_encryption.rs:300-331): decode its constant term
            row = bsk[gi, idx, 0, k1 - 1].reshape(-1)
            ph = int(o.glwe_phase(row)[0])
            m = ((ph + (1 << (63 - o.b))) >> (64 - o.b)) & 1
            assert m == (1 if idx == want else 0)


@pytest.fixture(scope="module")
def emul():
    import __graft_entry__ as g
    g.build()
    return C.CDLL(os.path.join(ROOT, "tests", "native", "libemul.so"))


@pytest.mark.parametrize("name", TOY)
def test_multibit_kernel_stages_on_cpu_match_oracle(emul, name):
    o = get_oracle(name)
    msgs = np.arange(o.pmod)
    cts = o.encrypt(o.encode(msgs), under="small", std=1e-9)
    lut = o.lut(_f(o))
    ref = o.pbs(cts, lut, sample_extract=False)
    out = np.zeros_like(ref)
    assert emul.emul_multibit_blind_rotate(o.n, o.k, o.N, o.b, o.l, o.bits, o.grouping, _p(o.bsk), _p(cts), len(msgs), _p(lut), _p(out)) == 0
    ph_ref = np.array([o.glwe_phase(g)[0] for g in ref])
    ph_out = np.array([o.glwe_phase(g)[0] for g in out])
    assert np.array_equal(o.decode(ph_out), np.array([_f(o)(int(m)) for m in msgs], np.uint64))
    assert np.array_equal(o.decode(ph_out), o.decode(ph_ref))
    assert _phase_diff(ph_out, ph_ref).max() < mb_phase_bound(o)


def mb_phase_bound(o):
    """Same law as conftest.phase_bound with n/G external products whose input is the full accumulator."""
    return phase_bound(o)


def test_multibit_host_argument_checks():
    """Python mirror rejects what the reference asserts (grouping factor must divide the LWE dimension; key size)."""
    import tfhe_rs_odd_b200 as t
    p = t.core_crypto.make_params(12, 1, 512, 12, 2, 4, 4, 64)

    class Ctx:
        pass
    with pytest.raises(t.B200TfheError, match="grouping factor"):
        t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(Ctx(), p, np.zeros(8, np.uint64), 5)
    with pytest.raises(t.B200TfheError, match="scalars"):
        t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(Ctx(), p, np.zeros(8, np.uint64), 2)
    p32 = t.core_crypto.make_params(12, 2, 512, 6, 3, 3, 4, 32)
    with pytest.raises(t.B200TfheError, match="64-bit"):
        t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(Ctx(), p32, np.zeros(8, np.uint32), 2)


# ---- GPU ---------------------------------------------------------------------------------------------------------------
def _keys(ctx, o):
    import tfhe_rs_odd_b200 as t
    p = params_of(o)
    bsk = t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, p, o.bsk, o.grouping)
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk)
    return p, bsk, ksk


@pytest.mark.gpu
@pytest.mark.parametrize("name,B", [("TOY_MB_G2", 9), ("TOY_MB_G3", 9), ("TOY_MB22_G2", 5), ("TOY_MB22_G3", 5),
                                    ("TOY_MB22_G2", 450), ("TOY_MB22_G3", 446)])
def test_gpu_multibit_pbs_matches_oracle(ctx, name, B):
    """B < 444 runs the one-ciphertext-per-CTA instance, B >= 444 the three-per-CTA one (ragged last CTA)."""
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    p, bsk, ksk = _keys(ctx, o)
    assert bsk.grouping_factor == o.grouping
    rng = np.random.default_rng(5)
    msgs = rng.integers(0, o.pmod, size=B)
    msgs[:min(B, o.pmod)] = np.arange(min(B, o.pmod))
    cts = o.encrypt(o.encode(msgs), under="small", std=1e-9)
    luts = np.stack([o.lut(_f(o)), o.lut(lambda x: x)])
    idx = (np.arange(B) % 2).astype(np.uint32)
    want = np.where(idx == 0, (3 * msgs + 1) % o.pmod, msgs).astype(np.uint64)
    got = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts, luts, bsk, lut_idx=idx)
    assert np.array_equal(o.decode(o.phase(got)), want)
    nref = min(B, 24)                                # the oracle is slow at N = 2048: check a prefix against it
    ref = o.pbs(cts[:nref], luts, idx[:nref])
    assert np.array_equal(o.decode(o.phase(got[:nref])), o.decode(o.phase(ref)))
    assert _phase_diff(o.phase(got[:nref]), o.phase(ref)).max() < mb_phase_bound(o)
    # GLWE output (blind rotation without sample extract) is the same ciphertext before extraction
    glwe = t.programmable_bootstrap_lwe_ciphertext_without_sample_extract(cts[:nref], luts, bsk, lut_idx=idx[:nref])
    ext = t.extract_lwe_sample_from_glwe_ciphertext(ctx, glwe, o.k, o.N, 64)
    # (a batch this small takes the two-kernel latency path where it is built, so compare with a bootstrap of the same batch)
    same = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts[:nref], luts, bsk, lut_idx=idx[:nref])
    assert np.array_equal(ext, same)
    assert _phase_diff(o.phase(same), o.phase(got[:nref])).max() < mb_phase_bound(o)
    # single-ciphertext form, empty batch
    one = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts[1], luts[0], bsk)
    assert o.decode(o.phase(one))[0] == _f(o)(int(msgs[1]))
    assert t.programmable_bootstrap_lwe_ciphertext(np.zeros((0, o.n + 1), np.uint64), luts, bsk).shape == (0, o.big + 1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["MULTI_BIT_2_2_GROUP_2", "MULTI_BIT_2_2_GROUP_3"])
def test_gpu_multibit_real_parameters_ks_pbs(ctx, name):
    """PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_{2,3}_KS_PBS through shortint ServerKey::apply_lookup_table with a
    multi-bit bootstrapping key: every message, real noise."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import shortint
    o = get_oracle(name)
    p, bsk, ksk = _keys(ctx, o)
    sk = shortint.ServerKey(ksk, bsk, 4, 4, "ks_pbs")
    B = 64
    msgs = np.arange(B) % 16
    cts = shortint.CiphertextList(o.encrypt(o.encode(msgs), under="big"), degree=15, noise_level=shortint.NOISE_NOMINAL)
    lut = sk.generate_lookup_table(lambda x: (x * x + 1) % 16)
    out = sk.apply_lookup_table(cts, lut)
    assert np.array_equal(o.decode(o.phase(out.data)), ((msgs * msgs + 1) % 16).astype(np.uint64))
    ref = o.ks_pbs(cts.data[:4], lut.acc)
    assert np.array_equal(o.decode(o.phase(ref)), o.decode(o.phase(out.data[:4])))
    assert _phase_diff(o.phase(out.data[:4]), o.phase(ref)).max() < mb_phase_bound(o)
    # a large batch on the three-per-CTA kernel: decrypt-checked against the plain function
    B = 1000
    msgs = np.random.default_rng(1).integers(0, 16, size=B)
    big = shortint.CiphertextList(o.encrypt(o.encode(msgs), under="big"), degree=15, noise_level=shortint.NOISE_NOMINAL)
    out = sk.apply_lookup_table(big, lut)
    assert np.array_equal(o.decode(o.phase(out.data)), ((msgs * msgs + 1) % 16).astype(np.uint64))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["TOY_MB22_G2", "TOY_MB22_G3", "MULTI_BIT_2_2_GROUP_2", "MULTI_BIT_2_2_GROUP_3"])
def test_gpu_multibit_latency_path_matches_fused_kernel(ctx, name):
    """Batches of at most `mb_lat_max` ciphertexts run mb_key_sum_kernel + the single-ciphertext kernel (n/G plain external
    products against per-ciphertext summed keys); larger ones the fused kernel.  Both compute the reference's
    GGSW_sum (.) acc per group and may differ only by f64 summation order."""
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    p, bsk, ksk = _keys(ctx, o)
    B = 7
    msgs = np.arange(B) % o.pmod
    cts = o.encrypt(o.encode(msgs), under="small", std=1e-9 if name.startswith("TOY") else None)
    luts = np.stack([o.lut(_f(o)), o.lut(lambda x: x)])
    idx = (np.arange(B) % 2).astype(np.uint32)
    want = np.where(idx == 0, (3 * msgs + 1) % o.pmod, msgs).astype(np.uint64)
    n0 = ctx.launch_count
    lat = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts, luts, bsk, lut_idx=idx)
    assert ctx.launch_count - n0 == 2                       # key sums, then the external products
    glwe_lat = t.programmable_bootstrap_lwe_ciphertext_without_sample_extract(cts, luts, bsk, lut_idx=idx)
    ctx.set_option("mb_lat_max", 0)
    try:
        n0 = ctx.launch_count
        fused = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts, luts, bsk, lut_idx=idx)
        assert ctx.launch_count - n0 == 1
    finally:
        ctx.set_option("mb_lat_max", 148)
    assert np.array_equal(o.decode(o.phase(lat)), want)
    assert np.array_equal(o.decode(o.phase(fused)), want)
    assert np.array_equal(t.extract_lwe_sample_from_glwe_ciphertext(ctx, glwe_lat, o.k, o.N, 64), lat)
    # same digits almost everywhere: the two outputs differ by f64 rounding unless a decomposition boundary was crossed
    assert _phase_diff(o.phase(lat), o.phase(fused)).max() < mb_phase_bound(o)
    nref = 2 if not name.startswith("TOY") else B
    ref = o.pbs(cts[:nref], luts, idx[:nref])
    assert np.array_equal(o.decode(o.phase(lat[:nref])), o.decode(o.phase(ref)))
    assert _phase_diff(o.phase(lat[:nref]), o.phase(ref)).max() < mb_phase_bound(o)


@pytest.mark.gpu
def test_gpu_multibit_rejects_bad_arguments(ctx):
    import tfhe_rs_odd_b200 as t
    o = get_oracle("TOY_MB_G2")
    p = params_of(o)
    with pytest.raises(t.B200TfheError, match="grouping"):
        t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, p, np.zeros(o.bsk.size * 8, np.uint64), 5)
    # a set the multi-bit kernel is not built for
    q = t.core_crypto.make_params(12, 3, 512, 18, 1, 4, 3, 64)
    with pytest.raises(t.B200TfheError, match="not built"):
        t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, q, np.zeros((12 // 2 * 4) * 16 * 512, np.uint64), 2)
    classic = get_oracle("TOY_U64")
    cb = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, params_of(classic), classic.bsk)
    assert cb.grouping_factor == 1
    with pytest.raises(t.B200TfheError, match="multi-bit"):
        t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(np.zeros(classic.n + 1, np.uint64), classic.lut(lambda x: x), cb)


@pytest.mark.gpu
def test_gpu_multibit_latency_path_boundary(ctx):
    """mb_lat_max is the largest batch on the two-kernel path: one ciphertext more runs the fused kernel; both sides decrypt, and the
    option can move the boundary."""
    import tfhe_rs_odd_b200 as t
    o = get_oracle("TOY_MB22_G3")
    p, bsk, ksk = _keys(ctx, o)
    B = 149
    msgs = np.arange(B) % o.pmod
    cts = o.encrypt(o.encode(msgs), under="small", std=1e-9)
    lut = o.lut(_f(o))
    want = np.array([_f(o)(int(m)) for m in msgs], np.uint64)
    for batch, launches in ((148, 2), (149, 1)):
        n0 = ctx.launch_count
        got = t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts[:batch], lut, bsk)
        assert ctx.launch_count - n0 == launches, batch
        assert np.array_equal(o.decode(o.phase(got)), want[:batch])
    ctx.set_option("mb_lat_max", 10)
    try:
        n0 = ctx.launch_count
        t.core_crypto.multi_bit_programmable_bootstrap_lwe_ciphertext(cts[:11], lut, bsk)
        assert ctx.launch_count - n0 == 1
    finally:
        ctx.set_option("mb_lat_max", 148)
