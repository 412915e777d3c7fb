"""Bit extraction, the first half of the WoP-PBS (reference core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs:66-225,
test it mirrors: wop_pbs/tests.rs:101-268 -- 5 bits including padding, every bit extracted, each decrypted with a 1-bit decomposer).
The reference's own toy set (N = 1024) has no team-FFT instance here (N = 2 R^2 only), so the test runs on the sets of the path."""
import numpy as np
import pytest

from conftest import get_oracle, params_of, phase_bound

U64 = 1 << 64


def _bits_msb_first(msgs, nbits):
    return np.array([[(int(m) >> (nbits - 1 - i)) & 1 for i in range(nbits)] for m in msgs])


def _diff_bound(o, nbits):
    """|phase_gpu - phase_oracle| of an extracted ciphertext after the first round: the running buffer holds the outputs of
    earlier bootstraps (two f64 FFT orders -> independent rounding draws, conftest.phase_bound), shifted left by up to nbits - 1
    bits, and the keyswitch of two different buffers draws its decomposition rounding (kN terms of 2^-(b_ks l_ks)) twice."""
    ks_var = o.k * o.N * 2.0 ** (-2 * o.b_ks * o.l_ks) / 12.0
    return phase_bound(o) * 2.0 ** nbits + 8.0 * (2.0 * ks_var) ** 0.5


def _decrypt_bits(o, out):
    """phase of every extracted ciphertext under the small key, rounded to one bit (SignedDecomposer(1, 1) of tests.rs:189)."""
    B, nb, _ = out.shape
    ph = o.phase(out.reshape(B * nb, o.n + 1), under="small").reshape(B, nb)
    bits = ((ph.astype(np.uint64) + np.uint64(1 << 62)) >> np.uint64(63)).astype(int) & 1
    return bits, ph


# ---- CPU: the oracle restatement against the definition ------------------------------------------------------------
@pytest.mark.parametrize("name,nbits", [("TOY_U64", 5), ("TOY_2_2", 4)])
def test_oracle_extract_bits_decrypts_every_bit(name, nbits):
    o = get_oracle(name)
    delta_log = 64 - nbits
    msgs = np.arange(1 << nbits, dtype=np.uint64)
    cts = o.encrypt(msgs << np.uint64(delta_log), under="big", seed=41)
    out = o.extract_bits(cts, delta_log, nbits)
    bits, ph = _decrypt_bits(o, out)
    assert np.array_equal(bits, _bits_msb_first(msgs, nbits))
    err = (ph - (bits.astype(np.uint64) << np.uint64(63))).astype(np.int64) / 2.0 ** 64
    assert np.abs(err).max() < 2.0 ** -6


def test_oracle_extract_bits_partial_window_and_first_round():
    """Fewer bits than the message holds: the window starts at delta_log; the first extracted ciphertext (last slot) is the
    keyswitch of the shifted input (mod.rs:160-184) and a single-bit extraction does no bootstrap at all."""
    o = get_oracle("TOY_U64")
    msgs = np.array([0b10110, 0b01101, 0b11111, 0], dtype=np.uint64)
    cts = o.encrypt(msgs << np.uint64(59), under="big", seed=43)
    out = o.extract_bits(cts, 59, 2)                      # bits 59 and 60 = the two low bits of the message
    bits, _ = _decrypt_bits(o, out)
    assert np.array_equal(bits, np.stack([(msgs >> np.uint64(1)) & np.uint64(1), msgs & np.uint64(1)], 1).astype(int))
    shifted = cts << np.uint64(64 - 59 - 1)
    assert np.array_equal(out[:, 1], o.keyswitch(shifted))
    assert np.array_equal(o.extract_bits(cts, 59, 1)[:, 0], o.keyswitch(shifted))


# ---- GPU: through the C ABI ----------------------------------------------------------------------------------------------
def _keys(ctx, o):
    import tfhe_rs_odd_b200 as t
    p = params_of(o)
    if o.grouping:
        bsk = t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, p, o.bsk, o.grouping)
    else:
        bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
    return t, bsk, t.LweKeyswitchKey.upload(ctx, p, o.ksk)


@pytest.mark.gpu
@pytest.mark.parametrize("name,nbits,B", [("TOY_U64", 5, 64), ("TOY_1_1", 3, 16), ("TOY_2_2", 4, 600), ("TOY_3_3", 6, 8)])
def test_gpu_extract_bits_matches_oracle(ctx, name, nbits, B):
    o = get_oracle(name)
    t, bsk, ksk = _keys(ctx, o)
    delta_log = 64 - nbits
    msgs = (np.arange(B, dtype=np.uint64) * np.uint64(7) + np.uint64(3)) % np.uint64(1 << nbits)
    cts = o.encrypt(msgs << np.uint64(delta_log), under="big", seed=47)
    got = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, delta_log, nbits)
    assert got.shape == (B, nbits, o.n + 1)
    bits, ph = _decrypt_bits(o, got)
    assert np.array_equal(bits, _bits_msb_first(msgs, nbits))                 # every bit of every message
    nref = min(B, 24)
    ref = o.extract_bits(cts[:nref], delta_log, nbits)
    assert np.array_equal(got[:nref, nbits - 1], ref[:, nbits - 1])           # first round: integer only, memcmp-exact
    _, ph_ref = _decrypt_bits(o, ref)
    d = (ph[:nref] - ph_ref).astype(np.int64) / 2.0 ** 64                     # later rounds go through bootstraps (f64 FFT)
    assert np.abs(d).max() < _diff_bound(o, nbits) < 0.125
    # a single ciphertext, and fewer bits than the window: same slots
    one = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts[5], bsk, ksk, delta_log, nbits)
    b1, _ = _decrypt_bits(o, one[None])
    assert np.array_equal(b1[0], _bits_msb_first(msgs[5:6], nbits)[0])
    two = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts[:4], bsk, ksk, delta_log, 2)
    assert np.array_equal(two[:, 1], got[:4, nbits - 1])                      # the lowest bit does not depend on the count
    b2, _ = _decrypt_bits(o, two)
    assert np.array_equal(b2, _bits_msb_first(msgs[:4] & np.uint64(3), 2))


@pytest.mark.gpu
def test_gpu_extract_bits_multibit_key_and_errors(ctx):
    """The key handle carries the grouping factor: the same entry runs on a multi-bit bootstrap key.  Argument validation
    mirrors the reference's assertions (mod.rs:86-118)."""
    o = get_oracle("TOY_MB22_G2")
    t, bsk, ksk = _keys(ctx, o)
    msgs = np.arange(16, dtype=np.uint64)
    cts = o.encrypt(msgs << np.uint64(60), under="big", seed=53)
    got = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 60, 4)
    bits, _ = _decrypt_bits(o, got)
    assert np.array_equal(bits, _bits_msb_first(msgs, 4))
    with pytest.raises(t.B200TfheError, match="exceeds"):
        t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 60, 5)
    with pytest.raises(t.B200TfheError):
        t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts[:, :-1], bsk, ksk, 60, 4)
    assert t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 60, 0).shape == (16, 0, o.n + 1)
    o2 = get_oracle("TOY_U64")
    _, bsk2, _ = _keys(ctx, o2)
    with pytest.raises(t.B200TfheError, match="length"):                       # host mirror: lwe_in size vs the bootstrap key (:107-112)
        t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk2, ksk, 60, 4)
    cts2 = o2.encrypt(np.zeros(2, np.uint64), under="big")
    with pytest.raises(t.B200TfheError, match="dimensions differ"):            # C ABI: ksk output vs bsk input dimension (:113-118)
        t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts2, bsk2, ksk, 60, 4)


@pytest.mark.gpu
def test_gpu_extract_bits_full_size_2_2(ctx):
    """PARAM_MESSAGE_2_CARRY_2 keys at full size: all 5 bits (padding included) of 296 ciphertexts decrypt to the message bits;
    the oracle runs the first 2 ciphertexts (4 bootstraps each)."""
    o = get_oracle("MESSAGE_2_CARRY_2")
    t, bsk, ksk = _keys(ctx, o)
    B, nbits = 296, 5
    msgs = (np.arange(B, dtype=np.uint64) * np.uint64(11)) % np.uint64(32)
    cts = o.encrypt(msgs << np.uint64(59), under="big", seed=59)
    got = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 59, nbits)
    bits, ph = _decrypt_bits(o, got)
    assert np.array_equal(bits, _bits_msb_first(msgs, nbits))
    ref = o.extract_bits(cts[:2], 59, nbits)
    assert np.array_equal(got[:2, nbits - 1], ref[:, nbits - 1])
    _, ph_ref = _decrypt_bits(o, ref)
    assert np.abs((ph[:2] - ph_ref).astype(np.int64) / 2.0 ** 64).max() < _diff_bound(o, nbits) < 0.125


# ---- circuit bootstrap + vertical packing (wop_pbs/mod.rs:243-853) ---------------------------------------------------------
def _wop(name):
    from oracle.oracle import WOPBS
    return WOPBS[name]


_PFPKSK = {}


def _pfpksk(o, name):
    if name not in _PFPKSK:
        pb, pl, _, _ = _wop(name)
        _PFPKSK[name] = o.gen_cbs_pfpksk_list(pb, pl)
    return _PFPKSK[name]


def _luts(o, nbits):
    """two 4-bit -> 4-bit tables as polynomials: coefficient j holds f(j) * 2^60"""
    lut = np.zeros((2, o.N), np.uint64)
    for j in range(1 << nbits):
        lut[0, j] = np.uint64(((j * j + 1) % 16) << 60)
        lut[1, j] = np.uint64(((15 - j) % 16) << 60)
    return lut


def _decode4(o, cts):
    ph = o.phase(cts.reshape(-1, o.big + 1)).reshape(cts.shape[:-1])
    return ((ph.astype(np.uint64) + np.uint64(1 << 59)) >> np.uint64(60)).astype(int) & 15, ph


def _sdiff(a, b):
    """signed torus difference of two 64-bit values as a Python int"""
    d = (int(a) - int(b)) % U64
    return d - U64 if d >= U64 // 2 else d


def test_oracle_pfpks_and_circuit_bootstrap():
    """pfpks with key r decrypts to -phase * S_r (r < k) / +phase (r = k) (lwe_wopbs.rs:128-157); the GGSW of a circuit
    bootstrap decrypts, level by level, to bit * 2^(64 - b*level) on the last row; vertical packing evaluates the tables."""
    o = get_oracle("TOY_WOP")
    wop = _wop("TOY_WOP")
    pf = _pfpksk(o, "TOY_WOP")
    key_len = (o.big + 1) * wop[1] * o.glwe_len
    ct = o.encrypt(np.array([3 << 60], dtype=np.uint64), under="big", seed=61)
    g = o.pfpks(pf[o.k * key_len:(o.k + 1) * key_len], wop[0], wop[1], ct)[0]
    ph = o.glwe_phase(g)
    assert abs(_sdiff(ph[0], 3 << 60)) < 2 ** 40 and np.abs(ph[1:].astype(np.int64)).max() < 2 ** 40
    msgs = np.array([5, 10, 15, 0, 7], dtype=np.uint64)
    bits = o.extract_bits(o.encrypt(msgs << np.uint64(60), under="big", seed=67), 60, 4)
    ggsw = o.circuit_bootstrap_boolean(pf, wop, bits[0, 1], 63)               # bit 1 of 5 = 0b0101 (MSB first) is 1
    for lvl in range(wop[3]):
        ph = o.glwe_phase(ggsw[lvl, o.k])
        assert abs(_sdiff(ph[0], 1 << (64 - wop[2] * (lvl + 1)))) < 2 ** 44
    out = o.cbs_vertical_packing(pf, wop, bits, _luts(o, 4))
    dec, ph = _decode4(o, out)
    assert np.array_equal(dec[:, 0], (msgs * msgs + 1) % 16) and np.array_equal(dec[:, 1], 15 - msgs)
    assert np.abs((ph - (dec.astype(np.uint64) << np.uint64(60))).astype(np.int64)).max() < 2.0 ** 64 / 256


def _wop_keys(ctx, o, name):
    import tfhe_rs_odd_b200 as t
    t_, bsk, ksk = _keys(ctx, o)
    wop = _wop(name)
    pf = t.core_crypto.LwePrivateFunctionalPackingKeyswitchKeyList.upload(ctx, params_of(o), wop[0], wop[1], _pfpksk(o, name))
    return t, bsk, ksk, pf, wop


@pytest.mark.gpu
def test_gpu_pfpks_bit_exact(ctx):
    o = get_oracle("TOY_WOP")
    t, bsk, ksk, pf, wop = _wop_keys(ctx, o, "TOY_WOP")
    rng = np.random.default_rng(71)
    B = 150
    cts = rng.integers(0, U64, size=(B, o.big + 1), dtype=np.uint64)
    cts[0, :] = 0
    cts[1, :] = U64 - 1
    cts[2, :] = 1 << 63
    cts[3, :] = (1 << 63) - 1
    cts[4, :] = np.uint64(1 << 33)                                            # exactly half of the last kept digit: rounding boundary
    key_len = (o.big + 1) * wop[1] * o.glwe_len
    host = _pfpksk(o, "TOY_WOP")
    for r in range(o.k + 1):
        got = t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, r, cts)
        ref = o.pfpks(host[r * key_len:(r + 1) * key_len], wop[0], wop[1], cts[:12])
        assert np.array_equal(got[:12], ref)
        one = t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, r, cts[7])
        assert np.array_equal(one, ref[7])
        assert np.array_equal(t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, r, cts[100:]), got[100:])
    with pytest.raises(t.B200TfheError, match="out of range"):
        t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, o.k + 1, cts)


@pytest.mark.gpu
@pytest.mark.parametrize("name,B", [("TOY_WOP", 16), ("WOPBS_MESSAGE_2_CARRY_2", 8)])
def test_gpu_wopbs_extract_then_circuit_bootstrap_vertical_packing(ctx, name, B):
    """The whole WoP-PBS of a 4-bit message (shortint/wopbs/mod.rs:547-830): extract_bits, then circuit bootstrap + vertical
    packing with two look-up tables.  Every message decrypts to table[message]; the oracle runs the first ciphertexts."""
    o = get_oracle(name)
    t, bsk, ksk, pf, wop = _wop_keys(ctx, o, name)
    nbits = 4
    msgs = (np.arange(B, dtype=np.uint64) * np.uint64(5) + np.uint64(1)) % np.uint64(16)
    cts = o.encrypt(msgs << np.uint64(60), under="big", seed=73)
    bits = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 60, nbits)
    b, _ = _decrypt_bits(o, bits)
    assert np.array_equal(b, _bits_msb_first(msgs, nbits))
    luts = _luts(o, nbits)
    out = t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(bits, luts, bsk, pf, wop[2], wop[3])
    assert out.shape == (B, 2, o.big + 1)
    dec, ph = _decode4(o, out)
    assert np.array_equal(dec[:, 0], (msgs * msgs + 1) % 16) and np.array_equal(dec[:, 1], 15 - msgs)
    err = np.abs((ph - (dec.astype(np.uint64) << np.uint64(60))).astype(np.int64)) / 2.0 ** 64
    nref = 2
    ref = o.cbs_vertical_packing(_pfpksk(o, name), wop, bits[:nref], luts)       # same extracted bits into the oracle
    dref, phr = _decode4(o, ref)
    assert np.array_equal(dref, dec[:nref])
    eref = np.abs((phr - (dref.astype(np.uint64) << np.uint64(60))).astype(np.int64)) / 2.0 ** 64
    assert err.max() < max(8 * eref.max(), 2.0 ** -9) < 1 / 64                   # same noise regime as the oracle's outputs
    one = t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(bits[3], luts[1], bsk, pf, wop[2], wop[3])
    d1, _ = _decode4(o, one)
    assert one.shape == (1, o.big + 1) and d1[0] == 15 - int(msgs[3])
    with pytest.raises(t.B200TfheError, match="too small"):
        big = np.zeros((1, 12, o.n + 1), np.uint64)
        t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(big, luts, bsk, pf, wop[2], wop[3])
    with pytest.raises(t.B200TfheError, match="power of two"):
        t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(
            bits, np.zeros((3, o.N), np.uint64), bsk, pf, wop[2], wop[3], number_of_luts=1)
    with pytest.raises(t.B200TfheError, match="no blind-rotation instance"):
        t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(bits, luts, bsk, pf, 4, 3)


def test_oracle_cmux_tree_tables_of_several_polynomials():
    """13 input bits, tables of 4 polynomials: 2 tree layers then 11 rotation steps (wop_pbs/mod.rs:812-853)."""
    o = get_oracle("TOY_WOP")
    wop = _wop("TOY_WOP")
    nb = 13
    msgs = np.array([5, 4097, 8191, 0, 6000], dtype=np.uint64)
    bits = o.extract_bits(o.encrypt(msgs << np.uint64(64 - nb), under="big", seed=89), 64 - nb, nb)
    idx = np.arange(1 << nb, dtype=np.uint64)
    f = (idx * idx + idx // np.uint64(7) + np.uint64(3)) % np.uint64(16)
    out = o.cbs_vertical_packing(_pfpksk(o, "TOY_WOP"), wop, bits, (f << np.uint64(60)).reshape(4, o.N), 4)
    dec, _ = _decode4(o, out)
    assert np.array_equal(dec[:, 0], f[msgs].astype(int))


@pytest.mark.gpu
@pytest.mark.parametrize("nbits,polys", [(12, 2), (13, 4), (3, 4)])
def test_gpu_vertical_packing_cmux_tree(ctx, nbits, polys):
    """Tables of several polynomials: the CMUX tree (wop_pbs/mod.rs:468-592) over the most significant bits, then the blind
    rotation over the others.  (3, 4): two tree layers and a single rotation step."""
    o = get_oracle("TOY_WOP")
    t, bsk, ksk, pf, wop = _wop_keys(ctx, o, "TOY_WOP")
    rng = np.random.default_rng(79 + nbits)
    msgs = rng.integers(0, 1 << nbits, 6, dtype=np.uint64)
    msgs[0], msgs[1] = 0, (1 << nbits) - 1
    cts = o.encrypt(msgs << np.uint64(64 - nbits), under="big", seed=83)
    bits = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(cts, bsk, ksk, 64 - nbits, nbits)
    idx = np.arange(polys * o.N, dtype=np.uint64)
    f = [(idx * idx + idx // np.uint64(7) + np.uint64(3)) % np.uint64(16), (idx ^ (idx >> np.uint64(3))) % np.uint64(16)]
    luts = np.stack([(v << np.uint64(60)).astype(np.uint64) for v in f]).reshape(2 * polys, o.N)
    out = t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(bits, luts, bsk, pf, wop[2], wop[3], number_of_luts=2)
    dec, ph = _decode4(o, out)
    # the nbits input bits address coefficient `msg` of a table of 2^nbits entries laid out over polys * N coefficients:
    # polynomial index = the log2(polys) most significant bits, rotation = the others
    n_br = nbits - int(np.log2(polys))
    pos = ((msgs >> np.uint64(n_br)) * np.uint64(o.N) + (msgs & np.uint64((1 << n_br) - 1))).astype(np.int64)
    assert np.array_equal(dec[:, 0], f[0][pos].astype(int)) and np.array_equal(dec[:, 1], f[1][pos].astype(int))
    ref = o.cbs_vertical_packing(_pfpksk(o, "TOY_WOP"), wop, bits[:2], luts, polys)
    dref, phr = _decode4(o, ref)
    assert np.array_equal(dref, dec[:2])
    err = np.abs((ph - (dec.astype(np.uint64) << np.uint64(60))).astype(np.int64)) / 2.0 ** 64
    eref = np.abs((phr - (dref.astype(np.uint64) << np.uint64(60))).astype(np.int64)) / 2.0 ** 64
    assert err.max() < max(8 * eref.max(), 2.0 ** -9) < 1 / 64


@pytest.mark.gpu
def test_gpu_wop_bootstrap_set_large_ragged_batch(ctx):
    """The bootstrap of the WoP-PBS set (N 2048, level 2, base 2^15) on a batch of several waves that is not a multiple of the
    ciphertexts per CTA (ghost ciphertext), two tables selected per ciphertext, against the oracle; GLWE and LWE epilogues agree."""
    import tfhe_rs_odd_b200 as t
    o = get_oracle("TOY_WOP")
    _, bsk, ksk = _keys(ctx, o)
    B = 3 * 148 + 157
    msgs = (np.arange(B) * 7 + 3) % o.pmod
    cts = o.encrypt(o.encode(msgs), under="small", seed=97)
    luts = np.stack([o.lut(lambda x: x), o.lut(lambda x: (5 * x + 2) % o.pmod)])
    idx = (np.arange(B) % 2).astype(np.uint32)
    want = np.where(idx == 0, msgs, (5 * msgs + 2) % o.pmod).astype(np.uint64)
    out = t.programmable_bootstrap_lwe_ciphertext(cts, luts, bsk, lut_idx=idx)
    ph = o.phase(out)
    assert np.array_equal(o.decode(ph) % o.pmod, want)
    nref = 16
    ref = o.pbs(cts[:nref], luts, idx[:nref])
    d = (ph[:nref] - o.phase(ref)).astype(np.int64) / 2.0 ** 64
    assert np.abs(d).max() < phase_bound(o)
    glwe = t.core_crypto.programmable_bootstrap_lwe_ciphertext_without_sample_extract(cts, luts, bsk, lut_idx=idx)
    ext = t.core_crypto.extract_lwe_sample_from_glwe_ciphertext(ctx, glwe, o.k, o.N, 64)
    assert np.array_equal(ext, out)                                                              # same kernel, other epilogue


# ---- host mirror of shortint::wopbs::WopbsKey (tfhe-rs-odd_b200/wopbs.py) ---------------------------------------------------
def test_wopbs_lut_tables_match_reference_formulas():
    """generate_lut (wopbs/mod.rs:282-295) and generate_lut_without_padding (:320-333), computed without a device."""
    import tfhe_rs_odd_b200 as t  # noqa: F401
    from tfhe_rs_odd_b200 import wopbs

    class _P:
        polynomial_size = 2048

    class _SK:
        params = _P()

    class _PF:
        decomp_level_count, decomp_base_log = 2, 15
    key = wopbs.WopbsKey(_SK(), _PF(), wopbs.WopbsParameters(2, 15, 3, 5, 4, 4))
    lut = key.generate_lut(lambda x: x * x % 4)
    assert lut.values.shape == (1, 2048) and lut.output_ciphertext_count == 1
    assert [int(v) >> 59 for v in lut.values[0, :16]] == [(i % 4) ** 2 % 4 for i in range(16)] and not lut.values[0, 16:].any()
    lut2 = key.generate_lut_without_padding(lambda x: 3 - x)
    assert [int(v) >> 60 for v in lut2.values[0, :16]] == [3 - i % 4 for i in range(16)]
    with pytest.raises(t.B200TfheError, match="does not match"):
        wopbs.WopbsKey(_SK(), _PF(), wopbs.WopbsParameters(3, 15, 3, 5, 4, 4))


@pytest.mark.gpu
def test_gpu_wopbs_key_applies_lut_to_every_message(ctx):
    """WopbsKey.wopbs (wopbs/mod.rs:434-451) and programmable_bootstrapping_without_padding (:472-495) over all 16 values of a
    2_2 block, as the reference's doc-tests do for one value."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import shortint, wopbs
    name = "TOY_WOP"
    o = get_oracle(name)
    _, bsk, ksk, pf, wop = _wop_keys(ctx, o, name)
    key = wopbs.WopbsKey(shortint.ServerKey(ksk, bsk, 4, 4), pf, wopbs.WopbsParameters(wop[1], wop[0], wop[3], wop[2], 4, 4))
    msgs = np.arange(16, dtype=np.uint64)
    f = lambda x: (x * x + 1) % 4                                                    # noqa: E731
    cts = shortint.CiphertextList(o.encrypt(msgs << np.uint64(59), under="big", seed=101), degree=15, noise_level=shortint.NOISE_NOMINAL)
    out = key.wopbs(cts, key.generate_lut(f))
    ph = o.phase(out.data)
    dec = ((ph.astype(np.uint64) + np.uint64(1 << 58)) >> np.uint64(59)).astype(int) & 31
    assert np.array_equal(dec, [f(int(m) % 4) for m in msgs]) and (out.degree == 3).all() and (out.noise_level == shortint.NOISE_NOMINAL).all()
    cts2 = shortint.CiphertextList(o.encrypt(msgs << np.uint64(60), under="big", seed=103), degree=15, noise_level=shortint.NOISE_NOMINAL)
    out2 = key.programmable_bootstrapping_without_padding(cts2, key.generate_lut_without_padding(f))
    dec2, _ = _decode4(o, out2.data)
    assert np.array_equal(dec2, [f(int(m) % 4) for m in msgs])


@pytest.mark.gpu
def test_gpu_cbs_vertical_packing_batch_slices(ctx):
    """Batches above 1024 inputs go through the device in slices (bounded workspace): same bits out as the slices run separately,
    every 2-bit input decrypts to its table entry."""
    o = get_oracle("TOY_WOP")
    t, bsk, ksk, pf, wop = _wop_keys(ctx, o, "TOY_WOP")
    B, nbits = 1024 + 77, 2
    msgs = (np.arange(B) * 3 + 1) % 4
    bitvals = _bits_msb_first(msgs, nbits).astype(np.uint64)
    bits = o.encrypt((bitvals.reshape(-1) << np.uint64(63)), under="small", seed=107).reshape(B, nbits, o.n + 1)
    lut = np.zeros((1, o.N), np.uint64)
    lut[0, :4] = (np.array([2, 0, 3, 1], dtype=np.uint64) << np.uint64(60))
    f = t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized
    out = f(bits, lut, bsk, pf, wop[2], wop[3])
    dec, _ = _decode4(o, out)
    assert np.array_equal(dec[:, 0], np.array([2, 0, 3, 1])[msgs])
    assert np.array_equal(out[1024:], f(bits[1024:], lut, bsk, pf, wop[2], wop[3]))
    assert np.array_equal(out[:8], f(bits[:8], lut, bsk, pf, wop[2], wop[3]))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["TOY_WOP_K3", "TOY_WOP_K3B"])
def test_gpu_wopbs_k3_n512_family(ctx, name):
    """The (k 3, N 512) WoP-PBS families (WOPBS_PARAM_MESSAGE_2_CARRY_1_KS_PBS / 4_CARRY_0: pbs 12/2, ks 2/4, pfks 12/2, cbs 3/4, and
    1_CARRY_2 / 3_CARRY_1 / 4_CARRY_1 / 6_CARRY_0: pbs 9/3, pfks 9/3, cbs 3/5) at a toy n:
    R = 16 instances of the bootstrap, of the per-ciphertext-GGSW blind rotation and of the CMUX layer; pfpks with a 2^12 base
    (two 6-bit sub-digits).  Every 3-bit message through one table, then 10 bits through a 2-polynomial table (one tree layer)."""
    o = get_oracle(name)
    t, bsk, ksk, pf, wop = _wop_keys(ctx, o, name)
    host = _pfpksk(o, name)
    rng = np.random.default_rng(109)
    rnd = rng.integers(0, U64, size=(9, o.big + 1), dtype=np.uint64)
    key_len = (o.big + 1) * wop[1] * o.glwe_len
    for r in (0, o.k):
        got = t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, r, rnd)
        assert np.array_equal(got, o.pfpks(host[r * key_len:(r + 1) * key_len], wop[0], wop[1], rnd))
    f = t.core_crypto.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized
    msgs = np.arange(8, dtype=np.uint64)
    bits = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(o.encrypt(msgs << np.uint64(61), under="big", seed=113), bsk, ksk, 61, 3)
    lut = np.zeros((1, o.N), np.uint64)
    lut[0, :8] = ((msgs * np.uint64(3) + np.uint64(1)) % np.uint64(8)) << np.uint64(61)
    out = f(bits, lut, bsk, pf, wop[2], wop[3])
    ph = o.phase(out.reshape(-1, o.big + 1))
    dec = ((ph.astype(np.uint64) + np.uint64(1 << 60)) >> np.uint64(61)).astype(int) & 7
    assert np.array_equal(dec, ((msgs * 3 + 1) % 8).astype(int))
    ref = o.cbs_vertical_packing(host, wop, bits[:2], lut)
    assert np.array_equal(((o.phase(ref.reshape(-1, o.big + 1)).astype(np.uint64) + np.uint64(1 << 60)) >> np.uint64(61)).astype(int) & 7, dec[:2])
    nb = 10                                                          # 2 polynomials of 512 entries: one CMUX layer + 9 rotation steps
    m10 = np.array([0, 1023, 512, 511, 700], dtype=np.uint64)
    bits10 = t.core_crypto.extract_bits_from_lwe_ciphertext_mem_optimized(o.encrypt(m10 << np.uint64(54), under="big", seed=127), bsk, ksk, 54, nb)
    b10, _ = _decrypt_bits(o, bits10)
    assert np.array_equal(b10, _bits_msb_first(m10, nb))
    idx = np.arange(2 * o.N, dtype=np.uint64)
    tab = (idx * np.uint64(5) + idx // np.uint64(9)) % np.uint64(8)
    out10 = f(bits10, (tab << np.uint64(61)).reshape(2, o.N), bsk, pf, wop[2], wop[3], number_of_luts=1)
    ph10 = o.phase(out10.reshape(-1, o.big + 1))
    dec10 = ((ph10.astype(np.uint64) + np.uint64(1 << 60)) >> np.uint64(61)).astype(int) & 7
    assert np.array_equal(dec10, tab[m10.astype(np.int64)].astype(int))
