"""Wire formats: bincode layouts of the reference's key / ciphertext types (tfhe-rs-odd_b200/serialization.py).
The reference holds no serialised fixtures, so the byte layouts are written out by hand here from the type definitions
(file:line in serialization.py) and the codec is checked against them; the Fourier-domain order is pinned against numpy."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from conftest import ROOT, get_oracle, params_of


def _ser():
    import tfhe_rs_odd_b200 as t  # noqa: F401
    from tfhe_rs_odd_b200 import serialization
    return serialization


def _q(*v):
    return b"".join(struct.pack("<Q", x) for x in v)


NATIVE64 = struct.pack("<QQ", 0, 0) + _q(64)        # CiphertextModulus{modulus: 0u128, scalar_bits: 64}


def test_keyswitch_key_bytes_by_hand():
    ser = _ser()
    # LweKeyswitchKey{data: Vec<u64> (len 12), decomp_base_log 3, decomp_level_count 2, output_lwe_size 3, modulus}
    data = np.arange(12, dtype=np.uint64) * np.uint64(0x0101010101010101)
    want = _q(12) + data.astype("<u8").tobytes() + _q(3, 2, 3) + NATIVE64
    assert ser.serialize_lwe_keyswitch_key(data, 3, 2, 3) == want
    k = ser.deserialize_lwe_keyswitch_key(want)
    assert np.array_equal(k["data"], data) and k["input_lwe_dimension"] == 2 and k["output_lwe_size"] == 3
    with pytest.raises(Exception, match="trailing"):
        ser.deserialize_lwe_keyswitch_key(want + b"\0")
    with pytest.raises(Exception, match="end of input"):
        ser.deserialize_lwe_keyswitch_key(want[:-1])
    bad = want[:-8] + _q(32)                                            # scalar_bits mismatch is an error in the reference
    with pytest.raises(Exception, match="64 bits"):
        ser.deserialize_lwe_keyswitch_key(bad)


def test_ciphertext_and_bootstrap_key_roundtrips():
    ser = _ser()
    rng = np.random.default_rng(0)
    ct = rng.integers(0, 2 ** 64, 11, dtype=np.uint64)
    assert ser.serialize_lwe_ciphertext(ct) == _q(11) + ct.tobytes() + NATIVE64
    assert np.array_equal(ser.deserialize_lwe_ciphertext(ser.serialize_lwe_ciphertext(ct)), ct)
    lst = rng.integers(0, 2 ** 64, (4, 5), dtype=np.uint64)
    assert np.array_equal(ser.deserialize_lwe_ciphertext_list(ser.serialize_lwe_ciphertext_list(lst)), lst)
    c32 = rng.integers(0, 2 ** 32, 7, dtype=np.uint32)
    assert np.array_equal(ser.deserialize_lwe_ciphertext(ser.serialize_lwe_ciphertext(c32, 32), 32), c32)
    g, poly = ser.deserialize_glwe_ciphertext(ser.serialize_glwe_ciphertext(lst.reshape(-1)[:16], 8))
    assert poly == 8 and np.array_equal(g, lst.reshape(-1)[:16])
    # LweBootstrapKey: n = 3, l = 2, glwe_size = 2, N = 4
    bsk = rng.integers(0, 2 ** 64, 3 * 2 * 2 * 2 * 4, dtype=np.uint64)
    b = ser.serialize_lwe_bootstrap_key(bsk, 2, 4, 7, 2)
    assert b == _q(bsk.size) + bsk.tobytes() + _q(2, 4, 7, 2) + NATIVE64
    k = ser.deserialize_lwe_bootstrap_key(b)
    assert k["input_lwe_dimension"] == 3 and np.array_equal(k["data"], bsk)
    # multi-bit: 2 groups of G = 2 -> 8 GGSWs
    mb = rng.integers(0, 2 ** 64, 8 * 2 * 2 * 2 * 4, dtype=np.uint64)
    k = ser.deserialize_lwe_multi_bit_bootstrap_key(ser.serialize_lwe_bootstrap_key(mb, 2, 4, 7, 2, grouping_factor=2))
    assert k["grouping_factor"] == 2 and k["input_lwe_dimension"] == 4 and k["ggsw_count"] == 8
    pk = rng.integers(0, 2 ** 64, 4 * 2 * 2 * 4, dtype=np.uint64)
    k = ser.deserialize_lwe_packing_keyswitch_key(ser.serialize_lwe_packing_keyswitch_key(pk, 3, 2, 2, 4))
    assert k["input_lwe_dimension"] == 4 and k["output_polynomial_size"] == 4
    s = ser.serialize_shortint_ciphertext(ct, 3, 1, 4, 4, "ks_pbs")
    assert s == _q(11) + ct.tobytes() + NATIVE64 + _q(3, 1, 4, 4) + struct.pack("<I", 0)
    d = ser.deserialize_shortint_ciphertext(s)
    assert d["degree"] == 3 and d["pbs_order"] == "ks_pbs" and np.array_equal(d["ct"], ct)


def test_fourier_key_bytes_by_hand():
    ser = _ser()
    # FourierLweBootstrapKey with n = 1, l = 1, glwe_size = 1 (degenerate), N = 4: ONE Fourier polynomial of 2 complex values
    f = np.array([[[1.5, -2.0], [0.25, 8.0]]])
    want = _q(3, 4, 1) + _q(2) + struct.pack("<dddd", 1.5, -2.0, 0.25, 8.0) + _q(1, 1, 9, 1)
    assert ser.serialize_fourier_lwe_bootstrap_key(f, 4, 1, 1, 9, 1) == want
    k = ser.deserialize_fourier_lwe_bootstrap_key(want)
    assert np.array_equal(k["fourier"], f) and k["decomp_base_log"] == 9 and k["grouping_factor"] == 1


def _numpy_fourier_key(o):
    """forward_as_torus of every key polynomial, natural frequency order (numpy's FFT; independent of oracle and device)."""
    N, M = o.N, o.N // 2
    polys = o.bsk.reshape(-1, N).astype(np.int64).astype(np.float64) * 2.0 ** -64
    x = (polys[:, :M] + 1j * polys[:, M:]) * np.exp(1j * np.pi * np.arange(M) / N)
    F = np.fft.fft(x, axis=1)
    return np.stack([F.real, F.imag], axis=-1)


@pytest.mark.parametrize("N,bits", [(512, 64), (2048, 64), (512, 32), (8192, 64)])
def test_device_fourier_layout_is_natural_order(N, bits):
    """What b200tfhe_bsk_upload_fourier relies on: the key conversion kernels write frequency f at position f
    (f = t/16 + 16 (t%16) + 256 p at position 256 p + t for N = 8192).  Executed on the CPU by the lane emulator."""
    import __graft_entry__ as g
    g.build()
    em = C.CDLL(os.path.join(ROOT, "tests", "native", "libemul.so"))
    rng = np.random.default_rng(1)
    M = N // 2
    poly = rng.integers(0, 2 ** bits, size=16 * N, dtype=np.uint64)
    out = np.zeros(2 * M)
    assert em.emul_key_forward(N, bits, poly.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)) == 0
    p = poly[:N].astype(np.int64) if bits == 64 else poly[:N].astype(np.uint32).astype(np.int32).astype(np.int64)
    ref = np.fft.fft((p[:M] + 1j * p[M:]) * 2.0 ** -bits * np.exp(1j * np.pi * np.arange(M) / N))
    got = out[0::2] + 1j * out[1::2]
    if N == 8192:
        pos = np.arange(M)
        t, pp = pos % 256, pos // 256
        ref = ref[t // 16 + 16 * (t % 16) + 256 * pp]
    assert np.abs(got - ref).max() < 1e-12


def _server_key_bytes(o, grouping=None):
    ser = _ser()
    return ser.serialize_shortint_server_key(
        o.ksk, o.b_ks, o.l_ks, o.n + 1, _numpy_fourier_key(o), o.N, o.n, o.k + 1, o.b, o.l,
        message_modulus=int(round(o.pmod ** 0.5)), carry_modulus=int(round(o.pmod ** 0.5)), max_degree=o.pmod - 1, max_noise_level=8,
        pbs_order="ks_pbs", grouping_factor=grouping)


def test_shortint_server_key_roundtrip_cpu():
    ser = _ser()
    o = get_oracle("TOY_U64")
    d = ser.deserialize_shortint_server_key(_server_key_bytes(o))
    assert d["pbs_order"] == "ks_pbs" and d["message_modulus"] == 2 and d["max_noise_level"] == 8
    assert np.array_equal(d["key_switching_key"]["data"], o.ksk)
    b = d["bootstrapping_key"]
    assert (b["input_lwe_dimension"], b["glwe_size"], b["polynomial_size"], b["decomp_base_log"], b["decomp_level_count"]) == (o.n, o.k + 1, o.N, o.b, o.l)
    assert b["fourier"].shape == (o.n * o.l * (o.k + 1) ** 2, o.N // 2, 2)
    mb = get_oracle("TOY_MB_G2")
    d = ser.deserialize_shortint_server_key(_server_key_bytes(mb, grouping=2))
    assert d["bootstrapping_key"]["grouping_factor"] == 2 and d["bootstrapping_key"]["deterministic_execution"] is False


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["TOY_U64", "TOY_2_2", "TOY_3_3", "TOY_MB22_G3"])
def test_gpu_server_key_from_bincode(ctx, name):
    """A serialised shortint ServerKey (keyswitch key + FOURIER bootstrap key computed with numpy, natural order) loads and
    bootstraps correctly: every message through apply_lookup_table, and the same decrypted values as the key converted on the device."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import shortint
    o = get_oracle(name)
    sk = shortint.ServerKey.from_bincode(ctx, _server_key_bytes(o, o.grouping or None))
    assert sk.bootstrapping_key.grouping_factor == (o.grouping or 1)
    msgs = np.arange(2 * o.pmod) % o.pmod
    cts = shortint.CiphertextList(o.encrypt(o.encode(msgs), under="big"), degree=o.pmod - 1, noise_level=shortint.NOISE_NOMINAL)
    lut = sk.generate_lookup_table(lambda x: (x * 3 + 2) % o.pmod)
    out = sk.apply_lookup_table(cts, lut)
    assert np.array_equal(o.decode(o.phase(out.data)), ((msgs * 3 + 2) % o.pmod).astype(np.uint64))
    p = params_of(o)
    if o.grouping:
        dev = t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, p, o.bsk, o.grouping)
    else:
        dev = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
    ref = shortint.ServerKey(sk.key_switching_key, dev, sk.message_modulus, sk.carry_modulus).apply_lookup_table(cts, lut)
    assert np.array_equal(o.decode(o.phase(ref.data)), o.decode(o.phase(out.data)))


def test_fourier_key_sanity_check_rejects_wrong_scaling_or_order():
    """shortint.ServerKey.from_bincode uploads the serialised Fourier key verbatim; before that it checks that a few polynomials invert to
    torus coefficients in [-1/2, 1/2) (the layout is not pinned by reference-produced bytes: fail loudly instead of bootstrapping garbage)."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200.shortint import _check_fourier_key
    N, M = 512, 256
    rng = np.random.default_rng(4)
    polys = rng.integers(0, 2 ** 64, size=(6, N), dtype=np.uint64).astype(np.int64).astype(np.float64) * 2.0 ** -64
    x = (polys[:, :M] + 1j * polys[:, M:]) * np.exp(1j * np.pi * np.arange(M) / N)
    F = np.fft.fft(x, axis=1)
    good = np.stack([F.real, F.imag], axis=-1)
    _check_fourier_key(good, N)                                        # natural order, 2^-64 scaling: accepted
    with pytest.raises(t.B200TfheError, match="order or scaling"):
        _check_fourier_key(good * 2.0 ** 64, N)                         # forward_as_integer instead of forward_as_torus
    with pytest.raises(t.B200TfheError, match="order or scaling"):
        _check_fourier_key(good * M, N)                                 # an un-normalised transform
