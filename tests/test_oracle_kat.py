"""CPU tests (-m "not gpu"): pin the oracle against every known-answer value and tolerance test the reference holds
for this path, so that it can be trusted as the checker of the CUDA path."""
import ctypes as C

import numpy as np
import pytest

from conftest import get_oracle
from oracle.oracle import Oracle, lib, _p


def test_closest_representable_kat():
    # doc-test KAT: core_crypto/commons/math/decomposition/decomposer.rs:95-96  (B = 2^4, l = 3, u32)
    assert lib().orc_closest_representable(C.c_uint64(1_340_987_234), 4, 3, 32) == 1_341_128_704


@pytest.mark.parametrize("bits,b,l", [(32, 4, 3), (64, 23, 1), (64, 15, 2), (64, 3, 5), (32, 6, 3), (32, 3, 4)])
def test_decomposition_recomposes_and_digit_range(bits, b, l):
    # decomposer.rs:135-143 doc-test + decomposition/tests.rs: digits in [-B/2, B/2], sum digit_i * q/B^i == closest_representable
    rng = np.random.default_rng(1)
    mask = (1 << bits) - 1
    for x in [0, mask, 1 << (bits - 1)] + [int(v) for v in rng.integers(0, 2 ** bits, 300, dtype=np.uint64)]:
        d = np.zeros(l, np.uint64)
        lib().orc_decompose(C.c_uint64(x), b, l, bits, _p(d))
        rec = 0
        for t in range(l):                      # t = 0 is level l
            level = l - t
            dv = int(d[t])
            sv = dv - (1 << bits) if dv >> (bits - 1) else dv
            assert -(1 << (b - 1)) <= sv <= (1 << (b - 1))
            rec += sv << (bits - b * level)
        assert rec & mask == lib().orc_closest_representable(C.c_uint64(x), b, l, bits)


def test_modswitch_range_and_values():
    # fft_impl/common.rs:20-43: result in [0, 2N]
    assert lib().orc_modswitch(C.c_uint64(0), 11, 64) == 0
    assert lib().orc_modswitch(C.c_uint64(2 ** 64 - 1), 11, 64) == 4096
    assert lib().orc_modswitch(C.c_uint64(1 << 63), 11, 64) == 2048
    assert lib().orc_modswitch(C.c_uint64(2 ** 32 - 1), 9, 32) == 1024


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("N", [32, 256, 2048, 8192])
def test_fft_roundtrip_tolerance(bits, N):
    # fft/tests.rs:9-80 test_roundtrip: u32 exact, u64 error < 2^14 ulp
    rng = np.random.default_rng(N + bits)
    poly = rng.integers(0, 2 ** bits, N, dtype=np.uint64)
    f = np.zeros(N, np.float64)
    lib().orc_fft_forward_torus(N, _p(poly), bits, _p(f))
    out = np.zeros(N, np.uint64)
    lib().orc_fft_add_backward_torus(N, _p(out), bits, _p(f))
    d = out - poly
    if bits == 32:
        assert np.array_equal(out & np.uint64(0xFFFFFFFF), poly)
    else:
        assert np.abs(d.astype(np.int64)).max() < (1 << 14)


@pytest.mark.parametrize("N", [64, 512, 2048])
def test_fft_product_vs_schoolbook(N):
    # fft/tests.rs:82-222 test_product: torus poly x 16-bit integer poly vs convolution_naive, u64 threshold 1 << (64 - (52-16-log2 N))
    rng = np.random.default_rng(N)
    a = rng.integers(0, 2 ** 64, N, dtype=np.uint64)
    b = rng.integers(0, 2 ** 16, N, dtype=np.uint64)
    ref = np.zeros(N, np.uint64)
    lib().orc_negacyclic_schoolbook(N, _p(a), _p(b), _p(ref), 64)
    fa, fb = np.zeros(N), np.zeros(N)
    lib().orc_fft_forward_torus(N, _p(a), 64, _p(fa))
    lib().orc_fft_forward_integer(N, _p(b), 64, _p(fb))
    M = N // 2      # oracle Fourier buffers are structure-of-arrays: M real parts then M imaginary parts
    ca, cb = fa[:M] + 1j * fa[M:], fb[:M] + 1j * fb[M:]
    pr = ca * cb
    prod = np.ascontiguousarray(np.concatenate([pr.real, pr.imag]))
    out = np.zeros(N, np.uint64)
    lib().orc_fft_add_backward_torus(N, _p(out), 64, _p(prod))
    thr = 1 << (64 - (52 - 16 - (N.bit_length() - 1)))
    assert np.abs((out - ref).astype(np.int64)).max() < thr


def test_monomial_ops_and_sample_extract_against_definition():
    # polynomial_algorithms.rs:315-354,425-490 ; glwe_sample_extraction.rs:126-146 -- against a big-int negacyclic definition
    N = 16
    rng = np.random.default_rng(5)
    p = rng.integers(0, 2 ** 64, N, dtype=np.uint64)

    def mul_monomial(poly, d):            # poly * X^d mod (X^N + 1), d may be negative
        out = [0] * N
        for j, c in enumerate(poly):
            e = (j + d) % (2 * N)
            out[e % N] = (out[e % N] + (int(c) if e < N else -int(c))) % 2 ** 64
        return np.array(out, dtype=np.uint64)
    for deg in [0, 1, 5, N - 1, N, N + 3, 2 * N - 1, 2 * N]:
        out = np.zeros(N, np.uint64)
        lib().orc_monomial_div(_p(out), _p(p), N, C.c_uint64(deg), 64)
        assert np.array_equal(out, mul_monomial(p, -deg)), deg
        lib().orc_monomial_mul_sub(_p(out), _p(p), N, C.c_uint64(deg), 64)
        assert np.array_equal(out, mul_monomial(p, deg) - p), deg
    k = 2
    glwe = rng.integers(0, 2 ** 64, (k + 1) * N, dtype=np.uint64)
    lwe = np.zeros(k * N + 1, np.uint64)
    lib().orc_sample_extract0(_p(lwe), _p(glwe), k, N, 64)
    assert lwe[k * N] == glwe[k * N]
    for q in range(k):
        assert lwe[q * N] == glwe[q * N]
        assert np.array_equal(lwe[q * N + 1:(q + 1) * N], (np.uint64(0) - glwe[q * N + 1:(q + 1) * N])[::-1])


@pytest.mark.parametrize("name", ["TOY_U64", "TOY_2_2", "TOY_1_1"])
def test_keyswitch_and_pbs_decrypt_every_message(name):
    # algorithms/test/lwe_keyswitch.rs:8-109 and test/lwe_programmable_bootstrapping.rs:70-160: every message, identity LUT
    o = get_oracle(name)
    msgs = np.arange(o.pmod).repeat(2)
    cts = o.encrypt(o.encode(msgs), "big")
    ks = o.keyswitch(cts)
    assert np.array_equal(o.decode(o.phase(ks, "small")) % o.pmod, msgs)
    out = o.ks_pbs(cts, o.lut(lambda x: x))
    assert np.array_equal(o.decode(o.phase(out, "big")) % o.pmod, msgs)
    out = o.ks_pbs(cts, o.lut(lambda x: (x * x + 1) % o.pmod))
    assert np.array_equal(o.decode(o.phase(out, "big")) % o.pmod, (msgs * msgs + 1) % o.pmod)


def test_boolean_truth_tables():
    # boolean/server_key/tests.rs: all six binary gates, PBS -> KS order of DEFAULT_PARAMETERS
    o = get_oracle("TOY_U32")
    a = np.array([0, 0, 1, 1], bool); b = np.array([0, 1, 0, 1], bool)
    T, F = 1 << 29, 7 << 29
    ca = o.encrypt(np.where(a, T, F).astype(np.uint64), "small", seed=1)
    cb = o.encrypt(np.where(b, T, F).astype(np.uint64), "small", seed=50)
    truth = {0: a & b, 1: ~(a & b), 2: a | b, 3: ~(a | b), 4: a ^ b, 5: ~(a ^ b)}
    for gate, want in truth.items():
        out = o.pbs_ks(o.gate_prologue(gate, ca, cb), o.boolean_lut())
        assert np.array_equal(o.decode_bool(o.phase(out, "small")), want), gate


@pytest.mark.parametrize("name", ["TOY_2_2", "TOY_U64", "TOY_U32", "TOY_1_1"])
def test_oracle_external_product_vs_exact_integer_product(name):
    """The oracle's add_external_product (f64 FFT, the reference's operation order) against the EXACT result: digits x the
    standard-domain key by schoolbook negacyclic products in wrapping integer arithmetic, no FFT (the reference's own
    convolution_naive, fft/tests.rs:83-100).  Threshold: the reference's (fft/tests.rs:166-167) summed over the (k+1) l products,
    and the tighter per-CMUX bound the GPU tests use (conftest.fft_bound)."""
    from conftest import fft_bound, get_oracle
    o = get_oracle(name)
    rng = np.random.default_rng(0xE1)
    for gi in (0, o.n - 1):
        acc = rng.integers(0, 2 ** o.bits, size=o.glwe_len, dtype=np.uint64)
        glwe = rng.integers(0, 2 ** o.bits, size=o.glwe_len, dtype=np.uint64)
        got = o.add_external_product(acc, gi, glwe)[0]
        exact = o.exact_external_product(acc, gi, glwe)
        d = (got - exact)
        if o.bits == 32:
            d = (d & np.uint64(0xFFFFFFFF)).astype(np.uint32).astype(np.int32).astype(np.int64)
        else:
            d = d.astype(np.int64)
        log2n = o.N.bit_length() - 1
        ref_threshold = (o.k + 1) * o.l * 2.0 ** max(0, o.bits - (52 - o.b - log2n))
        assert np.abs(d).max() <= ref_threshold
        assert np.abs(d).max() <= fft_bound(o), (np.abs(d).max(), fft_bound(o))
    # the vectorised decomposition used above is the KAT-pinned one
    x = rng.integers(0, 2 ** o.bits, size=64, dtype=np.uint64)
    digs = o.decompose_all(x)
    for i, v in enumerate(x):
        out = np.zeros(o.l, np.uint64)
        lib().orc_decompose(C.c_uint64(int(v)), o.b, o.l, o.bits, _p(out))
        assert np.array_equal(digs[:, i], out)
