"""The f64 stages pinned ON THE DEVICE at FFT precision (VERDICT r1 "what's weak" #1).

(a) stand-alone add_external_product_assign / cmux_assign entry points vs the oracle, every parameter set;
(b) ONE CMUX (an LWE input with a single non-zero mask element) through EVERY blind-rotation kernel vs the oracle's
    blind_rotate_assign: per-coefficient difference within the f64 rounding of one external product;
(c) a NOISELESS gadget key (GGSW_i = s_i * G, valid under any secret key): the blind rotation is then an exact monomial
    rotation up to FFT rounding, so the output rounded to the decomposition grid must equal LUT * X^(-b_hat + sum s_i a_hat_i)
    BIT FOR BIT, with mask elements that modulus-switch to the edge rotations 0, 1, N-1, N, 2N-1 -- the integer
    stages a3/a4/a5 (modulus switch, monomial div, rotate-and-subtract) checked exactly on hardware for every kernel.
"""
import numpy as np
import pytest

from conftest import fft_bound, get_oracle, params_of

pytestmark = pytest.mark.gpu

BR_AUTO, BR_GENERIC, BR_TMEM, BR_LAT, BR_GENERIC_1CT, BR_X16, BR_PAIR = 0, 1, 2, 3, 4, 6, 7
BR_LAT_1SM = 13        # kernel 3 with lat_cluster_max = 0: one CTA per ciphertext (default for small batches: a two-CTA cluster)
BR_X16B = 600 + 4      # kernel 6 with x16_variant 4 ("all transforms first")
BR_X16C = 600 + 12     # the same with the bootstrap key streamed through a TMA ring in shared memory
# (oracle set, kernels that are built for it)
KERNELS = [
    ("TOY_2_2", [BR_GENERIC, BR_TMEM, BR_LAT, BR_LAT_1SM, BR_GENERIC_1CT]),
    ("TOY_1_1", [BR_GENERIC, BR_PAIR]),   # BR_PAIR: six ciphertexts per CTA, two per warp (blind_rotate_pair.cu)
    ("TOY_U64", [BR_GENERIC]),
    ("TOY_U32", [BR_GENERIC, BR_X16, BR_X16B, BR_X16C]),  # the boolean shape (k 2, N 512, l 3) on the 2^32 torus; x16 = half-warp teams + tensor memory
    ("TOY_3_3", [BR_GENERIC]),          # N = 8192: br_launch falls through to blind_rotate_big_kernel
    ("TOY_WOP", [BR_GENERIC]),          # level 2, base 2^15
]
CASES = [(n, k) for n, ks in KERNELS for k in ks]


def _dt(o):
    return np.uint64 if o.bits == 64 else np.uint32


def _sdiff(a, b, bits):
    d = a.astype(np.uint64) - b.astype(np.uint64)
    if bits == 32:
        return (d & np.uint64(0xFFFFFFFF)).astype(np.uint32).astype(np.int32).astype(np.int64)
    return d.astype(np.int64)


class _kernel:
    def __init__(self, ctx, which):
        self.ctx, self.which = ctx, which

    def __enter__(self):
        if self.which == BR_LAT_1SM:
            self.ctx.set_option("br_kernel", BR_LAT)
            self.ctx.set_option("lat_cluster_max", 0)
            return
        self.ctx.set_option("br_kernel", self.which // 100 if self.which >= 100 else self.which)
        self.ctx.set_option("x16_variant", self.which % 100 if self.which >= 100 else 0)

    def __exit__(self, *a):
        self.ctx.set_option("br_kernel", BR_AUTO)
        self.ctx.set_option("x16_variant", -1)      # back to the default organisation
        self.ctx.set_option("lat_cluster_max", 74)


def _bsk(ctx, o):
    import tfhe_rs_odd_b200 as t
    return t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, params_of(o), o.bsk.astype(_dt(o)))


@pytest.mark.parametrize("name", ["TOY_2_2", "TOY_1_1", "TOY_U64", "TOY_U32", "TOY_3_3", "TOY_WOP", "TOY_WOP_K3", "TOY_WOP_K3B"])
def test_external_product_matches_oracle_to_fft_precision(ctx, name):
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    bsk = _bsk(ctx, o)
    rng = np.random.default_rng(0xE0)
    B = 5                                                     # odd: the boolean shape runs two ciphertexts per CTA
    out0 = rng.integers(0, 2 ** o.bits, size=(B, o.glwe_len), dtype=np.uint64)
    glwe = rng.integers(0, 2 ** o.bits, size=(B, o.glwe_len), dtype=np.uint64)
    glwe[1, :] = 0                                            # all-zero digits: adds exactly nothing
    glwe[2, :] = np.uint64((1 << (o.bits - o.b * o.l - 1)))    # the rounding boundary of closest_representable
    idx = np.array([0, 3, o.n - 1, 1, 2], np.uint32)
    ref = o.add_external_product(out0, idx, glwe)
    got = out0.astype(_dt(o)).copy()
    t.add_external_product_assign(got, bsk, glwe.astype(_dt(o)), idx)
    d = np.abs(_sdiff(got, ref, o.bits))
    assert d.max() <= fft_bound(o), (name, d.max(), fft_bound(o))
    assert np.array_equal(got[1].astype(np.uint64), out0[1]), "zero input must add exactly zero"
    # the reference's own acceptance criterion for this stage: against the EXACT integer product (no FFT on either side)
    exact = o.exact_external_product(out0[0], int(idx[0]), glwe[0])
    assert np.abs(_sdiff(got[0].astype(np.uint64), exact, o.bits)).max() <= fft_bound(o)
    # default index (NULL) = GGSW 0
    got0 = out0[:1].astype(_dt(o)).copy()
    t.add_external_product_assign(got0, bsk, glwe[:1].astype(_dt(o)))
    assert np.array_equal(got0[0], got[0])


@pytest.mark.parametrize("name", ["TOY_2_2", "TOY_U32", "TOY_3_3"])
def test_cmux_matches_oracle_and_selects(ctx, name):
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    bsk = _bsk(ctx, o)
    rng = np.random.default_rng(0xC0)
    B = 4
    ct0 = rng.integers(0, 2 ** o.bits, size=(B, o.glwe_len), dtype=np.uint64)
    ct1 = rng.integers(0, 2 ** o.bits, size=(B, o.glwe_len), dtype=np.uint64)
    idx = np.arange(B, dtype=np.uint32)
    ref0, ref1 = o.cmux(ct0, ct1, idx)
    g0, g1 = ct0.astype(_dt(o)).copy(), ct1.astype(_dt(o)).copy()
    t.cmux_assign(g0, g1, bsk, idx)
    assert np.array_equal(g1.astype(np.uint64), ref1), "ct1 <- ct1 - ct0 is an integer stage: bit-exact"
    assert np.abs(_sdiff(g0, ref0, o.bits)).max() <= fft_bound(o)
    # semantics: the GGSW of key bit s selects ct1 (s = 1) or ct0 (s = 0), up to the decomposition rounding + key noise
    for b in range(B):
        want = ct1[b] if int(o.small_key[b]) else ct0[b]
        ph_got, ph_want = o.glwe_phase(g0[b].astype(np.uint64)), o.glwe_phase(want)
        err = np.abs(_sdiff(ph_got, ph_want, o.bits)).max() / 2.0 ** o.bits
        assert err < 2.0 ** -(o.b * o.l) * o.N, (b, err)      # ~ rounding to b*l bits over (k+1)N/2 key coefficients


def test_external_product_argument_checks(ctx):
    import tfhe_rs_odd_b200 as t
    o = get_oracle("TOY_2_2")
    bsk = _bsk(ctx, o)
    out = np.zeros((2, o.glwe_len), np.uint64)
    with pytest.raises(t.B200TfheError, match="ggsw_idx out of range"):
        t.add_external_product_assign(out, bsk, out.copy(), np.array([0, o.n], np.uint32))
    with pytest.raises(t.B200TfheError):
        t.add_external_product_assign(out, bsk, np.zeros((3, o.glwe_len), np.uint64))
    t.add_external_product_assign(np.zeros((0, o.glwe_len), np.uint64), bsk, np.zeros((0, o.glwe_len), np.uint64))   # empty batch


@pytest.mark.parametrize("name,kernel", CASES)
def test_single_cmux_through_every_blind_rotation_kernel(ctx, name, kernel):
    """One effective CMUX: every mask element but one is zero (a zero element rotates by 0 and adds exactly nothing; the oracle skips
    it as bootstrap.rs:285 does), so GPU vs oracle differ by the f64 rounding of ONE external product, not by decryption margins."""
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    bsk = _bsk(ctx, o)
    rng = np.random.default_rng(0x51 + kernel)
    B = 9
    cts = np.zeros((B, o.n + 1), np.uint64)
    pos = rng.integers(0, o.n, size=B)
    cts[np.arange(B), pos] = rng.integers(1, 2 ** o.bits, size=B, dtype=np.uint64)
    cts[:, o.n] = rng.integers(0, 2 ** o.bits, size=B, dtype=np.uint64)       # body: the initial monomial division
    cts[0, o.n] = 0
    luts = rng.integers(0, 2 ** o.bits, size=(2, o.glwe_len), dtype=np.uint64)   # full random GLWE accumulators (mask and body)
    idx = (np.arange(B) % 2).astype(np.uint32)
    ref = o.blind_rotate(cts, luts, idx)
    with _kernel(ctx, kernel):
        got = t.programmable_bootstrap_lwe_ciphertext_without_sample_extract(cts.astype(_dt(o)), luts.astype(_dt(o)), bsk, idx)
    d = np.abs(_sdiff(got, ref, o.bits))
    assert d.max() <= fft_bound(o), (name, kernel, d.max(), fft_bound(o))


def _gadget_key(o, bits_s):
    """Noiseless GGSW list: GGSW_i = s_i * G, i.e. s_i * q / B^level on coefficient 0 of polynomial r of row (level, r)."""
    k1 = o.k + 1
    bsk = np.zeros((o.n, o.l, k1, k1, o.N), np.uint64)
    for lvl in range(o.l):
        g = np.uint64(1 << (o.bits - o.b * (lvl + 1)))
        for r in range(k1):
            bsk[:, lvl, r, r, 0] = bits_s.astype(np.uint64) * g
    return bsk


def _rot(poly, a, N, bits):
    """poly * X^a mod (X^N + 1), a in Z, exact."""
    a %= 2 * N
    m = np.uint64((1 << bits) - 1 if bits < 64 else 0xFFFFFFFFFFFFFFFF)
    out = np.empty_like(poly)
    for c in range(N):
        s = (c - a) % (2 * N)
        v = poly[s % N]
        out[c] = ((~v + np.uint64(1)) & m) if s >= N else v
    return out


@pytest.mark.parametrize("name,kernel", CASES + [("MESSAGE_2_CARRY_2", BR_TMEM), ("MESSAGE_2_CARRY_2", BR_LAT), ("MESSAGE_2_CARRY_2", BR_LAT_1SM), ("BOOLEAN_DEFAULT", BR_X16), ("BOOLEAN_DEFAULT", BR_X16B), ("BOOLEAN_DEFAULT", BR_X16C), ("MESSAGE_1_CARRY_1", BR_PAIR)])
def test_noiseless_key_blind_rotation_is_an_exact_rotation(ctx, name, kernel):
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name) if name not in ("MESSAGE_2_CARRY_2", "BOOLEAN_DEFAULT", "MESSAGE_1_CARRY_1") else _shape_only(name)
    N, n, bits = o.N, o.n, o.bits
    rng = np.random.default_rng(0x90 + kernel)
    s = rng.integers(0, 2, size=n)
    s[:6] = 1
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, params_of(o), _gadget_key(o, s).astype(_dt(o)))
    B = 6
    log2N = N.bit_length() - 1
    unit = 1 << (bits - log2N - 1)                             # one step of the modulus switch to 2N
    ahat = rng.integers(0, 2 * N, size=(B, n + 1))
    ahat[0, :6] = [0, 1, N - 1, N, 2 * N - 1, N + 1]           # the edge rotations, on mask elements whose key bit is 1
    ahat[1, :] = 0
    ahat[2, n] = 2 * N - 1
    ahat[3, n] = N
    m = (1 << bits) - 1
    # a value that switches to a_hat: a_hat * unit plus a sub-step offset in (-unit/2, unit/2) (both rounding directions)
    off = rng.integers(-(unit // 2) + 1, unit // 2, size=(B, n + 1))
    cts = ((ahat.astype(object) * unit + off.astype(object)) & m).astype(np.uint64)
    # accumulator: every coefficient on the decomposition grid (multiples of q / B^l), mask and body polynomials non-trivial
    grid = np.uint64(1 << (bits - o.b * o.l))
    lut = (rng.integers(0, 1 << (o.b * o.l), size=o.glwe_len, dtype=np.uint64) * grid)
    with _kernel(ctx, kernel):
        got = t.programmable_bootstrap_lwe_ciphertext_without_sample_extract(cts.astype(_dt(o)), lut.astype(_dt(o)), bsk)
    got = got.astype(np.uint64).reshape(B, o.k + 1, N)
    lutp = lut.reshape(o.k + 1, N)
    half = np.uint64(1 << (bits - o.b * o.l - 1))
    mm = np.uint64(m)
    for b in range(B):
        total = -int(ahat[b, n]) + int(np.sum(ahat[b, :n] * s))
        for j in range(o.k + 1):
            want = _rot(lutp[j], total, N, bits)
            rounded = (((got[b, j] + half) & mm) // grid) * grid & mm
            assert np.array_equal(rounded, want), (name, kernel, b, j)
            err = np.abs(_sdiff(got[b, j], want, bits)).max()
            assert err <= n * 64 * max(1.0, 2.0 ** (bits - 64 + 17)), (name, kernel, b, j, err)   # accumulated FFT rounding: far below the grid


class _Shape:
    pass


def _shape_only(name):
    """Parameter set without generated keys (the noiseless test builds its own key)."""
    from oracle.oracle import PARAMS
    t = PARAMS[name]
    o = _Shape()
    o.n, o.k, o.N, o.b, o.l, o.b_ks, o.l_ks, o.bits = t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7]
    o.glwe_len = (o.k + 1) * o.N
    return o
