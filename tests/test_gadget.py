"""Gadget extras of the fork (reference tfhe/src/gadget/engine/bootstrapping.rs): packing keyswitch,
multi-value bootstrap, accumulator packing.  Integer stages are bit-exact against the oracle restatement (oracle/orc_gadget.c);
the multi-value bootstrap is additionally decrypt-checked against the plain functions for every message."""
import numpy as np
import pytest

from conftest import get_oracle, params_of

U64 = 1 << 64


def _gadget():
    import tfhe_rs_odd_b200 as t  # noqa: F401
    from tfhe_rs_odd_b200 import gadget
    return gadget


# ---- CPU: oracle and host tables ----------------------------------------------------------------------------------------
def test_oracle_packing_keyswitch_decrypts():
    o = get_oracle("TOY_U64")
    pksk = o.gen_pksk()
    msgs = np.arange(4)
    g = o.packing_keyswitch(pksk, o.encrypt(o.encode(msgs), under="big"))
    for m, x in zip(msgs, g):
        ph = o.glwe_phase(x)
        assert int(o.decode(ph[:1])[0]) == m                                   # constant coefficient carries the message
        assert np.abs(ph[1:].astype(np.int64)).max() < 2.0 ** 40               # every other coefficient is noise only


def test_oracle_glwe_mul_plain_extract_matches_definition():
    o = get_oracle("TOY_U64")
    rng = np.random.default_rng(2)
    N, k = o.N, o.k
    glwe = rng.integers(0, U64, (k + 1) * N, dtype=np.uint64)
    v = np.zeros(N, np.uint64)
    v[[3, 100, 511]] = [5, U64 - 2, 7]
    got = o.glwe_mul_plain_extract(glwe, v)[0, 0]
    polys = [[int(x) for x in glwe[j * N:(j + 1) * N]] for j in range(k + 1)]
    vv = [int(x) for x in v]

    def prod_coef(a, c):
        s = 0
        for t in (3, 100, 511):
            d = c - t
            s += vv[t] * (a[d] if d >= 0 else -a[N + d])
        return s % U64
    want = []
    for j in range(k):
        want += [prod_coef(polys[j], 0)] + [(-prod_coef(polys[j], N - t)) % U64 for t in range(1, N)]
    want.append(prod_coef(polys[k], 0))
    assert [int(x) for x in got] == want


def test_mvb_tables_rebuild_the_accumulator():
    """Noise-free identity behind mvb_bootstrap: (common factor * X^{-mu}) * v_i, coefficient 0, equals the accumulator of
    (enc_in after f_i -> enc_out) read at mu, for mu in the middle of every window."""
    gadget = _gadget()
    import tfhe_rs_odd_b200 as t
    params = t.core_crypto.make_params(16, 1, 512, 12, 2, 4, 4, 64)
    N = 512
    p, pp = 5, 7
    enc_in = gadget.Encoding.new_canonical(5, [0, 1, 2, 3, 4], p)
    enc_out = gadget.Encoding.new_canonical(5, [0, 1, 2, 3, 4], pp)
    f = [3, 0, 4, 4, 1]
    inter = enc_in.apply_lut_to_encoding(lambda x: f[x])
    v = [int(x) for x in gadget.create_vi_for_mvb(params, inter, enc_out)]
    acc = [int(x) for x in gadget.accumulator_glwe(params, inter, enc_out)[N:]]
    unit = U64 // pp
    for z in range(p):
        mu = (2 * N * z) // p                     # modulus-switched phase of the encoding of z (window centre)
        # T = unit * (1 + X + ... + X^{N-1}); coefficient 0 of T * X^a (a in [0, 2N)) is +unit for a = 0 or a > N, else -unit
        s = 0
        for pos, val in enumerate(v):
            if val == 0:
                continue
            a = (pos - mu) % (2 * N)
            sign = 1 if a == 0 else (-1 if a <= N else 1)
            s += sign * val * unit
        want_idx = mu % N
        want = acc[want_idx] if mu < N else (-acc[want_idx]) % U64
        assert abs(((s - want + U64 // 2) % U64) - U64 // 2) <= 4 * pp, (z, s % U64, want)


def test_encoding_mirror_basics():
    gadget = _gadget()
    e = gadget.Encoding.new_canonical(3, [0, 1, 2], 5)
    assert e.is_canonical() and e.get_modulus() == 5 and e.inverse_encoding(2) == 2 and e.inverse_encoding(4) is None
    g = e.apply_lut_to_encoding(lambda x: [1, 1, 0][x])
    assert g.parts[1] == frozenset({0, 1}) and g.parts[0] == frozenset({2}) and not g.is_canonical()
    assert gadget.create_accumulator(e, gadget.Encoding.new_canonical(3, [0, 1, 2], 7)) == [0, 0, 1, 0, 2]
    with pytest.raises(Exception):
        gadget.Encoding(2, [{0}, {9}], 5)


# ---- GPU ------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name,B", [("TOY_U64", 37), ("TOY_1_1", 5), ("TOY_U64", 300)])
def test_gpu_packing_keyswitch_bit_exact(ctx, name, B):
    gadget = _gadget()
    o = get_oracle(name)
    p = params_of(o)
    host = o.gen_pksk()
    pksk = gadget.LwePackingKeyswitchKey.upload(ctx, p, host)
    rng = np.random.default_rng(4)
    cts = rng.integers(0, U64, size=(B, o.big + 1), dtype=np.uint64)
    cts[0, :] = 0
    cts[1, :] = U64 - 1
    cts[2, :] = 1 << 63
    nref = min(B, 40)
    ref = o.packing_keyswitch(host, cts[:nref])
    got = gadget.keyswitch_lwe_ciphertext_into_glwe_ciphertext(pksk, cts)
    assert np.array_equal(got[:nref], ref)
    one = gadget.keyswitch_lwe_ciphertext_into_glwe_ciphertext(pksk, cts[3])
    assert np.array_equal(one, ref[3])
    if B > nref:                                  # rows beyond the oracle prefix: same kernel on a sub-batch
        assert np.array_equal(gadget.keyswitch_lwe_ciphertext_into_glwe_ciphertext(pksk, cts[nref:]), got[nref:])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["TOY_U64", "TOY_1_1", "TOY_2_2"])
def test_gpu_glwe_mul_plain_extract_bit_exact(ctx, name):
    gadget = _gadget()
    o = get_oracle(name)
    p = params_of(o)
    rng = np.random.default_rng(6)
    glwe = rng.integers(0, U64, size=(7, o.glwe_len), dtype=np.uint64)
    polys = np.zeros((3, o.N), np.uint64)
    polys[0, rng.choice(o.N, 5, replace=False)] = rng.integers(1, 7, 5)
    polys[1, [0, o.N - 1]] = [U64 - 1, 3]
    polys[2] = rng.integers(0, U64, o.N, dtype=np.uint64)        # dense: same code path, nnz = N
    assert np.array_equal(gadget.glwe_mul_plain_extract(ctx, p, glwe, polys), o.glwe_mul_plain_extract(glwe, polys))


@pytest.mark.gpu
@pytest.mark.parametrize("pmod,count", [(3, 3), (5, 2), (7, 7)])
def test_gpu_pack_into_accumulator_bit_exact(ctx, pmod, count):
    gadget = _gadget()
    import tfhe_rs_odd_b200 as t
    o = get_oracle("TOY_U64")
    p = params_of(o)
    host = o.gen_pksk()
    pksk = gadget.LwePackingKeyswitchKey.upload(ctx, p, host)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
    bs = gadget.Bootstrapper(bsk, None, pksk)
    rng = np.random.default_rng(8)
    B = 3
    enc = gadget.Encoding.new_canonical(pmod, list(range(pmod)), pmod)
    vals = rng.integers(0, pmod, (B, count))
    lwes = np.stack([o.encrypt(np.array([enc.encode(int(v)) for v in row], np.uint64), under="big", seed=100 + 17 * b)
                     for b, row in enumerate(vals)])
    got = bs.pack_into_new_accumulator(lwes, pmod)
    for b in range(B):
        assert np.array_equal(got[b], o.pack_into_accumulator(host, lwes[b], pmod))
    assert np.array_equal(bs.pack_into_new_accumulator(lwes[1], pmod), got[1])
    # semantics: window kk of the packed accumulator decrypts to element kk (even: value kk/2; odd: minus value (p+1)/2+(kk-1)/2)
    ph = o.glwe_phase(got[0])
    size = o.N // pmod
    for kk in range(1, pmod):
        src = kk // 2 if kk % 2 == 0 else (pmod + 1) // 2 + (kk - 1) // 2
        want = 0 if src >= count else (int(vals[0, src]) if kk % 2 == 0 else (-int(vals[0, src])) % pmod)
        c = size // 2 + (kk - 1) * size + size // 2
        assert int(round(int(ph[c]) * pmod / U64)) % pmod == want
    with pytest.raises(t.B200TfheError, match="odd"):
        bs.pack_into_new_accumulator(lwes[0], 4)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["TOY_U64", "TOY_2_2"])
def test_gpu_multi_value_bootstrap(ctx, name):
    """mvb_bootstrap: one blind rotation, three functions.  (a) decrypts to f_i(m) for every m; (b) the integer
    post-processing is bit-identical to the oracle's applied to the same rotated GLWE; (c) decrypted parity with the oracle's
    own full pipeline."""
    gadget = _gadget()
    import tfhe_rs_odd_b200 as t
    o = get_oracle(name)
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk)
    bs = gadget.Bootstrapper(bsk, ksk, None, "ks_pbs")
    pin, pout = 5, 7
    enc_in = gadget.Encoding.new_canonical(pin, list(range(pin)), pin)
    encs_out = [gadget.Encoding.new_canonical(pin, [0, 1, 2, 3, 4], pout), gadget.Encoding.new_canonical(pin, [0, 2, 4, 6, 1], pout),
                gadget.Encoding.new_canonical(pin, [3, 1, 4, 0, 5], pout)]
    luts = [[3, 0, 4, 4, 1], [0, 1, 2, 3, 4], [2, 2, 2, 0, 1]]
    msgs = np.arange(2 * pin) % pin
    pts = np.array([enc_in.encode(int(m)) for m in msgs], np.uint64)
    small = o.encrypt(pts, under="small", std=1e-9)
    out = bs.mvb_bootstrap(small, enc_in, encs_out, luts, keyswitch_first=False)
    assert out.shape == (len(msgs), 3, o.big + 1)
    for b, m in enumerate(msgs):
        for i in range(3):
            assert encs_out[i].decode_phase(o.phase(out[b, i])[0]) == luts[i][m], (b, i)
    # (b) exact integer parity of the post-processing on the GPU's own rotated GLWE
    acc = gadget.common_factor_accumulator(p, encs_out[0])
    polys = np.stack([gadget.create_vi_for_mvb(p, enc_in.apply_lut_to_encoding(lambda x, t_=tuple(l): t_[x]), e) for e, l in zip(encs_out, luts)])
    glwe = t.programmable_bootstrap_lwe_ciphertext_without_sample_extract(small, acc, bsk)
    assert np.array_equal(out, o.glwe_mul_plain_extract(glwe, polys))
    # (c) the oracle's own pipeline decrypts to the same values
    ref = o.glwe_mul_plain_extract(o.pbs(small[:4], acc, sample_extract=False), polys)
    for b in range(4):
        for i in range(3):
            assert encs_out[i].decode_phase(o.phase(ref[b, i])[0]) == encs_out[i].decode_phase(o.phase(out[b, i])[0])
    # Big key choice: inputs under the big key, keyswitched on the device first
    big = o.encrypt(pts, under="big")
    out2 = bs.mvb_bootstrap(big, enc_in, encs_out, luts)
    for b, m in enumerate(msgs):
        for i in range(3):
            assert encs_out[i].decode_phase(o.phase(out2[b, i])[0]) == luts[i][m]
    # single-function bootstrap through the same encodings (Bootstrapper::bootstrap)
    one = bs.bootstrap(big, enc_in.apply_lut_to_encoding(lambda x: luts[0][x]), encs_out[0])
    assert [encs_out[0].decode_phase(ph) for ph in o.phase(one)] == [luts[0][m] for m in msgs]
