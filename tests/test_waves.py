"""Device-resident waves: the OpGraph scheduler (CPU), a radix addition with sequential carry
propagation and the homomorphic Trivium keystream (GPU), all checked against plain arithmetic / the eSTREAM vectors the
reference's own tests use (apps/trivium/src/trivium/test.rs:80-275)."""
import ctypes as C

import numpy as np
import pytest

from conftest import get_oracle, params_of


class _FakePool:
    capacity, n_luts = 64, 4


def test_wave_op_layout_matches_header():
    from tfhe_rs_odd_b200.waves import WaveOp
    assert C.sizeof(WaveOp) == 72
    assert WaveOp.scale.offset == 24 and WaveOp.plaintext.offset == 56 and WaveOp.lut.offset == 64


def test_opgraph_levels_respect_hazards():
    from tfhe_rs_odd_b200.waves import OpGraph
    from tfhe_rs_odd_b200 import B200TfheError
    g = OpGraph(_FakePool())
    assert g.op(10, [(0, 1), (1, 1)], lut=0) == 1            # independent of everything
    assert g.op(11, [(2, 2), (3, 1)], lut=1) == 1
    assert g.op(12, [(10, 1), (11, 1)], lut=0) == 2          # RAW on 10 and 11
    assert g.op(0, [(5, 1)]) == 1                            # WAR: slot 0 is read in wave 1 -> may be overwritten at its end
    assert g.op(10, [(6, 1)]) == 2                           # WAW / WAR: 10 is written in wave 1 and read in wave 2
    assert g.op(13, [(10, 1)], lut=2) == 3                   # sees the wave-2 value of slot 10
    assert g.op(14, [(7, 1), (7, 2)]) == 1                   # duplicate sources merge into one term
    assert g.ops[-1][2] == [(7, 3)]
    assert [len(w) for w in g.waves()] == [4, 2, 1]
    with pytest.raises(B200TfheError):
        g.op(15, [(i, 1) for i in range(6)])                 # more than five terms
    with pytest.raises(B200TfheError):
        g.op(15, [(i, 2 ** 40) for i in range(5)])           # five terms: one multiplier has to fit 32 bits
    g.op(15, [(0, 2 ** 40), (1, 1), (2, 2 ** 41), (3, -(2 ** 40)), (4, 2 ** 42)])
    assert g.ops[-1][2][-1] == (1, 1)                        # ... and it is moved to the fifth place
    with pytest.raises(B200TfheError):
        g.op(64, [(0, 1)])                                   # slot out of range
    with pytest.raises(B200TfheError):
        g.op(15, [(0, 1)], lut=4)                            # accumulator out of range


def test_plain_trivium_model_matches_estream_vectors():
    from tfhe_rs_odd_b200.trivium import trivium_reference_bits, hex_lsb_first
    key, iv = [0] * 80, [0] * 80
    h = hex_lsb_first(trivium_reference_bits(key, iv, 512 * 8))
    assert h[:128] == ("FBE0BF265859051B517A2E4E239FC97F563203161907CF2DE7A8790FA1B2E9CDF75292030268B7382B4C1A759AA2599A"
                       "285549986E74805903801A4CB5A5D4F2")
    assert h[448 * 2:448 * 2 + 32] == "68450EB0910A98EF1853E0FC1BED8AB6"
    key[7] = 1
    assert hex_lsb_first(trivium_reference_bits(key, iv, 64 * 8))[:40] == "38EB86FF730D7A9CAF8DF13A4420540DBB7B6514"


def _server_key(ctx, o):
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import shortint
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk)
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk)
    mm = int(round(o.pmod ** 0.5))
    return p, ksk, bsk, shortint.ServerKey(ksk, bsk, mm, mm, "ks_pbs")


def _bits_of_hex(h):
    """kreyvium/test.rs:135-156: byte i of the string -> bits 8i .. 8i+7, least significant first."""
    return [(int(h[i:i + 2], 16) >> j) & 1 for i in range(0, len(h), 2) for j in range(8)]


def test_plain_kreyvium_model_matches_reference_vectors():
    """apps/trivium/src/kreyvium/test.rs:80-168 (values from renaud1239/Kreyvium)."""
    from tfhe_rs_odd_b200.trivium import hex_lsb_first, kreyvium_reference_bits
    z = [0] * 128
    one = [1] + [0] * 127
    assert hex_lsb_first(kreyvium_reference_bits(z, z, 64)) == "26DCF1F4BC0F1922"
    assert hex_lsb_first(kreyvium_reference_bits(one, z, 64)) == "4FD421D4DA3D2C8A"
    assert hex_lsb_first(kreyvium_reference_bits(z, one, 64)) == "C9217BA0D762ACA1"
    key, iv = _bits_of_hex("0053A6F94C9FF24598EB000000000000"), _bits_of_hex("0D74DB42A91077DE45AC000000000000")
    assert hex_lsb_first(kreyvium_reference_bits(key, iv, 64)) == "D1F0303482061111"


@pytest.mark.gpu
def test_kreyvium_bool_keystream_on_device(ctx):
    """KreyviumStream<FheBool> (apps/trivium/src/kreyvium/kreyvium.rs): 15 gate bootstraps per step in three waves; stream 0 is the
    reference's test vector 4, stream 1 random; 128 keystream bits so that both halves of the rotating key / IV registers are used."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200.trivium import BOOL_FALSE, BOOL_TRUE, KreyviumStreamBool, hex_lsb_first, kreyvium_reference_bits
    o = get_oracle("BOOLEAN_DEFAULT")
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk.astype(np.uint32))
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk.astype(np.uint32))
    rng = np.random.default_rng(5)
    S = 2
    keys, ivs = rng.integers(0, 2, (S, 128)), rng.integers(0, 2, (S, 128))
    keys[0], ivs[0] = _bits_of_hex("0053A6F94C9FF24598EB000000000000"), _bits_of_hex("0D74DB42A91077DE45AC000000000000")
    pts = np.where(keys.reshape(-1) == 1, BOOL_TRUE, BOOL_FALSE).astype(np.uint64)
    key_cts = o.encrypt(pts, "small", seed=0x7719).astype(np.uint32).reshape(S, 128, -1)
    kr = KreyviumStreamBool(ctx, p, ksk, bsk, o.boolean_lut().astype(np.uint32), key_cts, ivs)
    out = np.concatenate([kr.next_64(), kr.next_64()], axis=1)
    assert all(w == (64 * S * g, 64 * S * g) for batch in kr.waves_run for w, g in zip(batch, (6, 5, 4)))
    for s in range(S):
        bits = o.decode_bool(o.phase(out[s].astype(np.uint64), "small")).astype(np.int64)
        assert np.array_equal(bits, kreyvium_reference_bits(keys[s], ivs[s], 128)), s
    bits0 = o.decode_bool(o.phase(out[0, :64].astype(np.uint64), "small")).astype(np.int64)
    assert hex_lsb_first(list(bits0)) == "D1F0303482061111"
    kr.pool.close()


@pytest.mark.gpu
def test_radix_addition_with_carry_propagation(ctx):
    """16-bit radix integers (8 blocks of 2 message bits at PARAM_MESSAGE_2_CARRY_2): block-wise unchecked_add, then the
    sequential carry propagation of the reference (message_extract / carry_extract per block, server_key/mod.rs:539-623):
    8 dependent waves, nothing leaves the device in between."""
    from tfhe_rs_odd_b200.waves import CiphertextPool, OpGraph
    o = get_oracle("MESSAGE_2_CARRY_2")
    p, ksk, bsk, sk = _server_key(ctx, o)
    M, NB = 24, 8
    rng = np.random.default_rng(9)
    x, y = rng.integers(0, 1 << 16, M), rng.integers(0, 1 << 16, M)
    blocks = lambda v: np.stack([(v >> (2 * i)) & 3 for i in range(NB)], axis=1)        # [M][NB], least significant first
    pool = CiphertextPool(ctx, p, capacity=4 * M * NB, big=True)
    pool.set_luts(np.stack([sk.generate_lookup_table(lambda z: z % 4).acc, sk.generate_lookup_table(lambda z: z // 4).acc]))
    X, Y, S, CARRY = 0, M * NB, 2 * M * NB, 3 * M * NB
    pool.write(X, o.encrypt(o.encode(blocks(x).reshape(-1)), "big", seed=0x51))
    pool.write(Y, o.encrypt(o.encode(blocks(y).reshape(-1)), "big", seed=0x52000))
    g = OpGraph(pool)
    for m in range(M):
        for i in range(NB):
            e = m * NB + i
            terms = [(X + e, 1), (Y + e, 1)] + ([(CARRY + e - 1, 1)] if i else [])
            g.op(S + e, terms, lut=0)                        # message_extract(x_i + y_i + carry_{i-1})
            g.op(CARRY + e, terms, lut=1)                    # carry_extract(...)
    waves = g.run(ksk, bsk)
    assert len(waves) == NB and all(w == (2 * M, 2 * M) for w in waves)
    got = o.decode(o.phase(pool.read(S, M * NB), "big")).reshape(M, NB) % 4
    total = sum(got[:, i].astype(np.int64) << (2 * i) for i in range(NB))
    assert np.array_equal(total, (x + y) % (1 << 16))
    pool.close()


@pytest.mark.gpu
def test_trivium_keystream_on_device(ctx):
    """BASELINE configs[4]: homomorphic Trivium at PARAM_MESSAGE_1_CARRY_1, 2 streams with different keys / IVs; the
    1152 warm-up steps and 64 keystream bits are 19 x 2 device-resident waves; decrypted bits equal plain Trivium."""
    from tfhe_rs_odd_b200.trivium import TriviumStreamShortint, trivium_reference_bits
    o = get_oracle("MESSAGE_1_CARRY_1")
    p, ksk, bsk, sk = _server_key(ctx, o)
    rng = np.random.default_rng(3)
    S = 2
    keys, ivs = rng.integers(0, 2, (S, 80)), rng.integers(0, 2, (S, 80))
    keys[0, :] = 0; ivs[0, :] = 0                        # stream 0 = the all-zero eSTREAM vector of the reference's test 1
    key_cts = o.encrypt(o.encode(keys.reshape(-1)), "big", seed=0x7717).reshape(S, 80, -1)
    luts = np.stack([sk.generate_lookup_table(lambda z: (z >> 1) & z & 1).acc, sk.generate_lookup_table(lambda z: z % 2).acc])
    tr = TriviumStreamShortint(ctx, p, ksk, bsk, luts, key_cts, ivs, encode_bit=lambda b: int(o.encode([b])[0]))
    out = tr.next_64()
    assert all(w == (64 * S * 4, 64 * S * 4) for batch in tr.waves_run for w in batch) and len(tr.waves_run) == 19
    for s in range(S):
        bits = o.decode(o.phase(out[s], "big")) % 2
        assert np.array_equal(bits, trivium_reference_bits(keys[s], ivs[s], 64)), s
    tr.pool.close()


@pytest.mark.gpu
def test_trivium_bool_keystream_on_device(ctx):
    """The FheBool variant (apps/trivium/src/trivium/trivium_bool.rs): 14 boolean gate bootstraps per step in three waves at
    boolean DEFAULT_PARAMETERS; decrypted bits equal plain Trivium (stream 0 = the all-zero eSTREAM vector)."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200.trivium import BOOL_FALSE, BOOL_TRUE, TriviumStreamBool, trivium_reference_bits
    o = get_oracle("BOOLEAN_DEFAULT")
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, o.bsk.astype(np.uint32))
    ksk = t.LweKeyswitchKey.upload(ctx, p, o.ksk.astype(np.uint32))
    rng = np.random.default_rng(4)
    S = 2
    keys, ivs = rng.integers(0, 2, (S, 80)), rng.integers(0, 2, (S, 80))
    keys[0, :] = 0; ivs[0, :] = 0
    pts = np.where(keys.reshape(-1) == 1, BOOL_TRUE, BOOL_FALSE).astype(np.uint64)
    key_cts = o.encrypt(pts, "small", seed=0x7718).astype(np.uint32).reshape(S, 80, -1)
    tr = TriviumStreamBool(ctx, p, ksk, bsk, o.boolean_lut().astype(np.uint32), key_cts, ivs)
    out = tr.next_64()
    assert len(tr.waves_run) == 19 and all(w == (64 * S * g, 64 * S * g) for batch in tr.waves_run for w, g in zip(batch, (6, 4, 4)))
    for s in range(S):
        bits = o.decode_bool(o.phase(out[s].astype(np.uint64), "small")).astype(np.int64)
        assert np.array_equal(bits, trivium_reference_bits(keys[s], ivs[s], 64)), s
    tr.pool.close()


@pytest.mark.gpu
def test_five_term_linear_wave_is_bit_exact(ctx):
    """A wave operation takes up to five terms (the chunk of a radix column sum at MESSAGE_2_CARRY_2); the fifth one travels in the
    two formerly reserved words of b200tfhe_wave_op.  Linear operations (no bootstrap) are wrapping u64 arithmetic: bit-exact."""
    from tfhe_rs_odd_b200.waves import CiphertextPool, OpGraph
    o = get_oracle("MESSAGE_2_CARRY_2")
    p = params_of(o)
    pool = CiphertextPool(ctx, p, capacity=16, big=True)
    rng = np.random.default_rng(31)
    cts = rng.integers(0, 2 ** 64, (8, pool.len), dtype=np.uint64)
    pool.write(0, cts)
    g = OpGraph(pool)
    scales = [3, -5, 2 ** 40 + 1, -(2 ** 33), 7]
    g.op(8, [(i, c) for i, c in enumerate(scales)], plaintext=0x1234)
    g.op(9, [(i, 1) for i in range(3, 8)])
    g.op(10, [(0, 1), (1, 1), (2, 1), (3, 1)])               # four terms: the fifth multiplier stays zero
    g.run(None, None)
    got = pool.read(8, 3)
    want = np.zeros((3, pool.len), np.uint64)
    with np.errstate(over="ignore"):
        for i, c in enumerate(scales):
            want[0] += np.uint64(c % 2 ** 64) * cts[i]
        want[0, -1] += np.uint64(0x1234)
        want[1] = cts[3:8].sum(axis=0, dtype=np.uint64)
        want[2] = cts[0:4].sum(axis=0, dtype=np.uint64)
    assert np.array_equal(got, want)
    pool.close()


@pytest.mark.gpu
def test_radix_multiplication_on_device(ctx):
    """BASELINE configs[3] at full width: FheUint32 = 16 blocks at PARAM_MESSAGE_2_CARRY_2, schoolbook product in the
    structure of unchecked_mul_parallelized (radix_parallel/mul.rs:329-424): one wave of bivariate block products, then
    column-sum waves; result equals x * y mod 2^32 for every pair.  The circuit is compiled once into a wave plan and
    replayed by one C call (b200tfhe_wave_sequence_run)."""
    from tfhe_rs_odd_b200.waves import CiphertextPool, OpGraph
    from tfhe_rs_odd_b200 import radix
    o = get_oracle("MESSAGE_2_CARRY_2")
    p, ksk, bsk, sk = _server_key(ctx, o)
    M, NB = 4, 16
    rng = np.random.default_rng(21)
    x, y = rng.integers(0, 1 << 32, M), rng.integers(0, 1 << 32, M)
    x[0], y[0] = 0xFFFFFFFF, 0xFFFFFFFF
    x[1], y[1] = 0, 0x12345678
    blocks = lambda v: np.stack([(v >> (2 * i)) & 3 for i in range(NB)], axis=1)
    pool = CiphertextPool(ctx, p, capacity=M * 1200, big=True)
    pool.set_luts(np.stack(radix.radix_luts(sk)))
    X, Y, OUT = 0, M * NB, 2 * M * NB
    pool.write(X, o.encrypt(o.encode(blocks(x).reshape(-1)), "big", seed=0x61))
    pool.write(Y, o.encrypt(o.encode(blocks(y).reshape(-1)), "big", seed=0x62000))
    g = OpGraph(pool)
    alloc = radix.SlotAllocator(3 * M * NB, pool.capacity)
    for m in range(M):
        radix.mul(g, alloc, X + m * NB, Y + m * NB, OUT + m * NB, NB)
    plan = g.compile()
    waves = plan.run(ksk, bsk)
    assert waves[0] == (M * (136 + 120), M * (136 + 120))    # the block-product wave: nb(nb+1)/2 lsb + nb(nb-1)/2 msb terms = 256
    assert sum(w[1] for w in waves) == M * 547 and len(waves) == 17   # five-term column chunks (+ the final copy wave)
    got = o.decode(o.phase(pool.read(OUT, M * NB), "big")).reshape(M, NB) % 4
    total = sum(got[:, i].astype(np.int64) << (2 * i) for i in range(NB))
    assert np.array_equal(total, (x * y) % (1 << 32))
    # replaying the compiled plan on fresh operands gives the new product (the plan holds slots, not values)
    x2, y2 = rng.integers(0, 1 << 32, M), rng.integers(0, 1 << 32, M)
    pool.write(X, o.encrypt(o.encode(blocks(x2).reshape(-1)), "big", seed=0x63))
    pool.write(Y, o.encrypt(o.encode(blocks(y2).reshape(-1)), "big", seed=0x64000))
    plan.run(ksk, bsk)
    got = o.decode(o.phase(pool.read(OUT, M * NB), "big")).reshape(M, NB) % 4
    total = sum(got[:, i].astype(np.int64) << (2 * i) for i in range(NB))
    assert np.array_equal(total, (x2 * y2) % (1 << 32))
    pool.close()
