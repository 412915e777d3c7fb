"""Committed golden vectors (tests/golden/ks_pbs_golden.npz, written by tests/golden/make_golden.py).

CPU half (`-m "not gpu"`): the oracle, loaded with the STORED keys, still reproduces every stored output -- integer stages
bit for bit, PBS outputs by decrypted message and by phase within the analytic bound (the f64 FFT of a rebuilt oracle may
round differently on another host).  GPU half (`-m gpu`): the CUDA path through the C ABI on the same stored bytes.
"""
import os

import numpy as np
import pytest

from conftest import params_of, phase_bound

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ks_pbs_golden.npz")
SHORTINT_SETS = ["GOLD_2_2", "GOLD_U64"]


@pytest.fixture(scope="module")
def z():
    return np.load(GOLD)


def _oracle(z, name):
    from oracle.oracle import Oracle
    return Oracle(name).load_keys(z[f"{name}.small_key"], z[f"{name}.glwe_key"], z[f"{name}.ksk"], z[f"{name}.bsk"])


def _phase_close(o, got_phase, ref_phase, bits):
    d = (got_phase.astype(np.uint64) - ref_phase.astype(np.uint64))
    if bits == 32:
        d = (d & np.uint64(0xFFFFFFFF)).astype(np.uint32).astype(np.int32).astype(np.float64)
    else:
        d = d.astype(np.int64).astype(np.float64)
    return np.abs(d).max() / 2.0 ** bits < phase_bound(o)


# ---- CPU: oracle vs stored vectors ---------------------------------------------------------------------------------
@pytest.mark.parametrize("name", SHORTINT_SETS)
def test_oracle_reproduces_golden_shortint(z, name):
    from oracle.oracle import lib
    o = _oracle(z, name)
    cts = z[f"{name}.cts"]
    ks = o.keyswitch(cts)
    assert np.array_equal(ks, z[f"{name}.ks_out"])
    ms = np.array([[lib().orc_modswitch(int(v), int(np.log2(o.N)), o.bits) for v in row] for row in ks], np.uint32)
    assert np.array_equal(ms, z[f"{name}.modswitch"])
    assert np.array_equal(np.stack([o.lut(lambda x: x), o.lut(lambda x: (3 * x + 1) % o.pmod)]), z[f"{name}.luts"])
    out = o.ks_pbs(cts, z[f"{name}.luts"], z[f"{name}.lut_idx"])
    ph = o.phase(out, "big")
    assert np.array_equal(o.decode(ph) % o.pmod, z[f"{name}.ks_pbs_msgs"])
    assert _phase_close(o, ph, z[f"{name}.ks_pbs_phase"], o.bits)
    # rows 2.. are honest encryptions: the stored messages are LUT(msg)
    msgs, idx = z[f"{name}.msgs"], z[f"{name}.lut_idx"]
    exp = np.where(idx == 0, msgs, (3 * msgs + 1) % o.pmod)
    assert np.array_equal(z[f"{name}.ks_pbs_msgs"][2:], exp[2:])


def test_oracle_reproduces_golden_boolean(z):
    name = "GOLD_U32"
    o = _oracle(z, name)
    a, b = z[f"{name}.a"].astype(np.uint64), z[f"{name}.b"].astype(np.uint64)
    pro = o.gate_prologue(1, a, b)
    assert np.array_equal(pro.astype(np.uint32), z[f"{name}.nand_prologue"])
    out = o.pbs_ks(pro, o.boolean_lut())
    bits = o.decode_bool(o.phase(out, "small"))
    assert np.array_equal(bits, z[f"{name}.nand_bits"])
    assert np.array_equal(bits, (1 - (z[f"{name}.a_bits"] & z[f"{name}.b_bits"])).astype(bool))


# ---- GPU: CUDA path vs stored vectors ------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", SHORTINT_SETS)
def test_cuda_path_matches_golden_shortint(ctx, z, name):
    import tfhe_rs_odd_b200 as t
    o = _oracle(z, name)
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, z[f"{name}.bsk"])
    ksk = t.LweKeyswitchKey.upload(ctx, p, z[f"{name}.ksk"])
    cts = z[f"{name}.cts"]
    ks = t.keyswitch_lwe_ciphertext(ksk, cts)
    assert np.array_equal(ks, z[f"{name}.ks_out"])                                     # bit-exact
    ms = t.fast_pbs_modulus_switch(ctx, ks, o.N, 64)
    assert np.array_equal(np.asarray(ms).reshape(ks.shape).astype(np.uint32), z[f"{name}.modswitch"])   # bit-exact
    out = t.keyswitch_programmable_bootstrap(ksk, bsk, cts, z[f"{name}.luts"], lut_idx=z[f"{name}.lut_idx"])
    ph = o.phase(out, "big")
    assert np.array_equal(o.decode(ph) % o.pmod, z[f"{name}.ks_pbs_msgs"])            # decrypted values equal
    assert _phase_close(o, ph, z[f"{name}.ks_pbs_phase"], o.bits)


@pytest.mark.gpu
def test_cuda_path_matches_golden_boolean(ctx, z):
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import boolean
    name = "GOLD_U32"
    o = _oracle(z, name)
    p = params_of(o)
    bsk = t.convert_standard_lwe_bootstrap_key_to_fourier(ctx, p, z[f"{name}.bsk"])
    ksk = t.LweKeyswitchKey.upload(ctx, p, z[f"{name}.ksk"])
    out = boolean.gate_batch(ksk, bsk, 1, z[f"{name}.a"], z[f"{name}.b"], pbs_first=True)
    bits = o.decode_bool(o.phase(out.astype(np.uint64), "small"))
    assert np.array_equal(bits, z[f"{name}.nand_bits"])
    ref_ph = o.phase(z[f"{name}.nand_out"].astype(np.uint64), "small")
    assert _phase_close(o, o.phase(out.astype(np.uint64), "small"), ref_ph, 32)


# ---- multi-bit PBS and the gadget engine's integer operations (tests/golden/next_rows_golden.npz) -----------
GOLD_NEXT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "next_rows_golden.npz")


@pytest.fixture(scope="module")
def zn():
    return np.load(GOLD_NEXT)


def _oracle_pk(zn):
    from oracle.oracle import Oracle
    o = Oracle("GOLD_PK")
    o.glwe_key = zn["GOLD_PK.glwe_key"].astype(np.uint64)
    return o


def test_oracle_reproduces_golden_multibit(zn):
    name = "GOLD_MB"
    o = _oracle(zn, name)
    assert o.grouping == 2 and o.bsk.size == (o.n // 2 * 4) * o.l * (o.k + 1) ** 2 * o.N
    out = o.ks_pbs(zn[f"{name}.cts"], zn[f"{name}.luts"], zn[f"{name}.lut_idx"])
    ph = o.phase(out, "big")
    assert np.array_equal(o.decode(ph) % o.pmod, zn[f"{name}.ks_pbs_msgs"])
    assert _phase_close(o, ph, zn[f"{name}.ks_pbs_phase"], 64)
    msgs, idx = zn[f"{name}.msgs"], zn[f"{name}.lut_idx"]
    assert np.array_equal(zn[f"{name}.ks_pbs_msgs"], np.where(idx == 0, msgs, (3 * msgs + 1) % o.pmod))


def test_oracle_reproduces_golden_gadget_integer_ops(zn):
    o = _oracle_pk(zn)
    pksk, lwes = zn["GOLD_PK.pksk"], zn["GOLD_PK.lwes"]
    glwes = o.packing_keyswitch(pksk, lwes)
    assert np.array_equal(glwes, zn["GOLD_PK.pks_out"])
    assert np.array_equal(o.glwe_mul_plain_extract(glwes, zn["GOLD_PK.polys"]), zn["GOLD_PK.mul_extract_out"])
    assert np.array_equal(o.pack_into_accumulator(pksk, lwes[2:4], 3), zn["GOLD_PK.packed_p3_count2"])
    assert np.array_equal(o.pack_into_accumulator(pksk, lwes[:5], 5), zn["GOLD_PK.packed_p5_count5"])


@pytest.mark.gpu
def test_cuda_path_matches_golden_multibit(ctx, zn):
    import tfhe_rs_odd_b200 as t
    name = "GOLD_MB"
    o = _oracle(zn, name)
    p = params_of(o)
    bsk = t.core_crypto.convert_standard_lwe_multi_bit_bootstrap_key_to_fourier(ctx, p, zn[f"{name}.bsk"], 2)
    ksk = t.LweKeyswitchKey.upload(ctx, p, zn[f"{name}.ksk"])
    out = t.keyswitch_programmable_bootstrap(ksk, bsk, zn[f"{name}.cts"], zn[f"{name}.luts"], lut_idx=zn[f"{name}.lut_idx"])
    ph = o.phase(out, "big")
    assert np.array_equal(o.decode(ph) % o.pmod, zn[f"{name}.ks_pbs_msgs"])
    assert _phase_close(o, ph, zn[f"{name}.ks_pbs_phase"], 64)


@pytest.mark.gpu
def test_cuda_path_matches_golden_gadget_integer_ops(ctx, zn):
    """Bit for bit against the stored vectors: packing keyswitch (tensor-core path, N = 64: ragged against every tile size),
    GLWE x plaintext polynomial + extraction, pack_into_new_accumulator."""
    import tfhe_rs_odd_b200 as t
    from tfhe_rs_odd_b200 import gadget
    o = _oracle_pk(zn)
    p = params_of(o)
    pksk = gadget.LwePackingKeyswitchKey.upload(ctx, p, zn["GOLD_PK.pksk"])
    lwes = zn["GOLD_PK.lwes"]
    glwes = gadget.keyswitch_lwe_ciphertext_into_glwe_ciphertext(pksk, lwes)
    assert np.array_equal(glwes, zn["GOLD_PK.pks_out"])
    assert np.array_equal(gadget.glwe_mul_plain_extract(ctx, p, zn["GOLD_PK.pks_out"], zn["GOLD_PK.polys"]), zn["GOLD_PK.mul_extract_out"])

    class _Key:                     # Bootstrapper only needs params / ctx of the bootstrap key for the packing entry point
        params, ctx = p, None
    k = _Key()
    k.ctx = ctx
    bs = gadget.Bootstrapper(k, None, pksk)
    assert np.array_equal(bs.pack_into_new_accumulator(lwes[2:4], 3), zn["GOLD_PK.packed_p3_count2"])
    assert np.array_equal(bs.pack_into_new_accumulator(lwes[:5], 5), zn["GOLD_PK.packed_p5_count5"])


# ---- WoP-PBS row: private functional packing keyswitch with the circuit-bootstrap key list (tests/golden/wopbs_golden.npz) ----
GOLD_WOP = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wopbs_golden.npz")


@pytest.fixture(scope="module")
def zw():
    return np.load(GOLD_WOP)


def test_oracle_reproduces_golden_pfpks(zw):
    from oracle.oracle import Oracle
    o = Oracle("GOLD_PK")
    o.glwe_key = zw["GOLD_PK.glwe_key"].astype(np.uint64)
    pb, pl = int(zw["pf_base_log"]), int(zw["pf_level"])
    pf, lwes = zw["GOLD_PK.pfpksk_list"], zw["GOLD_PK.lwes"]
    key_len = (o.big + 1) * pl * o.glwe_len
    assert pf.size == (o.k + 1) * key_len
    for r in range(o.k + 1):
        assert np.array_equal(o.pfpks(pf[r * key_len:(r + 1) * key_len], pb, pl, lwes), zw["GOLD_PK.pfpks_out"][r])
    # meaning of the stored vectors (lwe_wopbs.rs:128-157): an encryption of 5 * 2^60 goes to -m * S_r for r < k and to +m for r = k
    m, S = 5 << 60, o.glwe_key.reshape(o.k, o.N)
    ph = zw["GOLD_PK.enc5_phases"]
    for r in range(o.k + 1):
        want = np.zeros(o.N, dtype=object)
        if r < o.k:
            want[:] = [(-m * int(s)) % (1 << 64) for s in S[r]]
        else:
            want[0] = m
        d = np.array([((int(a) - int(b) + (1 << 63)) % (1 << 64)) - (1 << 63) for a, b in zip(ph[r], want)], dtype=np.float64)
        assert np.abs(d).max() < 2.0 ** 50


@pytest.mark.gpu
def test_cuda_path_matches_golden_pfpks(ctx, zw):
    """Bit for bit: the tensor-core GEMM over exactly split base-2^15 digits against the stored outputs of the reference loop."""
    import tfhe_rs_odd_b200 as t
    from oracle.oracle import Oracle
    o = Oracle("GOLD_PK")
    p = params_of(o)
    pf = t.core_crypto.LwePrivateFunctionalPackingKeyswitchKeyList.upload(ctx, p, int(zw["pf_base_log"]), int(zw["pf_level"]),
                                                                         zw["GOLD_PK.pfpksk_list"])
    for r in range(o.k + 1):
        got = t.core_crypto.private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext(pf, r, zw["GOLD_PK.lwes"])
        assert np.array_equal(got, zw["GOLD_PK.pfpks_out"][r])
