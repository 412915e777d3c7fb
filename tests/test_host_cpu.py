"""CPU tests (-m "not gpu"): the C-ABI library loads and exports every declared symbol; host logic; the CUDA kernel's
mathematics executed lane-by-lane on the CPU (tests/native/emul.cpp compiles the very same stage functions) vs the oracle."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, get_oracle, phase_bound
from oracle.oracle import _p


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    return True


def test_library_exports_every_header_symbol(built):
    import tfhe_rs_odd_b200 as t
    hdr = open(os.path.join(ROOT, "include", "b200tfhe.h")).read()
    declared = set(re.findall(r"\b(b200tfhe_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 40
    lib = C.CDLL(t.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/b200tfhe.h but not exported"
    assert declared == set(t._native.EXPORTED_SYMBOLS), declared ^ set(t._native.EXPORTED_SYMBOLS)


def test_no_cpu_fallback_without_gpu(built):
    import torch
    import tfhe_rs_odd_b200 as t
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(t.B200TfheError, match="no CPU fallback"):
        t.Context(0)


def test_sass_is_sm100a_and_product_never_touches_oracle(built):
    import tfhe_rs_odd_b200 as t
    out = subprocess.run(["cuobjdump", "-lelf", t.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    pkg = os.path.join(ROOT, "tfhe-rs-odd_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                assert "oracle" not in open(os.path.join(dp, f)).read().replace("# oracle-free", ""), f


def test_shortint_lookup_table_matches_oracle_fill_accumulator(built):
    from tfhe_rs_odd_b200 import shortint
    from tfhe_rs_odd_b200.core_crypto import make_params

    class FakeKey:
        def __init__(self, p):
            self.params = p
    for name in ["MESSAGE_2_CARRY_2", "MESSAGE_1_CARRY_1", "MESSAGE_3_CARRY_3"]:
        from oracle.oracle import Oracle
        o = Oracle(name)
        sk = shortint.ServerKey.__new__(shortint.ServerKey)
        sk.params = make_params(o.n, o.k, o.N, o.b, o.l, o.b_ks, o.l_ks, 64)
        m = int(round(o.pmod ** 0.5))
        sk.message_modulus, sk.carry_modulus = m, o.pmod // m
        f = lambda x: (5 * x + 3) % o.pmod
        lut = sk.generate_lookup_table(f)
        assert np.array_equal(lut.acc, o.lut(f))
        assert lut.degree == max(f(i) for i in range(o.pmod))


def test_shortint_trivial_pbs_is_host_side(built):
    # trivial ciphertexts never reach the GPU (server_key/mod.rs:763-781): works without a device
    from tfhe_rs_odd_b200 import shortint
    from tfhe_rs_odd_b200.core_crypto import make_params
    sk = shortint.ServerKey.__new__(shortint.ServerKey)
    sk.params = make_params(742, 1, 2048, 23, 1, 3, 5, 64)
    sk.message_modulus = sk.carry_modulus = 4
    sk.pbs_order = "ks_pbs"
    lut = sk.generate_lookup_table(lambda x: (x + 7) % 16)
    delta = (1 << 63) // 16
    data = np.zeros((20, 2049), np.uint64)
    vals = np.arange(20)                         # 16..19 have the padding bit set -> negated look-up
    data[:, -1] = vals.astype(np.uint64) * np.uint64(delta)
    cts = shortint.CiphertextList(data, degree=15, noise_level=shortint.NOISE_ZERO)
    out = sk.apply_lookup_table(cts, lut)
    want = np.where(vals < 16, ((vals + 7) % 16), (-(((vals % 16) + 7) % 16)) % 32)
    assert np.array_equal(out.data[:, -1] // np.uint64(delta), want.astype(np.uint64))
    assert (out.degree == lut.degree).all() and (out.noise_level == shortint.NOISE_ZERO).all()


@pytest.fixture(scope="module")
def emul(built):
    return C.CDLL(os.path.join(ROOT, "tests", "native", "libemul.so"))


def test_kernel_fft_roundtrip_on_cpu(emul):
    # same tolerance as the reference's test_roundtrip (fft/tests.rs:9-80): < 2^14 ulp of the u64 torus
    rng = np.random.default_rng(1)
    poly = rng.integers(0, 2 ** 64, 2048, dtype=np.uint64)
    out = np.zeros(2048, np.uint64)
    assert emul.emul_fft_roundtrip(2048, _p(poly), _p(out)) == 0
    assert np.abs((out - poly).astype(np.int64)).max() < (1 << 14)


@pytest.mark.parametrize("name", ["TOY_U64", "TOY_U32", "TOY_2_2", "TOY_1_1"])
def test_kernel_stages_on_cpu_match_oracle(emul, name):
    o = get_oracle(name)
    B = 4
    if o.pmod:
        msgs = np.arange(B) % o.pmod
        pts, lut = o.encode(msgs), o.lut(lambda x: (x + 1) % o.pmod)
    else:
        msgs = np.arange(B) % 2
        pts, lut = np.where(msgs, 1 << 29, 7 << 29).astype(np.uint64), o.boolean_lut()
    cts = o.encrypt(pts, "small")
    ref = o.pbs(cts, lut, sample_extract=False)
    out = np.zeros((B, o.glwe_len), np.uint64)
    assert emul.emul_blind_rotate(o.n, o.k, o.N, o.b, o.l, o.bits, _p(o.bsk), _p(cts), B, _p(lut), _p(out)) == 0
    ph_ref = np.array([o.glwe_phase(g)[0] for g in ref]); ph = np.array([o.glwe_phase(g)[0] for g in out])
    if o.pmod:
        assert np.array_equal(o.decode(ph) % o.pmod, (msgs + 1) % o.pmod)
        assert np.array_equal(o.decode(ph), o.decode(ph_ref))
    else:
        assert np.array_equal(o.decode_bool(ph), msgs.astype(bool))
    d = (ph - ph_ref)
    d = d.astype(np.int64) if o.bits == 64 else (d & np.uint64(0xFFFFFFFF)).astype(np.uint32).astype(np.int32)
    assert np.abs(d.astype(np.float64)).max() / 2.0 ** o.bits < phase_bound(o)


def test_kernel_single_cmux_matches_oracle_to_fft_precision(emul):
    # one non-zero mask element => exactly one external product: the two f64 FFTs must agree to ~2^-24 of the torus
    o = get_oracle("TOY_2_2")
    lut = o.lut(lambda x: x)
    ct = np.zeros((1, o.n + 1), np.uint64)
    ct[0, 3] = np.uint64(0x1234567890ABCDEF); ct[0, o.n] = np.uint64(0x0234567890ABCDEF)
    ref = o.pbs(ct, lut, sample_extract=False)
    out = np.zeros((1, o.glwe_len), np.uint64)
    assert emul.emul_blind_rotate(o.n, o.k, o.N, o.b, o.l, o.bits, _p(o.bsk), _p(ct), 1, _p(lut), _p(out)) == 0
    assert np.abs((out - ref).astype(np.int64)).max() < 2 ** 42      # 2^-22 of the torus


def test_three_pass_fft_matches_numpy(emul):
    # fft8.cuh (the latency kernel's 128-thread FFT, radix 8 x 16 x 8): forward vs numpy's DFT of the twisted input, and the inverse round trip
    rng = np.random.default_rng(11)
    x = rng.standard_normal(1024) + 1j * rng.standard_normal(1024)
    inp = np.empty(2048); inp[0::2] = x.real; inp[1::2] = x.imag
    out = np.empty(2048); coef = np.empty(2048)
    assert emul.emul_fft8_forward(_p(inp), _p(out)) == 0
    ref = np.fft.fft(x * np.exp(1j * np.pi * np.arange(1024) / 2048))
    assert np.abs((out[0::2] + 1j * out[1::2]) - ref).max() / np.abs(ref).max() < 2.0 ** -48
    assert emul.emul_fft8_inverse(_p(out), _p(coef)) == 0
    assert np.abs((coef[:1024] + 1j * coef[1024:]) - x).max() < 2.0 ** -46


def test_latency_kernel_stages_on_cpu_match_oracle(emul):
    # blind_rotate_lat.cu's stage sequence, with the bootstrap key converted by the standard routine
    o = get_oracle("TOY_2_2")
    B = 3
    msgs = np.arange(B) % o.pmod
    pts, lut = o.encode(msgs), o.lut(lambda x: (x + 1) % o.pmod)
    cts = o.encrypt(pts, "small")
    ref = o.pbs(cts, lut, sample_extract=False)
    out = np.zeros((B, o.glwe_len), np.uint64)
    assert emul.emul_blind_rotate_lat(o.n, _p(o.bsk), _p(cts), B, _p(lut), _p(out)) == 0
    ph_ref = np.array([o.glwe_phase(g)[0] for g in ref]); ph = np.array([o.glwe_phase(g)[0] for g in out])
    assert np.array_equal(o.decode(ph) % o.pmod, (msgs + 1) % o.pmod)
    assert np.abs((ph - ph_ref).astype(np.int64).astype(np.float64)).max() / 2.0 ** 64 < phase_bound(o)
    ct = np.zeros((1, o.n + 1), np.uint64)      # exactly one external product: FFT-precision agreement
    ct[0, 3] = np.uint64(0x1234567890ABCDEF); ct[0, o.n] = np.uint64(0x0234567890ABCDEF)
    ref1 = o.pbs(ct, o.lut(lambda x: x), sample_extract=False)
    out1 = np.zeros((1, o.glwe_len), np.uint64)
    assert emul.emul_blind_rotate_lat(o.n, _p(o.bsk), _p(ct), 1, _p(o.lut(lambda x: x)), _p(out1)) == 0
    assert np.abs((out1 - ref1).astype(np.int64)).max() < 2 ** 42


def test_single_level_digit_closed_form(emul):
    rng = np.random.default_rng(9)
    xs = list(rng.integers(0, 2 ** 64, 20000, dtype=np.uint64))
    for b in (23, 18):      # rounding and carry boundaries
        for top in (0, 1, (1 << (b - 1)) - 1, 1 << (b - 1), (1 << (b - 1)) + 1, (1 << b) - 1):
            for lo in (0, 1, (1 << (63 - b)) - 1, 1 << (63 - b), (1 << (63 - b)) + 1, (1 << (64 - b)) - 1):
                xs.append(np.uint64(((top << (64 - b)) | lo) & (2 ** 64 - 1)))
    xs = np.array(xs, dtype=np.uint64)
    assert emul.emul_check_single_level_digit(_p(xs), xs.size) == 0


def test_big_kernel_stages_on_cpu_match_oracle(emul):
    # N = 8192 (PARAM_MESSAGE_3_CARRY_3 shape, tiny n): the 256-thread stage sequence of blind_rotate_big.cu on the CPU
    o = get_oracle("TOY_3_3")
    msgs = np.array([1, 17, 40])
    cts = o.encrypt(o.encode(msgs), "small")
    lut = o.lut(lambda x: (x + 5) % o.pmod)
    ref = o.pbs(cts, lut, sample_extract=False)
    out = np.zeros((3, o.glwe_len), np.uint64)
    assert emul.emul_blind_rotate_big(o.n, o.k, o.N, o.b, o.l, o.bits, _p(o.bsk), _p(cts), 3, _p(lut), _p(out)) == 0
    ph = np.array([o.glwe_phase(g)[0] for g in out]); ph_ref = np.array([o.glwe_phase(g)[0] for g in ref])
    assert np.array_equal(o.decode(ph) % o.pmod, (msgs + 5) % o.pmod)
    assert np.array_equal(o.decode(ph), o.decode(ph_ref))
    assert np.abs((ph - ph_ref).astype(np.int64)).max() / 2.0 ** 64 < phase_bound(o)
    ct = np.zeros((1, o.n + 1), np.uint64)
    ct[0, 2] = np.uint64(0x1234567890ABCDEF); ct[0, o.n] = np.uint64(0x0234567890ABCDEF)
    ref = o.pbs(ct, lut, sample_extract=False)
    out = np.zeros((1, o.glwe_len), np.uint64)
    emul.emul_blind_rotate_big(o.n, o.k, o.N, o.b, o.l, o.bits, _p(o.bsk), _p(ct), 1, _p(lut), _p(out))
    assert np.abs((out - ref).astype(np.int64)).max() < 2 ** 36       # one CMUX: f64 FFT precision only


def test_constant_table_is_reproducible():
    """fft_reg.cuh cites tools/gen_tables.py for COS128_TAB: the committed generator must reproduce it bit for bit."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_tables.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_rust_sys_crate_matches_the_header_and_the_makefile():
    """bindings/rust/b200tfhe-sys: lib.rs is the generated transcription of include/b200tfhe.h (every prototype), build.rs compiles
    the same sources as csrc/Makefile."""
    import re
    import subprocess
    assert subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_rust_sys.py"), "--check"]).returncode == 0
    rs = open(os.path.join(ROOT, "bindings", "rust", "b200tfhe-sys", "src", "lib.rs")).read()
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "b200tfhe.h")).read(), flags=re.S)
    names = set(re.findall(r"\b(b200tfhe_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) > 80 and all(f"pub fn {n}(" in rs for n in names)
    mk = open(os.path.join(ROOT, "tfhe-rs-odd_b200", "csrc", "Makefile")).read()
    objs = re.search(r"^OBJS = (.*)$", mk, flags=re.M).group(1).split()
    brs = open(os.path.join(ROOT, "bindings", "rust", "b200tfhe-sys", "build.rs")).read()
    listed = re.findall(r'"([a-z0-9_]+)"', re.search(r"SOURCES: &\[&str\] = &\[(.*?)\];", brs, flags=re.S).group(1))
    assert sorted(listed) == sorted(o[:-2] for o in objs)
