"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE -- see oracle/tfhe_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Parameter sets restate /root/reference/tfhe/src/shortint/parameters/mod.rs:613-627,703-717,853-867
and boolean/parameters/mod.rs:123-134.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


class OrcParams(C.Structure):
    _fields_ = [("n", C.c_int), ("k", C.c_int), ("N", C.c_int), ("pbs_base_log", C.c_int),
                ("pbs_level", C.c_int), ("ks_base_log", C.c_int), ("ks_level", C.c_int),
                ("bits", C.c_int), ("lwe_std", C.c_double), ("glwe_std", C.c_double)]


# name -> (n, k, N, pbs_base_log, pbs_level, ks_base_log, ks_level, bits, lwe_std, glwe_std, msg_mod*carry_mod, order)
PARAMS = {
    "BOOLEAN_DEFAULT": (722, 2, 512, 6, 3, 3, 4, 32, 0.000013071021089943935, 0.00000004990272175010415, 0, "pbs_ks"),
    "MESSAGE_1_CARRY_1": (684, 3, 512, 18, 1, 4, 3, 64, 0.00002043784477291318, 0.0000000000034525330484572114, 4, "ks_pbs"),
    "MESSAGE_2_CARRY_2": (742, 1, 2048, 23, 1, 3, 5, 64, 0.000007069849454709433, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    "MESSAGE_3_CARRY_3": (864, 1, 8192, 15, 2, 3, 6, 64, 0.000000757998020150446, 0.0000000000000000002168404344971009, 64, "ks_pbs"),
    # toy sets for fast CPU tests (same role as the reference's FFT_WOPBS_N512_PARAMS toy set, test/mod.rs:276-289):
    # the polynomial sizes / decompositions of the named sets with a tiny LWE dimension
    "TOY_U64": (16, 1, 512, 12, 2, 4, 4, 64, 1e-9, 1e-15, 4, "ks_pbs"),
    "TOY_U32": (20, 2, 512, 6, 3, 3, 4, 32, 1e-7, 1e-8, 0, "pbs_ks"),
    "TOY_2_2": (12, 1, 2048, 23, 1, 3, 5, 64, 1e-9, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    "TOY_3_3": (6, 1, 8192, 15, 2, 3, 6, 64, 1e-9, 0.0000000000000000002168404344971009, 64, "ks_pbs"),
    "TOY_1_1": (12, 3, 512, 18, 1, 4, 3, 64, 1e-9, 0.0000000000034525330484572114, 4, "ks_pbs"),
    # multi-bit PBS sets (shortint/parameters/multi_bit.rs:115-132 group 2, :173-190 group 3) and toy versions
    "MULTI_BIT_2_2_GROUP_2": (818, 1, 2048, 22, 1, 5, 3, 64, 0.000002226459789930014, 0.0000000000000003152931493498455, 16, "ks_pbs"),
    "MULTI_BIT_2_2_GROUP_3": (888, 1, 2048, 21, 1, 7, 2, 64, 0.0000006125031601933181, 0.0000000000000003152931493498455, 16, "ks_pbs"),
    # WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS (shortint/parameters/parameters_wopbs_message_carry.rs:432-451) and a toy of the same shape
    "WOPBS_MESSAGE_2_CARRY_2": (769, 1, 2048, 15, 2, 6, 2, 64, 0.0000043131554647504185, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    "TOY_WOP": (12, 1, 2048, 15, 2, 6, 2, 64, 1e-9, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    # shape of WOPBS_PARAM_MESSAGE_2_CARRY_1_KS_PBS (:n 487, k 3, N 512, pbs 12/2, ks 2/4, pfks 12/2, cbs 3/4) with a toy n
    "TOY_WOP_K3": (12, 3, 512, 12, 2, 2, 4, 64, 1e-9, 1e-15, 8, "ks_pbs"),
    # shape of WOPBS_PARAM_MESSAGE_3_CARRY_1_KS_PBS (k 3, N 512, pbs 9/3, ks 2/4, pfks 9/3, cbs 3/5) with a toy n
    "TOY_WOP_K3B": (12, 3, 512, 9, 3, 2, 4, 64, 1e-9, 1e-15, 8, "ks_pbs"),
    # keyswitch bases wider than an int8 digit: 2^10 x 1 level (WOPBS_PARAM_MESSAGE_3_CARRY_3_KS_PBS), 2^15 x 2, and 2^12 x 2 on the 2^32 torus
    "TOY_KSW10": (12, 1, 2048, 23, 1, 10, 1, 64, 1e-9, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    "TOY_KSW15": (20, 1, 512, 12, 2, 15, 2, 64, 1e-9, 1e-15, 4, "ks_pbs"),
    "TOY_KSW12_U32": (20, 2, 512, 6, 3, 12, 2, 32, 1e-7, 1e-8, 0, "pbs_ks"),
    "TOY_MB_G2": (12, 1, 512, 12, 2, 4, 4, 64, 1e-9, 1e-15, 4, "ks_pbs"),
    "TOY_MB_G3": (12, 1, 512, 12, 2, 4, 4, 64, 1e-9, 1e-15, 4, "ks_pbs"),
    "TOY_MB22_G2": (12, 1, 2048, 22, 1, 5, 3, 64, 1e-9, 0.0000000000000003152931493498455, 16, "ks_pbs"),
    "TOY_MB22_G3": (12, 1, 2048, 21, 1, 7, 2, 64, 1e-9, 0.0000000000000003152931493498455, 16, "ks_pbs"),
    # the sets of the committed golden fixtures (tests/golden/): n = 4 keeps the stored keys small
    "GOLD_U64": (4, 1, 512, 12, 2, 4, 4, 64, 1e-9, 1e-15, 4, "ks_pbs"),
    "GOLD_U32": (4, 2, 512, 6, 3, 3, 4, 32, 1e-7, 1e-8, 0, "pbs_ks"),
    "GOLD_2_2": (4, 1, 2048, 23, 1, 3, 5, 64, 1e-9, 0.00000000000000029403601535432533, 16, "ks_pbs"),
    "GOLD_MB": (4, 1, 512, 12, 2, 4, 4, 64, 1e-9, 1e-15, 4, "ks_pbs"),        # multi-bit, grouping factor 2
    "GOLD_PK": (4, 1, 64, 12, 2, 4, 2, 64, 1e-9, 1e-15, 4, "ks_pbs"),          # packing keyswitch / gadget integer ops (tiny N)
}

# LweBskGroupingFactor of the multi-bit sets (absent = classic bootstrap key)
GROUPING = {"MULTI_BIT_2_2_GROUP_2": 2, "MULTI_BIT_2_2_GROUP_3": 3, "TOY_MB_G2": 2, "TOY_MB_G3": 3, "TOY_MB22_G2": 2, "TOY_MB22_G3": 3,
            "GOLD_MB": 2}


# (pfks_base_log, pfks_level, cbs_base_log, cbs_level) of the WoP-PBS sets
WOPBS = {"WOPBS_MESSAGE_2_CARRY_2": (15, 2, 5, 3), "TOY_WOP": (15, 2, 5, 3), "TOY_WOP_K3": (12, 2, 3, 4), "TOY_WOP_K3B": (9, 3, 3, 5)}


def build(force=False):
    if force or not os.path.exists(_LIB) or any(
            os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB)
            for f in os.listdir(_HERE) if f.endswith((".c", ".h"))):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None
_native = None     # path of a host-tuned build, see use_native_build()


def use_native_build():
    """For the TIMED CPU legs of bench.py only: compile the oracle for the host it runs on (-O3 -march=native, what BASELINE.md
    section 3 states for the CPU arm) into liboracle_native.so and make lib() load that one.  The portable build (x86-64-v3) is what
    travels with the repository and what the tests use.  Returns the compiler flags actually in effect."""
    global _native
    if _lib is not None:
        return "portable (library already loaded)"
    out = os.path.join(_HERE, "liboracle_native.so")
    try:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle_native.so"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        C.CDLL(out)
        _native = out
        return "-O3 -march=native -mprefer-vector-width=512"
    except (OSError, subprocess.CalledProcessError):
        return "-O3 -march=x86-64-v3 (native build failed)"


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_native or build())
        _lib.orc_closest_representable.restype = C.c_uint64
        _lib.orc_closest_representable.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int]
        _lib.orc_modswitch.restype = C.c_uint64
        _lib.orc_modswitch.argtypes = [C.c_uint64, C.c_int, C.c_int]
        _lib.orc_lwe_phase.restype = C.c_uint64
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


class Oracle:
    """One parameter set + one deterministic key set, in the reference's standard-domain layouts."""

    def __init__(self, name, seed=0xB2000001, nthreads=None):
        t = PARAMS[name]
        self.name = name
        self.p = OrcParams(*t[:10])
        self.n, self.k, self.N = t[0], t[1], t[2]
        self.l, self.b, self.l_ks, self.b_ks, self.bits = t[4], t[3], t[6], t[5], t[7]
        self.pmod, self.order = t[10], t[11]
        self.grouping = GROUPING.get(name, 0)
        # number of GGSW ciphertexts of the bootstrap key: n (classic) or n/G * 2^G (multi-bit)
        self.n_ggsw = self.n if not self.grouping else self.n // self.grouping * (1 << self.grouping)
        self.seed = seed
        self.nthreads = nthreads or os.cpu_count()
        self.big = self.k * self.N
        self.glwe_len = (self.k + 1) * self.N
        self.small_key = self.glwe_key = self.ksk = self.bsk = self.fbsk = None

    # ---- keys -------------------------------------------------------------------------------
    def keygen(self, fourier=True):
        L = lib()
        cache = os.path.join(os.environ.get("ORACLE_KEY_CACHE", "/tmp/b200tfhe_keys"), f"{self.name}_{self.n}_{self.k}_{self.N}_{self.b}_{self.l}_{self.b_ks}_{self.l_ks}_{self.bits}_{self.seed:x}.npz")
        if os.path.exists(cache):
            z = np.load(cache)
            self.small_key, self.glwe_key, self.ksk, self.bsk = z["s"], z["g"], z["ksk"], z["bsk"]
        else:
            self.small_key = np.zeros(self.n, np.uint64)
            self.glwe_key = np.zeros(self.big, np.uint64)
            L.orc_gen_binary_key(C.c_uint64(self.seed), self.n, _p(self.small_key))
            L.orc_gen_binary_key(C.c_uint64(self.seed + 1), self.big, _p(self.glwe_key))
            self.ksk = np.zeros(self.big * self.l_ks * (self.n + 1), np.uint64)
            self.bsk = np.zeros(self.n_ggsw * self.l * (self.k + 1) ** 2 * self.N, np.uint64)
            L.orc_gen_ksk(C.byref(self.p), _p(self.glwe_key), _p(self.small_key), C.c_uint64(self.seed + 2), _p(self.ksk), self.nthreads)
            if self.grouping:
                L.orc_gen_multibit_bsk(C.byref(self.p), self.grouping, _p(self.small_key), _p(self.glwe_key), C.c_uint64(self.seed + 3), _p(self.bsk), self.nthreads)
            else:
                L.orc_gen_bsk(C.byref(self.p), _p(self.small_key), _p(self.glwe_key), C.c_uint64(self.seed + 3), _p(self.bsk), self.nthreads)
            try:
                os.makedirs(os.path.dirname(cache), exist_ok=True)
                np.savez(cache, s=self.small_key, g=self.glwe_key, ksk=self.ksk, bsk=self.bsk)
            except OSError:
                pass
        if fourier:
            self._to_fourier()
        return self

    def _to_fourier(self):
        pf = OrcParams(*[getattr(self.p, f) for f, _ in OrcParams._fields_])
        pf.n = self.n_ggsw      # the conversion is per GGSW polynomial: same routine for classic and multi-bit keys
        self.fbsk = np.zeros(self.bsk.size, np.float64)
        lib().orc_bsk_to_fourier(C.byref(pf), _p(self.bsk), _p(self.fbsk), self.nthreads)

    def load_keys(self, small_key, glwe_key, ksk, bsk):
        """Install externally stored keys (golden fixtures) instead of generating them."""
        self.small_key, self.glwe_key, self.ksk, self.bsk = u64(small_key), u64(glwe_key), u64(ksk), u64(bsk)
        self._to_fourier()
        return self

    # ---- ciphertexts ------------------------------------------------------------------------
    def encrypt(self, plaintexts, under="big", seed=0xB2000002, std=None):
        key, dim = (self.glwe_key, self.big) if under == "big" else (self.small_key, self.n)
        if std is None:
            std = self.p.glwe_std if under == "big" else self.p.lwe_std
        pts = u64(plaintexts)
        out = np.zeros((pts.size, dim + 1), np.uint64)
        for i, pt in enumerate(pts):
            lib().orc_lwe_encrypt(dim, self.bits, _p(key), C.c_uint64(int(pt)), C.c_double(std), C.c_uint64(seed + i), _p(out[i]))
        return out

    def phase(self, cts, under="big"):
        key, dim = (self.glwe_key, self.big) if under == "big" else (self.small_key, self.n)
        cts = u64(cts).reshape(-1, dim + 1)
        return np.array([lib().orc_lwe_phase(dim, self.bits, _p(key), _p(c)) for c in cts], np.uint64)

    def glwe_phase(self, glwe):
        """body - sum_j A_j*S_j of a GLWE ciphertext, exact (binary key)."""
        glwe = u64(glwe).reshape(self.k + 1, self.N)
        ph = glwe[self.k].copy()
        tmp = np.zeros(self.N, np.uint64)
        for j in range(self.k):
            lib().orc_negacyclic_mul_binary(self.N, _p(u64(glwe[j])), _p(u64(self.glwe_key[j * self.N:(j + 1) * self.N])), _p(tmp), self.bits)
            ph = ph - tmp
        return ph & np.uint64((1 << self.bits) - 1 if self.bits < 64 else 0xFFFFFFFFFFFFFFFF)

    # shortint encode / decode: shortint/engine/mod.rs:101 (delta = 2^63 / p), client_key/mod.rs:281-301
    def delta(self):
        return (1 << (self.bits - 1)) // self.pmod

    def encode(self, msgs):
        return (u64(msgs) * np.uint64(self.delta()))

    def decode(self, phases):
        d = np.uint64(self.delta())
        ph = u64(phases)
        rounded = ph + ((ph & (d >> np.uint64(1))) << np.uint64(1))
        return (rounded // d) % np.uint64(2 * self.pmod)

    def decode_bool(self, phases):
        return u64(phases) < np.uint64(1 << 31)  # boolean/engine/mod.rs:353-375

    # ---- LUTs -------------------------------------------------------------------------------
    def lut(self, f):
        table = u64([f(i) for i in range(self.pmod)])
        out = np.zeros(self.glwe_len, np.uint64)
        lib().orc_fill_accumulator(C.byref(self.p), _p(table), self.pmod, _p(out))
        return out

    def boolean_lut(self):
        out = np.zeros(self.glwe_len, np.uint64)
        lib().orc_boolean_lut(C.byref(self.p), _p(out))
        return out

    # ---- the path ---------------------------------------------------------------------------
    def keyswitch(self, cts):
        cts = u64(cts).reshape(-1, self.big + 1)
        out = np.zeros((cts.shape[0], self.n + 1), np.uint64)
        for i in range(cts.shape[0]):
            lib().orc_keyswitch(C.byref(self.p), _p(self.ksk), _p(cts[i]), _p(out[i]))
        return out

    def _batch(self, fn, cts, luts, lut_idx, in_dim, out_dim):
        cts = u64(cts).reshape(-1, in_dim)
        luts = u64(luts).reshape(-1, self.glwe_len)
        B = cts.shape[0]
        idx = np.ascontiguousarray(lut_idx if lut_idx is not None else np.zeros(B), dtype=np.uint32)
        out = np.zeros((B, out_dim), np.uint64)
        fn(_p(cts), _p(luts), _p(idx), _p(out), B, self.nthreads)
        return out

    def ks_pbs(self, cts, luts, lut_idx=None):
        L = lib()
        if self.grouping:
            return self._batch(lambda *a: L.orc_multibit_batch(C.byref(self.p), self.grouping, 0, _p(self.ksk), _p(self.fbsk), *a),
                               cts, luts, lut_idx, self.big + 1, self.big + 1)
        return self._batch(lambda *a: L.orc_ks_pbs_batch(C.byref(self.p), _p(self.ksk), _p(self.fbsk), *a),
                           cts, luts, lut_idx, self.big + 1, self.big + 1)

    def pbs_ks(self, cts, luts, lut_idx=None):
        L = lib()
        return self._batch(lambda *a: L.orc_pbs_ks_batch(C.byref(self.p), _p(self.ksk), _p(self.fbsk), *a),
                           cts, luts, lut_idx, self.n + 1, self.n + 1)

    def pbs(self, cts, luts, lut_idx=None, sample_extract=True):
        L = lib()
        od = self.big + 1 if sample_extract else self.glwe_len
        if self.grouping:
            return self._batch(lambda *a: L.orc_multibit_batch(C.byref(self.p), self.grouping, 2 if sample_extract else 3, None, _p(self.fbsk), *a),
                               cts, luts, lut_idx, self.n + 1, od)
        return self._batch(lambda *a: L.orc_pbs_batch(C.byref(self.p), _p(self.fbsk), *a, 1 if sample_extract else 0),
                           cts, luts, lut_idx, self.n + 1, od)

    def blind_rotate(self, cts, luts, lut_idx=None):
        """blind_rotate_assign over a list: the rotated GLWE accumulators (no sample extraction)."""
        return self.pbs(cts, luts, lut_idx, sample_extract=False)

    def ggsw_fourier(self, index):
        """Fourier GGSW `index` of the bootstrap key, [l][k+1][k+1][N doubles]."""
        sz = self.l * (self.k + 1) ** 2 * self.N
        return self.fbsk[index * sz:(index + 1) * sz]

    def add_external_product(self, out, ggsw_index, glwe):
        """add_external_product_assign (fft64/crypto/ggsw.rs:477-598): returns out + ggsw[ggsw_index] (x) glwe, one GGSW index per ciphertext."""
        out, glwe = u64(out).reshape(-1, self.glwe_len).copy(), u64(glwe).reshape(-1, self.glwe_len)
        idx = np.broadcast_to(np.asarray(ggsw_index, dtype=np.int64), (out.shape[0],))
        for b in range(out.shape[0]):
            g = np.ascontiguousarray(self.ggsw_fourier(int(idx[b])))
            lib().orc_add_external_product(C.byref(self.p), _p(out[b]), _p(g), _p(np.ascontiguousarray(glwe[b])))
        return out

    def cmux(self, ct0, ct1, ggsw_index):
        """cmux_assign (ggsw.rs:766-777): ct1 -= ct0; ct0 += ggsw (x) ct1.  Returns the new (ct0, ct1)."""
        ct0, ct1 = u64(ct0).reshape(-1, self.glwe_len), u64(ct1).reshape(-1, self.glwe_len)
        d = ct1 - ct0
        if self.bits == 32:
            d &= np.uint64(0xFFFFFFFF)
        return self.add_external_product(ct0, ggsw_index, d), d

    def decompose_all(self, x):
        """Signed decomposition of every element (commons/math/decomposition/decomposer.rs:99-119, iter.rs:120-127), vectorised:
        returns [l][...] wrapping digits, index 0 = level l (the first the iterator emits)."""
        x = u64(x)
        m = np.uint64((1 << self.bits) - 1 if self.bits < 64 else 0xFFFFFFFFFFFFFFFF)
        b, L = self.b, self.l
        shift = np.uint64(self.bits - b * L)
        one = np.uint64(1)
        # closest_representable: round to the top b*l bits
        st = ((((x >> (shift - one)) + one) >> one) << shift) & m
        st = st >> shift
        out = []
        mb = np.uint64((1 << b) - 1)
        for _ in range(L):
            res = st & mb
            st = st >> np.uint64(b)
            carry = ((((res - one) & m) | st) & res) >> np.uint64(b - 1)
            st = (st + carry) & m
            out.append((res - (carry << np.uint64(b))) & m)
        return np.stack(out)

    def exact_external_product(self, out, ggsw_index, glwe):
        """out + ggsw (x) glwe in EXACT integer arithmetic mod 2^bits: digits x the STANDARD-domain key, schoolbook negacyclic
        products (the helper of the reference's own FFT test, fft/tests.rs:83-100) -- no FFT anywhere."""
        k1, N, L = self.k + 1, self.N, self.l
        out = u64(out).reshape(k1, N).copy()
        glwe = u64(glwe).reshape(k1, N)
        g = self.bsk.reshape(-1, L, k1, k1, N)[ggsw_index]
        digs = self.decompose_all(glwe)                      # [l][k1][N], index 0 = level l
        m = np.uint64((1 << self.bits) - 1 if self.bits < 64 else 0xFFFFFFFFFFFFFFFF)
        tmp = np.zeros(N, np.uint64)
        for t in range(L):
            lvl = L - 1 - t                                   # key levels stored 1..l
            for r in range(k1):
                d = np.ascontiguousarray(digs[t, r])
                for j in range(k1):
                    lib().orc_negacyclic_schoolbook(N, _p(d), _p(np.ascontiguousarray(g[lvl, r, j])), _p(tmp), self.bits)
                    out[j] = (out[j] + tmp) & m
        return out.reshape(-1)

    def gate_prologue(self, gate, a, b):
        a, b = u64(a).reshape(-1, self.n + 1), u64(b).reshape(-1, self.n + 1)
        out = np.zeros_like(a)
        for i in range(a.shape[0]):
            lib().orc_boolean_gate_prologue(gate, self.n + 1, _p(a[i]), _p(b[i]), _p(out[i]))
        return out

    # ---- gadget extras (gadget/engine/bootstrapping.rs) ----------------------------------------------
    def gen_pksk(self):
        """LwePackingKeyswitchKey big LWE key -> GLWE key with the keyswitch decomposition: [kN][l_ks][(k+1)N]."""
        pksk = np.zeros(self.big * self.l_ks * self.glwe_len, np.uint64)
        lib().orc_gen_pksk(C.byref(self.p), _p(self.glwe_key), _p(self.glwe_key), C.c_uint64(self.seed + 4), _p(pksk), self.nthreads)
        return pksk

    def packing_keyswitch(self, pksk, cts):
        cts = u64(cts).reshape(-1, self.big + 1)
        out = np.zeros((cts.shape[0], self.glwe_len), np.uint64)
        for i in range(cts.shape[0]):
            lib().orc_packing_keyswitch(C.byref(self.p), _p(pksk), _p(cts[i]), _p(out[i]))
        return out

    def glwe_mul_plain_extract(self, glwes, polys):
        glwes, polys = u64(glwes).reshape(-1, self.glwe_len), u64(polys).reshape(-1, self.N)
        out = np.zeros((glwes.shape[0], polys.shape[0], self.big + 1), np.uint64)
        for b in range(glwes.shape[0]):
            for i in range(polys.shape[0]):
                lib().orc_glwe_mul_plain_extract(C.byref(self.p), _p(glwes[b]), _p(polys[i]), _p(out[b, i]))
        return out

    def pack_into_accumulator(self, pksk, lwes, pmod):
        lwes = u64(lwes).reshape(-1, self.big + 1)
        out = np.zeros(self.glwe_len, np.uint64)
        lib().orc_pack_into_accumulator(C.byref(self.p), _p(pksk), _p(lwes), lwes.shape[0], int(pmod), _p(out))
        return out

    # ---- bit extraction (fft64/crypto/wop_pbs/mod.rs:66-225) -----------------------------------------
    def extract_bits(self, cts, delta_log, nbits):
        """[B][kN+1] under the big key -> [B][nbits][n+1] under the small key, most significant extracted bit first."""
        assert not self.grouping, "the reference's extract_bits takes a classic Fourier bootstrap key"
        cts = u64(cts).reshape(-1, self.big + 1)
        out = np.zeros((cts.shape[0], int(nbits), self.n + 1), np.uint64)
        for i in range(cts.shape[0]):
            lib().orc_extract_bits(C.byref(self.p), _p(self.ksk), _p(self.fbsk), _p(cts[i]), int(delta_log), int(nbits), _p(out[i]))
        return out

    # ---- circuit bootstrap + vertical packing (fft64/crypto/wop_pbs/mod.rs:243-853) ---------------------
    def gen_cbs_pfpksk_list(self, pf_base_log, pf_level):
        """k+1 private functional packing keyswitch keys, reference layout [k+1][kN+1][level 1..L][(k+1)N]."""
        out = np.zeros((self.k + 1) * (self.big + 1) * int(pf_level) * self.glwe_len, np.uint64)
        lib().orc_gen_cbs_pfpksk_list(C.byref(self.p), int(pf_base_log), int(pf_level), _p(self.glwe_key), C.c_uint64(self.seed + 5),
                                      _p(out), self.nthreads)
        return out

    def pfpks(self, key, pf_base_log, pf_level, cts):
        cts = u64(cts).reshape(-1, self.big + 1)
        out = np.zeros((cts.shape[0], self.glwe_len), np.uint64)
        for i in range(cts.shape[0]):
            lib().orc_pfpks(C.byref(self.p), int(pf_base_log), int(pf_level), _p(key), _p(cts[i]), _p(out[i]))
        return out

    def circuit_bootstrap_boolean(self, pfpksk_list, wop, bit_ct, delta_log):
        pb, pl, cb, cl = wop
        out = np.zeros((cl, self.k + 1, self.glwe_len), np.uint64)
        lib().orc_circuit_bootstrap_boolean(C.byref(self.p), _p(self.fbsk), _p(pfpksk_list), pb, pl, cb, cl, _p(u64(bit_ct)), int(delta_log), _p(out))
        return out

    def cbs_vertical_packing(self, pfpksk_list, wop, bits_in, luts, polys_per_lut=1):
        """bits_in [B][nbits][n+1] (MSB first, q/2-scaled), luts [n_out][polys_per_lut][N] -> [B][n_out][kN+1]."""
        pb, pl, cb, cl = wop
        bits_in = u64(bits_in)
        bits_in = bits_in.reshape(-1, bits_in.shape[-2], self.n + 1)
        luts = u64(luts).reshape(-1, int(polys_per_lut), self.N)
        out = np.zeros((bits_in.shape[0], luts.shape[0], self.big + 1), np.uint64)
        for b in range(bits_in.shape[0]):
            lib().orc_cbs_vertical_packing_tree(C.byref(self.p), _p(self.fbsk), _p(pfpksk_list), pb, pl, cb, cl, _p(bits_in[b]),
                                                bits_in.shape[1], _p(luts), luts.shape[0], int(polys_per_lut), _p(out[b]))
        return out
