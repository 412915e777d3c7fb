import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def ctx():
    import tfhe_rs_odd_b200 as t
    c = t.Context(0)   # raises on a box without a GPU / without the built library: no CPU fallback
    yield c
    c.close()


_ORACLES = {}


def get_oracle(name):
    """Session-cached oracle with generated keys (test infrastructure)."""
    from oracle.oracle import Oracle
    if name not in _ORACLES:
        _ORACLES[name] = Oracle(name).keygen()
    return _ORACLES[name]


def params_of(o):
    import tfhe_rs_odd_b200 as t
    return t.core_crypto.make_params(o.n, o.k, o.N, o.b, o.l, o.b_ks, o.l_ks, o.bits)


def phase_bound(o):
    """Bound on |phase_gpu - phase_oracle| (torus units) for one PBS.

    Two f64 FFT implementations with different operation order differ by ~2^-25 of the torus after ONE external
    product; from the first CMUX where that flips a rounding decision of the signed decomposition the two
    accumulators are different (equally valid) ciphertexts and every later rounding is independent.  The phase
    difference is then the difference of two independent draws of the decomposition-rounding noise, whose
    variance per PBS is at most  n * (1 + kN/2) * 2^(-2 b l) / 12  (n/2 for a balanced key).  We allow 8 sigma of the difference.
    """
    # n (not n/2): upper bound over the secret key's Hamming weight, which matters for the toy sets with n ~ 10
    var = o.n * (1.0 + o.k * o.N / 2.0) * 2.0 ** (-2 * o.b * o.l) / 12.0
    return 8.0 * (2.0 * var) ** 0.5


def fft_bound(o):
    """Per-coefficient bound, in units of the torus LSB, on the difference of two f64 evaluations (different operation order) of ONE
    external product: |digit| <= B/2 times key values in [-1/2, 1/2), a random walk over the (k+1) l N terms, relative error 2^-53
    per operation; 16x head-room for the ~log2(N) butterfly levels and the maximum over all coefficients.  That is 2^-22 of the
    torus (2^42 LSBs) at PARAM_MESSAGE_2_CARRY_2 -- measured maximum 2^-24.4 -- and at least one LSB.  The reference's own
    acceptance threshold for one FFT product (fft/tests.rs:166-167: 2^(BITS - (52 - magnitude - log2 N))) is 2^46 there."""
    terms = (o.k + 1) * o.l * o.N
    est = 2.0 ** (o.b - 1) * 0.5 * terms ** 0.5 * 2.0 ** -53
    return max(1.0, 16.0 * est * 2.0 ** o.bits)
