"""Host mirror of shortint::wopbs::WopbsKey (tfhe/src/shortint/wopbs/mod.rs) over ciphertext LISTS: the reference applies the WoP-PBS
to one shortint block at a time (extract_bits :547-628, then circuit_bootstrap_with_bits :756-823); here every call takes all blocks
and the two stages are one C-ABI call each (b200tfhe_extract_bits_batch_u64, b200tfhe_cbs_vertical_packing_batch_u64).
Same names, argument meaning and metadata behaviour as the reference; key generation stays on the client."""
import math

import numpy as np

from . import core_crypto as cc
from .core_crypto import B200TfheError
from .shortint import NOISE_NOMINAL, CiphertextList


class WopbsParameters:
    """The fields of shortint::parameters::WopbsParameters the server side reads (parameters/mod.rs; e.g.
    WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS, parameters_wopbs_message_carry.rs:432-451)."""

    def __init__(self, pfks_level, pfks_base_log, cbs_level, cbs_base_log, message_modulus, carry_modulus):
        self.pfks_level, self.pfks_base_log = int(pfks_level), int(pfks_base_log)
        self.cbs_level, self.cbs_base_log = int(cbs_level), int(cbs_base_log)
        self.message_modulus, self.carry_modulus = int(message_modulus), int(carry_modulus)


class ShortintWopbsLUT:
    """ShortintWopbsLUT / WopbsLUTBase (:54-150): a plaintext list of small_lut_size entries per output ciphertext."""

    def __init__(self, values, output_ciphertext_count=1):
        self.values = np.ascontiguousarray(values, dtype=np.uint64).reshape(int(output_ciphertext_count), -1)

    @property
    def output_ciphertext_count(self):
        return self.values.shape[0]

    @property
    def small_lut_size(self):
        return self.values.shape[1]


class WopbsKey:
    """WopbsKey{wopbs_server_key, pbs_server_key, cbs_pfpksk, ksk_pbs_to_wopbs, param} (:44-51), the `only_for_wopbs` shape (:235-252):
    the ciphertexts are already under the WoP-PBS parameters.  wopbs_server_key: a shortint.ServerKey (keyswitch + bootstrap key
    handles); cbs_pfpksk: core_crypto.LwePrivateFunctionalPackingKeyswitchKeyList."""

    def __init__(self, wopbs_server_key, cbs_pfpksk, param):
        self.wopbs_server_key, self.cbs_pfpksk, self.param = wopbs_server_key, cbs_pfpksk, param
        if cbs_pfpksk.decomp_level_count != param.pfks_level or cbs_pfpksk.decomp_base_log != param.pfks_base_log:
            raise B200TfheError("cbs_pfpksk decomposition does not match the WoP-PBS parameters")

    # ---- look-up tables (host tables, identical to the reference's) -------------------------------------------------------
    def _poly_size(self):
        return self.wopbs_server_key.params.polynomial_size

    def generate_lut(self, f, message_modulus=None, carry_modulus=None):
        """:282-295: f on the message bits, one bit of padding: entry i = f(i % message_modulus) << (64 - ceil(log2 basis) - 1)."""
        mm = message_modulus or self.param.message_modulus
        basis = mm * (carry_modulus or self.param.carry_modulus)
        delta = 64 - math.ceil(math.log2(basis)) - 1
        lut = np.zeros(self._poly_size(), np.uint64)
        for i in range(basis):
            lut[i] = np.uint64((int(f(i % mm)) << delta) % (1 << 64))
        return ShortintWopbsLUT(lut)

    def generate_lut_without_padding(self, f, message_modulus=None, carry_modulus=None):
        """:320-333: no bit of padding: << (64 - ceil(log2 basis))."""
        mm = message_modulus or self.param.message_modulus
        basis = mm * (carry_modulus or self.param.carry_modulus)
        delta = 64 - math.ceil(math.log2(basis))
        lut = np.zeros(self._poly_size(), np.uint64)
        for i in range(basis):
            lut[i] = np.uint64((int(f(i % mm)) << delta) % (1 << 64))
        return ShortintWopbsLUT(lut)

    # ---- the two stages ------------------------------------------------------------------------------------------------------
    def extract_bits(self, delta_log, ciphertexts, num_bits_to_extract):
        """:547-628 for every ciphertext of the list: [B][num_bits][n+1], most significant bit first."""
        sk = self.wopbs_server_key
        data = ciphertexts.data if isinstance(ciphertexts, CiphertextList) else ciphertexts
        return cc.extract_bits_from_lwe_ciphertext_mem_optimized(data, sk.bootstrapping_key, sk.key_switching_key, int(delta_log),
                                                                 int(num_bits_to_extract))

    def circuit_bootstrapping_vertical_packing(self, vec_lut, extracted_bits_blocks):
        """:630-658 / circuit_bootstrap_with_bits :756-823: extracted bits [B][nbits][n+1] + one table per output ciphertext."""
        lut = vec_lut if isinstance(vec_lut, ShortintWopbsLUT) else ShortintWopbsLUT(vec_lut)
        N = self._poly_size()
        if lut.small_lut_size % N:
            raise B200TfheError("look-up table size must be a multiple of the polynomial size")
        return cc.circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized(
            extracted_bits_blocks, lut.values.reshape(-1, N), self.wopbs_server_key.bootstrapping_key, self.cbs_pfpksk,
            self.param.cbs_base_log, self.param.cbs_level, number_of_luts=lut.output_ciphertext_count)

    def _extract_bits_circuit_bootstrapping(self, cts, lut, delta_log, nb_bit_to_extract):
        """:825-862; metadata of the result: degree message_modulus - 1, nominal noise."""
        bits = self.extract_bits(delta_log, cts, nb_bit_to_extract)
        out = self.circuit_bootstrapping_vertical_packing(lut, bits)
        out = out.reshape(-1, out.shape[-2], out.shape[-1])[:, 0, :]
        return CiphertextList(out, degree=self.param.message_modulus - 1, noise_level=NOISE_NOMINAL)

    def wopbs(self, cts, lut):
        """:434-451: ciphertexts with one bit of padding."""
        basis = self.param.message_modulus * self.param.carry_modulus
        delta = (1 << 63) // basis
        return self._extract_bits_circuit_bootstrapping(cts, lut, delta.bit_length() - 1, int(math.log2(basis)))

    def programmable_bootstrapping_without_padding(self, cts, lut):
        """:472-495: ciphertexts without padding bit (delta doubled)."""
        basis = self.param.message_modulus * self.param.carry_modulus
        delta = (1 << 63) // basis * 2
        return self._extract_bits_circuit_bootstrapping(cts, lut, delta.bit_length() - 1, int(math.log2(basis)))
