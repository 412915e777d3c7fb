"""Named parameter sets of the path (values restated from the reference):
  shortint/parameters/mod.rs:613-627 (1_1), :703-717 (2_2), :853-867 (3_3); boolean/parameters/mod.rs:123-134.
Tuple: (n, k, N, pbs_base_log, pbs_level, ks_base_log, ks_level, torus_bits, message_modulus*carry_modulus, pbs_order)."""
from ._native import Params

PBS_ORDER_KS_PBS = "ks_pbs"   # EncryptionKeyChoice::Big  -> KeyswitchBootstrap
PBS_ORDER_PBS_KS = "pbs_ks"   # EncryptionKeyChoice::Small -> BootstrapKeyswitch

_TABLE = {
    "BOOLEAN_DEFAULT_PARAMETERS": (722, 2, 512, 6, 3, 3, 4, 32, 0, PBS_ORDER_PBS_KS),
    "PARAM_MESSAGE_1_CARRY_1_KS_PBS": (684, 3, 512, 18, 1, 4, 3, 64, 4, PBS_ORDER_KS_PBS),
    "PARAM_MESSAGE_2_CARRY_2_KS_PBS": (742, 1, 2048, 23, 1, 3, 5, 64, 16, PBS_ORDER_KS_PBS),
    "PARAM_MESSAGE_3_CARRY_3_KS_PBS": (864, 1, 8192, 15, 2, 3, 6, 64, 64, PBS_ORDER_KS_PBS),
}
_TABLE["PARAM_MESSAGE_1_CARRY_1"] = _TABLE["PARAM_MESSAGE_1_CARRY_1_KS_PBS"]
_TABLE["PARAM_MESSAGE_2_CARRY_2"] = _TABLE["PARAM_MESSAGE_2_CARRY_2_KS_PBS"]
_TABLE["PARAM_MESSAGE_3_CARRY_3"] = _TABLE["PARAM_MESSAGE_3_CARRY_3_KS_PBS"]


def get(name):
    t = _TABLE[name]
    return Params(*t[:8])


def plaintext_modulus(name):
    return _TABLE[name][8]


def pbs_order(name):
    return _TABLE[name][9]


def names():
    return sorted(_TABLE)
