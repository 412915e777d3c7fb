"""Import shim: the package directory is `tfhe-rs-odd_b200/` (hyphenated); this exposes it as `tfhe_rs_odd_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tfhe-rs-odd_b200")
_spec = importlib.util.spec_from_file_location("tfhe_rs_odd_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["tfhe_rs_odd_b200"] = _mod
_spec.loader.exec_module(_mod)
