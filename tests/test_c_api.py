"""A plain C client of include/b200tfhe.h (tests/c_api/test_shortint_ks_pbs.c, modelled on the reference's tfhe/c_api_tests/*.c):
compiled with gcc against the shipped shared library; run on the GPU it uploads keys, keyswitches, bootstraps and decrypts."""
import os
import subprocess

import pytest

from conftest import ROOT

SRC = os.path.join(ROOT, "tests", "c_api", "test_shortint_ks_pbs.c")
EXE = os.path.join(ROOT, "tests", "c_api", "test_shortint_ks_pbs")


def _build():
    import __graft_entry__ as g
    g.build()
    pkg, orc = os.path.join(ROOT, "tfhe-rs-odd_b200"), os.path.join(ROOT, "oracle")
    cmd = ["gcc", "-O1", "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", orc, SRC, "-o", EXE,
           "-L", pkg, "-lb200tfhe", "-L", orc, "-loracle", f"-Wl,-rpath,{pkg}", f"-Wl,-rpath,{orc}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_c_client_compiles_and_links_against_the_library():
    _build()
    r = subprocess.run([EXE, "--link-only"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.startswith("link ok"), r.stderr


@pytest.mark.gpu
def test_c_client_keyswitch_and_bootstrap_on_the_gpu():
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok:")
