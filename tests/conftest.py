import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    """The product .so files are built in-tree by __graft_entry__.build(); if a fresh checkout lacks
    them (they are git-ignored), build once here so the tests exercise the real native code."""
    lib = os.path.join(ROOT, "tinyfk_b200", "libtinyfk_b200.so")
    mods = [f for f in os.listdir(os.path.join(ROOT, "tinyfk_b200")) if f.startswith("_tinyfk") and f.endswith(".so")]
    if not os.path.exists(lib) or not mods:
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tinyfk_b200", "csrc"), "all"], check=True)
    from oracle import refapi

    if not os.path.exists(refapi.port_lib_path(False)) or not os.path.exists(refapi.port_lib_path(True)):
        refapi.build_port()
    if not refapi.have_ref(False) or not refapi.have_ref(True):
        refapi.build_ref()  # only possible where /root/reference exists; otherwise tests fall back to the port


_ensure_built()


@pytest.fixture(scope="session")
def has_ref():
    from oracle import refapi

    return refapi.have_ref(False) and refapi.have_ref(True)
