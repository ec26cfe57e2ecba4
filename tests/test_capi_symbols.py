"""The C-ABI library loads without a GPU and exports every symbol include/tinyfk_b200.h declares;
host-only entries work, compute entries fail loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for f in os.listdir(os.path.join(ROOT, "include")):
        if f.endswith(".h"):
            text = open(os.path.join(ROOT, "include", f)).read()
            names |= set(re.findall(r"TFK_API\s+[\w\s\*]+?\b(tfk_\w+)\s*\(", text))
    return names


def test_every_declared_symbol_is_exported():
    declared = _declared()
    assert len(declared) >= 25 and declared == set(capi.SYMBOLS)
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert b"sm_100a" in capi.lib().tfk_version()


def test_library_contains_sm100a_kernels_only():
    out = os.popen("cuobjdump -lelf %s 2>/dev/null" % capi.LIB_PATH).read()
    if out.strip():
        assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_compute_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    m = capi.Model(W.urdf_text("kuka.urdf"))
    j, l = m.get_joint_ids(W.KUKA_JOINTS), m.get_link_ids(["lbr_iiwa_link_7"])
    with pytest.raises(capi.TfkError, match="(?i)cuda"):
        m.solve_fk(np.zeros((4, 7)), l, j, 2, 0, True)
    assert capi.kernel_launch_count() == 0


def test_product_never_references_the_oracle():
    """the product tree must not import, link or load anything under oracle/ or tests/"""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "tinyfk_b200")):
        if "build" in base:
            continue
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", ".hpp", "Makefile")):
                text = open(os.path.join(base, f), errors="ignore").read()
                if re.search(r"^\s*(import|from)\s+oracle|#\s*include\s*[\"<][^\n]*(oracle|emu)|(CDLL|dlopen)\([^\n]*(oracle|_ref|emu)"
                             r"|-l(fk_oracle|tinyfk_ref)|\.\./oracle", text, re.M):
                    bad.append(os.path.join(base, f))
    assert not bad, bad
