"""The kernels' libm ON THE DEVICE against the host build of the same header that the pinned reference links, bit for bit.

tinyfk_b200/csrc/fkmath.h defines sin / cos / asin / atan2 as sequences of IEEE operations (+ - * / sqrt fma); the strict FP64
parity claim rests on those sequences giving the same bits under nvcc (-fmad=false) and gcc.  The kernel parity tests establish that
on the arguments forward kinematics produces; this test feeds millions of arbitrary ones -- every quadrant and reduction count,
the split points of asin and atan2, huge / tiny / subnormal / non-finite values -- through a probe kernel (tests/cuda/libm_probe.cu,
built here with the strict translation unit's flags) and through oracle/build/libstrict_libm.so."""
import ctypes
import math
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import refapi

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
_dp = ctypes.POINTER(ctypes.c_double)


def _p(a):
    return a.ctypes.data_as(_dp)


@pytest.fixture(scope="module")
def probe(tmp_path_factory):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("no nvcc on this box: the probe kernel cannot be built")
    out = str(tmp_path_factory.mktemp("probe") / "libm_probe.so")
    subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-fmad=false", "-std=c++17", "-shared",
                    "-Xcompiler", "-fPIC", "-cudart", "static", "-o", out, os.path.join(HERE, "cuda", "libm_probe.cu")], check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope="module")
def host():
    path = os.path.join(os.path.dirname(refapi.port_lib_path()), "libstrict_libm.so")
    if not os.path.exists(path):
        refapi.build_port()
    return ctypes.CDLL(path)


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    nan = np.isnan(a)
    return np.array_equal(nan, np.isnan(b)) and np.array_equal(a[~nan].view(np.uint64), b[~nan].view(np.uint64))


def _specials():
    v = [0.0, 1.0, 0.5, 0.3, 0.25, 0.75, 0.975, 0.4375, 0.6875, 1.1875, 2.4375, 1e-300, 5e-324, 2.2250738585072014e-308, 1e300,
         1.7976931348623157e308, 2.0 ** 1022, 2.0 ** -900, 2.0 ** -27, 2.0 ** -29, 7.450580596923828125e-9, np.inf, np.nan,
         math.pi, math.pi / 2, math.pi / 4, 3 * math.pi / 4, 1647099.0, 1646592.0, 1647098.9, 1e6, 2.7e11, 2.0 ** 38, 2.0 ** 40, 2.0 ** 66,
         1.0 - 2 ** -53, 1.0 + 2 ** -52, 0.5 - 2 ** -54, 0.5 + 2 ** -53, 0.7853981633974483, 0.7853981633974484]
    v = np.array(v)
    return np.concatenate([v, -v])


def _angles(rng, n):
    sp = _specials()
    return np.concatenate([rng.uniform(-3.2, 3.2, n), rng.uniform(-0.8, 0.8, n // 4), rng.uniform(-100, 100, n // 4),
                           rng.uniform(-1.65e6, 1.65e6, n // 4), rng.uniform(-3e11, 3e11, n // 16),
                           rng.normal(size=n // 4) * 10.0 ** rng.integers(-300, 3, n // 4),
                           np.concatenate([k * math.pi / 2 + rng.uniform(-1e-9, 1e-9, 64) for k in range(-200, 201)]), sp])


def test_sincos_device_equals_host(probe, host):
    rng = np.random.default_rng(21)
    x = _angles(rng, 2_000_000)
    s0, c0, s1, c1 = (np.empty_like(x) for _ in range(4))
    host.tfkm_sincos_v(_p(x), _p(s0), _p(c0), ctypes.c_long(x.size))
    assert probe.probe_sincos(_p(x), _p(s1), _p(c1), ctypes.c_long(x.size)) == 0
    assert _same(s0, s1) and _same(c0, c1)
    # two arguments per call (what the specialised walk of the opt-in modes uses), rare inputs next to ordinary ones
    xb = np.roll(x, 4321)
    sa, ca, sb, cb = (np.empty_like(x) for _ in range(4))
    assert probe.probe_sincos2(_p(x), _p(xb), _p(sa), _p(ca), _p(sb), _p(cb), ctypes.c_long(x.size)) == 0
    assert _same(s0, sa) and _same(c0, ca) and _same(np.roll(s0, 4321), sb) and _same(np.roll(c0, 4321), cb)


def test_asin_device_equals_host(probe, host):
    rng = np.random.default_rng(22)
    n = 2_000_000
    x = np.concatenate([rng.uniform(-1, 1, n), rng.uniform(0.49, 0.51, n // 8) * rng.choice([-1, 1], n // 8),
                        1.0 - 10.0 ** rng.uniform(-16, -1, n // 8), rng.normal(size=n // 8) * 10.0 ** rng.integers(-300, 0, n // 8),
                        rng.uniform(-1.5, 1.5, 1000), _specials()])
    a0, a1 = np.empty_like(x), np.empty_like(x)
    host.tfkm_asin_v(_p(x), _p(a0), ctypes.c_long(x.size))
    assert probe.probe_asin(_p(x), _p(a1), ctypes.c_long(x.size)) == 0
    assert _same(a0, a1)


def test_atan2_device_equals_host(probe, host):
    rng = np.random.default_rng(23)
    n = 2_000_000
    sp = _specials()
    mx = rng.uniform(0.5, 2.0, n // 4)
    mn = mx * np.concatenate([rng.uniform(0.49, 0.51, n // 8), np.full(n // 16, 0.5), rng.uniform(0, 1, n // 16)])
    swap = rng.integers(0, 2, n // 4).astype(bool)
    ye = np.where(swap, mx, mn) * rng.choice([-1.0, 1.0], n // 4)
    xe = np.where(swap, mn, mx) * rng.choice([-1.0, 1.0], n // 4)
    e = rng.integers(-1070, 1020, n // 4)
    y = np.concatenate([rng.normal(size=n), rng.normal(size=n // 2) * 10.0 ** rng.integers(-300, 300, n // 2), ye, np.ldexp(ye, e),
                        np.ldexp(ye, rng.integers(-1000, 1000, n // 4)), np.repeat(sp, sp.size)])
    x = np.concatenate([rng.normal(size=n), rng.normal(size=n // 2) * 10.0 ** rng.integers(-300, 300, n // 2), xe, np.ldexp(xe, e),
                        xe, np.tile(sp, sp.size)])
    o0, o1 = np.empty_like(x), np.empty_like(x)
    host.tfkm_atan2_v(_p(y), _p(x), _p(o0), ctypes.c_long(x.size))
    assert probe.probe_atan2(_p(y), _p(x), _p(o1), ctypes.c_long(x.size)) == 0
    assert _same(o0, o1)


def test_fkmath_h_functions_as_device_code(probe, host):
    """the rare-input branches of the kernels call the fkmath.h functions themselves on the device (big-argument reduction,
    atan2's rescaling and special values, asin at and beyond 1)"""
    rng = np.random.default_rng(24)
    sp = _specials()
    x = np.concatenate([_angles(rng, 200_000), rng.uniform(-1.2, 1.2, 200_000), np.tile(sp, sp.size)])
    m = x.size - sp.size * sp.size
    y = np.concatenate([rng.normal(size=m) * 10.0 ** rng.integers(-300, 300, m), np.repeat(sp, sp.size)])
    s0, c0, as0, at0, s1, c1, as1, at1 = (np.empty_like(x) for _ in range(8))
    host.tfkm_sincos_v(_p(x), _p(s0), _p(c0), ctypes.c_long(x.size))
    host.tfkm_asin_v(_p(x), _p(as0), ctypes.c_long(x.size))
    host.tfkm_atan2_v(_p(y), _p(x), _p(at0), ctypes.c_long(x.size))
    assert probe.probe_plain(_p(y), _p(x), _p(s1), _p(c1), _p(as1), _p(at1), ctypes.c_long(x.size)) == 0
    assert _same(s0, s1) and _same(c0, c1) and _same(as0, as1) and _same(at0, at1)
