"""Portable libm (tinyfk_b200/csrc/fkmath.h): accuracy against mpmath and special values.  The same
source runs on the GPU (nvcc -fmad=false); bit-equality of the two sides is what the GPU parity tests
establish end to end."""
import ctypes
import math
import os

import mpmath as mp
import numpy as np
import pytest

from oracle import refapi

_dp = ctypes.POINTER(ctypes.c_double)


@pytest.fixture(scope="module")
def lib():
    path = os.path.join(os.path.dirname(refapi.port_lib_path()), "libstrict_libm.so")
    if not os.path.exists(path):
        refapi.build_port()
    return ctypes.CDLL(path)


def _call1(lib, name, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    o = np.empty_like(x)
    getattr(lib, name)(x.ctypes.data_as(_dp), o.ctypes.data_as(_dp), ctypes.c_long(x.size))
    return o


def _max_ulp(got, exact_fn, args):
    mp.mp.dps = 40
    worst = 0.0
    for g, a in zip(got, args):
        ex = exact_fn(a)
        u = math.ulp(float(ex)) if float(ex) != 0 else 5e-324
        worst = max(worst, abs(float((mp.mpf(float(g)) - ex) / u)))
    return worst


@pytest.mark.parametrize("lo,hi", [(-0.7854, 0.7854), (-3.1416, 3.1416), (-100, 100), (-1e6, 1e6)])
def test_sin_cos_below_one_ulp(lib, lo, hi):
    x = np.random.default_rng(1).uniform(lo, hi, 3000)
    assert _max_ulp(_call1(lib, "tfkm_sin_v", x), lambda a: mp.sin(mp.mpf(float(a))), x) < 1.0
    assert _max_ulp(_call1(lib, "tfkm_cos_v", x), lambda a: mp.cos(mp.mpf(float(a))), x) < 1.0


def test_sin_cos_near_multiples_of_half_pi(lib):
    rng = np.random.default_rng(2)
    x = np.concatenate([k * math.pi / 2 + rng.uniform(-1e-6, 1e-6, 50) for k in range(-20, 21)])
    assert _max_ulp(_call1(lib, "tfkm_sin_v", x), lambda a: mp.sin(mp.mpf(float(a))), x) < 1.0
    assert _max_ulp(_call1(lib, "tfkm_cos_v", x), lambda a: mp.cos(mp.mpf(float(a))), x) < 1.0


def test_big_arguments_stay_accurate_up_to_2_38(lib):
    rng = np.random.default_rng(3)
    x = rng.uniform(1.6e6, 2.7e11, 2000) * rng.choice([-1, 1], 2000)
    mp.mp.dps = 40
    s, c = _call1(lib, "tfkm_sin_v", x), _call1(lib, "tfkm_cos_v", x)
    assert max(abs(float(mp.sin(mp.mpf(float(a))) - mp.mpf(float(v)))) for a, v in zip(x, s)) < 2.3e-16
    assert max(abs(float(mp.cos(mp.mpf(float(a))) - mp.mpf(float(v)))) for a, v in zip(x, c)) < 2.3e-16


def test_sincos_equals_separate_calls(lib):
    x = np.random.default_rng(4).uniform(-50, 50, 5000)
    s, c = np.empty_like(x), np.empty_like(x)
    lib.tfkm_sincos_v(x.ctypes.data_as(_dp), s.ctypes.data_as(_dp), c.ctypes.data_as(_dp), ctypes.c_long(x.size))
    assert np.array_equal(s, _call1(lib, "tfkm_sin_v", x)) and np.array_equal(c, _call1(lib, "tfkm_cos_v", x))


def test_asin_below_one_ulp(lib):
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(-1, 1, 3000), rng.uniform(0.97, 0.99999, 500), -rng.uniform(0.97, 0.99999, 500)])
    assert _max_ulp(_call1(lib, "tfkm_asin_v", x), lambda a: mp.asin(mp.mpf(float(a))), x) < 1.0
    sp = _call1(lib, "tfkm_asin_v", np.array([0.0, -0.0, 1.0, -1.0, 1.5, np.nan]))
    assert sp[0] == 0 and np.signbit(sp[1]) and sp[2] == math.pi / 2 and sp[3] == -math.pi / 2 and np.isnan(sp[4:]).all()


def test_atan2_below_two_ulp_and_special_values(lib):
    rng = np.random.default_rng(6)
    y = rng.normal(size=3000) * 10.0 ** rng.integers(-8, 8, 3000)
    x = rng.normal(size=3000) * 10.0 ** rng.integers(-8, 8, 3000)
    o = np.empty_like(x)
    lib.tfkm_atan2_v(y.ctypes.data_as(_dp), x.ctypes.data_as(_dp), o.ctypes.data_as(_dp), ctypes.c_long(x.size))
    assert _max_ulp(o, lambda a: mp.atan2(mp.mpf(float(a[0])), mp.mpf(float(a[1]))), list(zip(y, x))) < 2.0
    sy = np.array([0.0, -0.0, 0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.inf, 1.0, 1.0, 0.0, -0.0, 1e-300, 1e300])
    sx = np.array([1.0, 1.0, -1.0, -1.0, 0.0, 0.0, np.inf, -np.inf, 1.0, np.inf, -np.inf, 0.0, -0.0, 1e300, 1e-300])
    so = np.empty_like(sx)
    lib.tfkm_atan2_v(sy.ctypes.data_as(_dp), sx.ctypes.data_as(_dp), so.ctypes.data_as(_dp), ctypes.c_long(sx.size))
    g = np.arctan2(sy, sx)
    assert np.allclose(so, g, rtol=0, atol=3e-16) and (np.signbit(so) == np.signbit(g)).all()
