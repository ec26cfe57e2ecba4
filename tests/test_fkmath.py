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
def test_sin_cos_below_one_and_a_half_ulp(lib, lo, hi):
    """no tail word in the reduced argument (fkmath.h, "sin / cos"): its half ulp adds to the kernels' own error"""
    x = np.random.default_rng(1).uniform(lo, hi, 3000)
    assert _max_ulp(_call1(lib, "tfkm_sin_v", x), lambda a: mp.sin(mp.mpf(float(a))), x) < 1.5
    assert _max_ulp(_call1(lib, "tfkm_cos_v", x), lambda a: mp.cos(mp.mpf(float(a))), x) < 1.5


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


def test_atan2_one_division_core_at_its_edges(lib):
    """the split at mn / mx = 1/2 (both sides of it, every quadrant and both orders of |x|, |y|), magnitudes that the fkmath.h
    function rescales before the core (beyond 2^1022, below 2^-900, subnormals), exponents hundreds of binades apart; the
    device variant (fkmath_simt.h) on the same arguments, bit for bit"""
    rng = np.random.default_rng(16)
    n = 4000
    mx = rng.uniform(0.5, 2.0, n)
    ratio = np.concatenate([rng.uniform(0.49, 0.51, n // 2), np.full(n // 4, 0.5), rng.uniform(0.0, 1.0, n // 4)])
    mn = mx * ratio
    swap = rng.integers(0, 2, n).astype(bool)
    y = np.where(swap, mx, mn) * rng.choice([-1.0, 1.0], n)
    x = np.where(swap, mn, mx) * rng.choice([-1.0, 1.0], n)
    e = rng.integers(-1070, 1020, n)
    ys, xs = np.ldexp(y, e), np.ldexp(x, e)                       # same ratio at every magnitude, subnormals included
    yw = np.ldexp(y, rng.integers(-1000, 1000, n))                 # ... and far apart
    Y, X = np.concatenate([y, ys, yw]), np.concatenate([x, xs, x])
    keep = (X != 0) | (Y != 0)
    Y, X = Y[keep], X[keep]
    o0, o1 = np.empty_like(X), np.empty_like(X)
    lib.tfkm_atan2_v(Y.ctypes.data_as(_dp), X.ctypes.data_as(_dp), o0.ctypes.data_as(_dp), ctypes.c_long(X.size))
    lib.tfkm_atan2_simt_v(Y.ctypes.data_as(_dp), X.ctypes.data_as(_dp), o1.ctypes.data_as(_dp), ctypes.c_long(X.size))
    assert _bits_equal(o0, o1)
    nz = (X != 0) & (Y != 0)  # mpmath has no signed zeros; those cases are in the special-value table above
    sub = (np.abs(X) < 2.3e-308) | (np.abs(Y) < 2.3e-308)  # a subnormal operand carries fewer bits: its own rounding dominates
    ok = nz & ~sub
    assert _max_ulp(o0[ok], lambda a: mp.atan2(mp.mpf(float(a[0])), mp.mpf(float(a[1]))), list(zip(Y[ok], X[ok]))) < 2.0
    assert np.allclose(o0[nz & sub], np.arctan2(Y[nz & sub], X[nz & sub]), rtol=1e-15, atol=5e-324)


def _special_doubles():
    return np.array([0.0, -0.0, 1.0, -1.0, 0.5, -0.5, 0.3, 0.78125, 0.7853981633974483, 0.7853981633974484, 0.975,
                     0.4375, 0.6875, 1.1875, 2.4375, 1e-300, -1e-300, 5e-324, 1e300, np.inf, -np.inf, np.nan,
                     7.450580596923828125e-9, 1.862645149230957e-9, 1647099.0, 1646592.0, 2.0 ** 66, 2.0 ** -29,
                     math.pi, -math.pi, math.pi / 2, 3 * math.pi / 2, 1e6, -1e6, 2.7e11, 1.0 - 2 ** -53, 1.0 + 2 ** -52])


def _bits_equal(a, b):
    return np.array_equal(np.asarray(a).view(np.uint64), np.asarray(b).view(np.uint64))


def test_simt_variants_are_bit_identical(lib):
    """fkmath_simt.h (what the kernels run) against fkmath.h (what the pinned reference links): same bits,
    including signs of zeros and NaN payload-free specials."""
    rng = np.random.default_rng(11)
    sp = _special_doubles()
    n = 2_000_000
    x = np.concatenate([rng.uniform(-3.2, 3.2, n), rng.uniform(-0.8, 0.8, n // 4), rng.uniform(-1e6, 1e6, n // 4),
                        rng.uniform(-3e11, 3e11, 10000), rng.normal(size=n // 4) * 10.0 ** rng.integers(-12, 3, n // 4),
                        np.concatenate([k * math.pi / 2 + rng.uniform(-1e-9, 1e-9, 200) for k in range(-40, 41)]), sp])
    s0, c0, s1, c1 = (np.empty_like(x) for _ in range(4))
    lib.tfkm_sincos_v(x.ctypes.data_as(_dp), s0.ctypes.data_as(_dp), c0.ctypes.data_as(_dp), ctypes.c_long(x.size))
    lib.tfkm_sincos_simt_v(x.ctypes.data_as(_dp), s1.ctypes.data_as(_dp), c1.ctypes.data_as(_dp), ctypes.c_long(x.size))
    nn = ~np.isnan(s0)
    assert _bits_equal(s0[nn], s1[nn]) and _bits_equal(c0[nn], c1[nn]) and np.isnan(s1[~nn]).all() and np.isnan(c1[~nn]).all()

    # two arguments at once (tfk_sincos2_simt: what the plan-specialised walk calls for pairs of joints), every
    # argument paired with an unrelated one so that rare inputs meet ordinary ones on both sides
    xb = np.roll(x, 12345)
    sa, ca, sb, cb = (np.empty_like(x) for _ in range(4))
    lib.tfkm_sincos2_simt_v(x.ctypes.data_as(_dp), xb.ctypes.data_as(_dp), sa.ctypes.data_as(_dp), ca.ctypes.data_as(_dp),
                            sb.ctypes.data_as(_dp), cb.ctypes.data_as(_dp), ctypes.c_long(x.size))
    assert _bits_equal(s0[nn], sa[nn]) and _bits_equal(c0[nn], ca[nn]) and np.isnan(sa[~nn]).all() and np.isnan(ca[~nn]).all()
    nb = np.roll(nn, 12345)
    assert _bits_equal(np.roll(s0, 12345)[nb], sb[nb]) and _bits_equal(np.roll(c0, 12345)[nb], cb[nb])
    assert np.isnan(sb[~nb]).all() and np.isnan(cb[~nb]).all()

    x = np.concatenate([rng.uniform(-1, 1, n), rng.uniform(0.9, 1.0, n // 4), -rng.uniform(0.4, 0.6, n // 4),
                        rng.normal(size=n // 4) * 10.0 ** rng.integers(-12, 0, n // 4), sp])
    a0, a1 = _call1(lib, "tfkm_asin_v", x), _call1(lib, "tfkm_asin_simt_v", x)
    nn = ~np.isnan(a0)
    assert _bits_equal(a0[nn], a1[nn]) and np.isnan(a1[~nn]).all()

    y = np.concatenate([rng.normal(size=n), rng.normal(size=n // 2) * 10.0 ** rng.integers(-20, 20, n // 2),
                        rng.uniform(-1, 1, n // 2), np.repeat(sp, sp.size)])
    x = np.concatenate([rng.normal(size=n), rng.normal(size=n // 2) * 10.0 ** rng.integers(-20, 20, n // 2),
                        rng.uniform(0.4, 2.5, n // 2) * rng.choice([-1, 1], n // 2), np.tile(sp, sp.size)])
    # interval edges of atan: |y/x| right at the thresholds
    for thr in (0.4375, 0.6875, 1.1875, 2.4375):
        xe = rng.uniform(0.5, 2.0, 2000)
        y = np.concatenate([y, xe * thr, np.nextafter(xe * thr, 0), np.nextafter(xe * thr, 9)])
        x = np.concatenate([x, xe, xe, xe])
    o0, o1 = np.empty_like(x), np.empty_like(x)
    lib.tfkm_atan2_v(y.ctypes.data_as(_dp), x.ctypes.data_as(_dp), o0.ctypes.data_as(_dp), ctypes.c_long(x.size))
    lib.tfkm_atan2_simt_v(y.ctypes.data_as(_dp), x.ctypes.data_as(_dp), o1.ctypes.data_as(_dp), ctypes.c_long(x.size))
    nn = ~np.isnan(o0)
    assert _bits_equal(o0[nn], o1[nn]) and np.isnan(o1[~nn]).all()


def test_shared_reciprocal_division_is_ieee_division():
    """MathPortable::div_ (fk_math.h: one correctly rounded reciprocal + two FMA residual corrections per
    quotient) against the hardware division, bit for bit -- random operands, denominators within a few
    ulp of 1 (quaternion norms), the constant 1e-7, tiny / huge / zero / non-finite operands."""
    import emuapi
    L = emuapi.lib()
    L.emu_div.argtypes = [_dp, _dp, _dp, ctypes.c_long]
    rng = np.random.default_rng(12)
    n = 4_000_000
    one = 1.0 + rng.integers(-64, 64, n) * 2.0 ** -53
    sp = _special_doubles()
    a = np.concatenate([rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), rng.normal(size=n) * 1e-8,
                        rng.normal(size=n) * 10.0 ** rng.integers(-300, 300, n), np.repeat(sp, sp.size)])
    b = np.concatenate([one, rng.uniform(0.5, 2.0, n), np.full(n, 1e-7),
                        rng.normal(size=n) * 10.0 ** rng.integers(-300, 300, n), np.tile(sp, sp.size)])
    o = np.empty_like(a)
    L.emu_div(a.ctypes.data_as(_dp), b.ctypes.data_as(_dp), o.ctypes.data_as(_dp), ctypes.c_long(a.size))
    with np.errstate(all="ignore"):
        ref = a / b
    nn = ~np.isnan(ref)
    # an exact-zero quotient may differ in sign only where the numerator is an exact zero
    same = o[nn].view(np.uint64) == ref[nn].view(np.uint64)
    assert (same | ((o[nn] == 0) & (ref[nn] == 0))).all() and np.isnan(o[~nn]).all()
