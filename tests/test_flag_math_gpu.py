"""The branch-free strict FP64 arithmetic of tinyfk_b200/csrc/fkmath_flag.h (first pass of the plan-specialised strict
kernels) against what it must equal, element by element on the GPU through the C-ABI test hook tfk_debug_flag_math:
every result that does not carry the rare-input flag is BIT-IDENTICAL to the ordinary device function of the strict
kernels (and, for 1/x, x/y and sqrt, to IEEE-754 as numpy computes it on the host); ordinary inputs -- including the
exact zeros of a robot at its zero configuration -- do not raise the flag."""
import numpy as np
import pytest

from tinyfk_b200 import capi

pytestmark = pytest.mark.gpu
N = 1 << 21


def _same_bits(a, b):
    return (a.view(np.int64) == b.view(np.int64)) | (np.isnan(a) & np.isnan(b))


def _rand_doubles(rng, n, emin, emax):
    """random sign, uniformly random exponent in [emin, emax], random 52-bit significand"""
    e = rng.integers(emin + 1023, emax + 1023 + 1, n, dtype=np.int64)
    m = rng.integers(0, 1 << 52, n, dtype=np.int64)
    s = rng.integers(0, 2, n, dtype=np.int64)
    return ((s << 63) | (e << 52) | m).view(np.float64)


def _edge_significands(e=0):
    m = np.array([0, 1, 2, (1 << 52) - 1, (1 << 52) - 2, 1 << 51, (1 << 51) - 1, (1 << 51) + 1, 0xFFFFFFFF, 0xFFFFF00000000,
                  (1 << 52) - (1 << 32), 0x5555555555555, 0xAAAAAAAAAAAAA], dtype=np.int64)
    return (((e + 1023) << 52) | m).view(np.float64)


def test_reciprocal_is_correctly_rounded():
    rng = np.random.default_rng(1)
    x = np.concatenate([_rand_doubles(rng, N, -500, 500), rng.uniform(0.5, 2.0, N), _edge_significands(), -_edge_significands(7),
                        np.array([1.0, -1.0, 3.0, 1e-200, 1e200, 0.0, np.inf, np.nan, 1e-300, 1e300, 5e-324])])
    out, ref, fl = capi.debug_flag_math(0, x)
    ok = fl == 0
    with np.errstate(all="ignore"):
        ieee = 1.0 / x
    assert _same_bits(out[ok], ieee[ok]).all() and _same_bits(out[ok], ref[ok]).all()
    assert _same_bits(ref, ieee).all()  # __drcp_rn itself, everywhere
    inrange = (np.abs(x) >= 2.0 ** -511) & (np.abs(x) < 2.0 ** 513)
    assert not fl[inrange].any() and fl[~inrange].all()
    # significand all ones (1 - 2^-53 is what the norm of a unit quaternion rounds to) and its neighbours, every exponent
    # of the contract, both signs: the one case the last Markstein step does not cover by itself (csrc/fkmath_flag.h)
    e = np.arange(-511, 513, dtype=np.int64) + 1023
    for m in ((1 << 52) - 1, (1 << 52) - 2, (1 << 52) - (1 << 32), (1 << 52) - (1 << 32) - 1, 0xFFFFFFFF, 0, 1):
        v = ((e << 52) | m).view(np.float64)
        v = np.concatenate([v, -v])
        out, ref, fl = capi.debug_flag_math(0, v)
        assert not fl.any() and _same_bits(out, 1.0 / v).all(), hex(m)


def test_division_is_ieee_division():
    rng = np.random.default_rng(2)
    a = np.concatenate([_rand_doubles(rng, N, -300, 300), rng.uniform(-4, 4, N), np.zeros(8), -np.zeros(8), _edge_significands(3)])
    b = np.concatenate([_rand_doubles(rng, N, -300, 300), rng.uniform(0.1, 4, N) * rng.choice([-1.0, 1.0], N),
                        rng.uniform(-3, 3, 16), _edge_significands(-2)[::-1]])
    out, ref, fl = capi.debug_flag_math(1, a, b)
    with np.errstate(all="ignore"):
        ieee = a / b
    ok = fl == 0
    assert _same_bits(ref, ieee).all()
    assert _same_bits(out[ok], ieee[ok]).all()  # incl. the sign of a zero quotient
    ordinary = ((np.abs(a) >= 2.0 ** -399) & (np.abs(a) < 2.0 ** 401) | (a == 0)) & (np.abs(b) >= 2.0 ** -511) & (np.abs(b) < 2.0 ** 513)
    assert not fl[ordinary].any()
    assert ok.mean() > 0.99
    # out-of-contract operands are reported
    a2 = np.array([1e-200, 1.0, np.nan, 1.0, np.inf, 1.0, 1e300])
    b2 = np.array([3.0, 0.0, 1.0, np.inf, 2.0, 1e-300, 3.0])
    assert capi.debug_flag_math(1, a2, b2)[2].all()


def test_square_root_is_correctly_rounded():
    rng = np.random.default_rng(3)
    x = np.abs(np.concatenate([_rand_doubles(rng, N, -900, 900), rng.uniform(0, 4, N), _edge_significands(), _edge_significands(1),
                               np.array([1.0, 4.0, 2.0, 0.25, 1e-290, 1e300])]))
    out, ref, fl = capi.debug_flag_math(2, x)
    ok = fl == 0
    ieee = np.sqrt(x)
    assert _same_bits(ref, ieee).all()
    assert _same_bits(out[ok], ieee[ok]).all() and not fl.any()
    bad = np.array([0.0, -1.0, np.inf, np.nan, 1e-310, 2.0 ** -980])
    assert capi.debug_flag_math(2, bad)[2].all()


def test_sincos_equals_the_simt_version():
    rng = np.random.default_rng(4)
    k = np.arange(-4000, 4000)
    near = (k * (np.pi / 2))[:, None] * (1 + np.array([0, 1e-17, -1e-16, 3e-13, -2e-9, 1e-6]))[None, :]
    x = np.concatenate([rng.uniform(-np.pi, np.pi, N), rng.uniform(-1.6e6, 1.6e6, N), rng.uniform(-1e-8, 1e-8, 1000), near.ravel(),
                        np.array([0.0, -0.0, np.pi / 4, -np.pi / 4, 0.7853981633974483, 0.7853981633974484, 1e-300, 2.0 ** -27,
                                  2.0 ** -28, 0.3, 0.78125, 1647098.9, 1647099.0, 1e7, 1e22, np.inf, np.nan])])
    out, ref, fl = capi.debug_flag_math(3, x)
    ok = np.tile(fl == 0, 2)
    assert _same_bits(out[ok], ref[ok]).all()
    n = x.shape[0]
    assert fl[-5:].all()  # huge, inf, nan
    assert fl[:N].mean() < 1e-4 and not fl[2 * N:2 * N + 1000].any()  # ordinary angles: (almost) never
    assert not fl[n - 17:n - 6].any()  # zeros, pi/4 boundary, tiny: ordinary


def test_asin_equals_the_simt_version():
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(-1, 1, N), rng.uniform(-1e-7, 1e-7, 1000), 1 - 10.0 ** rng.uniform(-17, -1, 100000),
                        -(1 - 10.0 ** rng.uniform(-17, -1, 100000)), rng.uniform(0.97, 0.98, 100000), rng.uniform(0.49, 0.51, 100000),
                        np.array([0.0, -0.0, 0.5, -0.5, 0.975, -0.975, 2.0 ** -27, 2.0 ** -28, 1e-300, 5e-324, np.nextafter(1.0, 0), 0.99999,
                                  -0.99999, 1.0, -1.0, 1.5, np.nan, np.inf])])
    out, ref, fl = capi.debug_flag_math(4, x)
    ok = fl == 0
    assert _same_bits(out[ok], ref[ok]).all()
    assert fl[-5:].all() and not fl[np.abs(x) < 1].any()  # |x| >= 1 and nan are reported, nothing else is


def test_atan2_equals_the_simt_version():
    rng = np.random.default_rng(6)
    y = np.concatenate([rng.uniform(-2, 2, N), _rand_doubles(rng, N // 4, -40, 40), rng.uniform(-1, 1, 1000)])
    x = np.concatenate([rng.uniform(-2, 2, N), _rand_doubles(rng, N // 4, -40, 40), np.ones(1000)])
    zs = np.array([0.0, -0.0, 0.0, -0.0, 0.0, -0.0, 0.0, -0.0, 1.0, -1.0, 1.0, -1.0, 0.3, -0.3, 1e-300, 2.0 ** -30])
    zx = np.array([1.0, 1.0, -1.0, -1.0, 0.0, 0.0, -0.0, -0.0, 0.0, 0.0, -0.0, -0.0, 1.0, 1.0, 1.0, 1.0])
    y, x = np.concatenate([y, zs]), np.concatenate([x, zx])
    out, ref, fl = capi.debug_flag_math(5, y, x)
    ok = fl == 0
    assert _same_bits(out[ok], ref[ok]).all()
    # ordinary pairs, x == 1.0 and the exact zeros are not reported; (1e-300, 1.0) is (exponents more than 2^60 apart)
    assert not fl[:N].any() and not fl[-1016:-2].any() and fl[-2] and not fl[-1]
    assert ok.mean() > 0.6  # the wide-exponent block holds many pairs more than 2^60 apart
    # the zeros follow the IEEE conventions: atan2(+-0, x > 0) = +-0, atan2(+-0, x < 0) = +-pi, atan2(y, +-0) = +-pi/2
    assert _same_bits(out[-16:-4], np.array([0.0, -0.0, np.pi, -np.pi, 0.0, -0.0, np.pi, -np.pi, np.pi / 2, -np.pi / 2, np.pi / 2,
                                             -np.pi / 2])).all()
    bad_y = np.array([np.nan, 1.0, np.inf, 1.0, 1e200])
    bad_x = np.array([1.0, np.nan, 1.0, np.inf, 1e-200])
    assert capi.debug_flag_math(5, bad_y, bad_x)[2].all()
