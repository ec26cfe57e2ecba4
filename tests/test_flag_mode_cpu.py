"""The branch-free strict arithmetic (csrc/fkmath_flag.h, MathPortableFlag) run on the CPU through tests/emu -- the very code
the plan-specialised strict kernels run as their first pass, with the two MUFU seeds emulated.  Contract: a configuration
that does not raise the rare-input flag gets the strict result BIT FOR BIT; ordinary inputs (random configurations, exact
zeros, the gimbal rows of a pose-only call) do not raise it; what the portable libm treats as rare does."""
import numpy as np
import pytest

from emuapi import MODE_STRICT64, MODE_STRICT64_FLAG, EmuModel
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W


def _bits_equal(a, b):
    return np.array_equal(a.view(np.int64), b.view(np.int64)) or np.array_equal(a, b, equal_nan=True)


def _setup(cfg):
    emu = EmuModel(W.urdf_text(W.CONFIGS[cfg][0]))
    j, l, rot, base = W.setup(emu, cfg)
    prod = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0)
    jp, _, _, _ = W.setup(prod, cfg)
    return emu, j, l, rot, base, prod.get_joint_limits(jp)


@pytest.mark.parametrize("cfg,n", [("cfg2", 1500), ("cfg3", 1500), ("cfg4", 60), ("cfg5", 40)])
@pytest.mark.parametrize("two_phase", [False, True])
def test_ordinary_inputs_take_the_first_pass_and_equal_strict_bit_for_bit(cfg, n, two_phase):
    emu, j, l, rot, base, lim = _setup(cfg)
    q = W.sample_q(lim, base, n, seed=21, wide_pitch=True)
    q[3] = 0.0                      # the zero configuration: exact zeros everywhere (atan2(0, 1), asin(0), 0 / norm)
    q[4, :len(j)] = 0.0
    if base == W.BASE_FLOATING:     # exact gimbal rows: pitch = +-pi/2 takes the |sarg| >= 0.99999 form of getRPY
        q[5, len(j) + 4] = np.pi / 2
        q[6, len(j) + 4] = -np.pi / 2
    tp = 16 if two_phase else 0
    P0, J0 = emu.solve_fk(q, l, j, rot, base, True, mode=MODE_STRICT64 | tp)
    P1, J1 = emu.solve_fk(q, l, j, rot, base, True, mode=MODE_STRICT64_FLAG | tp)
    fl = emu.last_flags(n)
    # the exact gimbal rows are reported when the Jacobian is asked for: rpy_derivative needs sin/cos of pitch = +-pi/2
    # exactly, which is the rare second stage of the argument reduction (for random orientations |sin pitch| >= 0.99999
    # has probability 1e-5).  Nothing else is.
    # Same for PR2's zero configuration in cfg5: one of its 50 links looks backwards there, getRPY returns yaw = pi exactly
    # and rpy_derivative takes sin/cos of it.  Nothing else is reported.
    allowed = np.zeros(n, dtype=bool)
    if base == W.BASE_FLOATING:
        allowed[[3, 5, 6]] = True
    assert not fl[~allowed].any(), (cfg, np.nonzero(fl)[0][:8], fl[fl != 0][:8])
    ok = fl == 0
    nl, pd = len(l), W.pose_dim(rot)
    assert _bits_equal(P0[np.repeat(ok, nl)], P1[np.repeat(ok, nl)]) and _bits_equal(J0[np.repeat(ok, nl * pd)], J1[np.repeat(ok, nl * pd)])
    # pose only: the gimbal rows take the first pass as well
    P2, _ = emu.solve_fk(q, l, j, rot, base, False, mode=MODE_STRICT64_FLAG | tp)
    assert not emu.last_flags(n).any() and _bits_equal(P0, P2)


@pytest.mark.parametrize("cfg", ["cfg2", "cfg3"])
def test_rare_inputs_are_reported_and_everything_else_is_still_exact(cfg):
    emu, j, l, rot, base, lim = _setup(cfg)
    n = 400
    q = W.sample_q(lim, base, n, seed=22)
    rare = {10: 1e7, 20: np.inf, 30: np.nan, 40: -3.3e9, 50: 1647099.5, 60: 2.0 ** 40}
    for r, v in rare.items():
        q[r, r % q.shape[1]] = v
    near = 3 * np.pi * (1 + 2.0 ** -45)   # half of it sits within 2^-44 of 3 pi / 2: the reduction's rare second stage
    q[70, 0] = near
    P0, J0 = emu.solve_fk(q, l, j, rot, base, True, mode=MODE_STRICT64)
    P1, J1 = emu.solve_fk(q, l, j, rot, base, True, mode=MODE_STRICT64_FLAG)
    fl = emu.last_flags(n)
    for r in list(rare) + [70]:
        assert fl[r] != 0, r
    ok = fl == 0
    assert ok.sum() >= n - 8
    pd = W.pose_dim(rot)
    assert _bits_equal(P0[ok], P1[ok]) and _bits_equal(J0[np.repeat(ok, pd)], J1[np.repeat(ok, pd)])
