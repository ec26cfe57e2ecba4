"""precision="fp64_fast" (TFK_F64_FAST; SURVEY.md section 7 hard part 4): double arithmetic with FMA contraction, the
platform libm, reciprocal-multiply quotients and ANALYTIC base-rotation Jacobian columns.

Contract (include/tinyfk_b200.h), against the STOCK reference (oracle/_ref, glibc):
  * poses and analytic Jacobian entries (joint columns, base-translation columns): <= 1e-12 (relative above 1 for the
    1/cos(pitch)-amplified RPY rows, as for the strict mode);
  * base-rotation columns: the exact derivative.  The reference's eps = 1e-7 forward difference differs from it by its own
    truncation + cancellation error (~1e-7 |f''| + 1e-16/1e-7): compared at 5e-6, and compared at 1e-9 with a CENTRAL
    difference of the strict FK evaluated in extended step (h = 1e-5) -- the derivative itself.
Strict FP64 stays the default and the headline."""
import numpy as np
import pytest

import parity
from emuapi import MODE_FAST64, EmuModel
from oracle import refapi
from tinyfk_b200 import workloads as W

needs_ref = pytest.mark.skipif(not refapi.have_ref(False), reason="oracle/_ref not built")
N = {"cfg2": 512, "cfg3": 512, "cfg4": 96, "cfg5": 48}


def _check(P, J, Pr, Jr, q, solve_strict, n_joint, base, rot, what):
    pd = W.pose_dim(rot)
    finite = np.isfinite(Pr).all(axis=1)
    assert parity.rel_err(P[finite], Pr[finite]).max() <= parity.TOL_F64, what
    fd = parity.fd_column_mask(n_joint, base, J.shape[1])
    bad = np.repeat(parity.gimbal_rows(Pr, rot) | ~finite, pd)
    Ja, Jb = J[~bad], Jr[~bad]
    ea = parity.rel_err(Ja[:, ~fd], Jb[:, ~fd]).max() if (~fd).any() else 0.0
    assert ea <= parity.TOL_F64, "%s analytic columns %.3e" % (what, ea)
    if fd.any():
        ef = parity.rel_err(Ja[:, fd], Jb[:, fd]).max()
        assert ef <= 5e-6, "%s base-rotation columns vs the reference's forward difference %.3e" % (what, ef)
        # the derivative itself: central difference of the strict FK, h = 1e-5 (error ~ h^2 |f'''| / 6 + 1e-16 / h)
        h = 1e-5
        cols = np.nonzero(fd)[0]
        m = min(16, q.shape[0])
        for c in cols:
            qp, qm = q[:m].copy(), q[:m].copy()
            qp[:, c] += h
            qm[:, c] -= h
            Pp, Pm = solve_strict(qp), solve_strict(qm)
            d = (Pp - Pm) / (2 * h)
            if rot == W.ROT_RPY:  # angle wrap-around at +-pi
                d[:, 3:] = (np.mod(Pp[:, 3:] - Pm[:, 3:] + np.pi, 2 * np.pi) - np.pi) / (2 * h)
            got = J[:d.shape[0] * pd, c].reshape(d.shape[0], pd)
            ok = ~(parity.gimbal_rows(Pr, rot) | ~finite)[:d.shape[0]]
            e = parity.rel_err(got[ok], d[ok]).max()
            assert e <= 2e-8, "%s column %d vs central difference %.3e" % (what, c, e)


@needs_ref
@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_fast64_core_contract_on_cpu(cfg):
    xml = W.urdf_text(W.CONFIGS[cfg][0])
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=False)
    W.setup(o, cfg)
    jids, lids, rot, base = W.setup(e, cfg)
    q = W.sample_q(o.get_joint_limits(jids), base, N[cfg], seed=41)
    Pr, Jr = o.solve_fk(q, lids, jids, rot, base, True)
    P, J = e.solve_fk(q, lids, jids, rot, base, True, mode=MODE_FAST64)
    _check(P, J, Pr, Jr, q, lambda x: o.solve_fk(x, lids, jids, rot, base, False)[0], len(jids), base, rot, cfg)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg4", "cfg5"])
@pytest.mark.parametrize("jit", [0, 2])
def test_gpu_fast64_contract(cfg, jit):
    from tinyfk_b200 import capi

    xml = W.urdf_text(W.CONFIGS[cfg][0])
    m, o = capi.Model(xml, host_libm=1), refapi.RefModel(xml, strict=False)
    W.setup(o, cfg)
    jids, lids, rot, base = W.setup(m, cfg)
    m.set_jit_mode(jit)  # 0: generic kernels; 2: plan-specialised at first use
    n = N[cfg] * 4 + 3
    q = W.sample_q(o.get_joint_limits(jids), base, n, seed=42)
    Pr, Jr = o.solve_fk(q, lids, jids, rot, base, True)
    P, J = m.solve_fk(q, lids, jids, rot, base, True, precision=2)
    assert P.dtype == np.float64
    if jit == 2:
        assert "spec" in m.last_launch_info()["kernel"], m.last_launch_info()
    _check(P, J, Pr, Jr, q, lambda x: o.solve_fk(x, lids, jids, rot, base, False)[0], len(jids), base, rot, cfg)
    Ps, Js = m.solve_fk(q, lids, jids, rot, base, True, precision=0)  # strict and fast agree to 1e-12 where both are analytic
    fd = parity.fd_column_mask(len(jids), base, J.shape[1])
    ok = np.isfinite(Ps).all(axis=1) & ~parity.gimbal_rows(Ps, rot)
    assert parity.rel_err(P[ok], Ps[ok]).max() <= parity.TOL_F64
    okj = np.repeat(ok, W.pose_dim(rot))
    assert parity.rel_err(J[okj][:, ~fd], Js[okj][:, ~fd]).max() <= parity.TOL_F64


@pytest.mark.gpu
def test_gpu_fast64_facade_and_other_entries():
    import tinyfk_b200
    from tinyfk_b200 import BaseType, KinematicModel, RotationType

    kin = KinematicModel(tinyfk_b200.pr2_urdfpath())
    jids = kin.get_joint_ids(W.PR2_RARM)
    lid = kin.get_link_ids(["r_gripper_tool_frame"])[0]
    q = np.array(W.README_Q)
    P, J = kin.solve_fk(q, lid, jids, RotationType.RPY, BaseType.PLANER, True, precision="fp64_fast")
    Ps, Js = kin.solve_fk(q, lid, jids, RotationType.RPY, BaseType.PLANER, True)
    assert np.abs(P - Ps).max() < 1e-12 and np.abs(J[:, :9] - Js[:, :9]).max() < 1e-12
    assert np.abs(J[:, 9] - Js[:, 9]).max() < 5e-6
    # the centre-of-mass, pair and inertia entries accept the mode too
    c, Jc = kin.solve_com_fk(q[:7], jids, with_jacobian=True, precision="fp64_fast")
    cs, Jcs = kin.solve_com_fk(q[:7], jids, with_jacobian=True)
    assert np.abs(c - cs).max() < 1e-12 and np.abs(Jc - Jcs).max() < 1e-12
    pairs = [(lid, kin.get_link_ids(["l_gripper_tool_frame"])[0])]
    s, G = kin.compute_inter_link_sqdists(q[:7], pairs, jids, with_jacobian=True, precision="fp64_fast")
    ss, Gs = kin.compute_inter_link_sqdists(q[:7], pairs, jids, with_jacobian=True)
    assert np.abs(s - ss).max() < 1e-12 and np.abs(G - Gs).max() < 1e-12
    I = kin.compute_total_inertia_matrix(q[:7], jids, precision="fp64_fast")
    Is = kin.compute_total_inertia_matrix(q[:7], jids)
    assert np.abs(I - Is).max() < 1e-11 * np.abs(Is).max()
