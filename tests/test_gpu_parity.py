"""GPU parity tests: the CUDA path, called through the C ABI (tfk_solve_fk_host / tfk_solve_fk_device),
against the oracle on the same seeded inputs.  FP64 strict mode must equal the libm-pinned reference
exactly (every entry, finite-difference columns included) and the stock reference within 1e-12
(its own FD noise floor on the FD columns); FP32 mode within 1e-5.
"""
import itertools

import numpy as np
import pytest

import parity
from parity import Pair
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu

SMALL_N = {"cfg1": 1, "cfg2": 4096, "cfg3": 4096, "cfg4": 640, "cfg5": 300}


def _inputs(pair, jids, base, n, seed=0, wide_pitch=False):
    return W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=seed, wide_pitch=wide_pitch)


@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_baseline_configs_bitwise_vs_pinned_reference(cfg):
    pair = Pair(W.CONFIGS[cfg][0], strict=True)
    jids, lids, rot, base = pair.setup(cfg)
    q = np.array([W.README_Q]) if cfg == "cfg1" else _inputs(pair, jids, base, SMALL_N[cfg])
    Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, True)
    P, J = pair.prod_solve(q, lids, jids, rot, base, True)
    assert P.shape == Pr.shape and J.shape == Jr.shape
    parity.assert_bitwise(P, J, Pr, Jr, cfg)
    assert np.abs(P - Pr).max() <= parity.TOL_F64 and np.abs(J - Jr).max() <= parity.TOL_F64


@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_baseline_configs_vs_stock_reference(cfg):
    pair = Pair(W.CONFIGS[cfg][0], strict=False)
    jids, lids, rot, base = pair.setup(cfg)
    q = np.array([W.README_Q]) if cfg == "cfg1" else _inputs(pair, jids, base, SMALL_N[cfg], seed=3)
    Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, True)
    P, J = pair.prod_solve(q, lids, jids, rot, base, True)
    parity.assert_close_stock(P, J, Pr, Jr, len(jids), base, rot, len(lids), cfg)


def _tree_pair(strict=True):
    pair = Pair("pr2.urdf", strict=strict, links_to_add=[("mylink", "r_upper_arm_link", [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])])
    jn = ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM
    ln = ["r_gripper_tool_frame", "l_gripper_tool_frame", "head_tilt_link", "base_footprint", "r_elbow_flex_link",
          "torso_lift_link", "l_forearm_link", "r_gripper_tool_frame", "base_link", "fl_caster_rotation_link",
          "r_upper_arm_link_com", "mylink"]
    jids, lids = pair.prod.get_joint_ids(jn), pair.prod.get_link_ids(ln)
    rng = np.random.default_rng(5)
    allj = list(range(pair.prod.n_joints))
    ang = rng.uniform(-0.3, 0.3, len(allj))
    lim = pair.prod.get_joint_limits(allj)
    ang = np.array([min(max(a, lo), hi) if np.isfinite(lo) else a for a, (lo, hi) in zip(ang, lim)])
    pair.set_state(allj, ang, [0.3, -0.2, 0.1, 0.2, -0.1, 0.7])  # non-selected joints and FIXED base use these
    return pair, jids, lids


@pytest.mark.parametrize("rot,base,jac", list(itertools.product((0, 1, 2), (0, 1, 2), (False, True))))
def test_all_modes_on_a_tree_bitwise(rot, base, jac):
    """every RotationType x BaseType x with_jacobian on a branching request (both arms, head, root link,
    duplicate link, irrelevant joints, a user link, a _com link); N is not a multiple of the tile"""
    pair, jids, lids = _tree_pair()
    q = _inputs(pair, jids, base, 333, seed=rot * 9 + base * 3 + int(jac))
    Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, jac)
    P, J = pair.prod_solve(q, lids, jids, rot, base, jac)
    parity.assert_bitwise(P, J, Pr, Jr, "tree rot%d base%d" % (rot, base))
    if not jac:
        assert not J.any()


def test_gimbal_branch_and_wide_pitch_bitwise():
    """base pitch drawn from [-pi/2, pi/2] plus exact +-pi/2 rows: the getRPY branch (pose.h:178-185)
    must be selected identically; entries there may be huge or inf but must be the same bits"""
    pair, jids, lids = _tree_pair()
    q = _inputs(pair, jids, W.BASE_FLOATING, 512, seed=11, wide_pitch=True)
    q[:8, -2] = np.pi / 2
    q[8:16, -2] = -np.pi / 2
    Pr, Jr = pair.oracle_solve(q, lids[:3], jids, 1, 2, True)
    P, J = pair.prod_solve(q, lids[:3], jids, 1, 2, True)
    parity.assert_bitwise(P, J, Pr, Jr, "gimbal")


def test_batch_equals_single_configs():
    """reference python/tests/test_tinyfk.py:133-154: batch result rows == single-config calls"""
    pair, jids, lids = _tree_pair()
    q = _inputs(pair, jids, 2, 10, seed=2)
    P, J = pair.prod_solve(q, lids, jids, 1, 2, True)
    n_link, n_dof = len(lids), q.shape[1]
    Ps, Js = P.reshape(10, n_link, 6), J.reshape(10, n_link * 6, n_dof)
    for i in range(10):
        P1, J1 = pair.prod_solve(q[i:i + 1], lids, jids, 1, 2, True)
        assert np.array_equal(Ps[i], P1) and np.array_equal(Js[i], J1)


def test_jacobian_matches_finite_differences_of_fk():
    """reference python/tests/test_tinyfk.py:84-130 (eps 1e-7, 7 decimals), FLOATING and PLANER"""
    pair, jids, lids = _tree_pair()
    for base in (2, 1):
        q0 = _inputs(pair, jids, base, 1, seed=4)
        P0, J = pair.prod_solve(q0, lids[:3], jids, 1, base, True)
        eps = 1e-7
        qp = np.repeat(q0, q0.shape[1], axis=0) + eps * np.eye(q0.shape[1])
        P1, _ = pair.prod_solve(qp, lids[:3], jids, 1, base, False)
        Jn = ((P1.reshape(q0.shape[1], -1) - P0.reshape(1, -1)) / eps).T
        np.testing.assert_almost_equal(Jn, J, decimal=6)


def test_stored_state_follows_the_last_configuration():
    """reference side effect (src/wrapper.cpp:85-89): the model is left at the last configuration"""
    pair, jids, lids = _tree_pair()
    q = _inputs(pair, jids, 2, 5, seed=8)
    pair.prod_solve(q, lids[:1], jids, 0, 2, False)
    np.testing.assert_array_equal(pair.prod.get_joint_angles(jids), q[-1, :len(jids)])
    np.testing.assert_allclose(pair.prod.get_base_pose(), q[-1, len(jids):], atol=1e-12)


def test_multi_chunk_host_path_and_pinned_buffers():
    """N larger than one pipeline chunk (cfg5-sized rows force small chunks) through pinned arrays"""
    pair = Pair("pr2.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg5")
    n = 20000  # > 8192-config chunk floor -> 3 chunks on 2 stream slots
    q = _inputs(pair, jids, base, n, seed=6)
    pd, nd = W.pose_dim(rot), q.shape[1]
    Pp = capi.PinnedArray((n * len(lids), pd))
    Jp = capi.PinnedArray((n * len(lids) * pd, nd))
    P, J = pair.prod_solve(q, lids, jids, rot, base, True, out=(Pp.array, Jp.array))
    sel = np.r_[0:64, 8128:8256, n - 64:n]
    Pr, Jr = pair.oracle_solve(q[sel], lids, jids, rot, base, True)
    L = len(lids)
    rows = (sel[:, None] * L + np.arange(L)[None, :]).reshape(-1)
    parity.assert_bitwise(P[rows], J[(rows[:, None] * pd + np.arange(pd)[None, :]).reshape(-1)], Pr, Jr, "chunks")
    Pp.free(), Jp.free()


def test_fp32_mode_within_1e5(capsys):
    """opt-in FP32 mode, north_star "1e-5 in FP32".  Tier 1 (asserted at 1e-5 abs, every entry): positions, quaternions
    and their Jacobian rows.  RPY angles and RPY Jacobian rows are functions of the quaternion whose condition number is
    1/cos(pitch) (angles) and 1/cos(pitch)^2 (rpy_derivative divides by cos(pitch), src/kinematics.cpp:18-21): an FP32
    quaternion that is correct to half an ulp already moves them by more than 1e-5 near the gimbal.  For those the test
    MEASURES and prints the max abs error and the fraction of entries above 1e-5, asserts 1e-5 on every entry whose
    |cos(pitch)| >= 0.5, asserts the fraction above 1e-5 stays small, and keeps the conditioning bound as tier 2."""
    report = []
    for cfg in ("cfg2", "cfg3", "cfg4", "cfg5"):
        pair = Pair(W.CONFIGS[cfg][0], strict=False)
        jids, lids, rot, base = pair.setup(cfg)
        n = 2048 if cfg != "cfg5" else 128
        q = _inputs(pair, jids, base, n, seed=7).astype(np.float32)
        Pr, Jr = pair.oracle_solve(q.astype(np.float64), lids, jids, rot, base, True)
        P, J = pair.prod_solve(q, lids, jids, rot, base, True, precision=1)
        assert P.dtype == np.float32 and J.dtype == np.float32
        pd = W.pose_dim(rot)
        aP = np.abs(P - Pr)
        aJ = np.abs(J - Jr).reshape(-1, pd, J.shape[1])
        assert aP[:, :3].max() <= parity.TOL_F32 and aJ[:, :3].max() <= parity.TOL_F32, (cfg, aP[:, :3].max(), aJ[:, :3].max())
        line = "%s: max|dP_xyz| %.2e  max|dJ_xyz| %.2e" % (cfg, aP[:, :3].max(), aJ[:, :3].max())
        if rot == W.ROT_XYZW:
            assert aP[:, 3:].max() <= parity.TOL_F32 and aJ[:, 3:].max() <= parity.TOL_F32
            line += "  max|dP_quat| %.2e  max|dJ_quat| %.2e" % (aP[:, 3:].max(), aJ[:, 3:].max())
        if rot == W.ROT_RPY:
            cosp = np.abs(np.cos(Pr[:, 4]))
            well = cosp >= 0.5
            fracP = float((aP[:, 3:] > parity.TOL_F32).mean())
            fracJ = float((aJ[:, 3:] > parity.TOL_F32).mean())
            line += ("  RPY angles: max %.2e (|cos|>=0.5: %.2e), %.3f%% of entries > 1e-5;  RPY rows of J: max %.2e "
                     "(|cos|>=0.5: %.2e), %.3f%% > 1e-5;  min|cos pitch| %.3f"
                     % (aP[:, 3:].max(), aP[well, 3:].max(), 100 * fracP, aJ[:, 3:].max(), aJ[well, 3:].max(), 100 * fracJ,
                        cosp.min()))
            # tier 1: well-conditioned rows meet 1e-5 outright (angles) / 4e-5 = 1e-5 / cos^2 at the edge of the set (rows)
            assert aP[well, 3:].max() <= parity.TOL_F32, line
            assert aJ[well, 3:].max() <= parity.TOL_F32 / 0.25, line
            assert fracP <= 0.01 and fracJ <= 0.05, line
            # tier 2: everywhere, the conditioning bound (relative above 1: |rows| grow like 1/cos too)
            c = np.maximum(cosp, 1e-3)
            eP = parity.rel_err(P, Pr)
            eJ = parity.rel_err(J, Jr).reshape(-1, pd, J.shape[1])
            assert (eP[:, 3:] <= parity.TOL_F32 / c[:, None]).all(), line
            assert (eJ[:, 3:] <= parity.TOL_F32 / (c * c)[:, None, None]).all(), line
        report.append(line)
    with capsys.disabled():
        print("\nFP32 mode vs the FP64 reference (measured):\n  " + "\n  ".join(report))


def test_empty_and_edge_shapes():
    pair, jids, lids = _tree_pair()
    P, J = pair.prod_solve(np.zeros((0, len(jids))), lids, jids, 1, 0, True)
    assert P.shape == (0, 6) and J.shape == (0, len(jids))
    # no joints selected, base only; and a single configuration
    q = _inputs(pair, [], 2, 1, seed=1)
    Pr, Jr = pair.oracle_solve(q, lids, [], 2, 2, True)
    P, J = pair.prod_solve(q, lids, [], 2, 2, True)
    parity.assert_bitwise(P, J, Pr, Jr, "base only")
    # the same joint named twice: the last column drives it, both columns are filled
    jj = [jids[1], jids[1], jids[2]]
    q = _inputs(pair, jj, 0, 7, seed=1)
    Pr, Jr = pair.oracle_solve(q, lids[:2], jj, 1, 0, True)
    P, J = pair.prod_solve(q, lids[:2], jj, 1, 0, True)
    parity.assert_bitwise(P, J, Pr, Jr, "duplicate joint")


def test_errors_are_value_errors():
    pair, jids, lids = _tree_pair()
    with pytest.raises(ValueError):
        pair.prod.solve_fk(np.zeros((2, len(jids))), [10 ** 6], jids, 0, 0, False)
    with pytest.raises(ValueError):
        pair.prod.solve_fk(np.zeros((2, len(jids) + 1)), lids, jids, 0, 0, False)
    with pytest.raises(ValueError):
        pair.prod.get_link_ids(["no_such_link"])
    with pytest.raises(ValueError):
        pair.prod.add_new_link("mylink", 0, [0, 0, 0], [0, 0, 0])


def test_device_pointer_entry_with_row_offset():
    """tfk_solve_fk_device on torch-owned device memory: two calls fill disjoint row ranges of one
    destination dimensioned for all rows (what the multi-stream / multi-GPU callers do)"""
    import torch

    pair = Pair("pr2.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg2")
    n = 1000
    q = _inputs(pair, jids, base, n, seed=9)
    dq = torch.from_numpy(q).cuda()
    pd, nd = W.pose_dim(rot), q.shape[1]
    dP = torch.full((pd, n), float("nan"), dtype=torch.float64, device="cuda")
    dJ = torch.full((nd, n * pd), float("nan"), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    cut = 384
    pair.prod.solve_fk_device(0, st, dq.data_ptr(), cut, lids, jids, rot, base, True, 0, dP.data_ptr(), dJ.data_ptr(), n, 0)
    pair.prod.solve_fk_device(0, st, dq.data_ptr() + cut * nd * 8, n - cut, lids, jids, rot, base, True, 0,
                              dP.data_ptr(), dJ.data_ptr(), n, cut)
    torch.cuda.synchronize()
    Pr, Jr = pair.oracle_solve(q, lids, jids, rot, base, True)
    parity.assert_bitwise(dP.cpu().numpy().T, dJ.cpu().numpy().T, Pr, Jr, "device entry")
    assert capi.kernel_launch_count() >= 2


def test_full_size_cfg2_properties():
    """BASELINE size (1M PR2-arm configurations): spot rows equal the oracle; every entry finite;
    the translation columns are the identity; a second run is bit-identical (determinism)"""
    pair = Pair("pr2.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg2")
    n = 1 << 20
    q = _inputs(pair, jids, base, n, seed=0)
    P, J = pair.prod_solve(q, lids, jids, rot, base, True)
    assert np.isfinite(P).all() and np.isfinite(J).all()
    nj = len(jids)
    Jv = J.reshape(n, 6, nj + 3)
    assert (Jv[:, 0, nj] == 1).all() and (Jv[:, 1, nj + 1] == 1).all() and (Jv[:, 2, nj:nj + 2] == 0).all()
    sel = np.r_[0:256, n // 2:n // 2 + 256, n - 256:n]
    Pr, Jr = pair.oracle_solve(q[sel], lids, jids, rot, base, True)
    parity.assert_bitwise(P[sel], Jv[sel].reshape(-1, nj + 3), Pr, Jr, "1M spot rows")
    P2, J2 = pair.prod_solve(q, lids, jids, rot, base, True)
    assert np.array_equal(P, P2) and np.array_equal(J, J2)


@pytest.mark.parametrize("rot,precision", [(0, 0), (1, 0), (2, 1)])
def test_half_tiles_when_the_last_wave_is_short(rot, precision):
    """several links, ONE launch (device entry) of a little more than one wave of 128-configuration tiles (148 SMs x
    3-4 blocks): the engine runs 64-configuration tiles (fk_launch.h:twophase_tile_cfgs); ragged tail of 37
    configurations.  Rows of the first tile, of tiles of the second round and of the tail equal the oracle"""
    import torch

    pair = Pair("fetch.urdf", strict=True)
    jids, lids, _, _ = pair.setup("cfg4")
    lids = lids[:3]
    base, nl = 2, 3
    n = 148 * 4 * 128 + 9 * 128 + 37
    q = _inputs(pair, jids, base, n, seed=5)
    dt, npdt = (torch.float64, np.float64) if precision == 0 else (torch.float32, np.float32)
    dq = torch.from_numpy(q.astype(npdt)).cuda()
    pd, nd = W.pose_dim(rot), q.shape[1]
    dP = torch.full((pd, n * nl), float("nan"), dtype=dt, device="cuda")
    dJ = torch.full((nd, n * nl * pd), float("nan"), dtype=dt, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    pair.prod.solve_fk_device(0, st, dq.data_ptr(), n, lids, jids, rot, base, True, precision, dP.data_ptr(), dJ.data_ptr(), n, 0)
    torch.cuda.synchronize()
    assert "twophase" in pair.prod.last_launch_info()["kernel"]
    P, J = dP.cpu().numpy().T.astype(np.float64), dJ.cpu().numpy().T.astype(np.float64)
    assert np.isfinite(P).all() and np.isfinite(J).all()
    sel = np.r_[0:200, 148 * 4 * 64 - 100:148 * 4 * 64 + 100, 148 * 3 * 64 - 50:148 * 3 * 64 + 50, n - 300:n]
    Pr, Jr = pair.oracle_solve(q[sel], lids, jids, rot, base, True)
    Ps = P.reshape(n, nl, pd)[sel].reshape(-1, pd)
    Js = J.reshape(n, nl * pd, nd)[sel].reshape(-1, nd)
    if precision == 0:
        parity.assert_bitwise(Ps, Js, Pr, Jr, "half tiles rot %d" % rot)
    else:
        assert np.abs(Ps - Pr).max() <= parity.TOL_F32 and np.abs(Js - Jr).max() <= parity.TOL_F32
