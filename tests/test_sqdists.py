"""compute_inter_link_sqdists (reference src/wrapper.cpp:189-237,
python/tinyfk/__init__.py:233-266).

Oracle = oracle/_ref: the reference's own get_link_pose / get_jacobian, with the three Eigen expressions of the
entry (diff.squaredNorm(), 2*(J1-J2)^T*diff) restated in oracle/ref_driver.cpp (Eigen is not installed, so Eigen's
internal summation order is unpinned: the contract for this entry is 1e-12).  With the libm-pinned
reference the product evaluates the same restated order, so the comparison below is exact anyway.
CPU tests run the kernels' per-thread code through tests/emu; GPU tests go through the C ABI."""
import numpy as np
import pytest

import emuapi
from emuapi import EmuModel
from oracle import refapi
from tinyfk_b200 import workloads as W

PAIR_LINKS = [("r_gripper_tool_frame", "l_gripper_tool_frame"), ("r_elbow_flex_link", "l_elbow_flex_link"),
              ("r_gripper_tool_frame", "head_tilt_link"), ("l_forearm_link", "base_link"),
              ("r_wrist_roll_link", "r_shoulder_pan_link"), ("torso_lift_link", "torso_lift_link")]
JOINTS = ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM

needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")


def _setup(m):
    jids = m.get_joint_ids(JOINTS)
    pairs = [tuple(m.get_link_ids(list(p))) for p in PAIR_LINKS]
    return list(jids), pairs


def _q(m, jids, base, n, seed):
    return W.sample_q(m.get_joint_limits(jids), base, n, seed=seed, wide_pitch=True)


@needs_ref
def test_oracle_sqdists_agree_with_its_own_fk_and_jacobians():
    o = refapi.RefModel(W.urdf_text("pr2.urdf"), strict=True)
    jids, pairs = _setup(o)
    q = _q(o, jids, 2, 7, 1)
    sq, G = o.inter_link_sqdists(q, pairs, jids, 2, True)
    lids = [p[0] for p in pairs] + [p[1] for p in pairs]
    P, J = o.solve_fk(q, lids, jids, 0, 2, True)
    n, k, L = q.shape[0], len(pairs), len(lids)
    P = P.reshape(n, L, 3)
    J = J.reshape(n, L, 3, -1)
    d = P[:, :k] - P[:, k:]
    assert np.allclose(sq.reshape(n, k), (d ** 2).sum(-1), rtol=0, atol=1e-14)
    g = 2 * np.einsum("nkrc,nkr->nkc", J[:, :k] - J[:, k:], d)
    assert np.allclose(G.reshape(n, k, -1), g, rtol=0, atol=1e-7)  # FD columns: 1e-9-level noise times 2*|d|


@needs_ref
@pytest.mark.parametrize("base", [0, 1, 2])
@pytest.mark.parametrize("jac", [False, True])
def test_pair_core_equals_pinned_reference_on_cpu(base, jac):
    xml = W.urdf_text("pr2.urdf")
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=True)
    jids, pairs = _setup(e)
    q = _q(o, jids, base, 16, 3 + base)
    sr, Gr = o.inter_link_sqdists(q, pairs, jids, base, jac)
    s, G = e.inter_link_sqdists(q, pairs, jids, base, jac)
    assert np.array_equal(s, sr)
    if jac:
        assert np.array_equal(G, Gr)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("base", [0, 1, 2])
def test_gpu_sqdists_bitwise_vs_pinned_reference(base):
    from tinyfk_b200 import capi
    xml = W.urdf_text("pr2.urdf")
    m, o = capi.Model(xml, host_libm=0), refapi.RefModel(xml, strict=True)
    jids, pairs = _setup(m)
    q = _q(o, jids, base, 1000, 11 + base)  # 1000: ragged last tile
    sr, Gr = o.inter_link_sqdists(q, pairs, jids, base, True)
    s, G = m.inter_link_sqdists(q, pairs, jids, base, True)
    assert s.shape == sr.shape and G.shape == Gr.shape and G.flags.f_contiguous
    assert np.array_equal(s, sr) and np.array_equal(G, Gr)
    s2, G2 = m.inter_link_sqdists(q, pairs, jids, base, False)
    assert np.array_equal(s2, sr) and G2.shape == (1, 1)
    # stock-glibc reference: 1e-12 on distances and analytic columns, FD noise floor on the base-rotation columns
    os_ = refapi.RefModel(xml, strict=False)
    ss, Gs = os_.inter_link_sqdists(q, pairs, jids, base, True)
    assert np.abs(s - ss).max() <= 1e-12
    nj = len(jids)
    assert np.abs(G[:, :nj] - Gs[:, :nj]).max() <= 1e-12
    assert np.abs(G - Gs).max() <= 1e-6


@pytest.mark.gpu
def test_gpu_sqdists_facade_and_numerical_gradient():
    """the reference's own test (python/tests/test_tinyfk.py:157-200): analytic gradient vs forward differences"""
    import tinyfk_b200
    from tinyfk_b200 import BaseType, KinematicModel
    kin = KinematicModel(tinyfk_b200.pr2_urdfpath())
    jids = kin.get_joint_ids(W.PR2_RARM + W.PR2_LARM)
    l1 = kin.get_link_ids(["r_shoulder_pan_link", "r_elbow_flex_link", "r_wrist_roll_link", "r_gripper_tool_frame"])
    l2 = kin.get_link_ids(["l_shoulder_pan_link", "l_elbow_flex_link", "l_wrist_roll_link", "l_gripper_tool_frame"])
    pairs = list(zip(l1, l2))
    rng = np.random.default_rng(0)
    for base in (BaseType.FIXED, BaseType.PLANER, BaseType.FLOATING):
        nb = {BaseType.FIXED: 0, BaseType.PLANER: 3, BaseType.FLOATING: 6}[base]
        q = rng.uniform(-0.5, 0.5, len(jids) + nb)
        v, Ja = kin.compute_inter_link_sqdists([q], pairs, jids, base, with_jacobian=True)
        assert v.shape == (4,) and Ja.shape == (4, len(q))
        eps = 1e-7
        Q = np.tile(q, (len(q), 1)) + eps * np.eye(len(q))
        v1, _ = kin.compute_inter_link_sqdists(Q, pairs, jids, base, with_jacobian=False)
        Jn = ((v1.reshape(len(q), 4) - v) / eps).T
        np.testing.assert_almost_equal(Jn, Ja, decimal=5)
    # FP32 mode within 1e-5
    q = rng.uniform(-0.5, 0.5, (64, len(jids) + 6))
    v64, J64 = kin.compute_inter_link_sqdists(q, pairs, jids, BaseType.FLOATING, with_jacobian=True)
    v32, J32 = kin.compute_inter_link_sqdists(q, pairs, jids, BaseType.FLOATING, with_jacobian=True, precision="fp32")
    assert v32.dtype == np.float32 and np.abs(v32 - v64).max() < 1e-5 and np.abs(J32 - J64).max() < 2e-5
