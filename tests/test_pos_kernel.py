"""fk_pos_kernel (tinyfk_b200/csrc/fk_kernels.cuh): several links with RotationType IGNORE -- collision-sphere
centres and their 3 x n position Jacobians, the shape of BASELINE cfg4 -- are produced one output entry per
thread (FkCore::pos_entry).  Compared with the oracle bit for bit: every BaseType, prismatic and irrelevant
joints, more Jacobian columns than one relevance-mask word, both record placements, ragged batch sizes."""
import numpy as np
import pytest

import emuapi
import parity
from emuapi import EmuModel
from oracle import refapi
from parity import Pair
from tinyfk_b200 import workloads as W



@pytest.fixture(autouse=True)
def _select_pos_kernel(monkeypatch):
    """the engine reads TFK_POS_KERNEL at every launch (engine.cpp: opt-in while the kernel is being tuned)"""
    monkeypatch.setenv("TFK_POS_KERNEL", "1")


needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")

PR2_LINKS = ["r_gripper_tool_frame", "l_gripper_tool_frame", "head_tilt_link", "base_footprint", "r_elbow_flex_link",
             "torso_lift_link", "l_forearm_link", "r_gripper_tool_frame", "fl_caster_rotation_link", "r_forearm_link",
             "l_gripper_l_finger_tip_link", "r_upper_arm_link_com", "laser_tilt_mount_link"]


@needs_ref
@pytest.mark.parametrize("base", [0, 1, 2])
def test_entry_per_thread_core_on_cpu_all_45_pr2_joints(base):
    """45 Jacobian columns (two mask words) through tests/emu, which runs FkCore::pos_entry / emit_base_rot_pos"""
    xml = W.urdf_text("pr2.urdf")
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=True)
    jids = list(range(e.n_joints))
    lids = e.get_link_ids(PR2_LINKS)
    q = W.sample_q([(-1.0, 1.0)] * len(jids), base, 6, seed=base, wide_pitch=True)
    Pr, Jr = o.solve_fk(q, lids, jids, 0, base, True)
    P, J = e.solve_fk(q, lids, jids, 0, base, True, mode=emuapi.MODE_STRICT64)
    parity.assert_bitwise(P, J, Pr, Jr, "pr2 all joints, base %d" % base)


@pytest.mark.gpu
@pytest.mark.parametrize("base", [0, 1, 2])
@pytest.mark.parametrize("jac", [False, True])
def test_gpu_fetch_spheres_bitwise(base, jac):
    """cfg4's request (8 joints incl. the prismatic torso, 30 sphere links) with every BaseType; 1000 is not a
    multiple of the 128-configuration tile"""
    pair = Pair("fetch.urdf", strict=True)
    jids, lids, rot, _ = pair.setup("cfg4")
    q = W.sample_q(pair.prod.get_joint_limits(jids), base, 1000, seed=10 + base, wide_pitch=True)
    Pr, Jr = pair.oracle_solve(q, lids, jids, 0, base, jac)
    P, J = pair.prod_solve(q, lids, jids, 0, base, jac)
    k = pair.prod.last_launch_info()["kernel"]
    assert k.startswith("fk_pos_kernel"), k
    parity.assert_bitwise(P, J if jac else P, Pr, Jr if jac else Pr, "fetch spheres base %d" % base)
    if not jac:
        assert not J.any()


@pytest.mark.gpu
@pytest.mark.parametrize("base", [0, 2])
def test_gpu_pr2_all_joints_bitwise(base):
    """all 45 PR2 joints (two relevance-mask words; records too large for shared memory -> scratch placement),
    links on both arms, head, casters, a duplicate and the root link"""
    pair = Pair("pr2.urdf", strict=True)
    jids = list(range(pair.prod.n_joints))
    lids = pair.prod.get_link_ids(PR2_LINKS)
    lim = [(max(lo, -1.0), min(hi, 1.0)) if np.isfinite(lo) else (-1.0, 1.0) for lo, hi in pair.prod.get_joint_limits(jids)]
    q = W.sample_q(lim, base, 517, seed=20 + base, wide_pitch=True)
    Pr, Jr = pair.oracle_solve(q, lids, jids, 0, base, True)
    P, J = pair.prod_solve(q, lids, jids, 0, base, True)
    assert pair.prod.last_launch_info()["kernel"] == "fk_pos_kernel"  # column-ordered joint records in the scratch
    parity.assert_bitwise(P, J, Pr, Jr, "pr2 all joints base %d" % base)


@pytest.mark.gpu
def test_gpu_two_links_one_config_and_fp32():
    pair = Pair("fetch.urdf", strict=True)
    jids, lids, _, _ = pair.setup("cfg4")
    q = W.sample_q(pair.prod.get_joint_limits(jids), 1, 1, seed=3)
    Pr, Jr = pair.oracle_solve(q, lids[:2], jids, 0, 1, True)
    P, J = pair.prod_solve(q, lids[:2], jids, 0, 1, True)
    parity.assert_bitwise(P, J, Pr, Jr, "two links, one configuration")
    q = W.sample_q(pair.prod.get_joint_limits(jids), 2, 700, seed=4)
    Pr, Jr = pair.oracle_solve(q, lids, jids, 0, 2, True)
    P, J = pair.prod_solve(q, lids, jids, 0, 2, True, precision=1)
    assert pair.prod.last_launch_info()["kernel"].startswith("fk_pos_kernel")
    assert np.abs(P - Pr).max() <= parity.TOL_F32 and np.abs(J - Jr).max() <= parity.TOL_F32
