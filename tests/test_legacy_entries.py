"""The reference's legacy pybind11 entries, called the way real tinyfk users call them:

  _tinyfk.RobotModel.solve_forward_kinematics(list-of-lists | ndarray, elink_ids, joint_ids, rot_type, with_base,
                                              with_jacobian, use_cache)           reference src/wrapper.cpp:60-124
  _tinyfk.RobotModel.compute_inter_link_squared_dists(qs, link_ids1, link_ids2, joint_ids, with_base, with_grads,
                                                      use_cache)                  reference src/wrapper.cpp:189-237
  _tinyfk.RobotModel.solve_com_forward_kinematics(qs, joint_ids, action_link_ids, forces, with_base, with_jacobian)
                                                                                  reference src/wrapper.cpp:126-179
  _tinyfk.RobotModel.compute_total_inertia_matrix(q, joint_ids)                   reference src/wrapper.cpp:181-187

as python/tinyfk/__init__.py:145-153, 189-195, 197-198, 257-266 does.  The oracle is the reference's own object code
(oracle/_ref, libm-pinned) behind the same loop; FP64 strict => bit for bit.
"""
import numpy as np
import pytest

import parity
from oracle import refapi
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")


def _robot_and_ref(urdf="pr2.urdf"):
    from tinyfk_b200 import _tinyfk

    xml = W.urdf_text(urdf)
    return _tinyfk.RobotModel(xml), refapi.RefModel(xml, strict=True)


def _rot_enum(rot):
    from tinyfk_b200 import _tinyfk

    return [_tinyfk.RotationType.IGNORE, _tinyfk.RotationType.RPY, _tinyfk.RotationType.XYZW][rot]


@needs_ref
@pytest.mark.parametrize("rot", [0, 1, 2])
@pytest.mark.parametrize("with_base", [False, True])
def test_seven_argument_solve_forward_kinematics(rot, with_base):
    """list-of-lists input (what KinematicModel.solve_fk hands the reference), with_base => 6-D floating base,
    use_cache True == use_cache False, ndarray input == list input"""
    robot, ref = _robot_and_ref()
    jids = robot.get_joint_ids(W.PR2_RARM)
    lids = robot.get_link_ids(["r_gripper_tool_frame", "r_upper_arm_link", "head_pan_link"])
    base = W.BASE_FLOATING if with_base else W.BASE_FIXED
    q = W.sample_q(ref.get_joint_limits(jids), base, 33, seed=11)
    as_lists = [list(map(float, row)) for row in q]
    P, J = robot.solve_forward_kinematics(as_lists, lids, jids, _rot_enum(rot), with_base, True, False)
    Pr, Jr = ref.solve_fk(q, lids, jids, rot, base, True)
    pd = W.pose_dim(rot)
    assert P.shape == (33 * 3, pd) and J.shape == (33 * 3 * pd, q.shape[1])
    assert P.flags.f_contiguous and J.flags.f_contiguous  # P.transpose() of a column-major matrix, wrapper.cpp:123
    parity.assert_bitwise(P, J, Pr, Jr, "legacy solve_forward_kinematics")
    P2, J2 = robot.solve_forward_kinematics(as_lists, lids, jids, _rot_enum(rot), with_base, True, True)  # use_cache
    assert np.array_equal(P, P2) and np.array_equal(J, J2)
    P3, J3 = robot.solve_forward_kinematics(q, lids, jids, _rot_enum(rot), with_base, True, False)  # ndarray
    assert np.array_equal(P, P3) and np.array_equal(J, J3)
    # without the Jacobian the reference returns zeros of the full shape (src/wrapper.cpp:81-82)
    P4, J4 = robot.solve_forward_kinematics(as_lists, lids, jids, _rot_enum(rot), with_base, False, False)
    assert np.array_equal(P, P4) and J4.shape == J.shape and not J4.any()
    # the model is left at the last configuration (src/wrapper.cpp:85-89)
    assert np.array_equal(robot.get_joint_angles(jids), q[-1, :len(jids)])
    if with_base:
        assert np.allclose(robot.get_base_pose(), q[-1, len(jids):], rtol=0, atol=1e-12)


def test_seven_argument_entry_dof_mismatch_is_a_value_error():
    """src/wrapper.cpp:74-78: with_base and rows that are not n_joints + 6 long -> std::invalid_argument -> ValueError"""
    robot, _ = _robot_and_ref()
    jids = robot.get_joint_ids(W.PR2_RARM)
    lids = robot.get_link_ids(["r_gripper_tool_frame"])
    with pytest.raises(ValueError, match="dof mismatch"):
        robot.solve_forward_kinematics([[0.0] * (len(jids) + 3)], lids, jids, _rot_enum(0), True, False, False)
    with pytest.raises(ValueError):  # fixed base, wrong row length: the reference reads out of bounds; here an error
        robot.solve_forward_kinematics([[0.0] * (len(jids) + 1)], lids, jids, _rot_enum(0), False, False, False)
    with pytest.raises(ValueError):  # ragged rows cannot form a 2-D array
        robot.solve_forward_kinematics([[0.0] * 7, [0.0] * 6], lids, jids, _rot_enum(0), False, False, False)


@needs_ref
@pytest.mark.parametrize("with_base", [False, True])
def test_seven_argument_compute_inter_link_squared_dists(with_base):
    robot, ref = _robot_and_ref()
    jids = robot.get_joint_ids(W.PR2_RARM + W.PR2_LARM)
    names1 = ["r_gripper_tool_frame", "r_forearm_link", "r_gripper_tool_frame"]
    names2 = ["l_gripper_tool_frame", "l_forearm_link", "head_pan_link"]
    l1, l2 = robot.get_link_ids(names1), robot.get_link_ids(names2)
    base = W.BASE_FLOATING if with_base else W.BASE_FIXED
    q = W.sample_q(ref.get_joint_limits(jids), base, 21, seed=12)
    as_lists = [list(map(float, row)) for row in q]
    sq, G = robot.compute_inter_link_squared_dists(as_lists, l1, l2, jids, with_base, True, False)
    sr, Gr = ref.inter_link_sqdists(q, list(zip(l1, l2)), jids, base, True)
    assert sq.shape == (21 * 3,) and G.shape == (21 * 3, q.shape[1])
    assert np.array_equal(sq, sr) and np.array_equal(G, Gr)
    sq2, G2 = robot.compute_inter_link_squared_dists(as_lists, l1, l2, jids, with_base, True, True)  # use_cache
    assert np.array_equal(sq, sq2) and np.array_equal(G, G2)
    sq3, G3 = robot.compute_inter_link_squared_dists(as_lists, l1, l2, jids, with_base, False, False)
    assert np.array_equal(sq, sq3) and G3.shape == (1, 1) and G3[0, 0] == 0  # src/wrapper.cpp:201-203
    with pytest.raises(ValueError):
        robot.compute_inter_link_squared_dists([[0.0] * 3], l1, l2, jids, with_base, True, False)


@needs_ref
def test_legacy_com_and_inertia_entries():
    robot, ref = _robot_and_ref("fetch.urdf")
    jids = robot.get_joint_ids(W.FETCH_JOINTS)
    q = W.sample_q(ref.get_joint_limits(jids), W.BASE_FLOATING, 9, seed=13)
    com, J = robot.solve_com_forward_kinematics([list(map(float, r)) for r in q], jids, [], [], True, True)
    cr, Jr = ref.solve_com_fk(q, jids, [], [], W.BASE_FLOATING, True)
    assert com.shape == (9, 3) and J.shape == (27, len(jids) + 6)
    assert np.array_equal(com, cr) and np.array_equal(J, Jr)
    m = robot.compute_total_inertia_matrix(list(map(float, q[0, :len(jids)])), jids)
    mr = ref.total_inertia_matrix(q[:1, :len(jids)], jids)[0]
    assert m.shape == (3, 3) and np.array_equal(m, mr)
    with pytest.raises(ValueError):
        robot.compute_total_inertia_matrix([0.0] * (len(jids) - 1), jids)
