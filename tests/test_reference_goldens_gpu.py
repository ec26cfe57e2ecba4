"""The reference's OWN golden vectors applied to the CUDA path directly (not via the oracle):

  * test/data/test_data.json (copied to tests/golden/reference_test_data.json): 10 PR2 link poses incl. the added
    "mylink" for one angle vector with a floating base -- checked by the reference in test/test_kinematics.cpp:94-173
    (isNear 1e-5... on x y z yaw pitch roll) and python/tests/test_tinyfk.py:61-81 (decimal=5);
  * the pinocchio constants of test/test_kinematics.cpp:198-222: total inertia matrix of the KUKA at q = 0, 1e-5;
  * the analytic-vs-numerical Jacobian check of test/test_kinematics.cpp:160-173 (1e-5) with a floating base.
"""
import json
import os

import numpy as np
import pytest

import tinyfk_b200
from tinyfk_b200 import BaseType, KinematicModel, RotationType
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _testdata():
    with open(os.path.join(GOLD, "reference_test_data.json")) as f:
        td = json.load(f)
    gt = np.array(td["pose_list"])[:, [0, 1, 2, 5, 4, 3]]  # stored as x y z yaw pitch roll (test_kinematics.cpp:150-157)
    return td, gt


def _pr2_with_mylink():
    kin = KinematicModel(tinyfk_b200.pr2_urdfpath())
    pid = kin.get_link_ids(["r_upper_arm_link"])[0]
    kin.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])  # test_kinematics.cpp:113-120
    return kin


@pytest.mark.parametrize("precision,tol", [("fp64", 1e-7), ("fp32", 1e-5)])
def test_solve_fk_reproduces_the_reference_golden_poses(precision, tol):
    td, gt = _testdata()
    kin = _pr2_with_mylink()
    lids, jids = kin.get_link_ids(td["link_names"]), kin.get_joint_ids(td["joint_names"])
    q = np.array(td["angle_vector"])
    P, J = kin.solve_fk(q, lids, jids, rot_type=RotationType.RPY, base_type=BaseType.FLOATING, with_jacobian=True,
                        precision=precision)
    assert P.shape == gt.shape
    err = np.abs(P - gt).max()
    assert err < tol, err
    if precision == "fp64":
        assert err < 1e-12  # what the strict FP64 path actually achieves against the stored decimals
    # every link on its own (fused single-link kernel) returns the row of the batch call (two-phase kernel)
    for k, lid in enumerate(lids):
        Pk, Jk = kin.solve_fk(q, lid, jids, rot_type=RotationType.RPY, base_type=BaseType.FLOATING, with_jacobian=True,
                              precision=precision)
        if precision == "fp64":
            assert np.array_equal(Pk[0], P[k]) and np.array_equal(Jk, J[6 * k:6 * k + 6])
        else:
            assert np.abs(Pk[0] - P[k]).max() < tol


def test_duplicate_link_name_raises_like_the_reference():
    kin = _pr2_with_mylink()
    pid = kin.get_link_ids(["r_upper_arm_link"])[0]
    with pytest.raises(Exception):  # test_kinematics.cpp:122-129
        kin.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])
    # relevance table follows the added link (test_kinematics.cpp:131-142)
    j = kin.get_joint_ids(["torso_lift_joint", "r_wrist_flex_joint"])
    l = kin.get_link_ids(["mylink", "head_pan_link", "fl_caster_rotation_link"])
    r = kin._robot.is_relevant
    assert r(l[0], j[0]) and not r(l[0], j[1]) and r(l[1], j[0]) and not r(l[2], j[0])


@pytest.mark.parametrize("rot", [RotationType.IGNORE, RotationType.RPY, RotationType.XYZW])
def test_analytic_jacobian_matches_numerical_jacobian_with_base(rot):
    """test/test_kinematics.cpp:160-173: central... forward differences (eps 1e-7) of solve_fk itself vs its Jacobian"""
    td, _ = _testdata()
    kin = _pr2_with_mylink()
    lid = kin.get_link_ids(td["link_names"])[-1]
    jids = kin.get_joint_ids(td["joint_names"])
    q0 = np.array(td["angle_vector"])
    eps = 1e-7
    Q = np.vstack([q0] + [q0 + eps * np.eye(q0.size)[k] for k in range(q0.size)])
    P, J = kin.solve_fk(Q, lid, jids, rot_type=rot, base_type=BaseType.FLOATING, with_jacobian=True)
    pd = P.shape[1]
    J_num = ((P[1:] - P[0]) / eps).T
    assert np.abs(J[:pd] - J_num).max() < 1e-5


def test_total_inertia_matrix_matches_the_pinocchio_constants():
    kin = KinematicModel(tinyfk_b200.kuka_urdfpath())
    imat_ref = np.array([[2.11765803e+00, 4.59516343e-05, -2.15483177e-04],
                         [4.59516343e-05, 2.08613831e+00, 2.17846493e-02],
                         [-2.15483177e-04, 2.17846493e-02, 1.14820264e-01]])  # test_kinematics.cpp:216-220
    jids = kin.get_joint_ids(W.KUKA_JOINTS)
    imat = kin.compute_total_inertia_matrix(np.zeros(len(jids)), jids)
    assert np.abs(imat - imat_ref).max() < 1e-5
    imat32 = kin.compute_total_inertia_matrix(np.zeros(len(jids)), jids, precision="fp32")
    assert np.abs(imat32 - imat_ref).max() < 1e-5
    # no joints named: the model's stored (zero) angles apply, as kin.get_total_inertia_matrix() in the C++ test
    imat0 = kin.compute_total_inertia_matrix(np.zeros(0), [])
    assert np.abs(imat0 - imat_ref).max() < 1e-5
