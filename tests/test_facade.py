"""Python facade (tinyfk_b200/__init__.py) mirrors tinyfk's: names, enums, shapes, pickling by replay
(reference python/tests/test_pickle.py, python/tests/test_tinyfk.py:43-58).  No GPU needed here."""
import copy
import pickle

import numpy as np
import pytest

import tinyfk_b200 as tinyfk
from tinyfk_b200 import BaseType, KinematicModel, RotationType
from tinyfk_b200 import workloads as W


def test_surface_and_enums():
    assert [e.name for e in BaseType] == ["FIXED", "PLANER", "FLOATING"] and [e.value for e in BaseType] == [0, 1, 2]
    assert [e.name for e in RotationType] == ["IGNORE", "RPY", "XYZW"]
    kin = KinematicModel(tinyfk.pr2_urdfpath())
    assert kin.root_link_name == "base_footprint" and abs(kin.total_mass - 266.732178) < 1e-6
    assert kin.get_joint_ids(W.PR2_RARM) == [39, 38, 40, 29, 30, 41, 42]
    assert kin.get_link_ids(["r_gripper_tool_frame"]) == [71]
    lim = kin.get_joint_limits(kin.get_joint_ids(W.PR2_RARM))
    assert lim[-1] == (None, None) and abs(lim[0][0] + 2.2853981634) < 1e-10
    with pytest.raises(ValueError):
        kin.get_link_ids(["nope"])
    with pytest.raises(AssertionError):  # reference python/tinyfk/__init__.py:133-142
        kin.solve_fk(np.zeros((2, 7)), 71, kin.get_joint_ids(W.PR2_RARM), base_type=BaseType.PLANER)
    with pytest.raises(RuntimeError):  # no CUDA device here: every compute entry fails loudly, none falls back to the CPU
        kin.compute_total_inertia_matrix(np.zeros(7), kin.get_joint_ids(W.PR2_RARM))


def test_pickle_and_deepcopy_replay():
    kin = KinematicModel(tinyfk.pr2_urdfpath())
    pickle.loads(pickle.dumps(kin))
    copy.copy(kin)
    pid = kin.get_link_ids(["r_upper_arm_link"])[0]
    kin.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])
    new_id = kin.get_link_ids(["mylink"])[0]
    jids = kin.get_joint_ids(W.PR2_RARM)
    q = np.array(W.README_Q)
    kin.set_q(jids, q, BaseType.PLANER)
    kin2 = copy.deepcopy(kin)
    assert kin2.get_link_ids(["mylink"])[0] == new_id
    np.testing.assert_almost_equal(kin2.get_q(jids, BaseType.PLANER), q)
    kin3 = copy.deepcopy(kin2)  # a copy of a copy still carries the history
    assert kin3.get_link_ids(["mylink"])[0] == new_id


@pytest.mark.gpu
def test_readme_example_and_trajectory_contract():
    """reference README.md:19-51 and python/tests/test_tinyfk.py:133-154"""
    kin = KinematicModel(tinyfk.pr2_urdfpath())
    jids = kin.get_joint_ids(W.PR2_RARM)
    lid = kin.get_link_ids(["r_gripper_tool_frame"])[0]
    P, J = kin.solve_fk(np.array(W.README_Q), lid, jids, rot_type=RotationType.RPY, base_type=BaseType.PLANER,
                        with_jacobian=True)
    assert P.shape == (1, 6) and J.shape == (6, 10) and P.flags.f_contiguous and J.flags.f_contiguous
    np.testing.assert_allclose(P[0], [0.50812707267784318, 0.84860531692602614, 0.75061268052719599,
                                      -2.0642683981753733, -0.21174995675381791, 1.493189858833309], atol=1e-12)
    links = kin.get_link_ids(["r_shoulder_pan_link", "r_elbow_flex_link", "base_link"])
    qs = [np.array(W.README_Q[:7] + [0.1, 0.2, 0.3, 0.1, 0.2, 0.3]) + np.random.randn(13) * 0.1 for _ in range(10)]
    P, J = kin.solve_fk(qs, links, jids, RotationType.RPY, BaseType.FLOATING, True)
    assert P.shape == (30, 6) and J.shape == (180, 13)
    Ps, Js = P.reshape(10, 3, 6), J.reshape(10, 18, 13)
    for i, q in enumerate(qs):
        P1, J1 = kin.solve_fk([q], links, jids, RotationType.RPY, BaseType.FLOATING, True)
        np.testing.assert_array_equal(Ps[i], P1)
        np.testing.assert_array_equal(Js[i], J1)
    P0, J0 = kin.solve_fk(qs, links, jids, RotationType.RPY, BaseType.FLOATING, False)
    np.testing.assert_array_equal(P0, P)
    assert J0.shape == J.shape and not J0.any()
