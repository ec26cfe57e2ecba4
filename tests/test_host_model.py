"""Host-side mirror of tinyfk::KinematicModel (tinyfk_b200/csrc/model.cpp) through the C ABI, no GPU:
id tables, link table bits, relevance, limits, state and error behaviour against the reference."""
import numpy as np
import pytest

from oracle import refapi
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

needs_ref = pytest.mark.skipif(not (refapi.have_ref(False) and refapi.have_ref(True)), reason="oracle/_ref not built")


@pytest.mark.parametrize("urdf,n_links,n_joints,root", [("pr2.urdf", 162, 45, "base_footprint"),
                                                          ("kuka.urdf", 16, 7, "lbr_iiwa_link_0"),
                                                          ("fetch.urdf", 45, 14, "base_link")])
def test_id_rule(urdf, n_links, n_joints, root):
    """links = sorted URDF names, then <name>_com in id order; joints = sorted movable names"""
    m = capi.Model(W.urdf_text(urdf))
    names = m.get_link_names()
    assert (m.n_links, m.n_joints, names[m.root_link_id]) == (n_links, n_joints, root)
    urdf_names = [n for n in names if not n.endswith("_com")]
    assert urdf_names == sorted(urdf_names) and names[:len(urdf_names)] == urdf_names
    com = names[len(urdf_names):]
    assert [c[:-4] for c in com] == sorted(c[:-4] for c in com) and all(c.endswith("_com") for c in com)
    assert m.get_joint_names() == sorted(m.get_joint_names())
    assert m.get_link_ids(names[::7]) == list(range(0, n_links, 7))


@needs_ref
@pytest.mark.parametrize("strict", [True, False])
@pytest.mark.parametrize("urdf", ["pr2.urdf", "kuka.urdf", "fetch.urdf"])
def test_link_table_bitwise_vs_reference(urdf, strict):
    r = refapi.RefModel(W.urdf_text(urdf), strict=strict)
    m = capi.Model(W.urdf_text(urdf), host_libm=0 if strict else 1)
    assert r.get_link_names() == m.get_link_names() and r.get_joint_names() == m.get_joint_names()
    assert r.root_link_id == m.root_link_id and r.total_mass == m.total_mass
    for i in range(r.n_links):
        a, b = r.link_info(i), m.link_info(i)
        assert a[0] == b[0] and a[2] == b[2] and (a[0] < 0 or a[1] == b[1])
        assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
    assert all(r.is_relevant(l, j) == m.is_relevant(l, j) for l in range(r.n_links) for j in range(r.n_joints))
    assert r.get_joint_limits(range(r.n_joints)) == m.get_joint_limits(range(m.n_joints))


def test_limits_and_relevance_pr2():
    """reference test/test_parse.cpp:6-40, python/tests/test_tinyfk.py:49-58, test/test_kinematics.cpp:131-142"""
    m = capi.Model(W.urdf_text("pr2.urdf"))
    pid = m.get_link_ids(["r_upper_arm_link"])[0]
    assert m.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1]) == 162
    lo, hi = m.get_joint_limits(m.get_joint_ids(["r_shoulder_pan_joint"]))[0]
    assert abs(lo + 2.2853981634) < 1e-10 and abs(hi - 0.714601836603) < 1e-10
    lo, hi = m.get_joint_limits(m.get_joint_ids(["r_wrist_roll_joint"]))[0]
    assert lo == -np.inf and hi == np.inf
    j = m.get_joint_ids(["torso_lift_joint", "r_wrist_flex_joint"])
    l = m.get_link_ids(["mylink", "head_pan_link", "fl_caster_rotation_link"])
    assert m.is_relevant(l[0], j[0]) and not m.is_relevant(l[0], j[1])
    assert m.is_relevant(l[1], j[0]) and not m.is_relevant(l[2], j[0])


def test_state_roundtrip_and_errors():
    m = capi.Model(W.urdf_text("pr2.urdf"))
    j = m.get_joint_ids(W.PR2_RARM)
    m.set_joint_angles(j, np.arange(7) * 0.1)
    assert np.array_equal(m.get_joint_angles(j), np.arange(7) * 0.1)
    m.set_base_pose([0.1, 0.2, 0.3, 0.3, 0.2, 0.1])
    np.testing.assert_allclose(m.get_base_pose(), [0.1, 0.2, 0.3, 0.3, 0.2, 0.1], atol=1e-15)
    with pytest.raises(ValueError, match="no joint named"):
        m.get_joint_ids(["nope"])
    with pytest.raises(ValueError, match="no link named"):
        m.get_link_ids(["nope"])
    m.add_new_link("x", 0, [0, 0, 0], [0, 0, 0])
    with pytest.raises(ValueError, match="already exists"):
        m.add_new_link("x", 0, [0, 0, 0], [0, 0, 0])
    with pytest.raises(ValueError):
        m.set_base_pose([0, 0, 0])
    with pytest.raises(ValueError):
        m.set_joint_angles([10 ** 6], [0.0])


@pytest.mark.parametrize("xml,exc", [
    ("", capi.TfkError),                                              # src/tinyfk.cpp:17-19 runtime_error
    ("<robot name='r'><link name='a'/><link name='a'/></robot>", ValueError),
    ("<robot name='r'></robot>", ValueError),
    ("<notrobot/>", ValueError),
    ("<robot name='r'><link name='a'/><link name='b'/></robot>", ValueError),      # two roots
    ("<robot name='r'><link name='a'/><link name='b'/><joint name='j' type='floating'><parent link='a'/>"
     "<child link='b'/></joint></robot>", ValueError),                             # src/tinyfk.cpp:61-63
    ("<robot name='r'><link name='a'/><link name='b'/><joint name='j' type='revolute'><parent link='a'/>"
     "<child link='b'/></joint></robot>", ValueError),                             # revolute without <limit>
    ("<robot name='r'><link name='a'><unclosed></link></robot>", ValueError),
    ("<robot name='r'><link name='a'>" + "<x>" * 100000 + "</x>" * 100000 + "</link></robot>", ValueError),  # nesting bound
    ("<robot name='r'><link name='a&#0;b'/></robot>", ValueError),          # NUL character reference
    ("<robot name='r'><link name='a&#x110000;'/></robot>", ValueError),     # beyond the last Unicode scalar value
])
def test_malformed_urdf_raises_instead_of_ub(xml, exc):
    with pytest.raises(exc):
        capi.Model(xml)


def test_minimal_urdf_defaults():
    xml = ("<?xml version='1.0'?><!-- c --><robot name='r'><link name='a'/><link name='b'/>"
           "<joint name='j' type='continuous'><parent link='a'/><child link='b'/></joint>"
           "<joint name='k' type='prismatic'><parent link='b'/><child link='c'/><axis/>"
           "<limit effort='1' velocity='1'/><origin xyz='1 2  3'/></joint><link name='c'/></robot>")
    m = capi.Model(xml)
    assert m.get_joint_names() == ["j", "k"] and m.get_link_names() == ["a", "b", "c"]
    p, t, jid, axis, origin = m.link_info(1)
    assert (p, t, jid) == (0, 2, 0) and list(axis) == [1, 0, 0] and list(origin) == [0, 0, 0, 0, 0, 0, 1]
    p, t, jid, axis, origin = m.link_info(2)
    assert (p, t, jid) == (1, 3, 1) and list(axis) == [0, 0, 0] and list(origin[:3]) == [1, 2, 3]
    assert m.get_joint_limits([1]) == [(0.0, 0.0)]


def test_plan_cache_is_keyed_on_the_signature_and_bounded():
    """Engine::plan_for: every *_host call writes the last configuration back into the model, which used to
    make the next call miss the cache and, when the stored state feeds the plan (alternating joint sets, as IK loops do),
    add an entry per call until a 64-entry flush.  Now: one entry per signature, replaced in place when its constants
    change, least recently used evicted at 64.  Runs without a GPU: n = 0 builds the plan and returns."""
    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    ja, jb = m.get_joint_ids(W.PR2_RARM), m.get_joint_ids(W.PR2_LARM)
    la = m.get_link_ids(["r_gripper_tool_frame", "l_gripper_tool_frame"])
    rng = np.random.default_rng(0)
    for k in range(100):
        # the signature over the right arm reads the stored angles of the left arm and vice versa
        m.set_joint_angles(jb, rng.uniform(-0.3, 0.3, 7))
        m.solve_fk(np.zeros((0, 7)), la, ja, 1, 0, True)
        m.set_joint_angles(ja, rng.uniform(-0.3, 0.3, 7))
        m.solve_fk(np.zeros((0, 7)), la, jb, 1, 0, True)
        assert m.plan_cache_size() == 2, k
    # an unchanged state keeps the model version (no rebuild), and state a plan does not read keeps its entry
    m.set_joint_angles(ja, m.get_joint_angles(ja))
    m.solve_fk(np.zeros((0, 7)), la, jb, 1, 0, True)
    assert m.plan_cache_size() == 2
    # many distinct signatures: bounded, least recently used evicted
    for lid in range(80):
        m.solve_fk(np.zeros((0, 7)), [lid], ja, 0, 0, False)
    assert m.plan_cache_size() == 64
