"""solve_com_fk (reference src/wrapper.cpp:126-179, src/kinematics.cpp:237-270,
python/tinyfk/__init__.py:160-195).

Oracle = oracle/_ref: the reference's own get_com / get_com_jacobian / get_jacobian (their Eigen expressions are
element-wise, so the storage-only Eigen stand-in evaluates them in the only possible order) driven by the loop of
wrapper.cpp:136-178 restated in oracle/ref_driver.cpp.  Strict FP64 is compared bit for bit.
CPU tests run the kernel's per-thread code through tests/emu; GPU tests go through the C ABI and the facade."""
import numpy as np
import pytest

from emuapi import EmuModel
from oracle import refapi
from tinyfk_b200 import workloads as W

JOINTS = ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM
ACTION = ["r_gripper_tool_frame", "l_gripper_tool_frame", "head_tilt_link"]
FORCES = [5.0, -3.0, 11.5]

needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")


def _q(o, jids, base, n, seed):
    return W.sample_q(o.get_joint_limits(jids), base, n, seed=seed, wide_pitch=True)


@needs_ref
def test_reference_com_is_the_mass_weighted_mean_of_its_com_links():
    """pins the restated loop: get_com against the reference's own solve_fk on the "_com" links"""
    o = refapi.RefModel(W.urdf_text("fetch.urdf"), strict=True)
    names = o.get_link_names()
    com_names = [n for n in names if n.endswith("_com")]
    assert len(com_names) == 20
    jids = o.get_joint_ids(W.FETCH_JOINTS)
    q = _q(o, jids, 2, 4, 5)
    com, J = o.solve_com_fk(q, jids, base_type=2, with_jacobian=True)
    lids = o.get_link_ids(com_names)
    P, Jl = o.solve_fk(q, lids, jids, 0, 2, True)
    P = P.reshape(4, len(lids), 3)
    # the COM lies inside the bounding box of the com links, and the translation block of J is the identity
    assert (com <= P.max(1) + 1e-12).all() and (com >= P.min(1) - 1e-12).all()
    nj = len(jids)
    for i in range(4):
        assert np.array_equal(J[3 * i:3 * i + 3, nj:nj + 3], np.eye(3))


@needs_ref
def test_reference_skips_the_cache_clear_with_a_fixed_base():
    """the quirk DESIGN.md documents and the product does NOT reproduce (src/wrapper.cpp:137-141)"""
    o = refapi.RefModel(W.urdf_text("kuka.urdf"), strict=True)
    jids = o.get_joint_ids(W.KUKA_JOINTS)
    q = _q(o, jids, 0, 3, 1)
    stale, _ = o.solve_com_fk(q, jids, clear_each=False)
    fresh, _ = o.solve_com_fk(q, jids, clear_each=True)
    assert np.array_equal(stale[1], stale[0]) and np.array_equal(stale[2], stale[0])
    assert not np.array_equal(fresh[1], fresh[0])
    assert np.array_equal(stale[0], fresh[0])


@needs_ref
@pytest.mark.parametrize("urdf,joints", [("pr2.urdf", JOINTS), ("fetch.urdf", W.FETCH_JOINTS), ("kuka.urdf", W.KUKA_JOINTS)])
@pytest.mark.parametrize("base", [0, 2])
@pytest.mark.parametrize("with_action", [False, True])
def test_com_core_equals_pinned_reference_on_cpu(urdf, joints, base, with_action):
    xml = W.urdf_text(urdf)
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=True)
    jids = list(e.get_joint_ids(joints))
    names = e.get_link_names()
    act = [len(names) // 3, len(names) // 2, 0] if with_action else []
    f = FORCES if with_action else []
    q = _q(o, jids, base, 6, 7 + base)
    cr, Jr = o.solve_com_fk(q, jids, act, f, base, True)
    c, J = e.solve_com_fk(q, jids, act, f, base, True)
    assert np.array_equal(c, cr) and np.array_equal(J, Jr)
    c2, _ = e.solve_com_fk(q, jids, act, f, base, False)
    assert np.array_equal(c2, cr)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("base", [0, 1, 2])
def test_gpu_com_bitwise_vs_pinned_reference(base):
    from tinyfk_b200 import capi
    xml = W.urdf_text("pr2.urdf")
    m, o = capi.Model(xml, host_libm=0), refapi.RefModel(xml, strict=True)
    jids = list(m.get_joint_ids(JOINTS))
    act = list(m.get_link_ids(ACTION))
    q = _q(o, jids, base, 300, 21 + base)  # ragged last tile
    for a, f in (([], []), (act, FORCES)):
        cr, Jr = o.solve_com_fk(q, jids, a, f, base, True)
        c, J = m.solve_com_fk(q, jids, a, f, base, True)
        assert c.shape == cr.shape and J.shape == Jr.shape == (900, len(jids) + (6 if base else 0))
        assert np.array_equal(c, cr) and np.array_equal(J, Jr)
        c2, J2 = m.solve_com_fk(q, jids, a, f, base, False)
        assert np.array_equal(c2, cr) and J2 is None
    assert m.last_launch_info()["kernel"] == "fk_com_kernel"
    # the model is left at the last configuration
    assert np.array_equal(np.asarray(m.get_joint_angles(jids)), q[-1, :len(jids)])
    # stock-glibc reference: 1e-12 on the COM and analytic columns, FD noise floor on the base-rotation columns
    cs, Js = refapi.RefModel(xml, strict=False).solve_com_fk(q, jids, act, FORCES, base, True)
    c, J = m.solve_com_fk(q, jids, act, FORCES, base, True)
    nj = len(jids)
    assert np.abs(c - cs).max() <= 1e-12 and np.abs(J[:, :nj] - Js[:, :nj]).max() <= 1e-12 and np.abs(J - Js).max() <= 1e-6


@pytest.mark.gpu
def test_gpu_com_facade_numerical_jacobian_and_fp32():
    import tinyfk_b200
    from tinyfk_b200 import BaseType, KinematicModel
    kin = KinematicModel(tinyfk_b200.fetch_urdfpath())
    jids = kin.get_joint_ids(W.FETCH_JOINTS)
    act = kin.get_link_ids(["gripper_link"])
    rng = np.random.default_rng(3)
    for base, nb in ((BaseType.FIXED, 0), (BaseType.PLANER, 3), (BaseType.FLOATING, 6)):
        q = rng.uniform(-0.4, 0.4, len(jids) + nb)
        com, J = kin.solve_com_fk([q], jids, act, [20.0], base, with_jacobian=True)
        n_dof = len(jids) + (6 if nb else 0)
        assert com.shape == (1, 3) and J.shape == (3, n_dof)
        eps = 1e-7
        Q = np.tile(q, (len(q), 1)) + eps * np.eye(len(q))
        c1, J0 = kin.solve_com_fk(Q, jids, act, [20.0], base, with_jacobian=False)
        assert J0.shape == (3 * len(q), n_dof) and not J0.any()
        Jn = ((c1 - com) / eps).T  # (3, len(q))
        cols = list(range(len(jids))) + ([len(jids), len(jids) + 1, len(jids) + 5] if nb == 3 else list(range(len(jids), n_dof)))
        np.testing.assert_almost_equal(Jn, J[:, cols], decimal=5)
    q = rng.uniform(-0.4, 0.4, (200, len(jids) + 6))
    c64, J64 = kin.solve_com_fk(q, jids, act, [20.0], BaseType.FLOATING, with_jacobian=True)
    c32, J32 = kin.solve_com_fk(q, jids, act, [20.0], BaseType.FLOATING, with_jacobian=True, precision="fp32")
    assert c32.dtype == np.float32 and np.abs(c32 - c64).max() < 1e-5 and np.abs(J32 - J64).max() < 2e-5
    # legacy 6-argument entry of _tinyfk.RobotModel
    c, J = kin._robot.solve_com_forward_kinematics(q[:3], jids, act, [20.0], True, True)
    assert np.array_equal(c, c64[:3]) and np.array_equal(J, J64.reshape(200, 3, -1)[:3].reshape(9, -1))


@pytest.mark.gpu
def test_gpu_com_device_entry_rejects_planer_and_bad_ids():
    from tinyfk_b200 import capi
    m = capi.Model(W.urdf_text("kuka.urdf"), host_libm=0)
    jids = list(m.get_joint_ids(W.KUKA_JOINTS))
    with pytest.raises(ValueError):
        m.solve_com_fk_device(0, 0, 0, 0, jids, [], [], 1, False, 0, 0, 0, 0, 0)
    with pytest.raises(ValueError):
        m.solve_com_fk(np.zeros((1, 7)), jids, [10 ** 6], [1.0])
