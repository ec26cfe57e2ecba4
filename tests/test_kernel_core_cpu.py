"""The kernels' per-thread code (tinyfk_b200/csrc/fk_core.h) run on the CPU by the test harness
tests/emu (NOT a product path) through the product's host model and plan flattening, against the
oracle.  Pins plan pruning / ordering / relevance masks / record layout / PLANER fusion without a GPU;
the GPU tests then only have to show that nvcc -fmad=false produces the same bits."""
import itertools

import numpy as np
import pytest

import emuapi
from emuapi import EmuModel
from oracle import refapi
from tinyfk_b200 import workloads as W


def _oracle_for(emu_setup, xml, strict):
    """(oracle, emu) with the same structure; RefModel when built, else PortModel from the product table"""
    if refapi.have_ref(strict):
        o = refapi.RefModel(xml, strict=strict)
        emu_setup(o)
        return o
    from tinyfk_b200 import capi
    m = capi.Model(xml, host_libm=0 if strict else 1)
    emu_setup(m)
    return refapi.PortModel(m.export_table(), strict=strict)


def _tree_setup(m):
    pid = m.get_link_ids(["r_upper_arm_link"])[0]
    m.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])


TREE_J = ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM
TREE_L = ["r_gripper_tool_frame", "l_gripper_tool_frame", "head_tilt_link", "base_footprint", "r_elbow_flex_link",
          "torso_lift_link", "l_forearm_link", "r_gripper_tool_frame", "base_link", "fl_caster_rotation_link",
          "r_upper_arm_link_com", "mylink"]


@pytest.mark.parametrize("rot,base,jac", list(itertools.product((0, 1, 2), (0, 1, 2), (False, True))))
def test_core_equals_pinned_reference_on_a_tree(rot, base, jac):
    xml = W.urdf_text("pr2.urdf")
    e = EmuModel(xml)
    _tree_setup(e)
    o = _oracle_for(_tree_setup, xml, True)
    jids, lids = e.get_joint_ids(TREE_J), e.get_link_ids(TREE_L)
    rng = np.random.default_rng(rot * 9 + base * 3 + int(jac))
    ang = rng.uniform(-0.3, 0.3, e.n_joints)
    bp = [0.3, -0.2, 0.1, 0.2, -0.1, 0.7]
    for m in (e, o):
        m.set_joint_angles(list(range(e.n_joints)), ang)
        m.set_base_pose(bp)
    lim = [(-1.0, 1.0)] * len(jids)
    q = W.sample_q(lim, base, 24, seed=rot + 3 * base, wide_pitch=True)
    Pr, Jr = o.solve_fk(q, lids, jids, rot, base, jac)
    P, J = e.solve_fk(q, lids, jids, rot, base, jac, mode=emuapi.MODE_STRICT64)
    assert ((P == Pr) | (np.isnan(P) & np.isnan(Pr))).all() and ((J == Jr) | (np.isnan(J) & np.isnan(Jr))).all()


@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_core_equals_committed_fixtures(cfg):
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_%s_pinned.npz" % cfg))
    e = EmuModel(W.urdf_text(W.CONFIGS[cfg][0]))
    jids, lids, rot, base = W.setup(e, cfg)
    P, J = e.solve_fk(g["q"], lids, jids, rot, base, True)
    assert np.array_equal(P, g["P"]) and np.array_equal(J, g["J"])


def test_core_fp32_and_native64_modes_are_close():
    xml = W.urdf_text("kuka.urdf")
    e = EmuModel(xml, system_libm=True)
    o = _oracle_for(lambda m: None, xml, False)
    jids, lids = e.get_joint_ids(W.KUKA_JOINTS), e.get_link_ids(["lbr_iiwa_link_7"])
    q = W.sample_q([(-2.0, 2.0)] * 7, 2, 200, seed=5)
    Pr, Jr = o.solve_fk(q, lids, jids, 2, 2, True)
    P, J = e.solve_fk(q, lids, jids, 2, 2, True, mode=emuapi.MODE_NATIVE64)
    assert np.abs(P - Pr).max() < 1e-13 and np.abs(J - Jr)[:, :10].max() < 1e-13 and np.abs(J - Jr).max() < 1e-6
    P, J = e.solve_fk(q, lids, jids, 2, 2, True, mode=emuapi.MODE_F32)
    assert np.abs(P - Pr).max() < 1e-5 and np.abs(J - Jr).max() < 1e-5


@pytest.mark.parametrize("cfg,n", [("cfg2", 64), ("cfg3", 64), ("cfg4", 12), ("cfg5", 6)])
@pytest.mark.parametrize("two_phase", [False, True])
def test_pair_layout_accessors_equal_the_plain_layout(cfg, n, two_phase):
    """FkCore<..., PAIRS = true> (the record layout of the several-link device kernels: whole 16-byte pairs per access,
    csrc/fk_core.h "record layout") run on the CPU with stride 1: every run of 3, 4 and 7 record elements starting on an
    even or an odd offset goes through the pair accessors and must give the plain layout's bits."""
    from emuapi import MODE_F32, MODE_STRICT64, EmuModel
    from tinyfk_b200 import capi
    emu = EmuModel(W.urdf_text(W.CONFIGS[cfg][0]))
    j, l, rot, base = W.setup(emu, cfg)
    prod = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0)
    jp, _, _, _ = W.setup(prod, cfg)
    q = W.sample_q(prod.get_joint_limits(jp), base, n, seed=31, wide_pitch=True)
    tp = 16 if two_phase else 0
    for mode in (MODE_STRICT64, MODE_F32):
        for jac in (True, False):
            P0, J0 = emu.solve_fk(q, l, j, rot, base, jac, mode=mode | tp)
            P1, J1 = emu.solve_fk(q, l, j, rot, base, jac, mode=mode | tp | 32)
            assert np.array_equal(P0, P1) and np.array_equal(J0, J1), (cfg, mode, jac)
