"""Pins the oracle before it is trusted (DESIGN.md "Oracle"):
  - oracle/_ref = the reference's own object code: reproduces the reference's golden vectors
    (test/data/test_data.json, used by test/test_kinematics.cpp:94-173 and python/tests/test_tinyfk.py:61-71)
    and the README example; its libm-pinned twin really is bound to fkmath.h;
  - oracle/fk_oracle.c (C restatement) == oracle/_ref bit for bit, and == the committed fixtures that were
    generated from _ref (tests/golden/make_golden.py), so the port can stand in where _ref is absent.
"""
import json
import os

import numpy as np
import pytest

from oracle import refapi
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
needs_ref = pytest.mark.skipif(not (refapi.have_ref(False) and refapi.have_ref(True)),
                               reason="oracle/_ref not built (needs /root/reference)")


def _testdata():
    with open(os.path.join(GOLD, "reference_test_data.json")) as f:
        td = json.load(f)
    gt = np.array(td["pose_list"])[:, [0, 1, 2, 5, 4, 3]]  # stored as x y z yaw pitch roll
    return td, gt


def _pr2_with_mylink(cls, **kw):
    m = cls(W.urdf_text("pr2.urdf"), **kw)
    pid = m.get_link_ids(["r_upper_arm_link"])[0]
    m.add_new_link("mylink", pid, [0.1, 0.1, 0.1], [0.3, 0.2, 0.1])
    return m


@needs_ref
@pytest.mark.parametrize("strict", [False, True])
def test_ref_reproduces_reference_golden_poses(strict):
    td, gt = _testdata()
    m = _pr2_with_mylink(refapi.RefModel, strict=strict)
    lids, jids = m.get_link_ids(td["link_names"]), m.get_joint_ids(td["joint_names"])
    P, _ = m.solve_fk([td["angle_vector"]], lids, jids, 1, 2, True)
    np.testing.assert_allclose(P, gt, rtol=0, atol=1e-7)  # the reference's own tolerance (7 decimals)
    assert np.abs(P - gt).max() < 1e-14                   # what it actually achieves


@needs_ref
def test_ref_readme_example():
    m = refapi.RefModel(W.urdf_text("pr2.urdf"))
    jids, lid = m.get_joint_ids(W.PR2_RARM), m.get_link_ids(["r_gripper_tool_frame"])[0]
    assert lid == 71 and jids == [39, 38, 40, 29, 30, 41, 42]
    P, J = m.solve_fk(W.README_Q, [lid], jids, 1, 1, True)
    np.testing.assert_allclose(P[0], [0.50812707267784318, 0.84860531692602614, 0.75061268052719599,
                                      -2.0642683981753733, -0.21174995675381791, 1.493189858833309], atol=1e-14)
    np.testing.assert_allclose(J[0, :9], [-0.84298458721470704, -0.026016505781018742, 0.23436381627257652,
                                          0.33428964230461938, 0.0036406922556169174, 0.17818746942806038, 0, 1, 0],
                               atol=1e-14)
    assert abs(J[0, 9] - (-0.64860533832700185)) < 1e-7  # finite-difference entry


@needs_ref
def test_pinned_ref_is_bound_to_the_portable_libm():
    import ctypes
    lib = ctypes.CDLL(os.path.join(os.path.dirname(refapi.port_lib_path()), "libstrict_libm.so"))
    x = np.random.default_rng(0).uniform(-3, 3, 2000)
    o = np.empty_like(x)
    dp = ctypes.POINTER(ctypes.c_double)
    lib.tfkm_sin_v(x.ctypes.data_as(dp), o.ctypes.data_as(dp), ctypes.c_long(x.size))
    differ = np.nonzero(o != np.sin(x))[0]
    assert differ.size > 0
    stock, pinned = refapi.RefModel(W.urdf_text("kuka.urdf"), False), refapi.RefModel(W.urdf_text("kuka.urdf"), True)
    for i in differ[:20]:
        assert stock.probe_sin(x[i]) == np.sin(x[i]) and pinned.probe_sin(x[i]) == o[i]


@needs_ref
@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_port_equals_ref_bitwise(cfg, strict):
    r = refapi.RefModel(W.urdf_text(W.CONFIGS[cfg][0]), strict=strict)
    jids, lids, rot, base = W.setup(r, cfg)
    p = refapi.PortModel(r.export_table(), strict=strict)
    q = W.sample_q(r.get_joint_limits(jids), base, 40, seed=1, wide_pitch=True)
    for jac in (False, True):
        Pr, Jr = r.solve_fk(q, lids, jids, rot, base, jac)
        Pp, Jp = p.solve_fk(q, lids, jids, rot, base, jac)
        assert np.array_equal(Pr, Pp) and np.array_equal(Jr, Jp)


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_port_equals_committed_reference_fixtures(cfg, strict):
    """works without /root/reference: link table from the product's parser, outputs vs fixtures made from _ref"""
    g = np.load(os.path.join(GOLD, "ref_%s_%s.npz" % (cfg, "pinned" if strict else "stock")))
    m = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0 if strict else 1)
    jids, lids, rot, base = W.setup(m, cfg)
    assert jids == list(g["joint_ids"]) and lids == list(g["link_ids"])
    p = refapi.PortModel(m.export_table(), strict=strict)
    P, J = p.solve_fk(g["q"], lids, jids, rot, base, True)
    if strict:
        assert np.array_equal(P, g["P"]) and np.array_equal(J, g["J"])
    else:  # stock fixtures carry the build container's glibc; another host's glibc may differ in the last bit
        assert np.abs(P - g["P"]).max() < 1e-12
        fd = np.zeros(J.shape[1], dtype=bool)
        fd[len(jids) + (2 if base == 1 else 3):] = base != 0
        assert np.abs(J - g["J"])[:, ~fd].max() < 1e-11 and np.abs(J - g["J"]).max() < 1e-6


@pytest.mark.parametrize("rot", [0, 1, 2])
def test_port_on_reference_testdata_fixture(rot):
    td, gt = _testdata()
    g = np.load(os.path.join(GOLD, "ref_testdata_rot%d_pinned.npz" % rot))
    m = _pr2_with_mylink(capi.Model, host_libm=0)
    lids, jids = m.get_link_ids(td["link_names"]), m.get_joint_ids(td["joint_names"])
    P, J = refapi.PortModel(m.export_table(), strict=True).solve_fk(g["q"], lids, jids, rot, 2, True)
    assert np.array_equal(P, g["P"]) and np.array_equal(J, g["J"])
    if rot == 1:
        np.testing.assert_allclose(P, gt, rtol=0, atol=1e-7)
