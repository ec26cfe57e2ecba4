"""compute_total_inertia_matrix (reference src/wrapper.cpp:181-187,
src/kinematics.cpp:272-317, python/tinyfk/__init__.py:197-198).

Oracle = the reference's own get_total_inertia_matrix in oracle/_ref.  Its Eigen 3x3 expressions run on the storage +
plain-loop stand-in oracle/eigen_stub (Eigen is not installed), so Eigen's internal summation order is unpinned: the
contract is 1e-12 relative.  Against the stand-in's order the product is bit-identical."""
import numpy as np
import pytest

from emuapi import EmuModel
from oracle import refapi
from tinyfk_b200 import workloads as W

needs_ref = pytest.mark.skipif(not refapi.have_ref(True), reason="oracle/_ref not built")
CASES = [("pr2.urdf", W.PR2_RARM + W.PR2_LARM), ("fetch.urdf", W.FETCH_JOINTS), ("kuka.urdf", W.KUKA_JOINTS)]


@needs_ref
@pytest.mark.parametrize("urdf,joints", CASES)
def test_inertia_core_equals_reference_on_cpu(urdf, joints):
    xml = W.urdf_text(urdf)
    e, o = EmuModel(xml), refapi.RefModel(xml, strict=True)
    jids = list(e.get_joint_ids(joints))
    q = W.sample_q(o.get_joint_limits(jids), 0, 8, seed=3)
    ref = o.total_inertia_matrix(q, jids)
    got = e.total_inertia_matrix(q, jids)
    assert np.array_equal(got, ref)
    # physical sanity of the oracle itself: symmetric positive definite, invariant trace bound
    for m in ref:
        assert np.allclose(m, m.T, rtol=0, atol=1e-12 * np.abs(m).max())
        assert (np.linalg.eigvalsh((m + m.T) / 2) > 0).all()


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("urdf,joints", CASES)
def test_gpu_inertia_vs_reference(urdf, joints):
    from tinyfk_b200 import capi
    xml = W.urdf_text(urdf)
    m, o = capi.Model(xml, host_libm=0), refapi.RefModel(xml, strict=True)
    jids = list(m.get_joint_ids(joints))
    q = W.sample_q(o.get_joint_limits(jids), 0, 200, seed=4)
    ref = o.total_inertia_matrix(q, jids)
    got = m.total_inertia_matrix(q, jids)
    assert got.shape == (200, 3, 3) and np.array_equal(got, ref)
    assert m.last_launch_info()["kernel"] == "fk_inertia_kernel"
    stock = refapi.RefModel(xml, strict=False).total_inertia_matrix(q, jids)
    assert np.abs(got - stock).max() <= 1e-12 * np.abs(stock).max()
    f32 = m.total_inertia_matrix(q, jids, precision=1)
    assert f32.dtype == np.float32 and np.abs(f32 - ref).max() <= 2e-5 * np.abs(ref).max()


@pytest.mark.gpu
def test_gpu_inertia_facade():
    import tinyfk_b200
    from tinyfk_b200 import KinematicModel
    kin = KinematicModel(tinyfk_b200.pr2_urdfpath())
    jids = kin.get_joint_ids(W.PR2_RARM)
    q = np.linspace(-0.5, 0.5, 7)
    m = kin.compute_total_inertia_matrix(q, jids)
    assert m.shape == (3, 3) and np.allclose(m, m.T, atol=1e-10)
    mb = kin.compute_total_inertia_matrix(np.stack([q, q * 0.5]), jids)
    assert mb.shape == (2, 3, 3) and np.array_equal(mb[0], m) and not np.array_equal(mb[1], m)
    assert np.array_equal(kin._robot.compute_total_inertia_matrix(list(q), jids), m)
    # every principal moment is bounded by total mass x (robot extent)^2
    assert 0 < np.trace(m) < kin.total_mass * 3 * 2.0 ** 2
