"""include/tinyfk_b200.hpp -- the header-only C++ view of the C ABI for callers that use tinyfk from C++ (the reference's
bench/bench_tinyfk.cpp drives tinyfk::KinematicModel::set_joint_angles / get_link_pose / get_jacobian directly).  The test
program tests/cpp/shim_check.cpp is compiled with g++ against the in-tree library; on a GPU its numbers must equal the Python
facade's bit for bit."""
import os
import subprocess

import numpy as np
import pytest

from tinyfk_b200 import workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
URDF = os.path.join(ROOT, "tinyfk_b200", "data", "pr2.urdf")


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("shim") / "shim_check")
    lib = os.path.join(ROOT, "tinyfk_b200")
    cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "shim_check.cpp"), "-L" + lib, "-ltinyfk_b200", "-Wl,-rpath," + lib, "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_the_header_compiles_links_and_resolves_ids_without_a_gpu(exe):
    r = subprocess.run([exe, URDF, "host-only"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    from tinyfk_b200 import capi
    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    j = m.get_joint_ids(W.PR2_RARM)
    l = m.get_link_ids(["r_gripper_tool_frame", "r_elbow_flex_link", "base_link"])
    new = m.add_new_link("mylink", l[0], [0.1, 0.2, 0.3], [0.3, 0.2, 0.1])
    assert r.stdout.split() == ["ids", str(j[0]), str(l[0]), str(l[2]), str(new)]


def _vals(line):
    return np.array([float.fromhex(x) for x in line.split()[1:]])


@pytest.mark.gpu
def test_cpp_calls_equal_the_python_facade_bit_for_bit(exe):
    r = subprocess.run([exe, URDF], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    from tinyfk_b200 import capi
    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    j = list(m.get_joint_ids(W.PR2_RARM))
    l = list(m.get_link_ids(["r_gripper_tool_frame", "r_elbow_flex_link", "base_link"]))
    new = m.add_new_link("mylink", l[0], [0.1, 0.2, 0.3], [0.3, 0.2, 0.1])
    ang = [0.564, 0.35, -0.74, -0.7, -0.7, -0.17, -0.63]
    base = [0.1, 0.2, 0.05, 0.02, -0.03, 0.3]
    q = np.array([ang + base])
    P, _ = m.solve_fk(q, l + [new], j, W.ROT_XYZW, W.BASE_FLOATING, False)
    for k in range(4):
        assert np.array_equal(_vals(lines[2 * k]), P[k, :3]) and np.array_equal(_vals(lines[2 * k + 1]), P[k, 3:])
    assert lines[8].split() == ["jac", "6", "13"]
    # get_jacobian(with_base) evaluates the STORED base pose: the header reads it back as (xyz, rpy) -- rpy recomputed from
    # the stored quaternion, equal to what was set to within rounding -- and sends it as the row's base part
    m.set_base_pose(base)
    stored = list(m.get_base_pose())
    assert np.abs(np.array(stored) - np.array(base)).max() < 1e-15
    _, J = m.solve_fk(np.array([ang + stored]), [l[0]], j, W.ROT_RPY, W.BASE_FLOATING, True)
    assert np.array_equal(_vals(lines[9]), J.T.ravel())     # column-major 6 x 13
    # (both sides now hold the base pose that call wrote back; the FIXED-base Jacobian below is evaluated at it)
    _, J2 = m.solve_fk(np.array([ang]), [new], j, W.ROT_IGNORE, W.BASE_FIXED, True)
    assert np.array_equal(_vals(lines[10]), J2.T.ravel())
    qs = np.array([[0.1 * (c + 1) - 0.37 * i for c in range(10)] for i in range(3)])
    Pb, Jb = m.solve_fk(qs, [l[0], new], j, W.ROT_XYZW, W.BASE_PLANER, True)
    assert np.array_equal(_vals(lines[11]), Pb.T.ravel()) and np.array_equal(_vals(lines[12]), Jb.T.ravel())
    assert lines[13] == "unknown link: invalid_argument"
