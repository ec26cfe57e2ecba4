"""The host entry's pipeline around the kernels (Engine::solve_host and friends), on a GPU:
  * destination pitches beyond cudaDevAttrMaxPitch (2 GiB - 1): one plain copy per column instead of cudaMemcpy2DAsync
    (cfg4 at 2^22 configurations needs 3.0 GB per Jacobian column);
  * Jacobian columns that are constants of the plan (all-zero columns of joints that move none of the links, the unit
    base-translation columns of src/kinematics.cpp:194-196) are filled on the host and never cross PCIe;
  * the device entry under CUDA-graph capture with the JIT in its default mode and a larger launch after the capture.
"""
import os

import numpy as np
import pytest

import parity
from parity import Pair
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _with_env(name, value, fn):
    old = os.environ.get(name)
    os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None:
            del os.environ[name]
        else:
            os.environ[name] = old


@pytest.mark.parametrize("cfg", ["cfg2", "cfg4", "cfg5"])
def test_per_column_copies_when_the_pitch_is_too_large(cfg):
    """TFK_MAX_PITCH (read once per model) lowers the limit so that small batches take the per-column path"""
    def run():
        pair = Pair(W.CONFIGS[cfg][0], strict=True)
        jids, lids, rot, base = pair.setup(cfg)
        n = 3001 if cfg != "cfg5" else 700
        q = W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=31)
        P, J = pair.prod_solve(q, lids, jids, rot, base, True)
        return pair, (jids, lids, rot, base), q, P, J
    pair, (jids, lids, rot, base), q, P, J = _with_env("TFK_MAX_PITCH", "4096", run)
    sel = np.r_[0:40, len(q) - 40:len(q)]
    Pr, Jr = pair.oracle_solve(q[sel], lids, jids, rot, base, True)
    L, pd = len(lids), W.pose_dim(rot)
    rows = (sel[:, None] * L + np.arange(L)[None, :]).reshape(-1)
    parity.assert_bitwise(P[rows], J[(rows[:, None] * pd + np.arange(pd)[None, :]).reshape(-1)], Pr, Jr, "per-column copies")
    P2, J2 = pair.prod_solve(q, lids, jids, rot, base, True)  # same model: limit cached -> same path; determinism
    assert np.array_equal(P, P2) and np.array_equal(J, J2)
    pair2 = Pair(W.CONFIGS[cfg][0], strict=True)              # fresh model, stock limit: strided copies
    pair2.setup(cfg)
    P3, J3 = pair2.prod_solve(q, lids, jids, rot, base, True)
    assert np.array_equal(P, P3) and np.array_equal(J, J3)


def test_cfg4_at_4m_configurations_through_the_host_entry():
    """BASELINE cfg4's nominal batch (2^22 configurations, 30 links): a Jacobian column is 3.0 GB, beyond the 2-D copy
    pitch limit.  One joint column keeps the host arrays at 3 + 3 GB.  Rows are compared with the device entry."""
    import torch

    free_kb = 0
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable"):
            free_kb = int(line.split()[1])
    if free_kb < 24 * 1024 * 1024:
        pytest.skip("needs ~8 GB of host memory")
    pair = Pair("fetch.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg4")
    jids = jids[1:2]  # shoulder_pan_joint
    n = 1 << 22
    q = W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=32)
    P, J = pair.prod.solve_fk(q, lids, jids, rot, base, True)
    assert J.shape == (n * 30 * 3, 1) and J.strides[1] > 2 ** 31
    sel = np.r_[0:300, n // 2 - 150:n // 2 + 150, n - 300:n]
    qs = np.ascontiguousarray(q[sel])
    m = len(sel)
    dq = torch.from_numpy(qs).cuda()
    dP = torch.empty((3, m * 30), dtype=torch.float64, device="cuda")
    dJ = torch.empty((1, m * 90), dtype=torch.float64, device="cuda")
    pair.prod.solve_fk_device(0, torch.cuda.current_stream().cuda_stream, dq.data_ptr(), m, lids, jids, rot, base, True, 0,
                              dP.data_ptr(), dJ.data_ptr(), m, 0)
    torch.cuda.synchronize()
    rows = (sel[:, None] * 30 + np.arange(30)[None, :]).reshape(-1)
    assert np.array_equal(P[rows], dP.cpu().numpy().T)
    assert np.array_equal(J[(rows[:, None] * 3 + np.arange(3)[None, :]).reshape(-1)], dJ.cpu().numpy().T)
    Pr, Jr = pair.oracle_solve(qs[:64], lids, jids, rot, base, True)
    parity.assert_bitwise(P[rows[:64 * 30]], J[:64 * 90], Pr, Jr, "4M rows")
    assert np.isfinite(P).all()


@pytest.mark.parametrize("base", [0, 1, 2])
def test_constant_columns_are_filled_on_the_host(base):
    """irrelevant joints (left-arm joints for a right-arm link) and base-translation columns: the default path (filled
    on the host, not copied) equals the copy-everything path (TFK_D2H_ALL=1) and the reference, float and double, with a
    ragged batch that spans several chunks"""
    pair = Pair("pr2.urdf", strict=True)
    jn = W.PR2_LARM[:2] + W.PR2_RARM + W.PR2_LARM[2:4]
    jids = pair.prod.get_joint_ids(jn)
    for lids, n in ((pair.prod.get_link_ids(["r_gripper_tool_frame"]), 50001),
                    (pair.prod.get_link_ids(["r_gripper_tool_frame", "r_elbow_flex_link", "r_forearm_link"]), 20011)):
        for rot in (0, 1, 2):
            q = W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=33 + rot)
            P, J = pair.prod_solve(q, lids, jids, rot, base, True)
            Pa, Ja = _with_env("TFK_D2H_ALL", "1", lambda: pair.prod_solve(q, lids, jids, rot, base, True))
            assert np.array_equal(P, Pa) and np.array_equal(J, Ja), (rot, base)
            assert not J[:, :2].any() and not J[:, 9:11].any()  # the left arm does not move the right arm
            P32, J32 = pair.prod_solve(q.astype(np.float32), lids, jids, rot, base, True, precision=1)
            Pb, Jb = _with_env("TFK_D2H_ALL", "1", lambda: pair.prod_solve(q.astype(np.float32), lids, jids, rot, base, True, precision=1))
            assert np.array_equal(P32, Pb) and np.array_equal(J32, Jb)
            sel = np.r_[0:20, n - 20:n]
            Pr, Jr = pair.oracle_solve(q[sel], lids, jids, rot, base, True)
            L, pd = len(lids), W.pose_dim(rot)
            rows = (sel[:, None] * L + np.arange(L)[None, :]).reshape(-1)
            parity.assert_bitwise(P[rows], J[(rows[:, None] * pd + np.arange(pd)[None, :]).reshape(-1)], Pr, Jr, "const columns")


def test_graph_capture_with_the_default_jit_mode_and_a_larger_launch_afterwards():
    """(1) the JIT may finish compiling in the middle of a capture -- the engine must not load a library or set
    function attributes then; (2) a later, larger launch on the same stream must not free the scratch a captured kernel
    node names (grow-only scratch)."""
    import torch

    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    assert m.get_jit_mode() == 1  # default: auto
    j = list(m.get_joint_ids(W.PR2_RARM))
    l = list(m.get_link_ids(["r_gripper_tool_frame", "r_elbow_flex_link", "r_wrist_roll_link"]))
    n, n_big, L = 256, 40000, 3
    dev = torch.device("cuda", 0)
    q = torch.from_numpy(W.sample_q(m.get_joint_limits(j), 1, n_big, seed=3)).to(dev)
    P = torch.zeros((3, n * L), dtype=torch.float64, device=dev)
    J = torch.zeros((10, n * L * 3), dtype=torch.float64, device=dev)
    Pb = torch.zeros((3, n_big * L), dtype=torch.float64, device=dev)
    Jb = torch.zeros((10, n_big * L * 3), dtype=torch.float64, device=dev)
    s = torch.cuda.Stream()

    def small(stream):
        m.solve_fk_device(0, stream.cuda_stream, q.data_ptr(), n, l, j, 0, 1, True, 0, P.data_ptr(), J.data_ptr(), n, 0)

    def big(stream):
        m.solve_fk_device(0, stream.cuda_stream, q.data_ptr(), n_big, l, j, 0, 1, True, 0, Pb.data_ptr(), Jb.data_ptr(), n_big, 0)

    with torch.cuda.stream(s):
        small(s)
    s.synchronize()
    ref = (P.clone(), J.clone())
    # the specialised kernel gets compiled now, but has never been launched: a capture must keep the generic kernel
    m.specialize(l, j, 0, 1, True, 0, wait=True, compile_only=True)
    P.zero_(), J.zero_()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        small(torch.cuda.current_stream())
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(P, ref[0]) and torch.equal(J, ref[1])
    # a much larger launch on the same stream grows the scratch; the captured node still names the old buffer
    with torch.cuda.stream(s):
        big(s)
    s.synchronize()
    P.zero_(), J.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(P, ref[0]) and torch.equal(J, ref[1])
    assert torch.equal(Pb[:, :n * L], ref[0])
    # growing inside a capture is refused with a clear error, not a CUDA failure
    s2 = torch.cuda.Stream()
    g2 = torch.cuda.CUDAGraph()
    with pytest.raises(ValueError, match="before capturing"):
        with torch.cuda.graph(g2, stream=s2):
            small(torch.cuda.current_stream())
