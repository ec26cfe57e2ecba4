"""tfk_solve_fk_device_noalloc (include/tinyfk_b200.h; tinyfk_b200/csrc/fk_noalloc.cu) -- no counterpart in the
reference, added: the several-link kernel with its scratch (records + per-link
tables) in a static device array, for callers that cannot call cudaMalloc.

Called through the C ABI (ctypes) and compared with the existing entry, tfk_solve_fk_device in TFK_F64_STRICT, bit for bit
(the tolerance of the strict-FP64 tests: identical numbers), on two calls in a row that SHARE the static array: a large
request first, then a smaller one with another signature, so the second finds the first one's leftovers everywhere.
torch owns the inputs and outputs here (the caller passes them); the entry itself must allocate nothing, which the test
observes through the device's free memory -- on REPEAT calls of a signature: the very first launch of a kernel instance
on a device lets the driver load its code (CUDA's lazy module loading, up to one 2 MiB granule, like any first kernel
launch anywhere), which is not an allocation made by the entry and is seen only in a fresh process.

This file sorts last on purpose (a failure here must not hide other tests), and passes both
alone in a fresh process and after the rest of the suite."""
import numpy as np
import pytest

import parity
from parity import Pair
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _run_both(pair, q, lids, jids, rot, base, jac, check_no_alloc=False):
    """the same device-resident inputs through tfk_solve_fk_device (strict FP64) and tfk_solve_fk_device_noalloc"""
    import torch

    n, nd = q.shape
    L, pd = len(lids), W.pose_dim(rot)
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(np.ascontiguousarray(q)).to(dev)
    outs = []
    st = torch.cuda.current_stream().cuda_stream
    for noalloc in (False, True):
        dP = torch.full((pd, n * L), float("nan"), dtype=torch.float64, device=dev)
        dJ = torch.full((nd, n * L * pd), float("nan"), dtype=torch.float64, device=dev) if jac else None
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        if noalloc:
            pair.prod.solve_fk_device_noalloc(0, st, dq.data_ptr(), n, lids, jids, rot, base, jac, dP.data_ptr(),
                                              dJ.data_ptr() if jac else 0, n, 0)
        else:
            pair.prod.solve_fk_device(0, st, dq.data_ptr(), n, lids, jids, rot, base, jac, 0, dP.data_ptr(),
                                      dJ.data_ptr() if jac else 0, n, 0)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info(0)[0]
        if noalloc and check_no_alloc:
            # cudaMalloc works in 2 MiB granules: any allocation by the entry would show (the records of this request
            # alone are tens of MB on the ordinary path)
            assert free0 - free1 < (1 << 20), "the allocation-free entry took %d bytes of device memory" % (free0 - free1)
        outs.append((dP.cpu().numpy().T, dJ.cpu().numpy().T if jac else None))
    return outs


def _tree(pair):
    jn = ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM
    ln = ["r_gripper_tool_frame", "l_gripper_tool_frame", "head_tilt_link", "base_footprint", "r_elbow_flex_link",
          "torso_lift_link", "l_forearm_link", "r_gripper_tool_frame", "base_link", "fl_caster_rotation_link"]
    return pair.prod.get_joint_ids(jn), pair.prod.get_link_ids(ln)


def test_noalloc_entry_equals_the_device_entry_on_two_calls_in_a_row():
    pair = Pair("pr2.urdf", strict=True)
    pair.prod.set_jit_mode(0)
    jids, lids = _tree(pair)
    limits = pair.prod.get_joint_limits(jids)
    # call 1: ten links, FLOATING base, RPY rows + Jacobian, several tiles per block, ragged tail
    q1 = W.sample_q(limits, W.BASE_FLOATING, 40000 + 77, seed=51)
    (P1, J1), (P1n, J1n) = _run_both(pair, q1, lids, jids, W.ROT_RPY, W.BASE_FLOATING, True)
    assert "twophase" in pair.prod.last_launch_info()["kernel"]
    parity.assert_bitwise(P1n, J1n, P1, J1, "no-alloc entry, call 1")
    assert np.isfinite(P1n).all()
    # call 2, right behind it on the same static array: another signature (one link, PLANER base, quaternion rows), fewer
    # configurations -- every byte it does not write itself is the first call's leftover
    jr = pair.prod.get_joint_ids(W.PR2_RARM)
    q2 = W.sample_q(pair.prod.get_joint_limits(jr), W.BASE_PLANER, 1000 + 3, seed=52)
    (P2, J2), (P2n, J2n) = _run_both(pair, q2, lids[:1], jr, W.ROT_XYZW, W.BASE_PLANER, True)
    parity.assert_bitwise(P2n, J2n, P2, J2, "no-alloc entry, call 2")
    # and against the reference itself on a few rows (the existing entry is pinned to it elsewhere)
    sel = np.r_[0:16, len(q2) - 16:len(q2)]
    Pr, Jr = pair.oracle_solve(q2[sel], lids[:1], jr, W.ROT_XYZW, W.BASE_PLANER, True)
    parity.assert_bitwise(P2n[sel], J2n.reshape(len(q2), 7, -1)[sel].reshape(-1, J2n.shape[1]), Pr, Jr, "no-alloc vs oracle")
    # calls 3 and 4: both requests again -- same bits as before (the array's contents in between do not matter), and, their
    # kernels being loaded and the symbol address cached by now, the device's free memory is left alone
    (_, _), (P3n, J3n) = _run_both(pair, q1, lids, jids, W.ROT_RPY, W.BASE_FLOATING, True, check_no_alloc=True)
    assert np.array_equal(P3n, P1n) and np.array_equal(J3n, J1n, equal_nan=True)
    (_, _), (P4n, J4n) = _run_both(pair, q2, lids[:1], jr, W.ROT_XYZW, W.BASE_PLANER, True, check_no_alloc=True)
    assert np.array_equal(P4n, P2n) and np.array_equal(J4n, J2n, equal_nan=True)


def test_noalloc_entry_without_jacobian_and_position_only():
    pair = Pair("fetch.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg4")
    q = W.sample_q(pair.prod.get_joint_limits(jids), base, 5000, seed=53)
    (P, _), (Pn, _) = _run_both(pair, q, lids, jids, rot, base, False)
    assert np.array_equal(P, Pn)
    (P, J), (Pn, Jn) = _run_both(pair, q, lids, jids, rot, base, True)
    parity.assert_bitwise(Pn, Jn, P, J, "no-alloc entry, cfg4")


def test_noalloc_entry_rejects_sizes_beyond_its_maximum():
    import torch

    pair = Pair("kuka.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg3")
    n = (1 << 16) + 1
    dq = torch.zeros((n, 7), dtype=torch.float64, device="cuda")
    dP = torch.zeros((7, n), dtype=torch.float64, device="cuda")
    dJ = torch.zeros((7, n * 7), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    with pytest.raises(ValueError, match="exceed the maximum of 65536"):
        pair.prod.solve_fk_device_noalloc(0, st, dq.data_ptr(), n, lids, jids, rot, base, True, dP.data_ptr(), dJ.data_ptr(), n, 0)
    # the maximum itself is accepted
    pair.prod.solve_fk_device_noalloc(0, st, dq.data_ptr(), n - 1, lids, jids, rot, base, True, dP.data_ptr(), dJ.data_ptr(), n, 0)
    torch.cuda.synchronize()
    assert torch.isfinite(dP[:, :n - 1]).all()
    with pytest.raises(ValueError):  # rows beyond the destination
        pair.prod.solve_fk_device_noalloc(0, st, dq.data_ptr(), 10, lids, jids, rot, base, True, dP.data_ptr(), dJ.data_ptr(), 5, 0)
