"""The engine's OWN multi-GPU split: tfk_solve_fk_host(devices=[...]) (include/tinyfk_b200.h; Engine::solve_host) cuts
one batch into contiguous per-device slices, each slice into chunks on that device's stream slots.  BASELINE cfg3
("16M configs ... sharded across 2/4/8 B200") and cfg5 ("8M ... FP64 and FP32 on 8xB200") go through it.

Checked on real devices (skips below 2): rows around every device and chunk boundary equal the reference bit for bit
(FP64) / within 1e-5 (FP32), the whole result equals the single-device result, and -- the reference's batch contract,
python/tests/test_tinyfk.py:133-154 -- rows of a batch equal single calls, across a device boundary.
Run with `gpurun --gpus 2` / `--gpus 8`; logs under profiles/.
"""
import numpy as np
import pytest

import parity
from parity import Pair
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _n_devices():
    try:
        return capi.device_count()
    except Exception:
        return 0


def _device_sets():
    n = _n_devices()
    sets = []
    if n >= 2:
        sets.append([0, 1])
    if n >= 3:
        sets.append([n - 1, 0, 1])  # any order, not starting at the current device
    if n >= 4:
        sets.append(list(range(n)))
    return sets


needs_two = pytest.mark.skipif(_n_devices() < 2, reason="needs at least two CUDA devices")


def _boundary_rows(n, n_dev, chunk_hint):
    per = (n + n_dev - 1) // n_dev
    rows = set(range(0, 6)) | set(range(n - 6, n))
    for g in range(1, n_dev):
        b = min(n, g * per)
        rows |= set(range(max(0, b - 6), min(n, b + 6)))
        c = min(n, (g - 1) * per + chunk_hint)  # first chunk boundary inside the previous slice
        rows |= set(range(max(0, c - 3), min(n, c + 3)))
    return np.array(sorted(rows))


@needs_two
@pytest.mark.parametrize("cfg,n_per_dev", [("cfg3", 150001), ("cfg5", 8192 + 541), ("cfg2", 70007), ("cfg4", 9000)])
@pytest.mark.parametrize("precision", [0, 1])
def test_device_list_split_equals_reference_and_single_device(cfg, n_per_dev, precision):
    for devs in _device_sets():
        pair = Pair(W.CONFIGS[cfg][0], strict=True)
        jids, lids, rot, base = pair.setup(cfg)
        n = n_per_dev * len(devs) - 3  # not a multiple of devices x chunk, ragged last slice
        limits = pair.prod.get_joint_limits(jids)
        q = W.sample_q(limits, base, n, seed=21)
        if precision == 1:
            q = q.astype(np.float32)
        P, J = pair.prod_solve(q, lids, jids, rot, base, True, precision=precision, devices=devs)
        P1, J1 = pair.prod_solve(q, lids, jids, rot, base, True, precision=precision, devices=[devs[0]])
        assert np.array_equal(P, P1) and np.array_equal(J, J1), (cfg, devs)
        sel = _boundary_rows(n, len(devs), 8192)
        Pr, Jr = pair.oracle_solve(q[sel].astype(np.float64), lids, jids, rot, base, True)
        L, pd = len(lids), W.pose_dim(rot)
        rows = (sel[:, None] * L + np.arange(L)[None, :]).reshape(-1)
        Ps, Js = P[rows], J[(rows[:, None] * pd + np.arange(pd)[None, :]).reshape(-1)]
        if precision == 0:
            parity.assert_bitwise(Ps, Js, Pr, Jr, "%s devices %s" % (cfg, devs))
        else:
            assert np.abs(Ps[:, :3] - Pr[:, :3]).max() <= parity.TOL_F32
            Jv, Jrv = Js.reshape(-1, pd, Js.shape[1]), Jr.reshape(-1, pd, Jr.shape[1])
            assert np.abs(Jv[:, :3] - Jrv[:, :3]).max() <= parity.TOL_F32
        # stored state = the LAST configuration of the whole batch, whichever device computed it (wrapper.cpp:85-89)
        assert np.array_equal(pair.prod.get_joint_angles(jids), q[-1, :len(jids)].astype(np.float64))


@needs_two
def test_batch_rows_equal_single_calls_across_a_device_boundary():
    """python/tests/test_tinyfk.py:133-154 (trajectory contract): solve_fk(batch)[i] == solve_fk(batch[i])"""
    import tinyfk_b200
    from tinyfk_b200 import BaseType, KinematicModel, RotationType

    kin = KinematicModel(tinyfk_b200.pr2_urdfpath())
    jids = kin.get_joint_ids(W.PR2_RARM)
    lids = kin.get_link_ids(["r_gripper_tool_frame", "r_forearm_link"])
    devs = list(range(min(_n_devices(), 8)))
    n = 16 * len(devs) + 5
    rng = np.random.default_rng(5)
    q = np.hstack([rng.uniform(-0.5, 0.5, (n, 7)), rng.uniform(-1, 1, (n, 3))])
    for rot in (RotationType.IGNORE, RotationType.RPY, RotationType.XYZW):
        P, J = kin.solve_fk(q, lids, jids, rot_type=rot, base_type=BaseType.PLANER, with_jacobian=True, devices=devs)
        pd = P.shape[1]
        per = (n + len(devs) - 1) // len(devs)
        for i in sorted({0, per - 1, per, per + 1, 2 * per - 1, n - 1} & set(range(n))):
            Pi, Ji = kin.solve_fk(q[i], lids, jids, rot_type=rot, base_type=BaseType.PLANER, with_jacobian=True)
            assert np.array_equal(P[2 * i:2 * i + 2], Pi) and np.array_equal(J[2 * pd * i:2 * pd * (i + 1)], Ji), (rot, i)


@needs_two
def test_bad_device_ordinal_is_a_value_error():
    pair = Pair("kuka.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg3")
    with pytest.raises(ValueError):
        pair.prod.solve_fk(np.zeros((4, 7)), lids, jids, rot, base, True, devices=[0, _n_devices()])
