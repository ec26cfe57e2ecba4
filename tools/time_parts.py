#!/usr/bin/env python
"""Device timing of the PR2-arm query in several shapes (pose only / FIXED base / PLANER / FLOATING, rot types)
to see what each part of the kernel costs.  usage: python tools/time_parts.py [n]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from tinyfk_b200 import capi  # noqa: E402
from tinyfk_b200 import workloads as W  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()


def run(urdf, jn, ln, rot, base, jac, prec=0, reps=10):
    m = capi.Model(W.urdf_text(urdf), host_libm=0)
    jids, lids = list(m.get_joint_ids(jn)), list(m.get_link_ids(ln))
    pd, n_in, L = W.pose_dim(rot), len(jids) + W.base_dof(base), len(lids)
    dt = torch.float32 if prec else torch.float64
    q = [torch.from_numpy(W.sample_q(m.get_joint_limits(jids), base, n, seed=r)).to(dev).to(dt) for r in range(2)]
    P = torch.empty((pd, n * L), dtype=dt, device=dev)
    J = torch.empty((max(n_in, 1), n * L * pd), dtype=dt, device=dev)

    def step(i):
        m.solve_fk_device(0, stream.cuda_stream, q[i % 2].data_ptr(), n, lids, jids, rot, base, jac, prec, P.data_ptr(),
                          J.data_ptr(), n, 0)
    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(reps):
        step(i)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    b = (n_in + L * pd * (1 + (n_in if jac else 0))) * (4 if prec else 8)
    return ms, b * n / ms / 1e6


for name, args in [
    ("pr2 arm IGNORE FIXED pose-only", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 0, 0, False)),
    ("pr2 arm RPY    FIXED pose-only", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 0, False)),
    ("pr2 arm IGNORE FIXED jac      ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 0, 0, True)),
    ("pr2 arm XYZW   FIXED jac      ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 2, 0, True)),
    ("pr2 arm RPY    FIXED jac      ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 0, True)),
    ("pr2 arm RPY    PLANER pose    ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 1, False)),
    ("pr2 arm IGNORE PLANER jac     ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 0, 1, True)),
    ("pr2 arm RPY    PLANER jac cfg2", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 1, True)),
    ("pr2 arm RPY    FLOAT  jac     ", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 2, True)),
    ("kuka    XYZW   FIXED jac  cfg3", ("kuka.urdf", W.KUKA_JOINTS, ["lbr_iiwa_link_7"], 2, 0, True)),
    ("kuka    XYZW   FIXED pose     ", ("kuka.urdf", W.KUKA_JOINTS, ["lbr_iiwa_link_7"], 2, 0, False)),
    ("pr2 arm RPY    PLANER jac fp32", ("pr2.urdf", W.PR2_RARM, ["r_gripper_tool_frame"], 1, 1, True, 1)),
]:
    ms, gbs = run(*args)
    print("%s  %8.4f ms  %8.1f GB/s  %.3e cfg/s" % (name, ms, gbs, n / ms * 1e3))
