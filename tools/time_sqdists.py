#!/usr/bin/env python
"""Device timing of compute_inter_link_sqdists on the PR2 dual-arm self-collision query.
usage: python tools/time_sqdists.py [n]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from tinyfk_b200 import capi  # noqa: E402
from tinyfk_b200 import workloads as W  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
jids = list(m.get_joint_ids(W.PR2_RARM + W.PR2_LARM))
R = ["r_shoulder_pan_link", "r_upper_arm_link", "r_elbow_flex_link", "r_forearm_link", "r_wrist_roll_link", "r_gripper_tool_frame"]
L = ["l" + s[1:] for s in R]
rid, lid = m.get_link_ids(R), m.get_link_ids(L)
for npair in (1, 6, 36):
    pairs = [(a, b) for a in rid for b in lid][:npair] if npair > 6 else list(zip(rid, lid))[:npair]
    for base in (0, 1, 2):
        for jac in (False, True):
            for prec in (0, 1):
                dt = torch.float32 if prec else torch.float64
                n_in = len(jids) + W.base_dof(base)
                q = [torch.from_numpy(W.sample_q(m.get_joint_limits(jids), base, n, seed=r)).to(dev).to(dt) for r in range(2)]
                sq = torch.empty(n * npair, dtype=dt, device=dev)
                G = torch.empty((n_in, n * npair), dtype=dt, device=dev)

                def step(i):
                    m.inter_link_sqdists_device(0, stream.cuda_stream, q[i % 2].data_ptr(), n, pairs, jids, base, jac, prec,
                                                sq.data_ptr(), G.data_ptr(), n, 0)
                for i in range(3):
                    step(i)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for i in range(10):
                    step(i)
                e1.record(stream)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 10
                b = (n_in + npair * (1 + (n_in if jac else 0))) * (4 if prec else 8)
                print("pairs %2d base %d jac %d %s  %8.4f ms  %7.1f GB/s  %.3e cfg/s  %s" % (
                    npair, base, jac, "fp32" if prec else "fp64", ms, b * n / ms / 1e6, n / ms * 1e3, m.last_launch_info()))

# a self-collision query of one arm: 7 joints, 6 arm links x 4 body links
jr = list(m.get_joint_ids(W.PR2_RARM))
body = m.get_link_ids(["base_link", "torso_lift_link", "head_pan_link", "l_shoulder_pan_link"])
pairs = [(a, b) for a in rid for b in body]
for base in (0, 1):
    for jac in (False, True):
        n_in = len(jr) + W.base_dof(base)
        q = [torch.from_numpy(W.sample_q(m.get_joint_limits(jr), base, n, seed=r)).to(dev) for r in range(2)]
        sq = torch.empty(n * len(pairs), dtype=torch.float64, device=dev)
        G = torch.empty((n_in, n * len(pairs)), dtype=torch.float64, device=dev)

        def step(i):
            m.inter_link_sqdists_device(0, stream.cuda_stream, q[i % 2].data_ptr(), n, pairs, jr, base, jac, 0, sq.data_ptr(),
                                        G.data_ptr(), n, 0)
        for i in range(3):
            step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(10):
            step(i)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        b = (n_in + len(pairs) * (1 + (n_in if jac else 0))) * 8
        print("one arm, pairs %2d base %d jac %d fp64  %8.4f ms  %7.1f GB/s  %.3e cfg/s  %s" % (
            len(pairs), base, jac, ms, b * n / ms / 1e6, n / ms * 1e3, m.last_launch_info()))
