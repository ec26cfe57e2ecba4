#!/usr/bin/env python
"""Device timing of solve_com_fk (centre of mass + Jacobian) on PR2 / Fetch.  usage: python tools/time_com.py [n]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from tinyfk_b200 import capi  # noqa: E402
from tinyfk_b200 import workloads as W  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 17
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
for urdf, jn in (("pr2.urdf", ["torso_lift_joint"] + W.PR2_RARM + W.PR2_LARM), ("fetch.urdf", W.FETCH_JOINTS)):
    m = capi.Model(W.urdf_text(urdf), host_libm=0)
    jids = list(m.get_joint_ids(jn))
    for base in (0, 2):
        for jac in (False, True):
            for prec in (0, 1):
                dt = torch.float32 if prec else torch.float64
                n_in = len(jids) + W.base_dof(base)
                q = [torch.from_numpy(W.sample_q(m.get_joint_limits(jids), base, n, seed=r)).to(dev).to(dt) for r in range(2)]
                com = torch.empty((3, n), dtype=dt, device=dev)
                J = torch.empty((max(n_in, 1), 3 * n), dtype=dt, device=dev)

                def step(i):
                    m.solve_com_fk_device(0, stream.cuda_stream, q[i % 2].data_ptr(), n, jids, [], [], base, jac, prec,
                                          com.data_ptr(), J.data_ptr(), n, 0)
                for i in range(3):
                    step(i)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for i in range(5):
                    step(i)
                e1.record(stream)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 5
                b = (n_in + 3 * (1 + (n_in if jac else 0))) * (4 if prec else 8)
                print("%-10s base %d jac %d %s  %8.4f ms  %7.1f GB/s  %.3e cfg/s  %s" % (
                    urdf, base, jac, "fp32" if prec else "fp64", ms, b * n / ms / 1e6, n / ms * 1e3, m.last_launch_info()))
