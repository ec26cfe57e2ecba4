#!/usr/bin/env python
"""Latency of small batches through tfk_solve_fk_device: one call per batch vs a CUDA graph of 64 calls (cfg1/cfg2 query:
PR2 arm, RPY, PLANER, pose + Jacobian).  usage: python tools/time_small_batches.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from tinyfk_b200 import capi  # noqa: E402
from tinyfk_b200 import workloads as W  # noqa: E402

dev = torch.device("cuda", 0)
m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
j, l = list(m.get_joint_ids(W.PR2_RARM)), list(m.get_link_ids(["r_gripper_tool_frame"]))
for jit in (0, 1):
    m.set_jit_mode(0)
    if jit:
        m.specialize(l, j, 1, 1, True, 0, wait=True)
        m.set_jit_mode(1)
    for n in (1, 32, 1024, 16384):
        q = torch.from_numpy(W.sample_q(m.get_joint_limits(j), 1, n, seed=0)).to(dev)
        P = torch.empty((6, n), dtype=torch.float64, device=dev)
        J = torch.empty((10, n * 6), dtype=torch.float64, device=dev)

        def run(stream):
            m.solve_fk_device(0, stream.cuda_stream, q.data_ptr(), n, l, j, 1, 1, True, 0, P.data_ptr(), J.data_ptr(), n, 0)
        s = torch.cuda.Stream()
        for _ in range(10):
            run(s)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(1000):
            run(s)
        torch.cuda.synchronize()
        t_call = (time.perf_counter() - t0) / 1000
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            cs = torch.cuda.current_stream()
            for _ in range(64):
                run(cs)
        g.replay()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            g.replay()
        torch.cuda.synchronize()
        t_graph = (time.perf_counter() - t0) / (20 * 64)
        print("jit %d  n %6d  per call %7.2f us (%.3e cfg/s)   in a 64-call graph %7.2f us (%.3e cfg/s)  %s"
              % (jit, n, t_call * 1e6, n / t_call, t_graph * 1e6, n / t_graph, m.last_launch_info()["kernel"][:24]))
