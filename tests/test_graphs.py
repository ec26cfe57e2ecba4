"""The device entry is asynchronous and, once a signature has been used, touches neither the allocator nor the host:
it can be captured into a CUDA graph (what a planner calling small batches in a tight loop wants)."""
import numpy as np
import pytest

from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W


@pytest.mark.gpu
@pytest.mark.parametrize("multi", [False, True])
def test_gpu_device_entry_is_graph_capturable(multi):
    import torch
    m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
    m.set_jit_mode(0)
    j = list(m.get_joint_ids(W.PR2_RARM))
    l = list(m.get_link_ids(["r_gripper_tool_frame"] + (["r_elbow_flex_link", "r_wrist_roll_link"] if multi else [])))
    n, K, L = 256, 6, len(l)
    dev = torch.device("cuda", 0)
    qs = [torch.from_numpy(W.sample_q(m.get_joint_limits(j), 1, n, seed=k)).to(dev) for k in range(K)]
    P = [torch.zeros((6, n * L), dtype=torch.float64, device=dev) for _ in range(K)]
    J = [torch.zeros((10, n * L * 6), dtype=torch.float64, device=dev) for _ in range(K)]

    def run(stream, k):
        m.solve_fk_device(0, stream.cuda_stream, qs[k].data_ptr(), n, l, j, 1, 1, True, 0, P[k].data_ptr(), J[k].data_ptr(), n, 0)

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for k in range(K):  # first use of the signature: plan upload, occupancy query, scratch allocation
            run(s, k)
    s.synchronize()
    ref = [(p.clone(), jj.clone()) for p, jj in zip(P, J)]
    for p, jj in zip(P, J):
        p.zero_()
        jj.zero_()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):  # the stream the signature was warmed up on: its scratch records exist already
        cs = torch.cuda.current_stream()
        for k in range(K):
            run(cs, k)
    assert not P[0].any()  # capture does not execute
    g.replay()
    torch.cuda.synchronize()
    for k in range(K):
        assert torch.equal(P[k], ref[k][0]) and torch.equal(J[k], ref[k][1])
    # new inputs, same graph
    qs[0].copy_(qs[1])
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(P[0], ref[1][0]) and torch.equal(J[0], ref[1][1])
