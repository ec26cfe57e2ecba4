import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from tinyfk_b200 import capi
from tinyfk_b200 import workloads as W
L = ctypes.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), sys.argv[1] if len(sys.argv) > 1 else "libspec_cfg2.so"))
L.spec_launch.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
n = 1 << 20
dev = torch.device("cuda", 0)
m = capi.Model(W.urdf_text("pr2.urdf"), host_libm=0)
jids, lids = list(m.get_joint_ids(W.PR2_RARM)), list(m.get_link_ids(["r_gripper_tool_frame"]))
q = [torch.from_numpy(W.sample_q(m.get_joint_limits(jids), 1, n, seed=r)).to(dev) for r in range(2)]
P0 = torch.empty((6, n), dtype=torch.float64, device=dev); J0 = torch.empty((10, n * 6), dtype=torch.float64, device=dev)
P1 = torch.zeros_like(P0); J1 = torch.zeros_like(J0)
st = torch.cuda.current_stream()
def gen(i): m.solve_fk_device(0, st.cuda_stream, q[i % 2].data_ptr(), n, lids, jids, 1, 1, True, 0, P0.data_ptr(), J0.data_ptr(), n, 0)
tile = L.spec_tile()
for grid in (148 * 512 // tile,):
    def spec(i):
        rc = L.spec_launch(q[i % 2].data_ptr(), n, P1.data_ptr(), J1.data_ptr(), grid, st.cuda_stream)
        assert rc == 0, rc
    gen(0); spec(0); torch.cuda.synchronize()
    print("grid", grid, "bitwise equal:", torch.equal(P0, P1), torch.equal(J0, J1), "smem", L.spec_smem())
    for name, f in (("generic", gen), ("spec", spec)):
        for i in range(3): f(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for i in range(20): f(i)
        e1.record(st); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print("%-8s %.4f ms  %.3e cfg/s" % (name, ms, n / ms * 1e3))
