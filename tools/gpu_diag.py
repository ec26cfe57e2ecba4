"""GPU-box diagnostics (not a test): FP32 error anatomy, PCIe bandwidth, e2e breakdown."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import parity
from parity import Pair
from tinyfk_b200 import capi, workloads as W

def fp32_anatomy():
    for cfg in ("cfg2", "cfg3", "cfg4", "cfg5"):
        pair = Pair(W.CONFIGS[cfg][0], strict=False)
        jids, lids, rot, base = pair.setup(cfg)
        n = 4096 if cfg != "cfg5" else 256
        q = W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=7).astype(np.float32)
        Pr, Jr = pair.oracle_solve(q.astype(np.float64), lids, jids, rot, base, True)
        P, J = pair.prod_solve(q, lids, jids, rot, base, True, precision=1)
        pd = W.pose_dim(rot)
        eP = parity.rel_err(P, Pr); eJ = parity.rel_err(J, Jr).reshape(-1, pd, J.shape[1])
        print(cfg, "P pos err %.2e" % eP[:, :3].max(), "P rot err %.2e" % (eP[:, 3:].max() if pd > 3 else 0),
              "J pos rows %.2e" % eJ[:, :3].max(), "J rot rows %.2e" % (eJ[:, 3:].max() if pd > 3 else 0))
        if rot == 1:
            cp = np.abs(np.cos(Pr[:, 4]))
            for thr in (0.0, 0.05, 0.1, 0.2, 0.5):
                sel = cp >= thr
                print("   |cos(pitch)|>=%.2f: n=%d P rot %.2e  J rot rows %.2e  J pos rows %.2e" % (thr, sel.sum(), eP[sel, 3:].max(), eJ[sel, 3:].max(), eJ[sel, :3].max()))
            fd = parity.fd_column_mask(len(jids), base, J.shape[1])
            print("   FD cols err %.2e, analytic cols err %.2e" % (eJ[:, :, fd].max(), eJ[:, :, ~fd].max()))

def pcie():
    for mb in (1, 8, 64, 512):
        n = mb * (1 << 20)
        h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
        for name, src, dst in (("H2D", h, d), ("D2H", d, h)):
            dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5): dst.copy_(src, non_blocking=True)
            torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
            print("pcie %s %4d MB: %.1f GB/s" % (name, mb, n / dt / 1e9))

def e2e_breakdown():
    pair = Pair("pr2.urdf", strict=True)
    jids, lids, rot, base = pair.setup("cfg2")
    n = 1 << 20
    q = W.sample_q(pair.prod.get_joint_limits(jids), base, n, seed=1)
    pq = capi.PinnedArray((n, 10), order="C"); pq.array[...] = q
    Pp = capi.PinnedArray((n, 6)); Jp = capi.PinnedArray((n * 6, 10))
    for label, qa, out in (("pinned in/out", pq.array, (Pp.array, Jp.array)), ("pageable in/out", q, None)):
        pair.prod.solve_fk(qa, lids, jids, rot, base, True, out=out)
        t0 = time.perf_counter()
        for _ in range(5): pair.prod.solve_fk(qa, lids, jids, rot, base, True, out=out)
        dt = (time.perf_counter() - t0) / 5
        print("e2e C-ABI host path, %s: %.2f ms per 1M configs = %.3e cfg/s (%.1f GB/s)" % (label, dt * 1e3, n / dt, n * 608 / dt / 1e9))
    t0 = time.perf_counter()
    for _ in range(5): pair.prod.solve_fk(pq.array, lids, jids, rot, base, False, out=(Pp.array, Jp.array))
    dt = (time.perf_counter() - t0) / 5
    print("pose only pinned: %.2f ms" % (dt * 1e3))

if __name__ == "__main__":
    which = sys.argv[1:] or ["fp32", "pcie", "e2e"]
    if "fp32" in which: fp32_anatomy()
    if "pcie" in which: pcie()
    if "e2e" in which: e2e_breakdown()
