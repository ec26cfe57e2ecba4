"""Out-of-bounds and determinism checks of every FK kernel family with guard regions of our own (compute-sanitizer is
closed on this GPU pool): P and J live in the middle of one device buffer filled with a sentinel bit pattern; after a launch
through the device entry of the C ABI the guards before, between and after them are untouched, rows outside
[row0, row0 + n) of arrays dimensioned for ld_cfg > n are untouched, every entry inside was written, and a second launch
reproduces the first bit for bit (the staged sink's warp-private strips are ordered by __syncwarp() pairs: a missing one
shows up as run-to-run differences).  Batch sizes leave a ragged tail tile and a ragged last warp."""
import numpy as np
import pytest

from parity import Pair
from tinyfk_b200 import workloads as W

pytestmark = pytest.mark.gpu
SENT = np.array([0x7FF8DEADBEEF1234], dtype=np.uint64)
GUARD = 8192  # reals

CASES = [  # cfg, links kept (None = all), rot override, base override, precision, jit, n
    ("cfg2", None, None, None, 0, 0, 1000 + 13),     # fused, generic
    ("cfg2", None, None, None, 0, 2, 1000 + 13),     # fused, plan-specialised (256-thread blocks)
    ("cfg2", None, None, 2, 1, 2, 777),              # FP32, FLOATING
    ("cfg3", None, None, None, 2, 2, 515),           # FP64 fast, XYZW
    ("cfg4", None, None, None, 0, 0, 300 + 5),       # two-phase generic, put_J2 (position rows)
    ("cfg4", None, None, None, 0, 2, 300 + 5),       # two-phase specialised, records in shared memory
    ("cfg4", 3, 2, 1, 1, 2, 148 * 5 * 32 + 32 * 9 + 7),  # FP32: 32-configuration tiles, more than one wave
    ("cfg5", 7, None, None, 0, 0, 130),              # RPY + FLOATING, generic, records in the L2 scratch
    ("cfg5", 7, None, None, 1, 2, 130),
    ("cfg5", None, None, None, 0, 0, 128 + 67),      # 50 links: joint records staged per warp (generic, strict)
    ("cfg5", None, None, None, 2, 2, 128 + 67),      # the same through the specialised FP64-fast kernel
    ("cfg5", 40, 2, 1, 1, 2, 100),                   # FP32, 40 links, XYZW, PLANER: records resident, link-record prefetch
]


@pytest.mark.parametrize("cfg,keep,rot_o,base_o,prec,jit,n", CASES)
def test_no_write_outside_the_outputs_and_runs_repeat(cfg, keep, rot_o, base_o, prec, jit, n):
    import torch

    pair = Pair(W.CONFIGS[cfg][0], strict=True)
    jids, lids, rot, base = pair.setup(cfg)
    lids = lids if keep is None else lids[:keep]
    rot = rot if rot_o is None else rot_o
    base = base if base_o is None else base_o
    m = pair.prod
    m.set_jit_mode(jit)
    lim = m.get_joint_limits(jids)
    q = W.sample_q(lim, base, n, seed=3, wide_pitch=True)
    dt, npdt = (torch.float32, np.float32) if prec == 1 else (torch.float64, np.float64)
    isz = 4 if prec == 1 else 8
    sent = SENT.view(np.float64)[0] if isz == 8 else SENT.view(np.float32)[0]
    nl, pd, nd = len(lids), W.pose_dim(rot), q.shape[1]
    ld, row0 = n + 96, 40          # arrays dimensioned for more rows than this launch fills
    szP, szJ = pd * ld * nl, nd * ld * nl * pd
    total = GUARD + szP + GUARD + szJ + GUARD
    host = np.full(total, sent, dtype=npdt)
    dq = torch.from_numpy(q.astype(npdt)).cuda()
    st = torch.cuda.current_stream().cuda_stream
    runs = []
    for rep in range(2):
        buf = torch.from_numpy(host.copy()).cuda()
        pP = buf.data_ptr() + GUARD * isz
        pJ = buf.data_ptr() + (GUARD + szP + GUARD) * isz
        m.solve_fk_device(0, st, dq.data_ptr(), n, lids, jids, rot, base, True, prec, pP, pJ, ld, row0)
        torch.cuda.synchronize()
        runs.append(buf.cpu().numpy())
    out = runs[0]
    bits = out.view(np.uint64 if isz == 8 else np.uint32)
    sbits = host.view(bits.dtype)[0]
    assert np.array_equal(runs[0].view(bits.dtype), runs[1].view(bits.dtype)), "two launches differ"
    for a, b in ((0, GUARD), (GUARD + szP, GUARD + szP + GUARD), (total - GUARD, total)):
        assert (bits[a:b] == sbits).all(), "guard region [%d, %d) was written" % (a, b)
    P = bits[GUARD:GUARD + szP].reshape(pd, ld, nl)
    J = bits[GUARD + szP + GUARD:total - GUARD].reshape(nd, ld, nl * pd)
    inside = np.zeros(ld, dtype=bool)
    inside[row0:row0 + n] = True
    assert (P[:, ~inside] == sbits).all() and (J[:, ~inside] == sbits).all(), "rows outside [row0, row0 + n) were written"
    assert not (P[:, inside] == sbits).any() and not (J[:, inside] == sbits).any(), "an output entry was not written"
    m.set_jit_mode(1)
