"""diagnostic (GPU): which warps of the specialised strict kernel are redone, and which line of fk_math.h reported
(build with TFK_JIT_DEFINES=...;TFK_SPEC_FLAG_MARK=1: the first P entry of a redone warp holds the line)"""
import collections, os, sys
os.environ["TFK_JIT_DEFINES"] = "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1;TFK_SPEC_FLAG=1;TFK_SPEC_FLAG_MARK=1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tinyfk_b200 import capi, workloads as W
for cfg in ("cfg2", "cfg3"):
    m = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0)
    j, l, rot, base = W.setup(m, cfg)
    n = 1 << 18
    q = W.sample_q(m.get_joint_limits(j), base, n, seed=0)
    m.set_jit_mode(1)
    m.specialize(l, j, rot, base, True, 0, wait=True)
    P, J = m.solve_fk(q, l, j, rot, base, True)
    print(cfg, m.last_launch_info()["kernel"])
    first = P[::32, 0]
    marked = first[(first > 100) & (first == np.round(first))]
    print("  warps:", n // 32, "redone:", marked.size, collections.Counter(marked.astype(int).tolist()).most_common(8))
