"""diagnostic (GPU): where the branch-free reciprocal / asin of csrc/fkmath_flag.h differ from IEEE / raise the flag"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tinyfk_b200 import capi
rng = np.random.default_rng(1)
N = 1 << 22
e = rng.integers(-500 + 1023, 500 + 1023, N, dtype=np.int64)
for name, m in (("random", rng.integers(0, 1 << 52, N, dtype=np.int64)), ("lo=ffffffff", (rng.integers(0, 1 << 20, N, dtype=np.int64) << 32) | 0xFFFFFFFF),
                ("lo>=fff00000", (rng.integers(0, 1 << 20, N, dtype=np.int64) << 32) | 0xFFF00000 | rng.integers(0, 1 << 20, N, dtype=np.int64)),
                ("lo<00100000", (rng.integers(0, 1 << 20, N, dtype=np.int64) << 32) | rng.integers(0, 1 << 20, N, dtype=np.int64))):
    x = ((e << 52) | m).view(np.float64)
    out, ref, fl = capi.debug_flag_math(0, x)
    ie = 1.0 / x
    bad = (out.view(np.int64) != ie.view(np.int64)) & (fl == 0)
    print(name, "mismatches", bad.sum(), "of", N, " ref(__drcp_rn) mismatches", (ref.view(np.int64) != ie.view(np.int64)).sum())
    for i in np.nonzero(bad)[0][:6]:
        print("   x=%s out=%s ieee=%s" % (hex(x.view(np.int64)[i]), hex(out.view(np.int64)[i]), hex(ie.view(np.int64)[i])))
x = np.concatenate([rng.uniform(-1, 1, N), 1 - 10.0 ** rng.uniform(-17, -1, 100000), np.array([np.nextafter(1.0, 0), 0.99999, 1 - 2.0 ** -30])])
out, ref, fl = capi.debug_flag_math(4, x)
print("asin flagged", fl.sum(), [float.hex(v) for v in x[fl != 0][:10]])
