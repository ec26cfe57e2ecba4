mkdir -p gpurun_out
export TFK_CACHE_DIR=/tmp/tfk_cache_ab
D="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
( TFK_JIT_DEFINES="$D" python -m pytest tests/test_jit.py tests/test_variants.py -m gpu -x -q ) 2>&1 | tail -4
: > gpurun_out/ab.log
for w in cfg2 cfg3; do
  python bench.py --workload $w --no-cpu-baseline --no-e2e >> gpurun_out/ab.log 2>&1
  TFK_JIT_DEFINES="$D" python bench.py --workload $w --no-cpu-baseline --no-e2e >> gpurun_out/ab.log 2>&1
done
python - <<PY
import json
for l in open("gpurun_out/ab.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["jit"])
    else: print(l[:300])
PY
