mkdir -p gpurun_out
( time python -m pytest tests/test_jit.py -m gpu -x -q ) 2>&1 | tail -12 > gpurun_out/t.log
cat gpurun_out/t.log
: > gpurun_out/b.log
for c in cfg4 cfg5; do
  python bench.py --workload $c --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
  python bench.py --workload $c --jit off --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
done
python bench.py --workload cfg4 --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --workload cfg5 --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["jit"][:40], d["config"]["launch"])
    else: print(l[:300])
PY
