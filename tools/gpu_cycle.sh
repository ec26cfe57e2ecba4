python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/t.log
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/b.log 2>&1
for c in cfg3 cfg4 cfg5; do python bench.py --workload $c --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1; done
python bench.py --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
cat gpurun_out/t.log
python - <<PY
import json
for l in open("gpurun_out/b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"])
    else: print(l[:300])
PY
python tools/time_sqdists.py > gpurun_out/sq.log 2>&1; tail -40 gpurun_out/sq.log
