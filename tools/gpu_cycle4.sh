mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) 2>&1 | tail -8 > gpurun_out/t.log
cat gpurun_out/t.log
: > gpurun_out/b.log
python bench.py --workload cfg4 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
TFK_POS_KERNEL=1 python bench.py --workload cfg4 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --workload cfg5 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --workload cfg4 --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"])
    else: print(l[:300])
PY
python tools/time_sqdists.py > gpurun_out/sq.log 2>&1; tail -12 gpurun_out/sq.log
python tools/time_com.py > gpurun_out/com.log 2>&1; tail -12 gpurun_out/com.log
