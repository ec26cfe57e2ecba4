mkdir -p gpurun_out
( time python -m pytest tests/test_pos_kernel.py tests/test_gpu_parity.py -m gpu -x -q ) 2>&1 | tail -8 > gpurun_out/t.log
cat gpurun_out/t.log
: > gpurun_out/b.log
TFK_POS_KERNEL=1 python bench.py --workload cfg4 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --workload cfg4 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
TFK_POS_KERNEL=1 python bench.py --workload cfg4 --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --workload cfg4 --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"])
    else: print(l[:300])
PY
TFK_POS_KERNEL=1 ncu --set full --clock-control none --import-source on -k regex:fk_pos -s 3 -c 1 -f -o gpurun_out/prof_cfg4_pos python bench.py --workload cfg4 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu4.log
