# round 2, cycle B: the new tests on one GPU + the D2H ceiling + e2e with/without constant-column fill
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/b_t.log 2>&1; tail -15 gpurun_out/b_t.log
python tools/d2h_bw.py > gpurun_out/b_d2h.log 2>&1; cat gpurun_out/b_d2h.log
python bench.py --no-cpu-baseline > gpurun_out/b_b.log 2>&1
TFK_D2H_ALL=1 python bench.py --no-cpu-baseline >> gpurun_out/b_b.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/b_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "e2e", d.get("e2e"))
    else: print(l[:300])
PY
