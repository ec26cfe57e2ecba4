# full verification + evidence cycle of a build (run under gpurun from the repo root); outputs in gpurun_out/
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) 2>&1 | tail -8 > gpurun_out/t.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bref.log 2>&1
python bench.py > gpurun_out/b.log 2>&1
for c in cfg3 cfg4 cfg5; do python bench.py --workload $c --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1; done
python bench.py --precision fp32 --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
python bench.py --jit off --no-cpu-baseline --no-e2e >> gpurun_out/b.log 2>&1
cat gpurun_out/t.log; tail -3 gpurun_out/smoke.log
python - <<PY
import json
for f in ("gpurun_out/bref.log", "gpurun_out/b.log"):
  for l in open(f):
    if l.startswith("{"):
        d=json.loads(l); print(d.get("impl","ours"), d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d.get("roofline",{}).get("frac",0), d["config"].get("launch"), "e2e", d.get("e2e",{}).get("value"), "cpu", d.get("cpu_baseline",{}).get("value"), d.get("clocks"), d.get("roofline",{}).get("fp64_pipe"))
    else: print(l[:300])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_cfg2.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tfk_spec -s 3 -c 1 -f -o gpurun_out/prof_cfg2_final python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -30
ncu --set full --clock-control none --import-source on -k regex:tfk_spec -s 3 -c 1 -f -o gpurun_out/prof_cfg4_final python bench.py --workload cfg4 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu4.log 2>&1
