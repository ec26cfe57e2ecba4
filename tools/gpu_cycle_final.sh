# full verification + evidence cycle of a build (run under gpurun from the repo root); outputs in gpurun_out/
mkdir -p gpurun_out; rm -f gpurun_out/f_*.log gpurun_out/f_*.txt gpurun_out/f_*.csv
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/f_t.log 2>&1; tail -4 gpurun_out/f_t.log
( time python -m pytest tests/test_zz_no_alloc.py -x -q ) > gpurun_out/f_noalloc_alone.log 2>&1; tail -3 gpurun_out/f_noalloc_alone.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/f_smoke.log; tail -3 gpurun_out/f_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f_bref.log 2>&1
( time python bench.py ) > gpurun_out/f_b.log 2>&1
python bench.py --jit off --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/f_b.log 2>&1
python - <<PY
import json
for f in ("gpurun_out/f_bref.log", "gpurun_out/f_b.log"):
  for l in open(f):
    if l.startswith("{"):
        d=json.loads(l); print(d.get("impl","ours"), d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d.get("roofline",{}).get("frac",0), "traffic", d.get("roofline",{}).get("traffic"), "e2e", d.get("e2e",{}).get("value"), "cpu", d.get("cpu_baseline",{}).get("value"), d.get("clocks",{}).get("samples"))
        for w in d.get("workloads", []): print("   ", w["workload"][:58], w["dtype"], "%.3e"%w["value"], "%.4f ms"%w["ms_per_step"], "frac %.3f"%w["roofline_frac"], w["kernel"][:22])
    elif l.startswith("real"): print(l.strip())
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/f_launches_cfg2.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-workloads > gpurun_out/f_ncu_l.log 2>&1
python tools/ncu_capture.py cfg2/fp64 cfg2/fp32 cfg2/fp64_fast cfg3/fp64 cfg3/fp32 cfg3/fp64_fast cfg4/fp64 cfg4/fp32 cfg4/fp64_fast cfg5/fp64 cfg5/fp32 cfg5/fp64_fast > gpurun_out/f_ncu.log 2>&1; tail -12 gpurun_out/f_ncu.log | cut -c1-160
for c in cfg2_fp64 cfg2_fp32 cfg2_fp64_fast cfg3_fp64 cfg3_fp32 cfg3_fp64_fast cfg4_fp64 cfg4_fp32 cfg4_fp64_fast cfg5_fp64 cfg5_fp32 cfg5_fp64_fast; do
  n=1048576; case $c in cfg3*) n=2097152;; cfg4*) n=524288;; cfg5*) n=65536;; esac
  python tools/ncu_lines.py gpurun_out/r2_$c.ncu-rep $n 45 > gpurun_out/f_lines_$c.txt 2>&1
done
rm -f gpurun_out/*.ncu-rep
