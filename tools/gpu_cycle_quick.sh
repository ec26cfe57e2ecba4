# quick A/B cycle: parity tests of the strict mode + every workload's bench line
mkdir -p gpurun_out; rm -f gpurun_out/q_*.log
( time python -m pytest tests/test_gpu_parity.py tests/test_jit.py tests/test_reference_goldens_gpu.py tests/test_fkmath.py -m gpu -x -q ) > gpurun_out/q_t.log 2>&1; tail -4 gpurun_out/q_t.log
( time python bench.py --no-cpu-baseline ) > gpurun_out/q_b.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/q_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d.get("roofline",{}).get("frac",0), "e2e", d.get("e2e",{}).get("value"))
        for w in d.get("workloads", []): print("   ", w["workload"][:58], w["dtype"], "%.3e"%w["value"], "%.4f ms"%w["ms_per_step"], "frac %.3f"%w["roofline_frac"], w["kernel"][:22])
    elif l.startswith("real"): print(l.strip())
PY
