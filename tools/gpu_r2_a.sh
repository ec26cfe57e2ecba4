# round 2, cycle A: numbers of the inherited build + the FP32 ncu captures round 1 never took
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/a_smi.log
for c in cfg2 cfg3 cfg4 cfg5; do for p in fp64 fp32; do
  python bench.py --workload $c --precision $p --no-cpu-baseline --no-e2e >> gpurun_out/a_b.log 2>&1
done; done
python - <<PY
import json
for l in open("gpurun_out/a_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:40], d["dtype"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"])
    else: print(l[:300])
PY
for c in cfg2 cfg4 cfg5; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:tfk_spec -s 3 -c 1 -f -o gpurun_out/r2a_${c}_fp32 python bench.py --workload $c --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_ncu_${c}.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:twophase -s 3 -c 1 -f -o gpurun_out/r2a_cfg5_fp64 python bench.py --workload cfg5 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_ncu_cfg5_64.log 2>&1
ls -la gpurun_out
