# cycle J: guard-region tests, the tile-size policy on every two-phase workload, ncu --set full of the dominant kernel of every
# workload / mode of this build (-> gpurun_out/ncu_traffic.json + one summary per capture)
mkdir -p gpurun_out; rm -f gpurun_out/j_*.log
( time python -m pytest tests/test_guards_gpu.py tests/test_gpu_parity.py tests/test_jit.py -m gpu -x -q ) > gpurun_out/j_t.log 2>&1; tail -4 gpurun_out/j_t.log
for cfgp in "cfg4 fp64" "cfg4 fp32" "cfg4 fp64_fast" "cfg5 fp64" "cfg5 fp32" "cfg5 fp64_fast" "cfg2 fp32" "cfg2 fp64_fast" "cfg3 fp32" "cfg3 fp64_fast"; do set -- $cfgp
  echo "== $1 $2" >> gpurun_out/j_b.log
  python bench.py --workload $1 --precision $2 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/j_b.log 2>&1
done
python - <<PY
import json
for l in open("gpurun_out/j_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
python tools/ncu_capture.py cfg2/fp64 cfg2/fp32 cfg2/fp64_fast cfg3/fp64 cfg4/fp64 cfg4/fp32 cfg5/fp64 cfg5/fp32 cfg5/fp64_fast > gpurun_out/j_ncu.log 2>&1; tail -12 gpurun_out/j_ncu.log | cut -c1-220
for c in cfg2_fp64 cfg2_fp32 cfg2_fp64_fast cfg3_fp64 cfg4_fp64 cfg4_fp32 cfg5_fp64 cfg5_fp32 cfg5_fp64_fast; do
  n=1048576; case $c in cfg3*) n=2097152;; cfg4*) n=524288;; cfg5*) n=65536;; esac
  python tools/ncu_lines.py gpurun_out/r2_$c.ncu-rep $n 45 > gpurun_out/j_lines_$c.txt 2>&1
done
rm -f gpurun_out/*.ncu-rep
