# cycle R: pair layout of the records in the plan-specialised two-phase kernels (TFK_REC_PAIRS): parity, then A/B
mkdir -p gpurun_out; rm -f gpurun_out/r_*.log
( time python -m pytest tests/test_jit.py tests/test_gpu_parity.py tests/test_guards_gpu.py tests/test_fp64_fast.py tests/test_reference_goldens_gpu.py -m gpu -x -q ) > gpurun_out/r_t.log 2>&1; tail -4 gpurun_out/r_t.log
for cfgp in "cfg4 fp64" "cfg4 fp32" "cfg4 fp64_fast" "cfg5 fp32" "cfg5 fp64_fast"; do set -- $cfgp
  for mode in off on; do
    echo "== $1 $2 pairs=$mode" >> gpurun_out/r_b.log
    if [ $mode = off ]; then export TFK_NO_REC_PAIRS=1; else unset TFK_NO_REC_PAIRS; fi
    python bench.py --workload $1 --precision $2 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/r_b.log 2>&1
  done
done
unset TFK_NO_REC_PAIRS
python - <<PY
import json
for l in open("gpurun_out/r_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
