# cycle P: cfg5 strict through the specialised two-phase kernel now that phase B stages the joint records; FP64-fast with the
# rolled walk (new default)
mkdir -p gpurun_out; rm -f gpurun_out/p_*.log
( time python -m pytest tests/test_jit.py tests/test_fp64_fast.py -m gpu -x -q ) > gpurun_out/p_t.log 2>&1; tail -3 gpurun_out/p_t.log
for so in 0 1; do
  echo "== cfg5 fp64 spec_strict_orient=$so" >> gpurun_out/p_b.log
  if [ $so = 1 ]; then export TFK_SPEC_STRICT_ORIENT=1; else unset TFK_SPEC_STRICT_ORIENT; fi
  python bench.py --workload cfg5 --precision fp64 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/p_b.log 2>&1
done
unset TFK_SPEC_STRICT_ORIENT
for wl in cfg4 cfg5; do
  echo "== $wl fp64_fast" >> gpurun_out/p_b.log
  python bench.py --workload $wl --precision fp64_fast --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/p_b.log 2>&1
done
python - <<PY
import json
for l in open("gpurun_out/p_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"], d["config"]["jit"][:60])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
