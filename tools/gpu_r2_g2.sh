mkdir -p gpurun_out; rm -f gpurun_out/g2_*.log
( time python -m pytest tests/test_flag_math_gpu.py tests/test_jit.py -m gpu -q ) > gpurun_out/g2_t.log 2>&1; tail -40 gpurun_out/g2_t.log
OLD="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
for d in "$OLD" "$OLD;TFK_SPEC_FLAG=1" "$OLD;TFK_SPEC_FLAG_ONLY=1"; do
  for wl in cfg2 cfg3; do
  echo "== $wl $d" >> gpurun_out/g2_b.log
  TFK_JIT_DEFINES="$d" python bench.py --workload $wl --precision fp64 --steps 50 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/g2_b.log 2>&1
  done
done
python - <<PY
import json
for l in open("gpurun_out/g2_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["block"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
