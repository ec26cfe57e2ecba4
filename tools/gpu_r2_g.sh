# cycle G: the branch-free first pass of the specialised strict kernels (fkmath_flag.h): unit tests of the arithmetic,
# kernel parity incl. rare rows, then A/B against the one-pass build and over block shapes
mkdir -p gpurun_out; rm -f gpurun_out/g_*.log
( time python -m pytest tests/test_flag_math_gpu.py tests/test_jit.py -m gpu -x -q ) > gpurun_out/g_t.log 2>&1; tail -15 gpurun_out/g_t.log
OLD="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
NEW="$OLD;TFK_SPEC_FLAG=1"
run() { # label, defines, shape, workload
  echo "== $1 $4 shape=${3:-default}" >> gpurun_out/g_b.log
  if [ -n "$3" ]; then export TFK_SPEC_SHAPE=$3; else unset TFK_SPEC_SHAPE; fi
  TFK_JIT_DEFINES="$2" python bench.py --workload $4 --precision fp64 --steps 50 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/g_b.log 2>&1
}
for wl in cfg2 cfg3; do
  run one-pass "$OLD" "" $wl
  run flag "$NEW" "" $wl
  run flag "$NEW" "128,4" $wl
  run flag "$NEW" "192,2" $wl
  run flag "$NEW" "384,1" $wl
  run flag-noinl "TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1;TFK_SPEC_FLAG=1" "" $wl
  run flag-nopair "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_FLAG=1" "" $wl
done
unset TFK_SPEC_SHAPE
python - <<PY
import json
for l in open("gpurun_out/g_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["block"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
