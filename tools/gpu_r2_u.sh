# cycle U: software prefetch knobs of the FP64-fast several-link kernels (TFK_FAST64_SPARE: bit 0 link record, bit 1 staging)
mkdir -p gpurun_out; rm -f gpurun_out/u_*.log
( time python -m pytest tests/test_jit.py tests/test_gpu_parity.py tests/test_guards_gpu.py tests/test_fp64_fast.py -m gpu -q ) > gpurun_out/u_t.log 2>&1; tail -3 gpurun_out/u_t.log
BASE="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
for wl in cfg5 cfg4; do for sp in 0 1 2 3; do
  echo "== $wl fp64_fast spare=$sp" >> gpurun_out/u_b.log
  TFK_JIT_DEFINES="$BASE;TFK_FAST64_SPARE=$sp" python bench.py --workload $wl --precision fp64_fast --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/u_b.log 2>&1
done; done
for wl in cfg5 cfg4; do
  echo "== $wl fp32" >> gpurun_out/u_b.log
  python bench.py --workload $wl --precision fp32 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/u_b.log 2>&1
done
python - <<PY
import json
for l in open("gpurun_out/u_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
