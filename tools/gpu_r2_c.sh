# round 2, cycle C: tests of the bulk-store build + fast modes; bench of every config x mode; A/B bulk on/off (spec kernels)
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/c_t.log 2>&1; tail -12 gpurun_out/c_t.log
rm -f gpurun_out/c_b.log
for c in cfg2 cfg3 cfg4 cfg5; do for p in fp64 fp32 fp64_fast; do
  python bench.py --workload $c --precision $p --no-cpu-baseline --no-e2e >> gpurun_out/c_b.log 2>&1
done; done
echo "--- bulk off" >> gpurun_out/c_b.log
export TFK_CACHE_DIR=/tmp/tfk_nobulk
for c in cfg2 cfg4 cfg5; do for p in fp64 fp32; do
  TFK_JIT_DEFINES="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1;TFK_BULK_STORE=0" python bench.py --workload $c --precision $p --no-cpu-baseline --no-e2e >> gpurun_out/c_b.log 2>&1
done; done
python - <<PY
import json
for l in open("gpurun_out/c_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"]["workload"][:12], d["config"]["mode"][:9], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:40], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    else: print(l[:300])
PY
