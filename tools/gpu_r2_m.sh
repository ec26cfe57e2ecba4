# cycle M: per-warp staging of the joint records in phase B of the two-phase kernel (cfg5-type requests): parity, then A/B
mkdir -p gpurun_out; rm -f gpurun_out/m_*.log
( time python -m pytest tests/test_gpu_parity.py tests/test_jit.py tests/test_guards_gpu.py tests/test_fp64_fast.py tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/m_t.log 2>&1; tail -4 gpurun_out/m_t.log
for prec in fp64 fp64_fast; do
  for ns in 1 0; do
    echo "== cfg5 $prec no_stage=$ns" >> gpurun_out/m_b.log
    if [ $ns = 1 ]; then export TFK_NO_STAGE=1; else unset TFK_NO_STAGE; fi
    python bench.py --workload cfg5 --precision $prec --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/m_b.log 2>&1
  done
done
unset TFK_NO_STAGE
python - <<PY
import json
for l in open("gpurun_out/m_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"], d["clocks"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
