# cycle N: cfg5 FP32 with the joint records staged per warp instead of resident in shared memory (more blocks per SM);
# ncu of the staged cfg5 strict kernel
mkdir -p gpurun_out; rm -f gpurun_out/n_*.log
for hot in 0 1; do
  echo "== cfg5 fp32 no_smem_hot=$hot" >> gpurun_out/n_b.log
  if [ $hot = 1 ]; then export TFK_NO_SMEM_HOT=1; else unset TFK_NO_SMEM_HOT; fi
  python bench.py --workload cfg5 --precision fp32 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/n_b.log 2>&1
  for tc in 64 128; do
  echo "== cfg5 fp32 no_smem_hot=$hot tile=$tc" >> gpurun_out/n_b.log
  TFK_TWOPHASE_TILE_CFGS=$tc python bench.py --workload cfg5 --precision fp32 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/n_b.log 2>&1
  done
done
unset TFK_NO_SMEM_HOT
python - <<PY
import json
for l in open("gpurun_out/n_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
python tools/ncu_capture.py cfg5/fp64 > gpurun_out/n_ncu.log 2>&1; tail -2 gpurun_out/n_ncu.log | cut -c1-200
python tools/ncu_lines.py gpurun_out/r2_cfg5_fp64.ncu-rep 65536 45 > gpurun_out/n_lines_cfg5_fp64.txt 2>&1
rm -f gpurun_out/*.ncu-rep
