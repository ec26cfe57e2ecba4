# cycle O: (1) rolled walk for the FP64-fast specialised two-phase kernels, (2) tile sizes for cfg5 with staged joint records
mkdir -p gpurun_out; rm -f gpurun_out/o_*.log
BASE="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
for wl in cfg4 cfg5; do
  for d in "$BASE" "$BASE;TFK_SPEC_ROLLED_WALK=1"; do
    echo "== $wl fp64_fast $d" >> gpurun_out/o_b.log
    TFK_JIT_DEFINES="$d" python bench.py --workload $wl --precision fp64_fast --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/o_b.log 2>&1
  done
done
for prec in fp64 fp64_fast; do
  for tc in 128 64 32; do
    echo "== cfg5 $prec tile=$tc" >> gpurun_out/o_b.log
    TFK_TWOPHASE_TILE_CFGS=$tc python bench.py --workload cfg5 --precision $prec --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/o_b.log 2>&1
  done
done
python - <<PY
import json
for l in open("gpurun_out/o_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
