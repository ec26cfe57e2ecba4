# cycle D: store-path matrix for the plan-specialised kernels: per-lane stores vs bulk stores with 1/2/3 strips per warp
mkdir -p gpurun_out; rm -f gpurun_out/d_b.log
for cfgp in "cfg2 fp32" "cfg2 fp64_fast" "cfg2 fp64" "cfg5 fp32" "cfg5 fp64_fast" "cfg4 fp32" "cfg4 fp64" "cfg3 fp32"; do set -- $cfgp
  for mode in "0 1" "1 1" "1 2" "1 3"; do set -- $cfgp $mode
    echo "== $1 $2 bulk=$3 nbuf=$4" >> gpurun_out/d_b.log
    TFK_BULK=$3 TFK_STG_NBUF=$4 python bench.py --workload $1 --precision $2 --no-cpu-baseline --no-e2e >> gpurun_out/d_b.log 2>&1
  done
done
python - <<PY
import json
for l in open("gpurun_out/d_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    else: print(l[:200].rstrip())
PY
