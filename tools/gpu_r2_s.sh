# cycle S: the generic two-phase kernels on the pair layout: full GPU suite, then A/B against a library built with
# -DTFK_GENERIC_PLAIN_RECORDS (tinyfk_b200/csrc/build/alt), jit off = generic kernels
mkdir -p gpurun_out; rm -f gpurun_out/s_*.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/s_t.log 2>&1; tail -4 gpurun_out/s_t.log
ALT=$PWD/tinyfk_b200/csrc/build/alt/libtinyfk_b200_plain.so
for cfgp in "cfg5 fp64 on" "cfg5 fp64 off" "cfg4 fp64 off" "cfg5 fp32 off" "cfg4 fp32 off"; do set -- $cfgp
  for lib in plain pairs; do
    echo "== $1 $2 jit=$3 generic records=$lib" >> gpurun_out/s_b.log
    if [ $lib = plain ]; then export TFK_B200_LIB=$ALT; else unset TFK_B200_LIB; fi
    python bench.py --workload $1 --precision $2 --jit $3 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/s_b.log 2>&1
  done
done
unset TFK_B200_LIB
python - <<PY
import json
for l in open("gpurun_out/s_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"], d["config"]["launch"]["smem_bytes"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
