# cycle Q: inlining knobs of the specialised strict fused kernel (cfg2, cfg3), one-pass build
mkdir -p gpurun_out; rm -f gpurun_out/q_*.log
for wl in cfg2 cfg3; do
for d in "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1" "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_PAIR=1" "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1" "TFK_SPEC_INLINE=1" "TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"; do
  echo "== $wl $d" >> gpurun_out/q_b.log
  TFK_JIT_DEFINES="$d" python bench.py --workload $wl --precision fp64 --steps 50 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/q_b.log 2>&1
done
done
python - <<PY
import json
for l in open("gpurun_out/q_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["block"], d["config"]["launch"]["blocks_per_sm"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
