# cycle I: full GPU suite on the current build, compute-sanitizer (memcheck, racecheck) over the tests that cover the staged
# sink (put_J, put_J2), half tiles, the ragged tail, every kernel variant and the allocation-free entry; tile-size A/B
mkdir -p gpurun_out; rm -f gpurun_out/i_*.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/i_all.log 2>&1; tail -4 gpurun_out/i_all.log
SUB="tests/test_gpu_parity.py::test_baseline_configs_bitwise_vs_pinned_reference tests/test_gpu_parity.py::test_all_modes_on_a_tree_bitwise tests/test_gpu_parity.py::test_half_tiles_when_the_last_wave_is_short tests/test_gpu_parity.py::test_device_pointer_entry_with_row_offset tests/test_variants.py tests/test_sqdists.py::test_gpu_sqdists_bitwise_vs_pinned_reference tests/test_com.py::test_gpu_com_bitwise_vs_pinned_reference tests/test_inertia.py tests/test_zz_no_alloc.py"
for tool in memcheck racecheck; do
  ( time timeout 900 compute-sanitizer --tool $tool --error-exitcode 99 --print-limit 20 python -m pytest $SUB -m gpu -x -q -p no:cacheprovider ) > gpurun_out/i_$tool.log 2>&1
  echo "$tool rc=$?" >> gpurun_out/i_$tool.log; tail -6 gpurun_out/i_$tool.log
done
for cfgp in "cfg5 fp64" "cfg5 fp32" "cfg5 fp64_fast" "cfg4 fp64" "cfg4 fp32"; do set -- $cfgp
  for tc in 128 64 32; do
    echo "== $1 $2 tile_cfgs=$tc" >> gpurun_out/i_b.log
    TFK_TWOPHASE_TILE_CFGS=$tc python bench.py --workload $1 --precision $2 --steps 30 --no-cpu-baseline --no-e2e --no-workloads >> gpurun_out/i_b.log 2>&1
  done
done
python - <<PY
import json
for l in open("gpurun_out/i_b.log"):
    if l.startswith("{"):
        d=json.loads(l); print("   ", "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], d["config"]["launch"]["kernel"][:30], d["config"]["launch"]["grid"], d["config"]["launch"]["blocks_per_sm"])
    elif l.startswith("==") or "rror" in l: print(l[:200].rstrip())
PY
