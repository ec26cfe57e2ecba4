# cycle H: why the two-pass build does not keep the first pass's speed: unit tests, then ncu of the cfg2 strict kernel
mkdir -p gpurun_out; rm -f gpurun_out/h_*.log
( time python -m pytest tests/test_flag_math_gpu.py -m gpu -q ) > gpurun_out/h_t.log 2>&1; tail -5 gpurun_out/h_t.log
python bench.py --workload cfg2 --precision fp64 --steps 50 --no-cpu-baseline --no-e2e --no-workloads > gpurun_out/h_b.log 2>&1; tail -c 600 gpurun_out/h_b.log
python tools/ncu_capture.py cfg2/fp64 > gpurun_out/h_ncu.log 2>&1; tail -3 gpurun_out/h_ncu.log
python tools/ncu_lines.py gpurun_out/r2_cfg2_fp64.ncu-rep 1048576 60 > gpurun_out/h_lines.txt 2>&1; head -50 gpurun_out/h_lines.txt
