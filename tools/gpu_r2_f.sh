# cycle F: the allocation-free entry -- alone in a fresh process, then after every other GPU test (pytest -x, as the round's end)
mkdir -p gpurun_out
( time python -m pytest tests/test_zz_no_alloc.py -x -q ) > gpurun_out/f_alone.log 2>&1; tail -25 gpurun_out/f_alone.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/f_all.log 2>&1; tail -8 gpurun_out/f_all.log
