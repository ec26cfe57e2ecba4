# A/B: FP64 fast mode with the portable libm (explicit fma) against libdevice
mkdir -p gpurun_out; rm -f gpurun_out/n_ab.log
D="TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1"
for w in cfg2 cfg3 cfg4 cfg5; do
  for v in A B; do
    if [ $v = A ]; then export TFK_JIT_DEFINES="$D"; else export TFK_JIT_DEFINES="$D;TFK_FAST_PORTABLE_LIBM=1"; fi
    python bench.py --workload $w --precision fp64_fast --no-cpu-baseline --no-e2e --no-workloads 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$w $v', '%.4f ms' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'], d['config'].get('kernel', ''))
" >> gpurun_out/n_ab.log
  done
done
cat gpurun_out/n_ab.log
