# A/B of TFK_JIT_DEFINES over every workload / mode (bench.py's workloads[])
mkdir -p gpurun_out; rm -f gpurun_out/n_ab.log
for v in "TFK_SPEC_INLINE=1;TFK_SPEC_SINCOS_CALL=1;TFK_SPEC_SINCOS_PAIR=1" "TFK_SPEC_INLINE=1"; do
  export TFK_JIT_DEFINES="$v"
  echo "== $v" >> gpurun_out/n_ab.log
  python bench.py --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['config']['workload'][:30], d['dtype'], '%.4f ms' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'])
        for w in d.get('workloads', []): print('   ', w['workload'][:40], w['dtype'], w.get('precision',''), '%.4f ms' % w['ms_per_step'], 'frac %.3f' % w['roofline_frac'], w['kernel'][:18])
" >> gpurun_out/n_ab.log
done
cat gpurun_out/n_ab.log
