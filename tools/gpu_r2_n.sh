mkdir -p gpurun_out; rm -f gpurun_out/n_ab.log
run() { # name workload precision
  python bench.py --workload $2 --precision $3 --no-cpu-baseline --no-e2e --no-workloads 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$2 $3 $1', '%.4f ms' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'], d['roofline'].get('kernel', d['config'].get('kernel','')))
" >> gpurun_out/n_ab.log
}
run base cfg5 fp64
TFK_SPEC_STRICT_ORIENT=1 run spec cfg5 fp64
TFK_SPEC_STRICT_ORIENT=1 TFK_TWOPHASE_TILE_CFGS=64 run spec64 cfg5 fp64
TFK_TWOPHASE_TILE_CFGS=64 run base64 cfg5 fp64
cat gpurun_out/n_ab.log
