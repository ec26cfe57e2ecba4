# multi-GPU cycle (N GPUs, `gpurun --gpus N -- bash tools/gpu_cycle_multi.sh N`): the engine's own device-list split on N real devices, bench.py under torchrun with N ranks (weak + strong
# readings, per-rank e2e, in-process e2e over devices 0..N-1); D2H=1 also measures the D2H ceiling per process layout
NG=${1:-8}
mkdir -p gpurun_out
( time python -m pytest tests/test_multi_device.py -x -q -rs ) > gpurun_out/l_t_${NG}.log 2>&1; tail -6 gpurun_out/l_t_${NG}.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NG --steps 50 --warmup 5 > gpurun_out/l_bN_${NG}.log 2>&1
if [ "${D2H:-0}" = 1 ]; then
  python tools/d2h_bw.py > gpurun_out/l_d2h_${NG}.log 2>&1
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 tools/d2h_bw.py >> gpurun_out/l_d2h_${NG}.log 2>&1
  grep "^{" gpurun_out/l_d2h_${NG}.log
  nvidia-smi topo -m > gpurun_out/l_topo_${NG}.log 2>&1; nproc >> gpurun_out/l_topo_${NG}.log
fi
python - <<PY
import json
for f in ("gpurun_out/l_bN_${NG}.log",):
  for l in open(f):
    if l.startswith("{"):
        d=json.loads(l); print(d.get("impl","ours"), "N", d["n_gpus"], "%.3e"%d["value"], "%.4f ms"%d["ms_per_step"], "frac", d.get("roofline",{}).get("frac"), "traffic", d.get("roofline",{}).get("traffic"), "clocks", d.get("clocks",{}).get("samples"), d.get("clocks",{}).get("reasons"))
        print("  e2e", d.get("e2e",{}).get("value"), "inproc", d.get("e2e_inprocess",{}).get("value"))
        for w in d.get("workloads", []): print("   ", w["workload"][:60], w["dtype"], w["scaling"][:12], "%.3e"%w["value"], "%.4f ms"%w["ms_per_step"], "frac %.3f"%w["roofline_frac"], w["kernel"][:24])
    elif "Error" in l or "error" in l: print(l[:300])
PY
