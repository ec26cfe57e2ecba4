#!/usr/bin/env python
"""Captures the dominant kernel of bench.py workloads with `ncu --set full` (one launch each, after warm-up) on the GPU box
and writes gpurun_out/ncu_traffic.json: per workload/mode the per-launch DRAM bytes, duration, FP64-pipe warp instructions
and the kernel-source hash of the build (bench.py:kernel_source_hash) -- bench.py reports `roofline.traffic` from
profiles/ncu_traffic.json only while that hash matches the build it is timing.

  python tools/ncu_capture.py cfg2/fp64 cfg4/fp64 cfg2/fp32 ...      (under gpurun; the .ncu-rep files stay in gpurun_out/)
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")


def raw_metrics(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        return {}
    return {h: (u, v) for h, u, v in zip(rows[0], rows[1], rows[2])}


def to_bytes(uv):
    u, v = uv
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


def main():
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "ncu_traffic.json")
    rec = {"entries": {}}
    if os.path.exists(path):
        rec = json.load(open(path))
    for item in sys.argv[1:]:
        cfg, prec = item.split("/")
        rep = os.path.join(OUT, "r2_%s_%s" % (cfg, prec))
        cmd = ["ncu", "--set", "full", "--clock-control", "none", "--import-source", "on", "-k", "regex:tfk_spec|fk_twophase|fk_fused",
               "-s", "3", "-c", "1", "-f", "-o", rep, sys.executable, os.path.join(ROOT, "bench.py"), "--workload", cfg,
               "--precision", prec, "--steps", "2", "--warmup", "3", "--no-cpu-baseline", "--no-e2e", "--no-workloads"]
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        if p.returncode != 0 or not os.path.exists(rep + ".ncu-rep"):
            print("capture of %s failed:\n%s" % (item, (p.stdout + p.stderr)[-2000:]))
            continue
        m = raw_metrics(rep + ".ncu-rep")
        n = bench.workload_n(cfg)
        e = {"kernel_source_hash": bench.kernel_source_hash(), "configs_per_launch": n,
             "dram_read_bytes": to_bytes(m["dram__bytes_read.sum"]), "dram_write_bytes": to_bytes(m["dram__bytes_write.sum"]),
             "source": "profiles/r2_full_%s_%s.txt (ncu --set full, one launch)" % (cfg, prec)}
        e["dram_bytes"] = e["dram_read_bytes"] + e["dram_write_bytes"]
        u, v = m.get("gpu__time_duration.sum", ("us", "0"))
        e["duration_us"] = float(v.replace(",", "")) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
        for name in ("sm__inst_executed_pipe_fp64.sum", "smsp__inst_executed_pipe_fp64.sum"):
            if name in m:
                e["fp64_warp_instructions_per_32_configs"] = float(m[name][1].replace(",", "")) / (n / 32.0)
                break
        e["kernel"] = "spec" if "tfk_spec" in (p.stdout + p.stderr) else "generic"
        e["registers_per_thread"] = m.get("launch__registers_per_thread", ("", ""))[1]
        rec["entries"][item] = e
        print(item, json.dumps(e))
        json.dump(rec, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
