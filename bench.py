#!/usr/bin/env python
"""bench.py -- PR2-arm FK + Jacobian throughput on B200 (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path (tfk_solve_fk_device, include/tinyfk_b200.h) over one batch of
synthetic configurations that is already resident in HBM: cfg2 = PR2 right arm 7-DoF + PLANER base,
r_gripper_tool_frame pose + Jacobian, RPY, FP64, 1 048 576 configurations PER GPU (weak scaling: the
batch shards trivially, no collective on the data path).  torch supplies device memory, the stream,
events and the rank plumbing; every FLOP is in libtinyfk_b200.so.

`value`      configs/s over all ranks, device-resident, CUDA events on the launching stream, max over ranks.
`e2e`        same metric through the reference-facing plugin call (_tinyfk.RobotModel.
             solve_forward_kinematics_batch: numpy in, numpy out) with HOST buffers; H2D + D2H inside.
`roofline`   algorithmic bytes (608 B/config) / measured kernel time vs the measured HBM peak.
`cpu_baseline` the reference's own C++ (oracle/_ref, built from /root/reference) on this host's cores.
--impl reference: the reference CPU implementation alone, same metric/config (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tinyfk_b200 import workloads as W  # noqa: E402

METRIC = "pr2_arm_fk_jacobian_configs_per_sec"
UNIT = "configs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--configs-per-gpu", type=int, default=0, help="override the workload's batch per GPU")
    ap.add_argument("--precision", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_n(cfg, override):
    if override:
        return override
    # per-GPU batch: the config's nominal size when it fits one GPU un-chunked, else a slice of it
    return {"cfg2": 1 << 20, "cfg3": 1 << 21, "cfg4": 1 << 19, "cfg5": 1 << 16}[cfg]


def workload_desc(cfg, n, precision):
    urdf, jn, ln, rot, base, _ = W.CONFIGS[cfg]
    n_link = 30 if ln is None else len(ln)
    return ("%s: %s, %d joints, %d link(s), rot=%s, base=%s, pose+Jacobian, %s, %d configs per GPU"
            % (cfg, urdf, len(jn), n_link, ["IGNORE", "RPY", "XYZW"][rot], ["FIXED", "PLANER", "FLOATING"][base],
               precision.upper(), n))


# ---- CPU reference arm ---------------------------------------------------------------------------
def cpu_reference_rate(cfg, n_threads, per_thread, repeats, seed=0):
    """Times the reference's own CPU implementation (oracle/_ref when built, else the C port) on
    `n_threads` host threads, one independent model per thread (the class is stateful and not
    thread-safe: include/tinyfk.hpp:67-75), `per_thread` configurations per thread per repeat.
    Returns (configs/s best of repeats, kind, list of per-repeat seconds)."""
    from oracle import refapi

    xml = W.urdf_text(W.CONFIGS[cfg][0])
    use_ref = refapi.have_ref(False)
    models, setups = [], None
    for _ in range(n_threads):
        if use_ref:
            m = refapi.RefModel(xml, strict=False)
            setups = W.setup(m, cfg)
        else:
            from tinyfk_b200 import capi
            pm = capi.Model(xml, host_libm=1)
            setups = W.setup(pm, cfg)
            m = refapi.PortModel(pm.export_table(), strict=False)
            m._limits = pm.get_joint_limits(setups[0])
        models.append(m)
    jids, lids, rot, base = setups
    limits = models[0].get_joint_limits(jids) if use_ref else models[0]._limits
    qs = [W.sample_q(limits, base, per_thread, seed=seed + t) for t in range(n_threads)]
    times = []
    for _ in range(repeats):
        barrier = threading.Barrier(n_threads + 1)
        done = []

        def work(t):
            barrier.wait()
            models[t].solve_fk(qs[t], lids, jids, rot, base, True)  # ctypes drops the GIL
            done.append(time.perf_counter())

        th = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
        for x in th:
            x.start()
        barrier.wait()
        t0 = time.perf_counter()
        for x in th:
            x.join()
        times.append(max(done) - t0)
    return n_threads * per_thread / min(times), ("reference" if use_ref else "port"), times


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = args.workload
    n = workload_n(cfg, args.configs_per_gpu)
    T = host_threads()
    per_thread = {"cfg2": 131072, "cfg3": 262144, "cfg4": 32768, "cfg5": 4096}[cfg]
    for _ in range(max(args.warmup, 1)):
        cpu_reference_rate(cfg, T, max(per_thread // 16, 256), 1)
    t_steps = []
    kind = "reference"
    for _ in range(args.steps):
        rate, kind, times = cpu_reference_rate(cfg, T, per_thread, 1)
        t_steps.append(times[0])
    ms = 1e3 * float(np.mean(t_steps))
    value = T * per_thread / (ms / 1e3)
    sample = "%d configs per step (%d threads x %d), one model per thread" % (T * per_thread, T, per_thread)
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_desc(cfg, n, "fp64"), "note": "CPU reference: bounded sample per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": T, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)
