#!/usr/bin/env python
"""bench.py -- PR2-arm FK + Jacobian throughput on B200 (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2] [--precision fp64]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path (tfk_solve_fk_device, include/tinyfk_b200.h) over one batch of
synthetic configurations that is already resident in HBM: cfg2 = PR2 right arm 7-DoF + PLANER base,
r_gripper_tool_frame pose + Jacobian, RPY, FP64, 1 048 576 configurations PER GPU (weak scaling: the
batch shards trivially, no collective on the data path).  torch supplies device memory, the stream,
events and the rank plumbing; every FLOP is in libtinyfk_b200.so.

`value`      configs/s over all ranks, device-resident, CUDA events on the launching stream, max over ranks (the K launches
             are enqueued while a spin kernel holds the stream: host scheduling hiccups are e2e's business, not this one's).
`e2e`        same metric through the reference-facing plugin call (_tinyfk.RobotModel.
             solve_forward_kinematics_batch: numpy in, numpy out) with HOST buffers; H2D + D2H inside.
`e2e_inprocess` (N > 1) the same call with devices=[0..N-1] made by rank 0 alone on N x the batch: the engine's own
             multi-GPU split (Engine::solve_host), while the other ranks wait.
`roofline`   algorithmic bytes (608 B/config) / measured kernel time vs the measured HBM peak.
`workloads`  the other BASELINE configurations and precision modes, each timed the same way (kernel-only) with its own
             roofline fraction; under torchrun also the STRONG readings (cfg2 at 1M total, cfg3 at 16M total).
`cpu_baseline` the reference's own C++ (oracle/_ref, built from /root/reference) on this host's cores.
--impl reference: the reference CPU implementation alone, same metric/config (rank 0 only); that arm never imports
the product package (workloads.py is loaded by path).
"""
import argparse
import hashlib
import importlib.util
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _load_by_path(name, rel):
    """workloads.py / sharding.py are plain numpy modules inside the package directory: loading them by path keeps the
    reference arm from importing tinyfk_b200/__init__.py (which loads the product's native libraries)"""
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


W = _load_by_path("tfk_workloads", os.path.join("tinyfk_b200", "workloads.py"))
sharding = _load_by_path("tfk_sharding", os.path.join("tinyfk_b200", "sharding.py"))

METRIC = "pr2_arm_fk_jacobian_configs_per_sec"
UNIT = "configs/s"
PRECISIONS = {"fp64": 0, "fp32": 1, "fp64_fast": 2}
MODE_NOTE = ["strict FP64: reference operation order, portable libm, no FMA contraction", "FP32 opt-in mode",
             "FP64 fast opt-in mode: FMA contraction, reciprocal quotients, analytic base columns (1e-12 contract)"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--configs-per-gpu", type=int, default=0, help="override the workload's batch per GPU")
    ap.add_argument("--precision", default="fp64", choices=sorted(PRECISIONS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="skip the other BASELINE configurations / modes")
    ap.add_argument("--jit", default="on", choices=["on", "off"],
                    help="on: specialise the kernel on the workload's plan before the warm-up "
                         "(tfk_specialize, part of model set-up); off: generic kernel")
    return ap.parse_args()


def workload_n(cfg, override=0):
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


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def kernel_source_hash():
    """identifies the kernel sources a build was made from (there is no .git on the GPU box): profiles/ncu_traffic.json
    entries written by tools/ncu_capture.py carry it, and a stale entry is reported as null"""
    d = os.path.join(ROOT, "tinyfk_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".h", ".cuh", ".cu", ".inc")):
            with open(os.path.join(d, f), "rb") as fh:
                h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def ncu_record(cfg, precision, kernel):
    """per-launch DRAM traffic (+ FP64 warp instructions per 32 configurations) of the dominant kernel from the committed
    `ncu --set full` capture of THIS build, or None when there is none / it is stale"""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            rec = json.load(f)
    except Exception:
        return None
    e = rec.get("entries", {}).get("%s/%s" % (cfg, precision))
    if not e or e.get("kernel_source_hash") != kernel_source_hash():
        return None
    if ("spec" in str(kernel)) != ("spec" in str(e.get("kernel", ""))):
        return None
    return e


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
        else:  # the reference could not be compiled here: the C restatement, driven by the product's link table
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


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = args.workload
    n = workload_n(cfg, args.configs_per_gpu)
    T = host_threads()
    per_thread = {"cfg2": 131072, "cfg3": 262144, "cfg4": 32768, "cfg5": 4096}[cfg]
    steps = args.steps
    if steps > 20:  # bounded sample: a step is ~0.3 s of all host cores at the nominal size; keep the run under a minute
        per_thread = max(2048, per_thread * 20 // steps)
    for _ in range(max(min(args.warmup, 3), 1)):
        cpu_reference_rate(cfg, T, max(per_thread // 16, 256), 1)
    t_steps = []
    kind = "reference"
    for _ in range(steps):
        rate, kind, times = cpu_reference_rate(cfg, T, per_thread, 1)
        t_steps.append(times[0])
    ms = 1e3 * float(np.mean(t_steps))
    value = T * per_thread / (ms / 1e3)
    sample = "%d configs per step (%d threads x %d), one model per thread" % (T * per_thread, T, per_thread)
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_desc(cfg, n, "fp64"), "note": "CPU reference: bounded sample per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": T, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ---- clocks ----------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (the quantities of the B200_PROFILING.md
    `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*` line, read through NVML) by a host
    thread that polls back to back (an NVML query takes a few hundred microseconds) while the region runs."""

    def __init__(self, gpu_index):
        self.samples, self.h, self.nv, self.t, self.on = [], None, None, None, False
        self.in_region, self.n_in_region, self.smax = None, 0, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(gpu_index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        except Exception:
            self.h = None
        # the first NVML queries of a process can take tens of milliseconds (lazy initialisation): pay for them here,
        # not inside the timed region
        for _ in range(3):
            self.sample()
        self.samples, self.n_in_region = [], 0

    def sample(self):
        if self.h is None:
            return
        nv = self.nv
        try:
            inside = self.in_region is not None and self.in_region()
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            if self.smax is None:
                self.smax = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            smax = self.smax
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            self.samples.append((float(sm), float(smax), int(r)))
            self.n_in_region += 1 if inside else 0
        except Exception:
            pass

    def start(self):
        self.on = True

        def loop():
            while self.on:
                self.sample()
                time.sleep(0.0002)
        self.t = threading.Thread(target=loop, daemon=True)
        self.t.start()

    def stop(self):
        self.on = False
        if self.t is not None:
            self.t.join()
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML unavailable"], "samples": 0}
        sm = [x[0] for x in self.samples]
        smax = [x[1] for x in self.samples]
        bits = 0
        for x in self.samples:
            bits |= x[2]
        # nvml.h: SwPowerCap 0x4, HwSlowdown 0x8, SwThermalSlowdown 0x20, HwThermalSlowdown 0x40
        names = [("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(n for n, b in names if bits & b), "samples": len(sm),
                "samples_in_timed_region": self.n_in_region,
                "note": "NVML polled back to back while the timed steps and an untimed tail of identical steps run (>= 60 ms of "
                        "the same load in all): an NVML query takes longer than a step"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- our arm ---------------------------------------------------------------------------------------
class _FacadeShim:
    """lets workloads.setup() drive a _tinyfk.RobotModel (list-returning methods)"""

    def __init__(self, robot):
        self.r = robot

    def get_joint_ids(self, names):
        return self.r.get_joint_ids(list(names))

    def get_link_ids(self, names):
        return self.r.get_link_ids(list(names))

    def add_new_link(self, name, parent, pos, rpy):
        self.r.add_new_link(name, int(parent), list(map(float, pos)), list(map(float, rpy)))


def constant_jacobian_columns(model, cfg, jids, lids, base):
    """columns of J the host entry fills on the CPU instead of copying them over PCIe (Engine::solve_host): joints that move
    none of the requested links, and the base-translation unit columns"""
    n_zero = sum(1 for j in jids if not any(model.is_relevant(l, j) for l in lids))
    return n_zero + {0: 0, 1: 2, 2: 3}[int(base)]


def measure_device_resident(ctx, cfg, precision, n, steps, warmup, jit=True, seed_base=0, clocks=None):
    """K timed passes of tfk_solve_fk_device over device-resident batches of `n` configurations (inputs rotate over
    distinct batches larger than L2), CUDA events on the launching stream.  Returns a dict with the per-rank total
    milliseconds, the per-step times, the launch description and the launch count."""
    import torch

    capi = ctx["capi"]
    prec = PRECISIONS[precision]
    np_dt, t_dt, rb = (np.float32, torch.float32, 4) if prec == 1 else (np.float64, torch.float64, 8)
    dev, local_rank = ctx["dev"], ctx["local_rank"]
    model = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0)
    jids, lids, rot, base = W.setup(model, cfg)
    pd, n_in, L = W.pose_dim(rot), len(jids) + W.base_dof(base), len(lids)
    bytes_per_cfg = W.algorithmic_bytes((len(jids), L, rot, base), rb)
    # inputs larger than L2: rotate over enough distinct input batches that their sum exceeds 126 MB;
    # the outputs (written once, never re-read) are far larger than L2 on their own
    n_rot = min(8, max(2, int(np.ceil(160e6 / max(n * n_in * rb, 1)))))
    limits = model.get_joint_limits(jids)
    host_q = [W.sample_q(limits, base, n, seed=seed_base + r).astype(np_dt) for r in range(n_rot)]
    d_q = [torch.from_numpy(h).to(dev) for h in host_q]
    d_P = torch.empty((pd, n * L), dtype=t_dt, device=dev)
    d_J = torch.empty((n_in, n * L * pd), dtype=t_dt, device=dev)
    stream = torch.cuda.current_stream()
    # plan specialisation (include/tinyfk_b200.h, tfk_specialize): set-up work like creating the model, outside
    # every timed region
    jit_note = "off"
    model.set_jit_mode(0)
    if jit:
        try:
            t0 = time.perf_counter()
            si = model.specialize(lids, jids, rot, base, True, prec, wait=True)
            model.set_jit_mode(1)
            jit_note = "specialised on the plan in %.1f s%s" % (time.perf_counter() - t0, " (disk cache)" if si["from_cache"] else "")
        except ValueError as e:  # not eligible: the generic kernel serves the plan
            jit_note = "not applicable: %s" % e
        except RuntimeError as e:  # NVRTC missing: the generic kernel serves the plan
            jit_note = "unavailable: %s" % str(e).splitlines()[0]

    def step(i):
        model.solve_fk_device(local_rank, stream.cuda_stream, d_q[i % n_rot].data_ptr(), n, lids, jids, rot, base,
                              True, prec, d_P.data_ptr(), d_J.data_ptr(), n, 0)

    wev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    wev[0].record(stream)
    for i in range(max(warmup, 3)):
        step(i)
    wev[1].record(stream)
    ctx["barrier"]()
    est_ms = max(wev[0].elapsed_time(wev[1]) / max(warmup, 3), 1e-3)
    launches0 = capi.kernel_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    if clocks is not None:
        clocks.in_region = lambda: not ev[-1].query()
        clocks.start()
    # The steps are enqueued while the stream is held by a spin kernel of torch's (not counted in `gpu_launches`): the K
    # launches and their events are already queued when the first timed event executes, so the timed region is K steps of
    # device work back to back -- a scheduling hiccup of the enqueuing host thread (8 ranks on a 32-vCPU VM: 1.3 ms seen on
    # one rank of a 10 ms region) is not part of the device-resident figure; host costs are what `e2e` measures.
    hold_ms = min(50.0, max(2.0, steps * 0.04))
    with torch.cuda.stream(stream):
        torch.cuda._sleep(int(hold_ms * 1.9e6))
    ev[0].record(stream)
    for i in range(steps):
        step(i)
        ev[i + 1].record(stream)
    launches = capi.kernel_launch_count() - launches0
    if clocks is not None:
        # The timed region is a few milliseconds and an NVML query takes about one: keep the GPU under the SAME load (the
        # same step, untimed, after the last event) until the sampler has seen >= 60 ms of it.  The events above are not
        # affected; `samples_in_timed_region` counts the samples taken before the last timed step finished.
        clocks.sample()
        for i in range(int(min(4000, max(0.0, 60.0 / est_ms - steps)))):
            step(steps + i)
    ctx["barrier"]()
    clk = clocks.stop() if clocks is not None else None
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    max_ms = ctx["max_over_ranks"](total_ms)
    per_rank = ctx["gather"]([total_ms / steps, float(np.median(step_ms)), float(np.max(step_ms))])
    info = model.last_launch_info()
    res = {"model": model, "ids": (jids, lids, rot, base), "host_q": host_q, "n_rot": n_rot, "n_in": n_in, "L": L, "pd": pd,
           "rb": rb, "np_dt": np_dt, "bytes_per_cfg": bytes_per_cfg, "max_ms": max_ms, "step_ms": step_ms, "info": info,
           "launches": launches, "jit_note": jit_note, "clocks": clk, "prec": prec, "per_rank": per_rank}
    del d_q, d_P, d_J
    return res


def roofline_of(cfg, precision, n, r, peak, peak_src):
    kern_ms = float(np.mean(r["step_ms"]))
    achieved = r["bytes_per_cfg"] * n / (kern_ms / 1e3) / 1e9
    rec = ncu_record(cfg, precision, r["info"].get("kernel"))
    out = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "traffic": rec["dram_bytes"] if rec else None,
           "traffic_source": (rec.get("source") if rec else "no ncu capture of this build (profiles/ncu_traffic.json)"),
           "peak_source": peak_src, "algorithmic_bytes_per_config": r["bytes_per_cfg"],
           "kernel": r["info"].get("kernel"), "kernel_ms": kern_ms}
    if precision == "fp64":
        out["note"] = ("bit-faithful FP64 (no FMA contraction, IEEE divisions, portable libm) is bound by instruction issue / "
                       "the FP64 pipe, not by HBM: DESIGN.md section 3")
    return out, rec


def measure_e2e(ctx, cfg, precision, n, r, steps, devices, jit_ok):
    """the same metric through the reference-facing plugin call with HOST buffers (pinned), H2D + D2H inside"""
    import torch
    from tinyfk_b200 import _tinyfk

    capi = ctx["capi"]
    jids, lids, rot, base = r["ids"]
    prec, rb, n_in, L, pd = r["prec"], r["rb"], r["n_in"], r["L"], r["pd"]
    robot = _tinyfk.RobotModel(W.urdf_text(W.CONFIGS[cfg][0]))
    W.setup(_FacadeShim(robot), cfg)
    robot.set_jit_mode(0)
    if jit_ok:
        robot.specialize(lids, jids, rot, base, True, prec, True, False)
        robot.set_jit_mode(1)
    n_total = n * len(devices)
    ne = min(n_total, max(4096, int(3e9 // max(r["bytes_per_cfg"], 1))))  # bound host memory: <= ~3 GB of output
    pin_q = capi.PinnedArray((ne, n_in), dtype=r["np_dt"], order="C")
    reps = -(-ne // r["host_q"][0].shape[0])
    pin_q.array[...] = np.concatenate([r["host_q"][k % r["n_rot"]] for k in range(reps)])[:ne]
    k_e2e = max(3, min(steps, 10))
    for _ in range(3):  # results are held across calls, as in the timed loop, so the pinned pool is warm
        P, J = robot.solve_forward_kinematics_batch(pin_q.array, lids, jids, rot, base, True, prec, devices, True)
    if len(devices) == 1:
        ctx["barrier"]()
    t0 = time.perf_counter()
    for _ in range(k_e2e):
        P, J = robot.solve_forward_kinematics_batch(pin_q.array, lids, jids, rot, base, True, prec, devices, True)
    for d in devices:
        torch.cuda.synchronize(d)
    dt = time.perf_counter() - t0
    chk = float(P[-1, 0]) + float(J[-1, -1])  # the result is read on the host
    assert np.isfinite(chk)
    n_const = constant_jacobian_columns(r["model"], cfg, jids, lids, base)
    out = {"seconds": dt, "configs": ne * k_e2e, "unit": UNIT, "h2d_bytes_per_step": int(ne * n_in * rb),
           "d2h_bytes_per_step": int(ne * L * pd * rb * (1 + n_in - n_const)),
           "host_filled_bytes_per_step": int(ne * L * pd * rb * n_const),
           "configs_per_step": int(ne), "steps": k_e2e,
           "api": "_tinyfk.RobotModel.solve_forward_kinematics_batch (numpy in, pinned numpy out)",
           "note": "%d of the %d Jacobian columns are constants of the plan (unit base-translation columns, joints that move no "
                   "requested link): filled by host threads, not copied" % (n_const, n_in)}
    del P, J
    pin_q.free()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    from tinyfk_b200 import capi  # raises if the CUDA library is not built: no fallback

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather(vals):  # every rank's list of floats, in rank order
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world == 1:
            return [t.tolist()]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [o.tolist() for o in out]

    ctx = {"capi": capi, "dev": dev, "local_rank": local_rank, "barrier": barrier, "max_over_ranks": max_over_ranks,
           "gather": gather}
    cfg, precision = args.workload, args.precision
    n = workload_n(cfg, args.configs_per_gpu)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    r = measure_device_resident(ctx, cfg, precision, n, args.steps, args.warmup, jit=args.jit == "on", seed_base=1000 * rank,
                                clocks=clocks)
    value = world * n * args.steps / (r["max_ms"] / 1e3)
    jit_ok = r["jit_note"].startswith("specialised")

    e2e = e2e_in = None
    if not args.no_e2e:
        e = measure_e2e(ctx, cfg, precision, n, r, args.steps, [local_rank], jit_ok)
        sec = max_over_ranks(e.pop("seconds"))
        e["value"] = world * e.pop("configs") / sec
        e["configs_per_step_per_gpu"] = e.pop("configs_per_step")
        e2e = e
        if world > 1:  # the engine's own multi-GPU split: one process, devices=[0..world-1], the other ranks wait
            barrier()
            if rank == 0:
                ei = measure_e2e(ctx, cfg, precision, n, r, min(args.steps, 5), list(range(world)), jit_ok)
                ei["value"] = ei.pop("configs") / ei.pop("seconds")
                ei["devices"] = list(range(world))
                e2e_in = ei
            barrier()

    # ---- the other BASELINE configurations and modes (kernel-only, same timing rules) ----------------------
    peak, peak_src = measured_peaks()
    others = []
    if not args.no_workloads and cfg == "cfg2" and precision == "fp64" and not args.configs_per_gpu:
        k = max(5, min(args.steps, 20))
        todo = [("cfg3", "fp64", workload_n("cfg3"), "weak"), ("cfg4", "fp64", workload_n("cfg4"), "weak"),
                ("cfg5", "fp64", workload_n("cfg5"), "weak"), ("cfg5", "fp32", workload_n("cfg5"), "weak"),
                ("cfg2", "fp32", workload_n("cfg2"), "weak"), ("cfg4", "fp32", workload_n("cfg4"), "weak"),
                ("cfg2", "fp64_fast", workload_n("cfg2"), "weak"), ("cfg3", "fp64_fast", workload_n("cfg3"), "weak")]
        if world > 1:  # the strong readings: BASELINE "1M batch at 1/2/4/8 B200" and cfg3 "16M sharded across 2/4/8"
            lo, hi = sharding.shard_bounds(1 << 20, world, rank)
            todo.append(("cfg2", "fp64", hi - lo, "strong: 1 048 576 configurations in total"))
            lo, hi = sharding.shard_bounds(1 << 24, world, rank)
            todo.append(("cfg3", "fp64", hi - lo, "strong: 16 777 216 configurations in total"))
        for c2, p2, n2, scaling in todo:
            r2 = measure_device_resident(ctx, c2, p2, n2, k, 3, jit=args.jit == "on", seed_base=1000 * rank + 77)
            rf, _ = roofline_of(c2, p2, n2, r2, peak, peak_src)
            total = n2 * world if scaling == "weak" else int(scaling.split(":")[1].split("conf")[0].replace(" ", ""))
            others.append({"workload": workload_desc(c2, n2, p2), "scaling": scaling, "dtype": "f32" if p2 == "fp32" else "f64",
                           "value": total * k / (r2["max_ms"] / 1e3), "unit": UNIT, "ms_per_step": r2["max_ms"] / k, "steps": k,
                           "roofline_frac": rf["frac"], "achieved_GBps": rf["achieved"], "kernel": rf["kernel"],
                           "launch": r2["info"], "jit": r2["jit_note"], "gpu_launches": int(r2["launches"])})
            del r2

    if rank == 0:
        info = r["info"]
        rf, rec = roofline_of(cfg, precision, n, r, peak, peak_src)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": r["max_ms"] / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if precision == "fp32" else "f64", "data": "synthetic",
            "config": {"workload": workload_desc(cfg, n, precision),
                       "l2": "inputs rotate over %d distinct batches (%.0f MB) > 126 MB L2; %.0f MB of outputs per step"
                             % (r["n_rot"], r["n_rot"] * n * r["n_in"] * r["rb"] / 1e6,
                                n * r["L"] * r["pd"] * r["rb"] * (1 + r["n_in"]) / 1e6),
                       "mode": MODE_NOTE[r["prec"]], "jit": r["jit_note"], "launch": info},
            "clocks": r["clocks"], "gpu_launches": int(r["launches"]), "roofline": rf,
            # per rank: [timed region / K, median step, longest step] in ms (`ms_per_step` is the max of the first column)
            "per_rank_ms_per_step": [[round(v, 6) for v in row] for row in r["per_rank"]],
        }
        out["config"]["timing"] = ("CUDA events on the launching stream around each of the K steps; the K launches are enqueued "
                                   "while a spin kernel holds the stream, so the region is device work back to back")
        clk = r["clocks"]
        if rec and rec.get("fp64_warp_instructions_per_32_configs") and clk and clk.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
            f64 = rec["fp64_warp_instructions_per_32_configs"]
            bound_ms = (n / 32.0) * f64 * 2.0 / (4 * sms * clk["sm_mhz"] * 1e6) * 1e3
            out["roofline"]["fp64_pipe"] = {"fp64_warp_instructions_per_32_configs": f64, "bound_ms": bound_ms,
                                            "frac": bound_ms / rf["kernel_ms"],
                                            "note": "2 cycles per FP64 warp instruction per scheduler, 4 schedulers x %d SMs at "
                                                    "the sampled SM clock" % sms}
        if e2e is not None:
            out["e2e"] = e2e
        if e2e_in is not None:
            out["e2e_inprocess"] = e2e_in
        if others:
            out["workloads"] = others
        if not args.no_cpu_baseline and world == 1:
            T = host_threads()
            per_thread = {"cfg2": 65536, "cfg3": 131072, "cfg4": 16384, "cfg5": 2048}[cfg]
            cpu_reference_rate(cfg, T, max(per_thread // 16, 256), 1)
            rate, kind, times = cpu_reference_rate(cfg, T, per_thread, 3)
            rate1, _, _ = cpu_reference_rate(cfg, 1, per_thread, 2)
            out["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": T, "kind": kind,
                                   "sample": "%d configs (%d threads x %d), best of 3, one model per thread"
                                             % (T * per_thread, T, per_thread),
                                   "single_thread_value": rate1}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
