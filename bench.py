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
    ap.add_argument("--jit", default="on", choices=["on", "off"],
                    help="on: specialise the single-link kernel on the workload's plan before the warm-up "
                         "(tfk_specialize, part of model set-up); off: generic kernel")
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


# ---- clocks ----------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (the quantities of the B200_PROFILING.md
    `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*` line, read through NVML)."""

    def __init__(self, gpu_index):
        """NVML from a host thread, one sample per millisecond while the timed region runs (the region is a few
        milliseconds long: an `nvidia-smi -lms` child would miss it), plus one forced sample while the GPU is busy."""
        self.samples, self.h, self.nv, self.t, self.on = [], None, None, None, False
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

    def sample(self):
        if self.h is None:
            return
        nv = self.nv
        try:
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            smax = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            self.samples.append((float(sm), float(smax), int(r)))
        except Exception:
            pass

    def start(self):
        import threading
        self.on = True

        def loop():
            while self.on:
                self.sample()
                time.sleep(0.001)
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
                "reasons": sorted(n for n, b in names if bits & b), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from the committed
# `ncu --set full` capture (profiles/), keyed by workload; None until a capture exists
# FP64 warp instructions per 32 configurations of the dominant kernel, from the same capture (executed-opcode histogram:
# DMUL + DADD + DFMA + DSETP); an FP64 instruction occupies its scheduler's FP64 pipe for 2 cycles on B200
# (tools/ubench/fp64_pipe.cu), which gives the kernel's FP64-pipe bound next to the HBM one
NCU_FP64_PER_WARP = {
    ("cfg2", "fp64", "spec"): 2140,     # profiles/r1_full_cfg2_strict_v10.txt (ncu: FP64 pipe 52 % active)
}
NCU_TRAFFIC_BYTES = {
    ("cfg2", "fp64"): 578.7e6,   # profiles/r1_full_cfg2_strict_v10.txt: 84.2 MB read + 494.5 MB written (algorithmic 637.5 MB)
    ("cfg4", "fp64"): 3500.7e6,  # profiles/r1_full_cfg4_twophase_v10.txt: 36.0 MB read + 3464.6 MB written (algorithmic 3430.9 MB)
    ("cfg5", "fp64"): 4030.8e6,  # profiles/r1_full_cfg5_twophase_v9.txt: 378.6 MB read + 3652.1 MB written (algorithmic 3471.3 MB)
}


# ---- our arm ---------------------------------------------------------------------------------------
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

    cfg = args.workload
    prec = 1 if args.precision == "fp32" else 0
    np_dt, t_dt, rb = (np.float32, torch.float32, 4) if prec else (np.float64, torch.float64, 8)
    n = workload_n(cfg, args.configs_per_gpu)
    model = capi.Model(W.urdf_text(W.CONFIGS[cfg][0]), host_libm=0)
    jids, lids, rot, base = W.setup(model, cfg)
    pd, n_in, L = W.pose_dim(rot), len(jids) + W.base_dof(base), len(lids)
    bytes_per_cfg = W.algorithmic_bytes((len(jids), L, rot, base), rb)

    # inputs larger than L2: rotate over enough distinct input batches that their sum exceeds 126 MB;
    # the outputs (written once, never re-read) are far larger than L2 on their own
    n_rot = max(2, int(np.ceil(160e6 / max(n * n_in * rb, 1))))
    n_rot = min(n_rot, 8)
    limits = model.get_joint_limits(jids)
    host_q = [W.sample_q(limits, base, n, seed=1000 * rank + r).astype(np_dt) for r in range(n_rot)]
    d_q = [torch.from_numpy(h).to(dev) for h in host_q]
    d_P = torch.empty((pd, n * L), dtype=t_dt, device=dev)
    d_J = torch.empty((n_in, n * L * pd), dtype=t_dt, device=dev)
    stream = torch.cuda.current_stream()

    # plan specialisation (include/tinyfk_b200.h, tfk_specialize): set-up work like creating the model, outside
    # every timed region; the same call is made on the e2e arm's model below
    jit_note = "off"
    model.set_jit_mode(0)
    if args.jit == "on":
        try:
            t0 = time.perf_counter()
            si = model.specialize(lids, jids, rot, base, True, prec, wait=True)
            model.set_jit_mode(1)
            jit_note = "specialised on the plan in %.1f s%s" % (time.perf_counter() - t0, " (disk cache)" if si["from_cache"] else "")
        except ValueError as e:  # tree too large for the unrolled build: the generic kernel serves the plan
            jit_note = "not applicable: %s" % e
        except RuntimeError as e:  # NVRTC missing: the generic kernel serves the plan
            jit_note = "unavailable: %s" % str(e).splitlines()[0]

    def step(i):
        model.solve_fk_device(local_rank, stream.cuda_stream, d_q[i % n_rot].data_ptr(), n, lids, jids, rot, base,
                              True, prec, d_P.data_ptr(), d_J.data_ptr(), n, 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local_rank) if rank == 0 else None
    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    launches0 = capi.kernel_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    if clocks is not None:
        clocks.start()
    ev[0].record(stream)
    for i in range(args.steps):
        step(i)
        ev[i + 1].record(stream)
    if clocks is not None:
        clocks.sample()  # the launches above are asynchronous: the GPU is inside the timed region now
    barrier()
    clk = clocks.stop() if clocks is not None else None
    launches = capi.kernel_launch_count() - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_ms = float(t.item())
    value = world * n * args.steps / (max_ms / 1e3)

    # ---- end to end through the plugin call, host buffers (pinned), H2D + D2H inside ---------------
    e2e = None
    if not args.no_e2e:
        import tinyfk_b200
        from tinyfk_b200 import _tinyfk

        robot = _tinyfk.RobotModel(W.urdf_text(W.CONFIGS[cfg][0]))
        W.setup(_FacadeShim(robot), cfg)
        robot.set_jit_mode(0)
        if args.jit == "on" and jit_note.startswith("specialised"):
            robot.specialize(lids, jids, rot, base, True, prec, True, False)
            robot.set_jit_mode(1)
        ne = min(n, max(4096, int(3e9 // max(bytes_per_cfg, 1))))  # bound host memory: <= ~3 GB of output
        pin_q = capi.PinnedArray((ne, n_in), dtype=np_dt, order="C")
        pin_q.array[...] = host_q[0][:ne]
        k_e2e = max(3, min(args.steps, 10))
        for _ in range(3):  # results are held across calls, as in the timed loop, so the pinned pool is warm
            P, J = robot.solve_forward_kinematics_batch(pin_q.array, lids, jids, rot, base, True, prec, [local_rank], True)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            P, J = robot.solve_forward_kinematics_batch(pin_q.array, lids, jids, rot, base, True, prec, [local_rank], True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        chk = float(P[-1, 0]) + float(J[-1, -1])  # the result is read on the host
        assert np.isfinite(chk)
        te = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": world * ne * k_e2e / float(te.item()), "unit": UNIT,
               "h2d_bytes_per_step": int(ne * n_in * rb), "d2h_bytes_per_step": int(ne * L * pd * rb * (1 + n_in)),
               "configs_per_step_per_gpu": int(ne), "steps": k_e2e,
               "api": "_tinyfk.RobotModel.solve_forward_kinematics_batch (numpy in, pinned numpy out)"}
        del P, J

    if rank == 0:
        peak, peak_src = measured_peaks()
        kern_ms = float(np.mean(step_ms))
        achieved = bytes_per_cfg * n / (kern_ms / 1e3) / 1e9
        info = model.last_launch_info()
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": max_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if prec else "f64", "data": "synthetic",
            "config": {"workload": workload_desc(cfg, n, args.precision),
                       "l2": "inputs rotate over %d distinct batches (%.0f MB) > 126 MB L2; %.0f MB of outputs per step"
                             % (n_rot, n_rot * n * n_in * rb / 1e6, n * L * pd * rb * (1 + n_in) / 1e6),
                       "mode": "strict FP64: reference operation order, portable libm, no FMA contraction" if not prec
                               else "FP32 opt-in mode",
                       "jit": jit_note, "launch": info},
            "clocks": clk, "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_BYTES.get((cfg, args.precision)),
                         "peak_source": peak_src, "algorithmic_bytes_per_config": bytes_per_cfg,
                         "kernel": info.get("kernel"), "kernel_ms": kern_ms,
                         "note": ("bit-faithful FP64 (no FMA contraction, IEEE divisions, portable libm) is bound by "
                                  "instruction issue / the FP64 pipe, not by HBM: DESIGN.md section 3") if not prec
                                 else "FP32 opt-in mode"},
        }
        key = (cfg, args.precision, "spec" if "spec" in str(info.get("kernel")) else "generic")
        if key in NCU_FP64_PER_WARP and clk and clk.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
            bound_ms = (n / 32.0) * NCU_FP64_PER_WARP[key] * 2.0 / (4 * sms * clk["sm_mhz"] * 1e6) * 1e3
            out["roofline"]["fp64_pipe"] = {"fp64_warp_instructions_per_32_configs": NCU_FP64_PER_WARP[key],
                                            "bound_ms": bound_ms, "frac": bound_ms / kern_ms,
                                            "note": "2 cycles per FP64 warp instruction per scheduler, 4 schedulers x %d SMs at "
                                                    "the sampled SM clock" % sms}
        if e2e is not None:
            out["e2e"] = e2e
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


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
