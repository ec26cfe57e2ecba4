#!/usr/bin/env python
"""Copies what tools/gpu_cycle_final.sh left in gpurun_out/ into profiles/ (tracked): the per-kernel ncu summaries with a
header naming the kernel-source hash they belong to, the per-launch traffic table bench.py reads, the launch list and the
bench lines.  Run in the build container after the cycle."""
import json
import os
import re
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def jsonl(src, dst):
    with open(dst, "w") as f:
        for l in open(src):
            if l.startswith("{"):
                f.write(l)


def main():
    rec = json.load(open(os.path.join(G, "ncu_traffic.json")))
    hashes = {e["kernel_source_hash"] for e in rec["entries"].values()}
    here = bench.kernel_source_hash()
    if hashes != {here}:
        raise SystemExit("captures belong to %s, the tree is %s: re-run the cycle" % (sorted(hashes), here))
    note = sys.argv[1] if len(sys.argv) > 1 else "final build"
    for c in ("cfg2", "cfg3", "cfg4", "cfg5"):
        for m in ("fp64", "fp32", "fp64_fast"):
            body = open(os.path.join(G, "f_lines_%s_%s.txt" % (c, m))).read()
            head = ("# ncu --set full --clock-control none --import-source on, one launch (the 4th) of the dominant kernel of\n"
                    "#   python bench.py --workload %s --precision %s --steps 2 --warmup 3   (tools/ncu_capture.py, "
                    "tools/gpu_cycle_final.sh; B200, CUDA 12.9)\n"
                    "# %s: kernel sources %s (bench.py:kernel_source_hash); summarised by tools/ncu_lines.py\n" % (c, m, note, here))
            open(os.path.join(P, "r2_full_%s_%s.txt" % (c, m)), "w").write(head + body)
    # FP64-pipe warp instructions per 32 configurations (bench.py's `roofline.fp64_pipe` bound): `--set full` has no per-pipe
    # instruction COUNT, so it comes from the opcode histogram of the same capture (tools/ncu_lines.py): the D* opcodes
    for key, e in rec["entries"].items():
        c, m = key.split("/")
        if m == "fp32":
            continue
        txt = open(os.path.join(G, "f_lines_%s_%s.txt" % (c, m))).read()
        hist = re.search(r"executed warp instructions per 32 configs: \d+\n(.*)\n", txt)
        if hist:
            ops = dict(kv.split("=") for kv in hist.group(1).split() if "=" in kv)
            e["fp64_warp_instructions_per_32_configs"] = float(sum(int(v) for k, v in ops.items() if k.startswith("D")))
            e["fp64_count_source"] = "D* opcodes (DFMA, DMUL, DADD, DSETP ...) of the capture's opcode histogram, top 24 opcodes"
    json.dump(rec, open(os.path.join(G, "ncu_traffic.json"), "w"), indent=1)
    shutil.copy(os.path.join(G, "ncu_traffic.json"), os.path.join(P, "ncu_traffic.json"))
    shutil.copy(os.path.join(G, "f_launches_cfg2.csv"), os.path.join(P, "r2_launches_cfg2_strict_final.csv"))
    jsonl(os.path.join(G, "f_b.log"), os.path.join(P, "r2_logs", "bench_n1_final.jsonl"))
    jsonl(os.path.join(G, "f_bref.log"), os.path.join(P, "r2_logs", "bench_reference_arm_final.jsonl"))
    print("profiles/ updated for kernel sources", here)


if __name__ == "__main__":
    main()
