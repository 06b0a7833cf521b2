#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, on the CPU box): key pipe utilisations, per-opcode instruction mix and
the most-stalled SASS instructions of the first kernel in the report.
    python tools/ncu_summary.py gpurun_out/x.ncu-rep [units_per_launch] [--top N]
"""
import collections
import csv
import io
import subprocess
import sys


def page(rep, which, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", which, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else None
    top_n = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    raw = page(rep, "raw")
    hdr, units_row, first = raw[0], raw[1], raw[2]
    want = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg", "sm__cycles_elapsed.avg.per_second",
            "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
            "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active"]
    print("kernel:", first[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
    for i, h in enumerate(hdr):
        if h in want:
            print(f"  {h} [{units_row[i]}] = {first[i]}")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in h]
    st = sorted(((float(first[i] or 0), h.replace("smsp__pcsamp_warps_issue_stalled_", "")) for i, h in stall_cols), reverse=True)
    tot = sum(v for v, _ in st) or 1
    print("  stall samples:", ", ".join(f"{n} {100 * v / tot:.1f}%" for v, n in st[:10]))

    src = page(rep, "source", ["--print-source", "sass"])
    h2 = src[1]
    idx = {h: i for i, h in enumerate(h2)}
    body = []
    for r in src[2:]:
        if len(r) < 10 or r[0] == "Kernel Name":
            break
        if r[0] != "Address":
            body.append(r)
    ops, smp = collections.Counter(), collections.Counter()
    for r in body:
        parts = r[1].split()
        op = parts[1] if parts[0].startswith("@") else parts[0]
        op = op if op.startswith("MUFU") else op.split(".")[0]
        ops[op] += int(r[idx["Instructions Executed"]])
        smp[op] += int(r[idx["# Samples"]])
    total = sum(ops.values())
    div = units or 1.0
    print(f"  instructions: {total} total" + (f", {total / div:.1f} per unit" if units else ""))
    for op, n in ops.most_common(22):
        print(f"    {op:12s} {n / div:10.1f}  samples {smp[op]}")
    stalls = [h for h in h2 if h.startswith("stall_") and "Not Issued" not in h]
    print(f"  top {top_n} instructions by samples (index, samples, executed, sass, stalls):")
    order = sorted(range(len(body)), key=lambda i: -int(body[i][idx["# Samples"]]))[:top_n]
    for i in sorted(order):
        r = body[i]
        s = {h[6:]: int(r[idx[h]]) for h in stalls if int(r[idx[h]]) > 0}
        s = dict(sorted(s.items(), key=lambda kv: -kv[1])[:4])
        print(f"    {i:5d} {r[idx['# Samples']]:>6s} {r[idx['Instructions Executed']]:>9s}  {r[1].strip()[:60]:60s} {s}")


if __name__ == "__main__":
    main()
