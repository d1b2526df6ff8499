#!/usr/bin/env python3
"""Summarise an ncu launch list (csv) and an `ncu --set full` report into a small text file for profiles/.
usage: tools/ncu_summary.py <launches.csv> <report.ncu-rep> <out.md> [title]"""
import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed.avg.per_cycle_active",
    "smsp__issue_active.avg.per_cycle_active",
    "smsp__inst_executed.sum",
    "sm__cycles_elapsed.avg", "smsp__cycles_elapsed.avg.per_second",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def main():
    launches, rep, out = sys.argv[1:4]
    title = sys.argv[4] if len(sys.argv) > 4 else rep
    lines = [f"# {title}", ""]
    rows = list(csv.reader(l for l in open(launches) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        tot[r[ki]] += v * scale
        cnt[r[ki]] += 1
    s = sum(tot.values())
    lines += ["## launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised:",
              "## compare shares, not absolutes)", "", "| kernel | launches | total ms | share |", "|---|---|---|---|"]
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        lines.append(f"| `{k[:100]}` | {cnt[k]} | {v:.3f} | {100 * v / s:.1f}% |")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    h, u = rr[0], rr[1]
    lines += ["", "## top kernel, `ncu --set full --clock-control none` (first captured launch)", ""]
    ni = h.index("Kernel Name")
    lines.append(f"kernel: `{rr[2][ni]}`")
    lines += ["", "| metric | unit | value |", "|---|---|---|"]
    for k in KEYS:
        if k in h:
            i = h.index(k)
            lines.append(f"| {k} | {u[i]} | {rr[2][i]} |")
    # stall reasons (warp states), if present
    lines += ["", "## warp-state / stall metrics present in the report", "", "| metric | value |", "|---|---|"]
    for i, name in enumerate(h):
        if "issue_stalled" in name and name.endswith("_per_warp_active.pct"):
            lines.append(f"| {name} | {rr[2][i]} |")
    for i, name in enumerate(h):
        if name.startswith("smsp__average_warp") or name.startswith("smsp__warps_issue_stalled"):
            lines.append(f"| {name} | {rr[2][i]} |")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()
