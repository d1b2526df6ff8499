#!/usr/bin/env python3
"""Dynamic instruction counts of a kernel from a committed `ncu --set full --import-source on` capture.

usage: tools/ncu_counts.py <report.ncu-rep> <kernel_sass.json> <kernel_counts.json> --workload c2 --log-n 24
           --perms-per-launch N --source profiles/<summary>.md [--summary-out profiles/<summary>.md --title "..."]

Reads the SASS source page (`ncu -i rep --page source --csv --print-source sass`): every SASS line with the number of
warp-level executions.  Sums them by opcode class and divides by the warp-permutations of the launch, which gives
the warp instructions PER PERMUTATION by pipe -- measured, not modelled.  The record is keyed by the kernel's
demangled name and carries the SASS fingerprint of the build it was captured from (tools/kernel_sass.py), so that
bench.py can tell whether the numbers belong to the library it is running.  Also writes a short markdown summary
(launch facts, pipe activity, stall reasons, DRAM traffic) for profiles/."""
import argparse
import collections
import hashlib
import csv
import json
import re
import subprocess


def classify(op):
    if op.startswith("IMAD.WIDE"):
        return "imad_wide"
    if op.startswith("IMAD.HI"):
        return "imad_hi"
    if op == "IMAD" or op.startswith("IMAD.U32") or op.startswith("IMAD.X"):
        return "imad"
    if op.startswith(("IMAD.MOV", "IMAD.IADD", "IMAD.SHL")):
        return "imad_other"
    if op.startswith(("DFMA", "DADD", "DMUL")):
        return "fp64"
    if op.startswith(("IMMA", "HMMA", "DMMA")) or (op.startswith("UTC") and "MMA" in op):
        return "mma"   # warp-level mma.sync, or tcgen05.mma (one warp-level execution = one instruction for the whole tile)
    if op.startswith(("LDS", "STS", "LDSM")):
        return "shared_ldst"
    if op.startswith(("LDG", "STG", "LDL", "STL", "LD.", "ST.")):
        return "global_ldst"
    if op.startswith(("LDC", "LDCU", "ULDC")):
        return "const_ld"
    if op.startswith("U") and not op.startswith("UNPACK"):
        return "uniform"
    return "alu_other"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("kernel_sass")
    ap.add_argument("out")
    ap.add_argument("--workload", required=True)
    ap.add_argument("--log-n", type=int, required=True)
    ap.add_argument("--perms-per-launch", type=float, required=True)
    ap.add_argument("--source", default=None)
    ap.add_argument("--summary-out", default=None)
    ap.add_argument("--title", default=None)
    a = ap.parse_args()

    src = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout.splitlines()
    kernel = list(csv.reader([src[0]]))[0][1]
    rows = list(csv.reader(src[1:]))
    hdr = rows[0]
    si, ei = hdr.index("Source"), hdr.index("Instructions Executed")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    by_class, stalls, total = collections.Counter(), collections.Counter(), 0
    opcodes = []
    for r in rows[1:]:
        if len(r) <= ei:
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[si])
        if not m:
            continue
        opcodes.append(m.group(2))
        try:
            n = int(float(r[ei]))
        except ValueError:
            continue
        by_class[classify(m.group(2))] += n
        total += n
        for i, h in stall_cols:
            try:
                stalls[h] += int(float(r[i]))
            except (ValueError, IndexError):
                pass
    warp_perms = a.perms_per_launch / 32.0
    per = {k: v / warp_perms for k, v in by_class.items()}
    per["total"] = total / warp_perms

    raw = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    h, u, v = rr[0], rr[1], rr[2]

    def metric(name, cast=float):
        if name in h:
            try:
                return cast(v[h.index(name)].replace(",", ""))
            except ValueError:
                return None
        return None

    def scaled(name):   # to plain units
        x = metric(name)
        if x is None:
            return None
        unit = u[h.index(name)]
        return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "s": 1, "ns": 1e-9,
                    "Ghz": 1e9, "Mhz": 1e6}.get(unit, 1)

    sass = json.load(open(a.kernel_sass))
    rec = dict(workload=a.workload, log_n=a.log_n, perms_per_launch=a.perms_per_launch, source=a.source or a.report,
               # fingerprint of the CAPTURED kernel's opcode sequence (what tools/kernel_sass.py computes for a build)
               opcode_sha256=hashlib.sha256("\n".join(opcodes).encode()).hexdigest(),
               matches_build_at_processing_time=(sass.get(kernel) or {}).get("opcode_sha256") ==
               hashlib.sha256("\n".join(opcodes).encode()).hexdigest(),
               warp_inst_per_perm={k: round(x, 2) for k, x in sorted(per.items())},
               registers_per_thread=metric("launch__registers_per_thread", lambda s: int(float(s))),
               block_size=metric("launch__block_size", lambda s: int(float(s))),
               grid_size=metric("launch__grid_size", lambda s: int(float(s))),
               kernel_ms=(scaled("gpu__time_duration.sum") or 0) * 1e3,
               sm_mhz=(scaled("sm__cycles_elapsed.avg.per_second") or 0) / 1e6, sm_count=148,
               dram_bytes_per_launch=(scaled("dram__bytes_read.sum") or 0) + (scaled("dram__bytes_write.sum") or 0),
               pipe_pct=dict(fmaheavy=metric("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                             fp64=metric("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                             alu=metric("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                             tensor=metric("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                             issue=metric("sm__inst_executed.avg.pct_of_peak_sustained_elapsed")))
    try:
        out = json.load(open(a.out))
    except (OSError, ValueError):
        out = {}
    out[kernel] = rec
    json.dump(out, open(a.out, "w"), indent=1, sort_keys=True)
    print(json.dumps({kernel: rec}, indent=1))

    if a.summary_out:
        tot_st = sum(stalls.values()) or 1
        lines = [f"# {a.title or kernel}", "", f"kernel: `{kernel}`", f"report: `{a.report}` (ncu --set full --clock-control none "
                 f"--import-source on, first captured launch)", "", "| fact | value |", "|---|---|"]
        for k in ("kernel_ms", "grid_size", "block_size", "registers_per_thread", "sm_mhz", "dram_bytes_per_launch"):
            lines.append(f"| {k} | {rec[k]} |")
        lines += ["", "## pipe activity (% of peak)", "", "| pipe | % |", "|---|---|"]
        lines += [f"| {k} | {x} |" for k, x in rec["pipe_pct"].items()]
        lines += ["", f"## warp instructions per permutation ({a.perms_per_launch:.0f} permutations in the launch), from the "
                  "per-instruction execution counts of the source page", "", "| class | per permutation |", "|---|---|"]
        lines += [f"| {k} | {x} |" for k, x in rec["warp_inst_per_perm"].items()]
        lines += ["", "## warp stall samples (source page, all samples)", "", "| reason | share |", "|---|---|"]
        lines += [f"| {k} | {100.0 * x / tot_st:.1f} % |" for k, x in stalls.most_common(10)]
        open(a.summary_out, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
