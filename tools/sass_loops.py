#!/usr/bin/env python3
"""Loop-level opcode histogram from `cuobjdump -sass`: for every backward branch of a kernel, the body
[target, branch] and its multiply mix (GPU-free audit of the hot loops).
usage: tools/sass_loops.py <object> <mangled-kernel-name>"""
import collections
import re
import subprocess
import sys


def main():
    obj, fun = sys.argv[1], sys.argv[2]
    out = subprocess.run(["cuobjdump", "-sass", "-fun", fun, obj], capture_output=True, text=True).stdout
    ins = []
    for line in out.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
    addr_idx = {a: i for i, (a, _, _) in enumerate(ins)}
    loops = []
    for i, (a, op, rest) in enumerate(ins):
        if op.startswith("BRA"):
            m = re.search(r"0x([0-9a-f]+)", rest)
            if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr_idx:
                loops.append((addr_idx[int(m.group(1), 16)], i))
    print(f"{len(ins)} instructions, {len(loops)} backward branches")
    for s, e in sorted(loops):
        h = collections.Counter(op for _, op, _ in ins[s:e + 1])
        wide = sum(v for k, v in h.items() if k.startswith("IMAD.WIDE"))
        hi = sum(v for k, v in h.items() if k.startswith("IMAD.HI"))
        lo = sum(v for k, v in h.items() if k == "IMAD" or k.startswith("IMAD.U32") or k.startswith("IMAD.X"))
        fp64 = sum(v for k, v in h.items() if k.startswith("DFMA") or k.startswith("DADD") or k.startswith("DMUL"))
        n = e - s + 1
        print(f"loop {ins[s][0]:#07x}..{ins[e][0]:#07x}: {n} instr, WIDE={wide} HI={hi} IMAD={lo} FP64={fp64} "
              f"int-pipe-cycles~{4 * wide + 4 * hi + 2 * lo} fp64-pipe-cycles~{2 * fp64}")
        print("    " + ", ".join(f"{k}:{v}" for k, v in h.most_common(12)))


if __name__ == "__main__":
    main()
