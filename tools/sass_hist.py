#!/usr/bin/env python3
"""Opcode histogram per kernel from `cuobjdump -sass` (GPU-free audit).
usage: tools/sass_hist.py <object-or-so> [kernel-name-regex]"""
import collections
import re
import subprocess
import sys

FMA_PIPE = ("IMAD", "FFMA", "FMUL", "FADD", "HFMA2")   # IMAD* (incl. IMAD.MOV/IADD/X/SHL) issue on the fma pipe


def main():
    obj = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    fn, hist = None, collections.defaultdict(collections.Counter)
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and fn:
            hist[fn][m.group(2)] += 1
    for fn, h in hist.items():
        if pat and not pat.search(fn):
            continue
        total = sum(h.values())
        wide = sum(v for k, v in h.items() if k.startswith("IMAD.WIDE"))
        hi = sum(v for k, v in h.items() if k.startswith("IMAD.HI"))
        imad = sum(v for k, v in h.items() if k == "IMAD" or k.startswith("IMAD.U32"))
        other_fma = sum(v for k, v in h.items() if k.startswith("IMAD")) - wide - hi - imad
        print(f"{fn}\n  total={total} IMAD.WIDE={wide} IMAD.HI={hi} IMAD(lo)={imad} other-IMAD*(mov/x/iadd/shl)={other_fma}")
        print("  " + ", ".join(f"{k}:{v}" for k, v in h.most_common(18)))


if __name__ == "__main__":
    main()
