#!/usr/bin/env python3
"""Static facts about the shipped sponge kernels, taken from their SASS at build time (no GPU needed).

usage: tools/kernel_sass.py <out.json> <object.o> [<object.o> ...]

For every `sponge_kernel<...>` entry in the given objects (cuobjdump -sass):
  * `sass_sha256`   fingerprint of the instruction stream (opcodes + operands, addresses and encodings
                    stripped); `opcode_sha256` the same over the opcode sequence alone, which is what an ncu report's
                    source page reproduces exactly (it prints absolute branch targets).  bench.py uses the latter to
                    decide whether the ncu-measured per-permutation instruction counts in profiles/
                    (tools/ncu_counts.py) were taken from THIS build of the kernel;
  * `instructions`  static instruction count and the multiply / FP64 / load mix of the whole kernel;
  * `loops`         every backward branch's body [target, branch] with the same mix -- the partial-round loop of
                    the variant-5 kernels is the largest body without an inner loop (straight-line code).
Per-object results are cached next to the object (keyed by size + mtime), so an incremental build only
disassembles what was recompiled."""
import collections
import hashlib
import json
import os
import re
import subprocess
import sys

INS = re.compile(r"\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)(.*?);")


def mix(ops):
    h = collections.Counter(ops)

    def cnt(pred):
        return sum(v for k, v in h.items() if pred(k))
    return dict(total=len(ops),
                imad_wide=cnt(lambda k: k.startswith("IMAD.WIDE")),
                imad_hi=cnt(lambda k: k.startswith("IMAD.HI")),
                imad=cnt(lambda k: k == "IMAD" or k.startswith("IMAD.U32") or k.startswith("IMAD.X")),
                imad_other=cnt(lambda k: k.startswith("IMAD.MOV") or k.startswith("IMAD.IADD") or k.startswith("IMAD.SHL")),
                fp64=cnt(lambda k: k.startswith(("DFMA", "DADD", "DMUL"))),
                iadd3=cnt(lambda k: k.startswith("IADD3")),
                lds=cnt(lambda k: k.startswith("LDS")),
                ldc=cnt(lambda k: k.startswith(("LDC", "LDCU"))),
                mov_sel=cnt(lambda k: k in ("MOV", "SEL") or k.startswith("IMAD.MOV")))


def analyse(obj):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    kernels, fn, ins = {}, None, []

    def flush():
        if fn and "sponge_kernel" in fn and ins:
            text = "\n".join(f"{op}{rest}" for _, op, rest in ins)
            addr_idx = {a: i for i, (a, _, _) in enumerate(ins)}
            loops = []
            for i, (a, op, rest) in enumerate(ins):
                if op.startswith("BRA"):
                    m = re.search(r"0x([0-9a-f]+)", rest)
                    if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr_idx:
                        loops.append((addr_idx[int(m.group(1), 16)], i))
            lrec = []
            for s, e in sorted(loops):
                inner = any(s <= s2 and e2 <= e and (s2, e2) != (s, e) for s2, e2 in loops)
                d = mix([op for _, op, _ in ins[s:e + 1]])
                d.update(first=f"{ins[s][0]:#x}", last=f"{ins[e][0]:#x}", has_inner_loop=inner)
                lrec.append(d)
            ops = "\n".join(op for _, op, _ in ins)
            kernels[fn] = dict(sass_sha256=hashlib.sha256(text.encode()).hexdigest(),
                               opcode_sha256=hashlib.sha256(ops.encode()).hexdigest(), instructions=mix([op for _, op, _ in ins]),
                               loops=lrec)

    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            fn, ins = m.group(1), []
            continue
        m = INS.match(line)
        if m and fn:
            ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
    flush()
    return kernels


def demangle(names):
    if not names:
        return {}
    try:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def main():
    out_path, objs = sys.argv[1], sys.argv[2:]
    merged = {}
    for obj in objs:
        st = os.stat(obj)
        cache = obj + ".sass.json"
        key = f"v2:{st.st_size}:{st.st_mtime_ns}"
        rec = None
        if os.path.exists(cache):
            try:
                rec = json.load(open(cache))
                if rec.get("key") != key:
                    rec = None
            except ValueError:
                rec = None
        if rec is None:
            rec = dict(key=key, kernels=analyse(obj))
            json.dump(rec, open(cache, "w"))
        merged.update(rec["kernels"])
    names = demangle(list(merged))
    result = {}
    for mangled, rec in merged.items():
        rec["mangled"] = mangled
        result[names.get(mangled, mangled)] = rec
    json.dump(result, open(out_path, "w"), indent=1, sort_keys=True)
    print(f"{out_path}: {len(result)} kernels")


if __name__ == "__main__":
    main()
