#!/usr/bin/env python
"""Opcode histogram of the innermost loops of a kernel's SASS. Usage: loop_hist.py OBJ_OR_SO MANGLED_SUBSTR"""
import re, subprocess, sys
obj, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", out)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    ins = []
    for line in f.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    print(name)
    for i, (addr, txt) in enumerate(ins):
        m = re.search(r"BRA (?:\w+ )?`?\(?\.?L_x_\d+\)?|BRA 0x([0-9a-f]+)", txt)
        if "BRA" in txt:
            m2 = re.search(r"0x([0-9a-f]+)", txt)
            if m2:
                tgt = int(m2.group(1), 16)
                if tgt < addr:
                    body = [t for a, t in ins if tgt <= a <= addr]
                    h = {}
                    for t in body:
                        op = t.split()[1] if t.startswith("@") else t.split()[0]
                        op = op.split(".")[0]
                        h[op] = h.get(op, 0) + 1
                    fp64 = sum(v for k, v in h.items() if k in ("DFMA", "DMUL", "DADD"))
                    fp32 = sum(v for k, v in h.items() if k in ("FFMA", "FMUL", "FADD"))
                    mufu = h.get("MUFU", 0)
                    if mufu: print(f"  loop {tgt:#x}..{addr:#x}: {len(body)} instrs, FP64 {fp64}, FP32 {fp32}, MUFU {mufu}, other {len(body) - fp64 - fp32 - mufu}; per pair (by MUFU): "
                          f"fp {(fp64 + fp32) / max(mufu, 1):.2f}, other {(len(body) - fp64 - fp32 - mufu) / max(mufu, 1):.2f}  ", dict(sorted(h.items(), key=lambda kv: -kv[1])))
