#!/usr/bin/env python
"""Summarise build/*.ptxas.log: one line per kernel (demangled name, registers, stack, spills, smem)."""
import re, subprocess, sys
from pathlib import Path
logs = sorted(Path(sys.argv[1] if len(sys.argv) > 1 else "pathtracing_b200/csrc/build").glob("*.ptxas.log"))
pat = sys.argv[2] if len(sys.argv) > 2 else ""
for log in logs:
    txt = log.read_text()
    ents = []
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'", txt):
        rest = txt[m.end():]
        nxt = rest.find("Compiling entry function")
        blk = rest[:nxt] if nxt >= 0 else rest
        fp = re.search(r"Function properties for " + re.escape(m.group(1)) + r"\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", blk)
        us = None
        for us in re.finditer(r"Used (\d+) registers", blk):
            pass
        if fp and us:
            ents.append((m.group(1), fp.group(1), fp.group(2), fp.group(3), us.group(1)))
    if not ents:
        continue
    names = subprocess.run(["c++filt"], input="\n".join(e[0] for e in ents), capture_output=True, text=True).stdout.splitlines()
    for n, e in zip(names, ents):
        n = re.sub(r"\(.*", "", n).replace("ptb::", "").replace("void ", "")
        if pat and not re.search(pat, n):
            continue
        print(f"{log.stem.split('.')[0]:14s} {n:70s} regs {e[4]:>3s} stack {e[1]:>4s} spill st/ld {e[2]}/{e[3]}")
