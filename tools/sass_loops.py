#!/usr/bin/env python
"""Length and opcode mix of the packed-fp32x2 step loops of every transient kernel in an object file -- a CPU-side check before GPU time:
   python tools/sass_loops.py nlos_gaussian_renderer_b200/lib/transient_narrow.o [more .o]
ptxas sometimes rotates the loop-carried accumulators through a dozen MOVs at the back edge (55 -> 65 instructions per 4 steps, +7 % on the
backward): a loop whose MOV count is not 0 - 2 (backward) is the thing to look for after any change near the candidate loop."""
import re
import subprocess
import sys
from collections import Counter

for obj in sys.argv[1:]:
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    fun, funs = None, {}
    for ln in sass.splitlines():
        m = re.match(r"\s+Function : (\S+)", ln)
        if m:
            fun = m.group(1)
            funs[fun] = []
        elif fun and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", ln):
            funs[fun].append(ln)
    for fun, ins in funs.items():
        if "transient" not in fun:
            continue
        addr = {int(re.match(r"\s+/\*([0-9a-f]+)\*/", ln).group(1), 16): i for i, ln in enumerate(ins)}
        spill = sum(("LDL" in ln or "STL" in ln) for ln in ins)
        print(f"{fun[:64]}: {len(ins)} instructions, {spill} LDL/STL")
        for i, ln in enumerate(ins):
            m = re.search(r"BRA\s+(0x[0-9a-f]+)", ln)
            if not m:
                continue
            tgt = int(m.group(1), 16)
            if tgt in addr and addr[tgt] < i:
                body = ins[addr[tgt]:i + 1]
                ops = [re.sub(r"@!?U?P\d\s+", "", x.split("*/")[1].strip()).split()[0] for x in body]
                c = Counter(o.split(".")[0] for o in ops)
                packed = c["FMUL2"] + c["FFMA2"] + c["FADD2"]
                if packed >= 8 and len(body) < 200:
                    mov = sum(1 for o in ops if o.startswith("MOV") or o.startswith("IMAD.MOV"))
                    print(f"   loop @{tgt:x}: {len(body)} instructions, {packed} packed, LDS {c['LDS']}, STS {c['STS']}, MOV {mov}")
