#!/usr/bin/env python3
"""Static SASS opcode histogram per kernel of the built objects (no GPU needed):
   tools/sass_static.py [object ...]   (default: the two objects of libmsk144b200.so)
Evidence for what the hot kernels are made of: packed FP32 (FFMA2 / FADD2 / FMUL2), warp collectives (REDUX / CREDUX,
VOTE, SHFL), three-input min/max (FMNMX3), asynchronous copies (LDGSTS), and -- by design, DESIGN.md section 4 -- no
tensor-core or TMA instructions (UTCxMMA, UTMALDG) in this FP32 SIMT path."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
objs = sys.argv[1:] or [os.path.join(ROOT, "msk144cudecoder_b200/csrc/build", f) for f in ("decoder.o", "channelizer.o")]
MARK = ["FFMA2", "FADD2", "FMUL2", "FFMA", "FMNMX3", "FMNMX", "REDUX", "CREDUX", "VOTE", "SHFL", "MUFU", "LDGSTS", "LDS", "STS",
        "LDG", "STG", "BAR", "ATOMG", "RED", "UTCHMMA", "UTCQMMA", "UTCIMMA", "UTMALDG", "HMMA", "LDL", "STL"]
for obj in objs:
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    name, hist = None, collections.Counter()

    def flush():
        if name and hist:
            tot = sum(hist.values())
            demangled = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
            print(f"===== {demangled}   ({tot} SASS instructions)")
            top = ", ".join(f"{k} {v}" for k, v in hist.most_common(14))
            print(f"  top: {top}")
            marks = {m: sum(v for k, v in hist.items() if k == m or k.startswith(m + ".")) for m in MARK}
            print("  of interest: " + ", ".join(f"{m} {marks[m]}" for m in MARK if marks[m]))
            absent = [m for m in ("UTCHMMA", "UTCQMMA", "UTCIMMA", "UTMALDG", "HMMA") if not marks[m]]
            print("  absent: " + ", ".join(absent))
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name, hist = m.group(1), collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and name:
            op = m.group(1)
            if op in ("LDS", "STS", "LDG", "STG", "MUFU", "REDUX", "CREDUX"):
                op += m.group(2)
            hist[op] += 1
    flush()
