#!/usr/bin/env python
"""SASS opcode histogram of liblutvq.so (cuobjdump -sass): whole library, then per kernel -> profiles/r02_sass_opcodes.txt.
Runs without a GPU."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "rvv-lut-vq_b200", "csrc", "liblutvq.so")
OUT = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_opcodes.txt")
INTEREST = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKPF", "UBLKCP", "HMMA", "ATOMG", "REDG", "PRMT", "LDS", "LDGSTS",
            "CREATEPOLICY", "ACQBULK", "SYNCS"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
total, per, cur = collections.Counter(), collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = demangle(m.group(1))
        per[cur] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur is not None:
        op = m.group(1)
        total[op] += 1
        per[cur][op] += 1
with open(OUT, "w") as f:
    f.write("# SASS opcode histogram of rvv-lut-vq_b200/csrc/liblutvq.so (sm_100a), cuobjdump -sass (tools/sass_histogram.py); whole library, then per kernel (top opcodes)\n")
    f.write("# tcgen05 = UTCHMMA / UTCBAR / LDTM + STTM (tensor-memory loads / stores); TMA tensor copy = UTMALDG; bulk async copy / prefetch = UBLKCP / UBLKPF;\n"
            "# legacy tensor core = HMMA; L2 policy = CREATEPOLICY\n")
    f.write("TOTAL " + str(dict(total.most_common(60))) + "\n")
    f.write("of interest: " + str({k: total.get(k, 0) for k in INTEREST}) + "\n")
    for name, c in per.items():
        extra = {k: c[k] for k in INTEREST[:8] if c.get(k)}
        f.write(f"{name} {sum(c.values())} {dict(c.most_common(14))}" + (f" | {extra}" if extra else "") + "\n")
print(open(OUT).read().splitlines()[4])
