#!/usr/bin/env python
"""Opcode counts per kernel of the in-tree library (`cuobjdump -sass`): the evidence that the tensor-core kernels
issue tcgen05 MMAs (UTCHMMA), use tensor memory (LDTM / STTM, UTCBAR) and bulk copies (UBLKCP)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "deeplearninginassetpricing_paperreplication_b200/libdlap_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
INTEREST = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "SYNCS", "HMMA", "FFMA", "DFMA", "DADD", "LDG", "STG",
            "LDS", "STS", "ATOM", "RED", "SHFL", "BAR", "MUFU", "ELECT")
kern, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()[:110]
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1).split(".")[0]
        counts[kern]["total"] += 1
        for k in INTEREST:
            if op.startswith(k):
                counts[kern][k] += 1
                break
for k, c in counts.items():
    if c["total"] == 0:
        continue
    print(k)
    print("    " + "  ".join(f"{n}={c[n]}" for n in ("total",) + INTEREST if c[n]))
