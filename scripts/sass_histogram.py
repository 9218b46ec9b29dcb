#!/usr/bin/env python
"""SASS opcode histogram of the built library, per kernel (cuobjdump -sass): the evidence for the TMA / mbarrier / packed
FP32x2 / multimem / cp.async claims in DESIGN.md.  usage: python scripts/sass_histogram.py > profiles/rNN_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "splatam_b200", "libsplat_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True, check=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        hist[kern][op.split(".")[0]] += 1
        for tag in ("LDGMC", "LDGSTS", "UBLKCP", "SYNCS", "REDG", "ATOMS", "FFMA2", "FMUL2", "FADD2", "MUFU"):
            if op.startswith(tag) and "." in op:
                hist[kern]["  " + op] += 1
KEY = ("FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL", "FADD", "MUFU", "LDS", "STS", "LDGSTS", "UBLKCP", "SYNCS", "REDG", "LDGMC",
       "STG", "LDG", "ATOMS", "ATOMG", "BAR", "SHFL", "VOTE", "FSETP", "FSEL", "FMNMX", "ISETP", "IMAD", "BRA")
print("SASS opcode histogram of", os.path.relpath(lib, ROOT), "(static instruction counts per kernel)")
for k, c in hist.items():
    if not c:
        continue
    tot = sum(v for o, v in c.items() if not o.startswith("  "))
    print("\n%s  [%d instructions]" % (k, tot))
    print("   " + "  ".join("%s %d" % (o, c[o]) for o in KEY if c.get(o)))
    detail = ["%s %d" % (o.strip(), v) for o, v in c.items() if o.startswith("  ")]
    if detail:
        print("   detail: " + "  ".join(sorted(detail)))
