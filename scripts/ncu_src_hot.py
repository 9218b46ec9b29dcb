#!/usr/bin/env python
"""Per-SASS-instruction executed counts of one kernel from an ncu `--page source --csv` export.
usage: ncu -i X.ncu-rep --page source --csv --kernel-name regex:K > k.csv ; python scripts/ncu_src_hot.py k.csv [min_frac]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
minf = float(sys.argv[2]) if len(sys.argv) > 2 else 0.002
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index('Address'), hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
body = [r for r in rows[2:] if len(r) > iex and r[iex].isdigit()]
tot = sum(int(r[iex]) for r in body)
smp = sum(int(r[ismp]) for r in body if r[ismp].isdigit())
print('total warp-instructions', tot, 'samples', smp)
for r in body:
    n = int(r[iex])
    if n > tot * minf:
        print(r[ia][-4:], '%10d %5.2f%% smp %5.2f%% ' % (n, 100.0 * n / tot, 100.0 * int(r[ismp] or 0) / max(smp, 1)), r[isrc][:100])
