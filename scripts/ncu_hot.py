#!/usr/bin/env python
"""Hot SASS lines of one kernel from `ncu --page source --csv` (executed share / stall-sample share)."""
import csv
import subprocess
import sys


def main(rep, kernel, skip="0", thresh="0.004"):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kernel,
                          '--launch-skip', skip, '--launch-count', '1'], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[1]
    data = []
    for r in rows[2:]:
        if r and r[0] == 'Kernel Name':
            break
        if len(r) == len(hdr):
            data.append(r)
    si, ei, sa, ta = (hdr.index(k) for k in ('Source', 'Instructions Executed', 'Warp Stall Sampling (All Samples)',
                                              'Avg. Threads Executed'))
    tot = sum(int(r[ei]) for r in data)
    tots = sum(int(r[sa]) for r in data)
    print('total warp-inst', tot, 'stall samples', tots, 'sass lines', len(data))
    th = float(thresh)
    for i, r in enumerate(data):
        e, s = int(r[ei]), int(r[sa])
        if e > tot * th or s > tots * th * 1.5:
            print("%4d %-64s exec=%5.2f%% smp=%5.2f%% thr=%s" % (i, r[si].strip()[:64], 100 * e / tot, 100 * s / tots, r[ta]))


if __name__ == '__main__':
    main(*sys.argv[1:])
