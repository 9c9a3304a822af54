#!/usr/bin/env python
"""Aggregate `ncu --page source --csv --print-source cuda,sass` per source line:
instructions executed and stall samples.  Usage: ncu_source_lines.py REP [top]"""
import csv
import subprocess
import sys
from collections import defaultdict


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source',
                          'cuda,sass'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                         text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    # find the header row
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
    hdr = rows[hi]
    c_line, c_src = 0, 1
    c_inst = hdr.index('Instructions Executed')
    c_samp = hdr.index('# Samples')
    c_file = None
    inst = defaultdict(float)
    samp = defaultdict(float)
    text = {}
    cur_file = ''
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            if r and r[0] == 'File Path':
                cur_file = r[1].split('/')[-1]
            continue
        try:
            ln = int(r[c_line])
        except ValueError:
            continue
        key = ln
        try:
            inst[key] += float(r[c_inst] or 0)
            samp[key] += float(r[c_samp] or 0)
        except ValueError:
            pass
        text.setdefault(key, r[c_src].strip()[:90])
    tot_i = sum(inst.values()) or 1
    tot_s = sum(samp.values()) or 1
    print(f'total inst {tot_i:.3e}  samples {tot_s:.0f}')
    for k in sorted(inst, key=lambda k: -inst[k])[:top]:
        print(f'{k:5d} inst {100*inst[k]/tot_i:5.1f}%  stall {100*samp[k]/tot_s:5.1f}%  {text[k]}')


if __name__ == '__main__':
    main()
