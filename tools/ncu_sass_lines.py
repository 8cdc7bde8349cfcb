#!/usr/bin/env python
"""Print the SASS attributed to a range of source lines: python tools/ncu_sass_lines.py cs.csv file.cuh 900 906"""
import csv, sys
path, f, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
hdr = None; fname = None; show = False
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] in ('File Path', 'File Name'): fname = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    if r[0]:
        show = (fname == f and lo <= int(r[0]) <= hi)
        if show: print('LINE', r[0], r[1][:100], r[7])
    elif show:
        print('      ', r[3][:80], r[7])
