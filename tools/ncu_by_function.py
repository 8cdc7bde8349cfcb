#!/usr/bin/env python
"""Executed warp instructions and stall samples of an ncu report, summed per source function (inlined code is
attributed to the function its source line lies in).   python tools/ncu_by_function.py report.ncu-rep"""
import collections, csv, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
pat = re.compile(r"^(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__|inline|__host__).*?\b(\w+)\s*\(")
starts = {}
def owner(fname, line):
    path = os.path.join(ROOT, "dexrobot_isaac_b200", "csrc", fname)
    if not os.path.exists(path):
        return fname
    if path not in starts:
        starts[path] = [(i, m.group(1)) for i, l in enumerate(open(path), 1) for m in [pat.match(l)] if m]
    name = "?"
    for ln, nm in starts[path]:
        if ln <= line:
            name = nm
    return name
inst, smp, stall = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
fname, hdr, seen = None, None, set()
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) != len(hdr) or not r[0].isdigit():
        continue
    key = (fname, int(r[0]))
    if key in seen:
        continue
    seen.add(key)
    d = {}
    for k, v in zip(hdr, r):
        d.setdefault(k, v)
    fn = owner(*key)
    inst[fn] += int(d.get("Instructions Executed") or 0)
    smp[fn] += int(d.get("# Samples") or 0)
    for k, v in d.items():
        if k.startswith("stall_") and "Not Issued" not in k and v not in ("", "0"):
            stall[fn][k[6:]] += int(v)
ti, ts = sum(inst.values()), sum(smp.values())
print(f"total warp instructions {ti}, stall samples {ts}")
for fn, n in smp.most_common(22):
    top = ", ".join(f"{k} {v}" for k, v in stall[fn].most_common(4))
    print(f"  {fn:22s} inst {inst[fn]:9d} ({100*inst[fn]/ti:4.1f} %)  samples {n:5d} ({100*n/ts:4.1f} %)  {top}")
