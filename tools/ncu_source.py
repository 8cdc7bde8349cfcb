#!/usr/bin/env python
"""Per-source-line hot spots of an ncu report (needs -lineinfo and --import-source on).

    python tools/ncu_source.py gpurun_out/prof.ncu-rep [top_n] [kernel-id]
Prints the lines with the most executed warp instructions and the most stall samples."""
import csv
import subprocess
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows, fname, hdr, seen_kernel = [], None, None, 0
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path" or r[0] == "File Name":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name" or r[0] == "Kernel Name":
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr) or not r[0]:
            continue  # rows with an empty line number are the SASS instructions under a source line
        d = {}
        for k, v in zip(hdr, r):
            d.setdefault(k, v)  # "Source" appears twice (CUDA line, SASS text): keep the first
        try:
            inst = int(d.get("Instructions Executed", "0") or 0)
            smp = int(d.get("# Samples", "0") or 0)
        except ValueError:
            continue
        if inst or smp:
            rows.append((fname, int(d["Line No"]), inst, smp, d["Source"].strip()[:110], d))
    # the report may hold several launches of the kernel: every launch repeats every line -> keep the first occurrence
    first = {}
    for r in rows:
        first.setdefault((r[0], r[1]), r)
    rows = list(first.values())
    ti, ts = sum(r[2] for r in rows), sum(r[3] for r in rows)
    print(f"total warp instructions {ti}, stall samples {ts}")
    print("---- by instructions executed")
    for r in sorted(rows, key=lambda r: -r[2])[:top]:
        print(f"{r[0]:22s}:{r[1]:4d} inst {r[2]:8d} ({100*r[2]/max(ti,1):4.1f}%) smp {r[3]:6d}  {r[4]}")
    print("---- by stall samples")
    keys = [k for k in rows[0][5] if k.startswith("stall_") and "Not Issued" not in k]
    for r in sorted(rows, key=lambda r: -r[3])[:top]:
        st = sorted(((int(r[5][k] or 0), k[6:]) for k in keys), reverse=True)[:3]
        print(f"{r[0]:22s}:{r[1]:4d} smp {r[3]:6d} ({100*r[3]/max(ts,1):4.1f}%) inst {r[2]:8d} {st}  {r[4][:70]}")


if __name__ == "__main__":
    main()
