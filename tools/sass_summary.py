#!/usr/bin/env python
"""Per-kernel SASS summary of the shipped library: instruction count, registers / spill stack / static shared memory, and
the memory-instruction mix the north star asks about (128-bit global loads / stores, LDGSTS, bulk and tensor TMA copies,
mbarrier / named-barrier synchronisation).   python tools/sass_summary.py > profiles/<round>_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "dexrobot_isaac_b200", "libdexstep.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout

usage = {}
name = None
for l in res.splitlines():
    m = re.search(r"Function (\S+):", l)
    if m:
        name = m.group(1)
        continue
    if name and "REG:" in l:
        usage[name] = {k: int(v) for k, v in re.findall(r"(REG|STACK|SHARED|LOCAL):(\d+)", l)}
        name = None

KEYS = ["LDG.128", "LDG.64", "LDG.32", "LDG.8/16", "STG.128", "STG.64", "STG.32", "STG.8/16", "LDGSTS", "UBLKCP (bulk copy)",
        "UTMALDG (tensor copy)", "UTMASTG", "SYNCS (mbarrier)", "BAR", "RED/ATOM", "LDS", "STS", "MUFU", "BRA"]


def classify(op):
    for base in ("LDG", "STG"):
        if op.startswith(base + "."):
            parts = op.split(".")
            if "128" in parts:
                return base + ".128"
            if "64" in parts:
                return base + ".64"
            if any(x in parts for x in ("U8", "S8", "U16", "S16")):
                return base + ".8/16"
            return base + ".32"
    if op.startswith("LDGSTS"):
        return "LDGSTS"
    if op.startswith("UBLKCP"):
        return "UBLKCP (bulk copy)"
    if op.startswith("UTMALDG"):
        return "UTMALDG (tensor copy)"
    if op.startswith("UTMASTG"):
        return "UTMASTG"
    if op.startswith("SYNCS"):
        return "SYNCS (mbarrier)"
    if op.startswith("BAR."):
        return "BAR"
    if op.split(".")[0] in ("RED", "ATOMG", "ATOM", "REDG"):
        return "RED/ATOM"
    if op.startswith("LDS"):
        return "LDS"
    if op.startswith("STS"):
        return "STS"
    if op.startswith("MUFU"):
        return "MUFU"
    if op.split(".")[0] == "BRA":
        return "BRA"
    return None


counts = collections.OrderedDict()
name = None
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        name = m.group(1)
        counts[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", l) if name else None
    if m:
        counts[name]["instructions"] += 1
        k = classify(m.group(1))
        if k:
            counts[name][k] += 1


def short(n):
    n = re.sub(r"_ZN\d+_GLOBAL__N__[0-9a-f]{8}_\d+_dexstep_cu_[0-9a-f]{8}", "", n)
    n = re.sub(r"^_Z", "", n)
    n = re.sub(r"^3pp2", "", n)
    m = re.match(r"(\d+)", n)
    if not m:
        return n
    k, d = int(m.group(1)), len(m.group(1))
    base, rest = n[d:d + k], n[d + k:]
    targs = re.findall(r"L[bi](\d+)E", rest.split("Ev")[0]) if rest.startswith("I") else []  # bool and int arguments
    return base + ("<" + ",".join(targs) + ">" if targs else "")


print(f"SASS summary of {os.path.relpath(lib, ROOT)} (sm_100a; cuobjdump -sass / --dump-resource-usage)\n")
print("template arguments: post_physics_kernel<BG,DICT,NORM>, post_physics_v2_kernel<BG,STD,TMA,OBS,G> (G = tiles per pass: 1 persistent, 2 pair), "
      "post_physics_v2s_kernel<BG,STD,OBS>\n")
keys = ["instructions"] + KEYS
for n, c in sorted(counts.items(), key=lambda kv: -kv[1]["instructions"]):
    u = usage.get(n, {})
    print(f"{short(n)}:  {c['instructions']} instr ({c['instructions'] * 16 // 1024} KB), {u.get('REG', '?')} regs, "
          f"{u.get('STACK', '?')} B spill stack, {u.get('SHARED', '?')} B static smem")
    print("    " + "  ".join(f"{k} {c[k]}" for k in keys[1:] if c[k]))
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print("\nwhole library: " + "  ".join(f"{k} {tot[k]}" for k in keys if tot[k]))
