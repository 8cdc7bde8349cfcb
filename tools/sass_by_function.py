#!/usr/bin/env python
"""SASS instruction count of one kernel attributed to the source function each instruction was inlined from.
usage: tools/sass_by_function.py [kernel-substring]   (needs nvdisasm; library built with -lineinfo)
Attribution: nvdisasm -g prints '//## File "...", line N' before each group; a line belongs to the __device__ function
whose definition precedes it in the file (function starts found with a regex over the source)."""
import collections, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
want = sys.argv[1] if len(sys.argv) > 1 else "post_physics_kernelILb1ELb0"
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "dexrobot_isaac_b200", "libdexstep.so")], cwd=tmp,
                      stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", cubin], cwd=tmp, capture_output=True, text=True).stdout.splitlines()

starts = {}  # file -> sorted [(line, name)]
def fn_starts(path):
    if path not in starts:
        out = []
        pat = re.compile(r"^(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__|inline|__host__).*?\b(\w+)\s*\(")
        for i, l in enumerate(open(path), 1):
            m = pat.match(l)
            if m and not l.startswith(" "):
                out.append((i, m.group(1)))
        starts[path] = out
    return starts[path]

def owner(path, line):
    name = "?"
    for ln, nm in fn_starts(path):
        if ln <= line:
            name = nm
        else:
            break
    return name

per_fn, per_line = collections.Counter(), collections.Counter()
inside, cur = False, ("?", 0)
for l in txt:
    if l.startswith(".text."):
        inside = want in l
    if l.startswith("\t.section") or l.startswith(".section"):
        inside = inside and ".text." in l
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    if re.match(r"^\s+/\*[0-9a-f]{4,}\*/\s", l):
        fn = owner(*cur) if os.path.exists(cur[0]) else "?"
        per_fn[fn] += 1
        per_line[(os.path.basename(cur[0]), cur[1])] += 1
tot = sum(per_fn.values())
print(f"{want}: {tot} instructions ({tot * 16 / 1024:.0f} KB)")
for fn, n in per_fn.most_common(25):
    print(f"  {n:6d}  {100 * n / tot:5.1f} %  {fn}")
print("top lines:")
for (f, ln), n in per_line.most_common(15):
    print(f"  {n:6d}  {f}:{ln}")
