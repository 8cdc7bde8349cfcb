#!/usr/bin/env python
"""Instruction-cache microbenchmark for B200 (run on the GPU box): cycles per instruction of straight-line FFMA code
of a given size, for the FIRST pass through the code in a launch (cold unless it survived from the previous launch)
and for the steady state of a loop over it, with 1 / 4 / 8 warps per SM running the same code, and with a second
kernel of `evict_kb` launched in between (the step alternates two kernels).

    python tools/microbench/icache.py > gpurun_out/icache.txt
"""
import os
import subprocess
import sys
import tempfile

import torch
from torch.utils.cpp_extension import load_inline  # noqa: F401  (not used: plain nvcc + ctypes keeps this self-contained)
import ctypes

SIZES_KB = [8, 16, 24, 32, 48, 64, 96, 128, 192, 256]


def gen(kb):
    n = kb * 1024 // 16  # SASS instructions (16 B each)
    body = []
    for i in range(n // 4):
        c = 1.0 + (i % 97) * 1e-3
        body.append(f"a0 = a0 * {c:.6f}f + b; a1 = a1 * {c + 1e-4:.6f}f + b; a2 = a2 * {c + 2e-4:.6f}f + b; a3 = a3 * {c + 3e-4:.6f}f + b;")
    return "\n".join(body)


src = ["#include <cuda_runtime.h>\n#include <stdint.h>\n"]
for kb in SIZES_KB:
    src.append(f"""
extern "C" __global__ void k{kb}(float* out, long long* clk, int iters, float b) {{
  float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
  long long t0 = clock64(), t1 = 0;
  for (int it = 0; it < iters; ++it) {{
    {gen(kb)}
    if (it == 0) t1 = clock64();
  }}
  long long t2 = clock64();
  if (threadIdx.x % 32 == 0) {{
    clk[(blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32) * 2] = t1 - t0;
    clk[(blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32) * 2 + 1] = t2 - t1;
  }}
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}}
extern "C" void run{kb}(float* out, long long* clk, int iters, int grid, int block, void* stream) {{
  k{kb}<<<grid, block, 0, (cudaStream_t)stream>>>(out, clk, iters, 0.5f);
}}
""")
d = tempfile.mkdtemp()
cu = os.path.join(d, "ic.cu")
open(cu, "w").write("".join(src))
so = os.path.join(d, "ic.so")
subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-fmad=true", "-shared", "-Xcompiler", "-fPIC",
                "--cudart", "shared", cu, "-o", so], check=True)
lib = ctypes.CDLL(so)
dev = torch.device("cuda:0")
out = torch.zeros(148 * 1024, device=dev)
clk = torch.zeros(148 * 32 * 2, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream


def run(kb, iters, warps, evict_kb=0, reps=5):
    f = getattr(lib, f"run{kb}")
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    first, steady = [], []
    for _ in range(reps):
        if evict_kb:
            g = getattr(lib, f"run{evict_kb}")
            g.argtypes = f.argtypes
            g(out.data_ptr(), clk.data_ptr(), 1, 148, 32 * warps, st)
        f(out.data_ptr(), clk.data_ptr(), iters, 148, 32 * warps, st)
        torch.cuda.synchronize()
        c = clk[: 148 * warps * 2].view(-1, 2).double()
        first.append(float(c[:, 0].median()))
        steady.append(float(c[:, 1].median()) / max(iters - 1, 1))
    n = kb * 1024 // 16
    return sorted(first)[len(first) // 2] / n, sorted(steady)[len(steady) // 2] / n


print("B200 instruction fetch: cycles per SASS instruction (FFMA x4 independent chains), median over SMs and 5 launches")
print(f"{'code KB':>8} | {'warps/SM':>8} | {'first pass':>10} | {'steady loop':>11} | first pass after a 128 KB kernel ran in between")
for kb in SIZES_KB:
    for warps in (1, 4, 8):
        a, b = run(kb, 6, warps)
        e, _ = run(kb, 6, warps, evict_kb=128 if kb != 128 else 192)
        print(f"{kb:8d} | {warps:8d} | {a:10.2f} | {b:11.2f} | {e:10.2f}")
