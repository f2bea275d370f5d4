#!/usr/bin/env python3
"""B200 instruction-cache microbenchmark: a warp loops over N straight-line FFMA instructions (16 B each); cycles per
instruction as a function of the loop's code size shows the capacity of each instruction-cache level and the cost of
running code that does not fit (the decode-chain kernel executes each phase's code once per phase: cold code).
usage: icache.py [gen|run]   gen: writes icache_gen.cu;  run: compiles (nvcc) and runs it on the GPU."""
import os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
SIZES = [512, 1024, 2048, 3072, 4096, 6144, 8192, 12288, 16384, 24576, 32768]

def gen():
    out = ["#include <cstdio>", "#include <cuda_runtime.h>"]
    for n in SIZES:
        out.append(f"__global__ void k{n}(float * o, long long * t, int iters, float a, float b) {{")
        out.append("  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;")
        out.append("  long long t0 = 0;")
        out.append("  for (int it = 0; it < iters + 1; ++it) {")
        out.append("    if (it == 1) t0 = clock64();")
        for i in range(n // 4):
            out.append("    asm volatile(\"fma.rn.f32 %0, %0, %4, %5; fma.rn.f32 %1, %1, %4, %5; fma.rn.f32 %2, %2, %4, %5; fma.rn.f32 %3, %3, %4, %5;\" : \"+f\"(x0), \"+f\"(x1), \"+f\"(x2), \"+f\"(x3) : \"f\"(a), \"f\"(b));")
        out.append("  }")
        out.append("  long long t1 = clock64();")
        out.append("  if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;")
        out.append("  o[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3;")
        out.append("}")
    out.append("int main() { float * o; long long * t; cudaMalloc(&o, 148 * 1024 * 4); cudaMalloc(&t, 148 * 8); long long h[148];")
    out.append("  const int iters = 20;")
    for warps in (1, 4, 16):
        for n in SIZES:
            out.append(f"  k{n}<<<148, {warps * 32}>>>(o, t, iters, 1.0001f, 0.5f); cudaMemcpy(h, t, 148 * 8, cudaMemcpyDeviceToHost);")
            out.append(f"  {{ long long s = 0; for (int i = 0; i < 148; ++i) s += h[i]; printf(\"warps %2d  loop %6d instrs (%4d KB): %.2f cycles/instr (warm loop), \", {warps}, {n}, {n} * 16 / 1024, (double) s / 148 / iters / {n}); }}")
            out.append(f"  k{n}<<<148, {warps * 32}>>>(o, t, 0, 1.0001f, 0.5f); cudaMemcpy(h, t, 148 * 8, cudaMemcpyDeviceToHost);")
            out.append(f"  {{ printf(\"\\n\"); }}")
    out.append("  return 0; }")
    open(os.path.join(HERE, "icache_gen.cu"), "w").write("\n".join(out))

if __name__ == "__main__":
    gen()
    if len(sys.argv) > 1 and sys.argv[1] == "run":
        exe = os.path.join(HERE, "icache_gen")
        if not os.path.exists(exe):
            subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-o", exe, os.path.join(HERE, "icache_gen.cu")])
        subprocess.check_call([exe])
