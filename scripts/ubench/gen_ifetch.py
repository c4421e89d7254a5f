"""Instruction-fetch micro-benchmark generator: straight-line FMA code of a chosen size inside a loop.
Measures issued warp-instructions per clock per SM as a function of code size and resident warps (instruction fetch)."""
import sys
sizes = [256, 1024, 2048, 4096, 8192, 16384, 32768]   # FMA instructions per loop body (16 B each)
print("#include <cstdio>\n#include <cuda_runtime.h>")
for n in sizes:
    print(f"__global__ void k{n}(float* out, int iters, float x, float y) {{")
    print("  float a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, a4 = 4, a5 = 5, a6 = 6, a7 = 7;")
    print("  for (int it = 0; it < iters; it++) {")
    for i in range(n // 8):
        # coefficients differ per instruction so the compiler cannot re-roll the body
        print(f"    a0 = fmaf(a0, x, {i}.25f); a1 = fmaf(a1, y, {i}.5f); a2 = fmaf(a2, x, {i}.75f); a3 = fmaf(a3, y, {i}.125f); a4 = fmaf(a4, x, {i}.375f); a5 = fmaf(a5, y, {i}.625f); a6 = fmaf(a6, x, {i}.875f); a7 = fmaf(a7, y, {i}.0625f);")
    print("  }")
    print("  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;\n}")
print("int main() {\n  float* out; cudaMalloc(&out, 148 * 1024 * 4 * 8);\n  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);")
print("  int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);")
print("  const int blocks_per_sm[] = {1, 1, 1, 1, 2, 4}; const int threads[] = {32, 128, 256, 512, 512, 512};")
print("  for (int c = 0; c < 6; c++) {")
for n in sizes:
    print(f"    {{ int iters = {max(2, 2_000_000 // n)}; k{n}<<<sms * blocks_per_sm[c], threads[c]>>>(out, 2, 1.0001f, 0.9999f); cudaDeviceSynchronize();")
    print(f"      cudaEventRecord(e0); k{n}<<<sms * blocks_per_sm[c], threads[c]>>>(out, iters, 1.0001f, 0.9999f); cudaEventRecord(e1); cudaEventSynchronize(e1);")
    print(f"      float ms; cudaEventElapsedTime(&ms, e0, e1); double winst = (double){n} * iters * blocks_per_sm[c] * threads[c] / 32.0;")
    print(f"      printf(\"body %6d instr (%4d KB)  warps/SM %2d  ms %8.3f  IPC/SM %.3f\\n\", {n}, {n * 16 // 1024}, blocks_per_sm[c] * threads[c] / 32, ms, winst / (ms * 1e-3 * 1.965e9)); }}")
print("  }\n  return 0;\n}")
